import sys, json, torch
sys.path.insert(0, "/root/repo")
sys.argv=["bench.py"]
import bench
dev=torch.device("cuda:0"); torch.cuda.set_device(dev)
r=bench.time_reductions(dev, 6548.8)
print(json.dumps({k:{kk:vv for kk,vv in v.items() if "sweep" in kk or "c4" in kk} for k,v in r.items()}, indent=1))
