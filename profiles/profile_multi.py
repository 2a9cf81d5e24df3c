"""Config 2's 13 VGG16 activation shapes (bs 256) through tp_reduce_rows_multi: Taylor, APoZ and
Sensitivity, twice each (wrap in `ncu --set full -k regex:reduce_rows_multi`)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from torchpruner_b200 import _lib
DEV = torch.device("cuda:0")
shapes = [(256, 64, 32, 32)] * 2 + [(256, 128, 16, 16)] * 2 + [(256, 256, 8, 8)] * 3 + \
    [(256, 512, 4, 4)] * 3 + [(256, 512, 2, 2)] * 3
zs = [torch.relu(torch.randn(s, device=DEV)) for s in shapes]
gs = [torch.randn(s, device=DEV) for s in shapes]
for _ in range(2):
    _lib.reduce_rows_multi("taylor", list(zip(zs, gs)))
    _lib.reduce_rows_multi("apoz", zs)
    _lib.reduce_rows_multi("abs_sum", gs)
torch.cuda.synchronize()
print("bytes: taylor", sum(8 * z.numel() + 4 * z.shape[0] * z.shape[1] for z in zs),
      "apoz/abs_sum", sum(4 * z.numel() + 4 * z.shape[0] * z.shape[1] for z in zs))
