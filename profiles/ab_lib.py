"""Run bench.py --core-only against an alternative build of the library (A/B measurements).
    python profiles/ab_lib.py path/to/lib.so [bench args]"""
import os, runpy, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, root)
from torchpruner_b200 import _lib
_lib.LIB_PATH = os.path.abspath(sys.argv[1])
sys.argv = [os.path.join(root, "bench.py"), "--core-only"] + sys.argv[2:]
runpy.run_path(sys.argv[0], run_name="__main__")
