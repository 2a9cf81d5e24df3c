"""Largest image (H*W) that gets position-major tiles: per-layer time of the conv kernel for
conv_pos_major_max_hw in {64, 256, 1024} (fp16x3 flavour unless `tf32` is given)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import profile_conv as pc
from torchpruner_b200 import _lib
for mx in (64, 256, 1024):
    _lib.set_tuning("conv_pos_major_max_hw", mx)
    for shp in pc.LAYERS[:5]:
        N = (pc.images_for_f16 if pc.F16 else pc.images_for)(*shp[:3])
        ms, tf = pc.run(*shp, N)
        print(mx, shp, N, round(ms, 3))
