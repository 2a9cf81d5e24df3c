import sys; sys.path.insert(0,"/root/repo"); sys.path.insert(0,"/root/repo/profiles")
import profile_conv as pc
from torchpruner_b200 import _lib
for mx in (64, 256, 1024):
    _lib.set_tuning("conv_pos_major_max_hw", mx)
    for shp in pc.LAYERS[:5]:
        N = pc.images_for(*shp[:3]); ms, tf = pc.run(*shp, N)
        print(mx, shp, N, round(ms,3))
