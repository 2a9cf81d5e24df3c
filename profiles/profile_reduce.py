"""Launch the row-reduction kernels on config-4's largest tensor and config-2's largest layer a
few times (to be wrapped in `ncu --set full -k regex:reduce_rows`)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from torchpruner_b200 import _lib  # noqa: E402

dev = torch.device("cuda:0")
for shape in ((256, 64, 112, 112), (256, 64, 32, 32), (256, 2048, 7, 7)):
    z = torch.relu(torch.randn(shape, device=dev))
    g = torch.randn(shape, device=dev)
    for _ in range(2):
        _lib.taylor_reduce(z, g)
        _lib.apoz_count(z)
    torch.cuda.synchronize()
    n, rows = z.numel(), shape[0] * shape[1]
    print(shape, "algorithmic bytes: taylor", 8 * n + 4 * rows, "apoz", 4 * n + 4 * rows)
