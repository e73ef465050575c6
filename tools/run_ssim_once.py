#!/usr/bin/env python
"""One fused L1 + SSIM loss (value + gradient) on a 1080p image pair, a few times: the target of an ncu capture
(ncu -k regex:fused_ssim --launch-skip 3 -c 1 python tools/run_ssim_once.py)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from gaussian_splatting_lightning_b200.training import fused_ssim, photometric_loss  # noqa: E402

a = torch.rand(1, 3, 1080, 1920, device="cuda", requires_grad=True)
b = torch.rand(1, 3, 1080, 1920, device="cuda")
for _ in range(5):
    loss = photometric_loss(a, b, 0.2)[0]
    loss.backward()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    fused_ssim(a, b)
e1.record()
torch.cuda.synchronize()
print("fused_ssim value+gradient: %.4f ms" % (e0.elapsed_time(e1) / 20))
