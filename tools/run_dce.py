#!/usr/bin/env python
"""Zero-DCE forward timing: python tools/run_dce.py precision frames steps [pair 0|1]"""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
prec, frames, steps = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
if len(sys.argv) > 4:
    os.environ["B200DCE_DCE_PAIR"] = sys.argv[4]
import torch  # noqa: E402

import etss_07_imageenhance_b200 as P  # noqa: E402

dev = torch.device("cuda:0")
m = P.ZeroDCE_RE(precision=prec).to(dev)
g = torch.Generator(device=dev).manual_seed(1)
x = torch.rand(frames, 3, 1080, 1920, device=dev, generator=g) * 0.5
for _ in range(3):
    out = m({"image": x})
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(steps):
    out = m({"image": x})
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / steps
px = frames * 1080 * 1920
print(f"zero_dce {prec} {frames}x1080p pair={os.environ.get('B200DCE_DCE_PAIR', '1')}: {ms:.3f} ms/forward = "
      f"{px * 158400 / ms / 1e9:.0f} TFLOP/s algorithmic, checksum {float(out['enhanced'].double().sum()):.6f}")
