#!/usr/bin/env python
"""Run the Zero-DCE forward a few times (for ncu / timing).  usage: run_dce.py [precision] [frames] [reps]"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
import etss_07_imageenhance_b200 as P

prec = sys.argv[1] if len(sys.argv) > 1 else "bf16"
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 4
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
dev = torch.device("cuda:0")
m = P.ZeroDCE_RE(precision=prec).to(dev)
g = torch.Generator(device=dev).manual_seed(1)
x = (0.02 + 0.3 * torch.rand(frames, 3, 1080, 1920, device=dev, generator=g)).contiguous()
for _ in range(3):
    out = m({"image": x})
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    out = m({"image": x})
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
px = frames * 1080 * 1920
print(f"{prec} frames={frames}: {ms:.3f} ms/forward, {px / 1e6 / (ms / 1e3):.0f} MP/s, "
      f"{px * 158400 / (ms / 1e3) / 1e12:.1f} TFLOP/s algorithmic")
