#!/usr/bin/env python
"""Time the RUAS path (CUDA events): python tools/run_ruas.py [frames] [reps] [denoise 0/1]"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import etss_07_imageenhance_b200 as P  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
den = bool(int(sys.argv[3])) if len(sys.argv) > 3 else True
dev = torch.device("cuda:0")
m = P.RUAS_RE(denoise=den).to(dev)
x = torch.rand(B, 3, 1080, 1920, device=dev) * 0.4
pk = m.packed_weights()
for _ in range(2):
    torch.ops.b200dce.ruas_forward(x, pk, den)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize()
e0.record()
for _ in range(reps):
    torch.ops.b200dce.ruas_forward(x, pk, den)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
print(f"ruas denoise={den} {B}x1080p: {ms:.3f} ms  {B * 1080 * 1920 / 1e6 / (ms / 1e3):.0f} MP/s")
