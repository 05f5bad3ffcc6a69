#!/usr/bin/env python
"""SCI forward timing + check against the CPU oracle on a crop: python tools/time_sci.py [frames H W [steps]]"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch  # noqa: E402

import etss_07_imageenhance_b200 as P  # noqa: E402
from oracle import oracle as O  # noqa: E402  (test tool)

F, H, W = (int(a) for a in sys.argv[1:4]) if len(sys.argv) >= 4 else (16, 1080, 1920)
STEPS = int(sys.argv[4]) if len(sys.argv) >= 5 else 200
dev = torch.device("cuda:0")
sd = O.make_sci_state_dict(59, "stress")
packed = P.ops.sci_pack(sd)
for (b, h, w) in [(2, 61, 135), (1, 24, 64), (1, 25, 65), (3, 50, 130), (1, 100, 258)]:
    xs = O.synthetic_lowlight(b, h, w, seed=5)
    il, r = torch.ops.b200dce.sci_forward(xs.to(dev), packed, True)
    ri, rr = O.sci_forward(xs, sd)
    print(f"{b}x{h}x{w}: max|illu - oracle| = {(il.cpu() - ri).abs().max().item():.3e}, "
          f"max|r - oracle| = {(r.cpu() - rr).abs().max().item():.3e}", flush=True)
x = O.synthetic_lowlight(1, H, W, seed=3).to(dev).expand(F, 3, H, W).contiguous()
for _ in range(5):
    il, r = torch.ops.b200dce.sci_forward(x, packed, True)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(STEPS):
    il, r = torch.ops.b200dce.sci_forward(x, packed, True)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / STEPS
print(f"sci {F}x{H}x{W}: {ms:.4f} ms/forward (incl. the op's output allocation) = {F * H * W / ms / 1e6:.1f} GP/s, "
      f"{F * H * W * 36 / ms / 1e6:.0f} GB/s algorithmic")
