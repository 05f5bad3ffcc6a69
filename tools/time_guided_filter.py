#!/usr/bin/env python
"""GCENet's post-network path (curve loop + guided filter) timing: python tools/time_guided_filter.py [frames H W]"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch  # noqa: E402

import etss_07_imageenhance_b200 as P  # noqa: E402,F401

F, H, W = (int(a) for a in sys.argv[1:4]) if len(sys.argv) >= 4 else (16, 1080, 1920)
dev = torch.device("cuda:0")
x = torch.rand(F, 3, H, W, device=dev) * 0.6
adj = torch.rand(F, 3, H, W, device=dev) * 2 - 1


def timed(fn, n=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(n):
        fn()
    t1.record()
    torch.cuda.synchronize()
    return t0.elapsed_time(t1) / n


import os  # noqa: E402

enh = torch.ops.b200dce.gce_curve(x, adj, 15, 2.6, False)[0]
bad = 0
for (b, c, cx, hh, ww, rad) in [(1, 3, 3, 1080, 1920, 3), (2, 3, 1, 67, 131, 3), (1, 1, 1, 17, 300, 7), (3, 3, 3, 40, 33, 5),
                                (1, 2, 2, 300, 2500, 9), (1, 1, 1, 100, 4096, 21), (2, 3, 3, 200, 1022, 5), (16, 3, 3, 300, 1920, 3), (8, 3, 1, 1080, 2500, 9)]:
    xs = torch.rand(b, cx, hh, ww, device=dev)
    ys = torch.rand(b, c, hh, ww, device=dev)
    res = []
    for mode in ("0", "2", ""):
        if mode:
            os.environ["B200DCE_GF_FUSED"] = mode
        else:
            os.environ.pop("B200DCE_GF_FUSED", None)
        res.append(torch.ops.b200dce.guided_filter(xs, ys, rad, 1e-4))
    eq = torch.equal(res[0], res[1]) and torch.equal(res[0], res[2])
    bad += not eq
    print(f"{b}x{c}({cx})x{hh}x{ww} radius {rad}: four passes {'==' if eq else '!='} six passes"
          + ("" if eq else f" (max |d| {(res[0] - res[1]).abs().max().item():.3e})"), flush=True)
for mode in ("0", "2", ""):
    if mode:
        os.environ["B200DCE_GF_FUSED"] = mode
    else:
        os.environ.pop("B200DCE_GF_FUSED", None)
    print(f"B200DCE_GF_FUSED={mode or 'unset'}: guided_filter {timed(lambda: torch.ops.b200dce.guided_filter(x, enh, 3, 1e-4)):.3f} ms", flush=True)
t_c = timed(lambda: torch.ops.b200dce.gce_curve(x, adj, 15, 2.6, False))
t_g = timed(lambda: torch.ops.b200dce.guided_filter(x, enh, 3, 1e-4))
px = F * H * W
print(f"{F} x {H} x {W}: gce_curve {t_c:.3f} ms ({px * 36 / t_c / 1e6:.0f} GB/s on 36 B/px), guided_filter {t_g:.3f} ms "
      f"({px * 36 / t_g / 1e6:.0f} GB/s on the algorithmic 36 B/px; four passes + two totals passes)")
sys.exit(1 if bad else 0)
