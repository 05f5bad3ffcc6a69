#!/usr/bin/env python
"""ZeroDCE-V forward: error of the tail against the CPU oracle and timing of core / tail:
    python tools/time_dcev.py [frames H W S]"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch  # noqa: E402

import etss_07_imageenhance_b200 as P  # noqa: E402
from oracle import oracle as O  # noqa: E402  (test tool)

F, H, W, S = (int(a) for a in sys.argv[1:5]) if len(sys.argv) >= 5 else (4, 1080, 1920, 256)
dev = torch.device("cuda:0")
sd = O.make_dce_state_dict(29, "stress", num_iters=15, in_channels=1)
m = P.ZeroDCEV(down_size=S)
m.load_state_dict(sd, strict=False)
m = m.to(dev)
x1 = O.synthetic_lowlight(1, H, W, seed=77)
ref = O.zero_dce_v_forward(x1, sd, down_size=S, num_iters=15)
out = m({"image": x1.to(dev)})
for k in ("image_v_fixed_lr", "image_v_fixed", "enhanced"):
    d = (out[k].cpu().double() - ref[k].double()).abs()
    print(f"{k}: max|d| = {d.max().item():.3e}, share of pixels > 1e-4: {(d > 1e-4).double().mean().item():.2e}")
x = O.synthetic_lowlight(1, H, W, seed=78).to(dev).expand(F, 3, H, W).contiguous()


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


core = m.forward_core({"image": x})
t_core = timed(lambda: m.forward_core({"image": x}))
t_tail = timed(lambda: torch.ops.b200dce.dcev_tail(x, core["image_v"], core["image_v_lr"], core["image_v_fixed_lr"]))
t_all = timed(lambda: m({"image": x}))
px = F * H * W
print(f"{F} x {H} x {W}, S = {S}: core {t_core:.3f} ms, tail {t_tail:.3f} ms ({px * 36 / t_tail / 1e6:.0f} GB/s on 36 B/px: "
      f"v 4 in, v_fixed 4 out | rgb 12 + v_fixed 4 in, enhanced 12 out), forward {t_all:.3f} ms = {px / t_all / 1e3:.0f} MP/s")
