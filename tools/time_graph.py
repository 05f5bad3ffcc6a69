#!/usr/bin/env python
"""Zero-DCE++ forward for small batches: eager `model(datapoint)` against a replayed CUDA graph of the same call
(wall clock per call with the device kept busy, i.e. what a one-frame-at-a-time loop sees):
    python tools/time_graph.py [H W]"""
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch  # noqa: E402

import etss_07_imageenhance_b200 as P  # noqa: E402

H, W = (int(a) for a in sys.argv[1:3]) if len(sys.argv) >= 3 else (1080, 1920)
dev = torch.device("cuda:0")
m = P.ZeroDCEpp_RE(scale_factor=12).to(dev)


def wall(fn, n=300):
    for _ in range(20):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e6


for B in (1, 2, 4, 16):
    x = torch.rand(B, 3, H, W, device=dev) * 0.5
    eager = wall(lambda: m({"image": x}))
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        m({"image": x})
    torch.cuda.current_stream().wait_stream(side)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        out = m({"image": x})
    replay = wall(g.replay)
    ref = m({"image": x})
    print(f"{B} x {H} x {W}: eager {eager:7.1f} us per call, graph replay {replay:7.1f} us "
          f"({B * H * W / replay:7.0f} MP/s), same bits: {torch.equal(out['enhanced'], ref['enhanced'])}", flush=True)
