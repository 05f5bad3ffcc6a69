#!/usr/bin/env python
"""Experiment: chunked Zero-DCE++ step with the low-res net of chunk c+1 overlapping the full-resolution pass of
chunk c on a second stream (captured into one CUDA graph).  usage: time_dcepp_overlap.py [frames]"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
import etss_07_imageenhance_b200 as P
from etss_07_imageenhance_b200._lib import check, lib

F = int(sys.argv[1]) if len(sys.argv) > 1 else 16
H, W, SF = 1080, 1920, 12
dev = torch.device("cuda:0")
L = lib()
m = P.ZeroDCEpp_RE(scale_factor=SF).to(dev)
packed = m.packed_weights()
x = torch.rand(F, 3, H, W, device=dev) * 0.4
enh, adj = torch.empty_like(x), torch.empty_like(x)
h, w = H // SF, W // SF
a_lr = torch.empty(F, 3, h, w, device=dev)
fb = 3 * H * W * 4          # bytes per frame
ab = 3 * h * w * 4


def build(chunk, nws):
    n = F // chunk
    wsb = L.b200_dcepp_workspace_bytes(chunk, H, W, SF)
    wss = [torch.empty(wsb, dtype=torch.uint8, device=dev) for _ in range(nws)]
    s0, s1 = torch.cuda.Stream(priority=-1), torch.cuda.Stream(priority=0)

    def run():
        cur = torch.cuda.current_stream()
        s0.wait_stream(cur)
        s1.wait_stream(cur)
        evs = []
        with torch.cuda.stream(s0):
            for c in range(n):
                ws = wss[c % nws]
                check(L.b200_dcepp_lowres(x.data_ptr() + c * chunk * fb, packed.data_ptr(), a_lr.data_ptr() + c * chunk * ab,
                                          chunk, H, W, SF, ws.data_ptr(), wsb, s0.cuda_stream), "lowres")
                e = torch.cuda.Event()
                e.record(s0)
                evs.append(e)
        with torch.cuda.stream(s1):
            for c in range(n):
                s1.wait_event(evs[c])
                check(L.b200_dcepp_apply(x.data_ptr() + c * chunk * fb, a_lr.data_ptr() + c * chunk * ab,
                                         enh.data_ptr() + c * chunk * fb, adj.data_ptr() + c * chunk * fb, chunk, H, W, h, w, 8,
                                         s1.cuda_stream), "apply")
        cur.wait_stream(s0)
        cur.wait_stream(s1)
    return run


def timeit(fn, n=200):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


ref = m({"image": x})
for chunk in (F, F // 2, F // 4, F // 8):
    if chunk < 1:
        continue
    run = build(chunk, 1)
    run()
    torch.cuda.synchronize()
    ok = torch.equal(enh, ref["enhanced"]) and torch.equal(adj, ref["adjust"])
    eager = timeit(run, 50)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            run()
        t = timeit(g.replay)
    print(f"frames={F} chunk={chunk:2d}: eager {eager:7.1f} us  graph {t:7.1f} us/step  ({F*H*W/t:8.0f} MP/s)  bit-equal={ok}")
