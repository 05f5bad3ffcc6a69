#!/usr/bin/env python
"""Where does the Zero-DCE++ step go?  lowres / apply / whole step, eager launches vs CUDA-graph replay."""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
import etss_07_imageenhance_b200 as P
from etss_07_imageenhance_b200._lib import check, lib

F, H, W, SF = 16, 1080, 1920, 12
dev = torch.device("cuda:0")
L = lib()
m = P.ZeroDCEpp_RE(scale_factor=SF).to(dev)
packed = m.packed_weights()
x = torch.rand(F, 3, H, W, device=dev) * 0.4
enh, adj = torch.empty_like(x), torch.empty_like(x)
h, w = H // SF, W // SF
wsb = L.b200_dcepp_workspace_bytes(F, H, W, SF)
ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
a_lr = torch.empty(F, 3, h, w, device=dev)

def lowres():
    check(L.b200_dcepp_lowres(x.data_ptr(), packed.data_ptr(), a_lr.data_ptr(), F, H, W, SF, ws.data_ptr(), wsb,
                              torch.cuda.current_stream().cuda_stream), "lowres")
def apply():
    check(L.b200_dcepp_apply(x.data_ptr(), a_lr.data_ptr(), enh.data_ptr(), adj.data_ptr(), F, H, W, h, w, 8,
                             torch.cuda.current_stream().cuda_stream), "apply")
def step():
    lowres(); apply()

def timeit(fn, n=100):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record()
    for _ in range(n): fn()
    e1.record(); t_enq = time.perf_counter() - t0
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n, t_enq / n * 1e3

for name, fn in (("lowres", lowres), ("apply", apply), ("step", step)):
    gpu, cpu = timeit(fn)
    print(f"{name:8s} eager: {gpu*1e3:8.1f} us/iter on the GPU timeline, CPU enqueue {cpu*1e3:8.1f} us/iter")
s = torch.cuda.Stream()
with torch.cuda.stream(s):
    step(); torch.cuda.synchronize()
    for name, fn in (("lowres", lowres), ("apply", apply), ("step", step)):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            fn()
        gpu, cpu = timeit(g.replay)
        print(f"{name:8s} graph: {gpu*1e3:8.1f} us/iter")
