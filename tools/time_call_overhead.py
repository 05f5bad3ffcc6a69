#!/usr/bin/env python
"""Where the host time of one small `model(datapoint)` call goes (Zero-DCE++, 1 x 3 x 96 x 96: the GPU work is a few
microseconds, the loop is host-bound):  python tools/time_call_overhead.py"""
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch  # noqa: E402

import etss_07_imageenhance_b200 as P  # noqa: E402
from etss_07_imageenhance_b200._lib import check, lib  # noqa: E402

dev = torch.device("cuda:0")
m = P.ZeroDCEpp_RE(scale_factor=12).to(dev)
x = torch.rand(1, 3, 96, 96, device=dev)
L = lib()
packed = m.packed_weights()
wsb = L.b200_dcepp_workspace_bytes(1, 96, 96, 12)
ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
enh, adj = torch.empty_like(x), torch.empty_like(x)
stream = torch.cuda.current_stream().cuda_stream


def wall(fn, n=2000):
    for _ in range(50):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e6


print(f"model(datapoint)                         {wall(lambda: m({'image': x})):7.1f} us per call")
print(f"torch.ops.b200dce.dcepp_forward          {wall(lambda: torch.ops.b200dce.dcepp_forward(x, packed, 12, 8, True)):7.1f}")
print(f"model.packed_weights() (key compare)     {wall(lambda: m.packed_weights()):7.1f}")
print(f"model._image(datapoint)                  {wall(lambda: m._image({'image': x})):7.1f}")
print(f"3 x torch.empty                          {wall(lambda: (torch.empty_like(x), torch.empty_like(x), torch.empty(wsb, dtype=torch.uint8, device=dev))):7.1f}")
print(f"b200_dcepp_forward through ctypes        {wall(lambda: check(L.b200_dcepp_forward(x.data_ptr(), packed.data_ptr(), enh.data_ptr(), adj.data_ptr(), 1, 96, 96, 12, 8, ws.data_ptr(), wsb, stream), 'fwd')):7.1f}")
g = P.GraphedForward(m, x)
print(f"GraphedForward replay                    {wall(lambda: g()):7.1f}")
