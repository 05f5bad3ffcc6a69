#!/usr/bin/env python
"""A few Zero-DCE++ low-resolution-net launches (16 x 1080p, sf=12) for ncu: python tools/run_lowres.py [n]."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
import etss_07_imageenhance_b200 as P
from etss_07_imageenhance_b200._lib import check, lib

F, H, W, SF = 16, 1080, 1920, 12
n = int(sys.argv[1]) if len(sys.argv) > 1 else 3
dev = torch.device("cuda:0")
L = lib()
m = P.ZeroDCEpp_RE(scale_factor=SF).to(dev)
packed = m.packed_weights()
x = torch.rand(F, 3, H, W, device=dev) * 0.4
wsb = L.b200_dcepp_workspace_bytes(F, H, W, SF)
ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
a_lr = torch.empty(F, 3, H // SF, W // SF, device=dev)
for _ in range(n):
    check(L.b200_dcepp_lowres(x.data_ptr(), packed.data_ptr(), a_lr.data_ptr(), F, H, W, SF, ws.data_ptr(), wsb,
                              torch.cuda.current_stream().cuda_stream), "lowres")
torch.cuda.synchronize()
print("ok", float(a_lr.abs().mean()))
