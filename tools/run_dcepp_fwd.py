#!/usr/bin/env python
"""Run b200_dcepp_forward N times (for ncu):  python tools/run_dcepp_fwd.py frames H W mode n [adjust]"""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
F, H, W = (int(a) for a in sys.argv[1:4])
os.environ["B200DCE_DCEPP_FUSE"] = sys.argv[4]
N = int(sys.argv[5])
ADJ = len(sys.argv) < 7 or sys.argv[6] != "0"
import torch  # noqa: E402

import etss_07_imageenhance_b200 as P  # noqa: E402

dev = torch.device("cuda:0")
m = P.ZeroDCEpp_RE(scale_factor=12, return_adjust=ADJ).to(dev)
g = torch.Generator(device=dev).manual_seed(1)
x = torch.rand(F, 3, H, W, device=dev, generator=g) * 0.6
for _ in range(N):
    out = m({"image": x})
torch.cuda.synchronize()
print("ok", float(out["enhanced"].mean()))
