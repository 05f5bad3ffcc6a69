#!/usr/bin/env python
"""A few ZeroDCE-V forwards (16 x 1080p, S = 256) for ncu: python tools/run_dcev.py [n]"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch  # noqa: E402

import etss_07_imageenhance_b200 as P  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2
dev = torch.device("cuda:0")
m = P.ZeroDCEV(down_size=256).to(dev)
x = torch.rand(16, 3, 1080, 1920, device=dev) * 0.6
for _ in range(n):
    out = m({"image": x})
torch.cuda.synchronize()
print("ok", float(out["enhanced"].mean()))
