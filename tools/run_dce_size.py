#!/usr/bin/env python
"""python tools/run_dce_size.py frames H W pair steps"""
import os, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
F, H, W = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
os.environ["B200DCE_DCE_PAIR"] = sys.argv[4]
steps = int(sys.argv[5])
import torch
import etss_07_imageenhance_b200 as P
dev = torch.device("cuda:0")
m = P.ZeroDCE_RE(precision="bf16").to(dev)
x = torch.rand(F, 3, H, W, device=dev) * 0.5
for _ in range(steps):
    out = m({"image": x})
torch.cuda.synchronize()
print("ok")
