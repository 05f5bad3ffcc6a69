#!/usr/bin/env python
"""Small Zero-DCE++ forward against the oracle (debug helper; run under compute-sanitizer when a kernel faults)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
import etss_07_imageenhance_b200 as P
from oracle import oracle as O

B, H, W, SF = (int(v) for v in (sys.argv[1:5] + ["2", "96", "120", "12"][len(sys.argv) - 1:]))
torch.manual_seed(0)
dev = torch.device("cuda:0")
m = P.ZeroDCEpp_RE(scale_factor=SF)
with torch.no_grad():
    for p in m.parameters():
        if p.dim() == 4 and p.shape[-1] * p.shape[-2] * p.shape[1] > 1:
            p.normal_(0, 0.25)
sd = {k: v.clone() for k, v in m.state_dict().items()}
x = torch.rand(B, 3, H, W) * 0.5
ref = O.zero_dcepp_forward(x, sd, scale_factor=SF)
out = m.to(dev)({"image": x.to(dev)})
torch.cuda.synchronize()
for k in ("adjust", "enhanced"):
    print(k, "max abs err", (out[k].cpu() - ref[k]).abs().max().item(), "ref max", ref[k].abs().max().item())
