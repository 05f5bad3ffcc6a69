#!/usr/bin/env python
"""Random shapes through b200_dcepp_forward: the fused launch (B200DCE_DCEPP_FUSE=1, 2) against the three-launch path
(=0), bit for bit, abort word checked, every case repeated (a scheduling race would show as a flaky difference):
    python tools/fuzz_dcepp_fused.py [cases [seed]]"""
import os
import random
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch  # noqa: E402

import etss_07_imageenhance_b200 as P  # noqa: E402
from etss_07_imageenhance_b200._lib import check, lib  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 40
rng = random.Random(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
dev = torch.device("cuda:0")
L = lib()
bad = 0
for case in range(N):
    sf = rng.choice([2, 4, 6, 12, 12, 12])
    B = rng.randint(1, 7)
    h, w = rng.randint(1, 110), rng.randint(1, 120)
    H, W = h * sf, w * sf
    if W % 4:
        W *= 2
        w *= 2
    torch.manual_seed(case)
    m = P.ZeroDCEpp_RE(scale_factor=sf).to(dev)
    for p in m.parameters():
        p.data.mul_(5.0)
    packed = m.packed_weights()
    x = torch.rand(B, 3, H, W, device=dev) * 0.7
    wsb = L.b200_dcepp_workspace_bytes(B, H, W, sf)
    outs = []
    for mode in ("0", "1", "2", "1", "2"):
        os.environ["B200DCE_DCEPP_FUSE"] = mode
        ws = torch.full((wsb,), 0x5A, dtype=torch.uint8, device=dev)   # poisoned workspace
        enh, adj = torch.empty_like(x), torch.empty_like(x)
        check(L.b200_dcepp_forward(x.data_ptr(), packed.data_ptr(), enh.data_ptr(), adj.data_ptr(), B, H, W, sf, 8,
                                   ws.data_ptr(), wsb, torch.cuda.current_stream().cuda_stream), "fwd")
        torch.cuda.synchronize()
        outs.append((enh, adj))
    ok = all(torch.equal(o[0], outs[0][0]) and torch.equal(o[1], outs[0][1]) for o in outs[1:])
    bad += not ok
    print(f"case {case:3d}: B={B} {H}x{W} sf={sf}: {'ok' if ok else 'MISMATCH'}", flush=True)
print(f"{N - bad} / {N} cases bit-identical")
sys.exit(1 if bad else 0)
