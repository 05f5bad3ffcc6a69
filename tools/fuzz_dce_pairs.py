#!/usr/bin/env python
"""Random shapes through the Zero-DCE bf16 forward: fused e_conv1 + e_conv2 / e_conv4 + e_conv5 launches (default)
against one launch per layer (B200DCE_DCE_FIRST_PAIR=0, B200DCE_DCE_PAIR=0), bit for bit, every case twice:
    python tools/fuzz_dce_pairs.py [cases [seed]]"""
import os
import random
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch  # noqa: E402

import etss_07_imageenhance_b200 as P  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 40
rng = random.Random(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
dev = torch.device("cuda:0")
torch.manual_seed(3)
m = P.ZeroDCE_RE(precision="bf16").to(dev)
for p in m.parameters():
    p.data.mul_(5.0)
bad = 0
for case in range(N):
    B = rng.randint(1, 4)
    H, W = rng.randint(1, 260), rng.randint(1, 700)
    if rng.random() < 0.2:
        H, W = rng.randint(200, 1100), rng.randint(400, 1950)
        B = 1
    x = torch.rand(B, 3, H, W, device=dev) * 0.7
    outs = []
    for fp, pr in (("0", "0"), ("1", "1"), ("1", "0"), ("0", "1"), ("1", "1")):
        os.environ["B200DCE_DCE_FIRST_PAIR"] = fp
        os.environ["B200DCE_DCE_PAIR"] = pr
        o = m({"image": x})
        torch.cuda.synchronize()
        outs.append((o["enhanced"].clone(), o["adjust"].clone()))
    ok = all(torch.equal(o[0], outs[0][0]) and torch.equal(o[1], outs[0][1]) for o in outs[1:])
    bad += not ok
    print(f"case {case:3d}: B={B} {H}x{W}: {'ok' if ok else 'MISMATCH'}", flush=True)
print(f"{N - bad} / {N} cases bit-identical")
sys.exit(1 if bad else 0)
