#!/usr/bin/env python
"""A/B of the rows-per-segment choice of the per-layer Zero-DCE launches (B200DCE_DCE_ROWS=40: round 1's fixed value;
unset: cost-based), alternating in one process:  python tools/ab_dce_rows.py precision frames [steps]"""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch  # noqa: E402

import etss_07_imageenhance_b200 as P  # noqa: E402

prec, F = sys.argv[1], int(sys.argv[2])
STEPS = int(sys.argv[3]) if len(sys.argv) > 3 else 20
dev = torch.device("cuda:0")
m = P.ZeroDCE_RE(precision=prec).to(dev)
x = torch.rand(F, 3, 1080, 1920, device=dev) * 0.5
ref = None
for rows in ("40", "", "40", "", "40", ""):
    if rows:
        os.environ["B200DCE_DCE_ROWS"] = rows
    else:
        os.environ.pop("B200DCE_DCE_ROWS", None)
    for _ in range(3):
        out = m({"image": x})
    torch.cuda.synchronize()
    if ref is None:
        ref = out["enhanced"].clone()
    same = torch.equal(out["enhanced"], ref)
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(STEPS):
        m({"image": x})
    t1.record()
    torch.cuda.synchronize()
    print(f"{prec} {F} x 1080p rows={rows or 'cost-based'}: {t0.elapsed_time(t1) / STEPS:.3f} ms, bit-equal to the first run: {same}", flush=True)
