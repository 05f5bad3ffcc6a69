#!/usr/bin/env python
"""Time dce_tc_first_pair_kernel alone (B200DCE_DCE_FIRST_PAIR=only stops the bf16 forward behind its first launch):
    python tools/time_first_pair.py [frames [steps]]"""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
os.environ["B200DCE_DCE_FIRST_PAIR"] = "only"
import torch  # noqa: E402

import etss_07_imageenhance_b200 as P  # noqa: E402
from etss_07_imageenhance_b200._lib import PREC_BF16, check, lib  # noqa: E402

F = int(sys.argv[1]) if len(sys.argv) > 1 else 4
STEPS = int(sys.argv[2]) if len(sys.argv) > 2 else 50
dev = torch.device("cuda:0")
L = lib()
packed = P.ZeroDCE_RE(precision="bf16").to(dev).packed_weights()
x = torch.rand(F, 3, 1080, 1920, device=dev) * 0.5
wsb = L.b200_dce_workspace_bytes(F, 1080, 1920, PREC_BF16)
ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
enh, adj = torch.empty_like(x), torch.empty(F, 24, 1080, 1920, device=dev)


def fwd():
    check(L.b200_dce_forward(x.data_ptr(), packed.data_ptr(), adj.data_ptr(), enh.data_ptr(), F, 1080, 1920, 8, PREC_BF16,
                             ws.data_ptr(), wsb, torch.cuda.current_stream().cuda_stream), "fwd")


for _ in range(5):
    fwd()
torch.cuda.synchronize()
t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0.record()
for _ in range(STEPS):
    fwd()
t1.record()
torch.cuda.synchronize()
us = t0.elapsed_time(t1) / STEPS * 1e3
px = F * 1080 * 1920
print(f"first pair, {F} x 1080p: {us:.1f} us per launch, {px * (12 + 128) / us / 1e6:.2f} TB/s algorithmic (x in, x1 + x2 out)")
