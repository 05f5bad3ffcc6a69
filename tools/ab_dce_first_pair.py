#!/usr/bin/env python
"""A/B of the Zero-DCE bf16 forward: e_conv1 + e_conv2 as one launch (default) vs two (B200DCE_DCE_FIRST_PAIR=0).
Calls the C-ABI with its own (poisoned) workspace and checks bit equality of x1, x2 and of both outputs on a set of
shapes (ragged heights, several segments / strips), then times 16 x 1080p in alternation.

    python tools/ab_dce_first_pair.py [steps]"""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch  # noqa: E402

import etss_07_imageenhance_b200 as P  # noqa: E402
from etss_07_imageenhance_b200._lib import PREC_BF16, check, lib  # noqa: E402
from oracle import oracle as O  # noqa: E402  (test tool: stress weights)

STEPS = int(sys.argv[1]) if len(sys.argv) > 1 else 20
dev = torch.device("cuda:0")
L = lib()
m = P.ZeroDCE_RE(precision="bf16")
m.load_state_dict(O.make_dce_state_dict(7, "stress"), strict=False)
m = m.to(dev)
packed = m.packed_weights()
g = torch.Generator(device=dev).manual_seed(3)


def fwd(x, mode, ws, enh, adj):
    os.environ["B200DCE_DCE_FIRST_PAIR"] = mode
    B, _, H, W = x.shape
    check(L.b200_dce_forward(x.data_ptr(), packed.data_ptr(), adj.data_ptr(), enh.data_ptr(), B, H, W, 8, PREC_BF16,
                             ws.data_ptr(), ws.numel(), torch.cuda.current_stream().cuda_stream), "fwd")


bad = 0
for (B, H, W) in [(1, 8, 8), (1, 40, 128), (2, 41, 124), (1, 37, 252), (3, 83, 256), (1, 120, 380), (2, 1080, 1920),
                  (1, 3, 4), (1, 200, 1000), (1, 45, 128 * 3 + 4)]:
    x = torch.rand(B, 3, H, W, device=dev, generator=g)
    wsb = L.b200_dce_workspace_bytes(B, H, W, PREC_BF16)
    fmap = (B * H * W * 64 + 255) // 256 * 256
    res = []
    for mode in ("0", "1"):
        ws = torch.full((wsb,), 0x7F, dtype=torch.uint8, device=dev)  # bf16 0x7F7F: a huge finite value
        enh, adj = torch.empty_like(x), torch.empty(B, 24, H, W, device=dev)
        fwd(x, mode, ws, enh, adj)
        torch.cuda.synchronize()
        x1 = ws[:B * H * W * 64].view(torch.int16).view(B, H, W, 32)
        x2 = ws[fmap:fmap + B * H * W * 64].view(torch.int16).view(B, H, W, 32)
        res.append((x1, x2, enh, adj))
    names = ("x1", "x2", "enhanced", "adjust")
    eq = [torch.equal(a, b) for a, b in zip(res[0], res[1])]
    bad += not all(eq)
    msg = ", ".join(f"{n} {'==' if e else '!='}" for n, e in zip(names, eq))
    print(f"{B}x{H}x{W}: {msg}", flush=True)
    for n, e, a, b in zip(names, eq, res[0], res[1]):
        if not e and n in ("x1", "x2"):
            d = (a != b).any(dim=3)
            idx = d.nonzero()
            print(f"   {n}: {idx.shape[0]} pixels differ; first {idx[:4].tolist()}, last {idx[-2:].tolist()}; "
                  f"columns {sorted(set(idx[:, 2].tolist()))[:12]}", flush=True)

x = torch.rand(16, 3, 1080, 1920, device=dev, generator=g) * 0.5
wsb = L.b200_dce_workspace_bytes(16, 1080, 1920, PREC_BF16)
ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
enh, adj = torch.empty_like(x), torch.empty(16, 24, 1080, 1920, device=dev)
for mode in ("0", "1", "0", "1", "0", "1"):
    for _ in range(3):
        fwd(x, mode, ws, enh, adj)
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(STEPS):
        fwd(x, mode, ws, enh, adj)
    t1.record()
    torch.cuda.synchronize()
    ms = t0.elapsed_time(t1) / STEPS
    print(f"first_pair={mode}: {ms:.3f} ms / 16 x 1080p = {16 * 1080 * 1920 * 158400 / ms / 1e9:.0f} TFLOP/s algorithmic", flush=True)
sys.exit(1 if bad else 0)
