#!/usr/bin/env python
"""Per-item timeline of the fused Zero-DCE++ launch (B200DCE_TRACE=1): which CTA ran which tile / block when, how long
its consumers waited for the producer.   python tools/trace_dcepp.py frames H W mode [out.npz]"""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
F, H, W = (int(a) for a in sys.argv[1:4])
os.environ["B200DCE_DCEPP_FUSE"] = sys.argv[4]
os.environ["B200DCE_TRACE"] = "1"
import numpy as np  # noqa: E402
import torch  # noqa: E402

import etss_07_imageenhance_b200 as P  # noqa: E402
from etss_07_imageenhance_b200._lib import check, lib  # noqa: E402

SF = 12
dev = torch.device("cuda:0")
L = lib()
m = P.ZeroDCEpp_RE(scale_factor=SF).to(dev)
packed = m.packed_weights()
x = torch.rand(F, 3, H, W, device=dev) * 0.6
enh, adj = torch.empty_like(x), torch.empty_like(x)
need = L.b200_dcepp_workspace_bytes(F, H, W, SF)
extra = 4 << 20
ws = torch.zeros(need + extra, dtype=torch.uint8, device=dev)
for _ in range(4):
    check(L.b200_dcepp_forward(x.data_ptr(), packed.data_ptr(), enh.data_ptr(), adj.data_ptr(), F, H, W, SF, 8,
                               ws.data_ptr(), need + extra, torch.cuda.current_stream().cuda_stream), "fwd")
torch.cuda.synchronize()
tr = ws[need:].view(torch.int32).cpu().numpy().view(np.uint32).reshape(-1, 4)
n = int(tr[0, 0])
rec = tr[1:n + 1].astype(np.int64)
kind, smid, cta = rec[:, 0] & 15, (rec[:, 0] >> 4) & 0xfff, rec[:, 0] >> 16
t0 = rec[:, 2].min()
wait, start, end = (rec[:, 1] - t0) / 1e3, (rec[:, 2] - t0) / 1e3, (rec[:, 3] - t0) / 1e3
print(f"{n} items, span {end.max():.1f} us")
for k in range(7):
    sel = kind == k
    if sel.any():
        print(f"kind {k} ({'block' if k == 6 else 'layer ' + str(k + 2)}): n={sel.sum():6d} first start {start[sel].min():7.1f} "
              f"last end {end[sel].max():7.1f} us | busy mean {np.mean(end[sel] - start[sel]):6.2f} us "
              f"(p50 {np.median(end[sel] - start[sel]):.2f}, p95 {np.percentile(end[sel] - start[sel], 95):.2f}) | "
              f"wait-for-producer mean {np.mean(start[sel] - wait[sel]):6.2f} us")
# utilisation timeline: fraction of CTAs busy with tiles / blocks / waiting per 10 us bin
edges = np.arange(0, end.max() + 10, 10.0)
ncta = len(np.unique(cta))
print(f"{ncta} CTAs; per 10 us bin: CTA-time share in tiles / blocks / waiting for the producer")
for a, b in zip(edges[:-1], edges[1:]):
    def overlap(s, e, sel):
        return np.clip(np.minimum(e[sel], b) - np.maximum(s[sel], a), 0, None).sum()
    tiles = overlap(start, end, kind < 6) / (ncta * (b - a))
    blocks = overlap(start, end, kind == 6) / (ncta * (b - a))
    waits = overlap(wait, start, kind >= 0) / (ncta * (b - a))
    print(f"  {a:6.0f}-{b:6.0f} us: tiles {tiles:5.2f} blocks {blocks:5.2f} wait {waits:5.2f}")
if len(sys.argv) > 5:
    np.savez_compressed(sys.argv[5], rec=rec)
