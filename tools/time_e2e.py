#!/usr/bin/env python
"""Sweep FramePipeline chunking for the host-fed (e2e) path: 16 x 1080p fp32 in / out of pinned host memory."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
import etss_07_imageenhance_b200 as P

F, H, W = 16, 1080, 1920
dev = torch.device("cuda:0")
m = P.ZeroDCEpp_RE(scale_factor=12).to(dev)
hin = torch.rand(F, 3, H, W).mul_(0.4).pin_memory()
hout = torch.empty(F, 3, H, W).pin_memory()
for wire in (sys.argv[1:] or ("f32", "u8")):
    if wire == "u8":
        hin8 = (hin.permute(0, 2, 3, 1) * 255).to(torch.uint8).contiguous().pin_memory()
        hout8 = torch.empty_like(hin8).pin_memory()
    for chunk, slots in ((1, 2), (1, 3), (1, 4), (1, 6), (2, 2), (2, 3), (2, 4), (2, 6), (4, 2), (4, 3), (4, 4), (4, 6), (4, 8), (2, 8), (8, 2), (8, 4)):
        pipe = P.FramePipeline(m, frames_per_chunk=chunk, H=H, W=W, wire=wire, slots=slots)
        a, b = (hin, hout) if wire == "f32" else (hin8, hout8)
        for _ in range(2):
            pipe.run(a, b)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(8):  # streaming form (the bench's e2e leg): the slot ring turns across steps
            pipe.run(a, b, wait=False)
        pipe.drain()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 8
        print(f"wire={wire} chunk={chunk:2d} slots={slots}: {ms:7.3f} ms/step  {F*H*W/1e6/(ms/1e3):8.0f} MP/s  "
              f"{a.numel()*a.element_size()/1e9/(ms/1e3):5.1f} GB/s each way", flush=True)
        del pipe
