import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[2]))
import torch
import etss_07_imageenhance_b200 as P
dev = torch.device("cuda:0")
for (B, H, W) in ((16, 1080, 1920), (8, 2160, 3840)):
    h, w = H // 12, W // 12
    x = torch.rand(B, 3, H, W, device=dev) * 0.4
    a = torch.rand(B, 3, h, w, device=dev) * 2 - 1
    for _ in range(3):
        torch.ops.b200dce.dcepp_apply(x, a, 8, True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(50):
        torch.ops.b200dce.dcepp_apply(x, a, 8, True)
    e1.record(); torch.cuda.synchronize()
    print(B, H, W, round(e0.elapsed_time(e1) / 50, 4), "ms")
    del x, a
