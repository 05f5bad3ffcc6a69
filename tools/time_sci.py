import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
import etss_07_imageenhance_b200 as P
dev = torch.device("cuda:0")
m = P.SCI_RE().to(dev)
x = torch.rand(16, 3, 1080, 1920, device=dev) * 0.4
for _ in range(3): m({"image": x})
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20): m({"image": x})
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 20
print(f"SCI 16x1080p: {ms:.3f} ms, {16*1080*1920/1e6/(ms/1e3):.0f} MP/s, {16*1080*1920*36/(ms/1e3)/1e9:.0f} GB/s algorithmic")
