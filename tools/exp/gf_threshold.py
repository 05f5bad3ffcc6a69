import os, sys
sys.path.insert(0, "/root/repo")
import torch
import etss_07_imageenhance_b200 as P
dev = torch.device("cuda:0")
def timed(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(n): fn()
    t1.record(); torch.cuda.synchronize()
    return t0.elapsed_time(t1) / n
for (B, H, W) in [(1,1080,1920),(2,1080,1920),(4,1080,1920),(6,1080,1920),(8,1080,1920),(12,1080,1920),(16,1080,1920),(4,2160,3840),(16,540,960)]:
    x = torch.rand(B,3,H,W,device=dev); y = torch.rand(B,3,H,W,device=dev)
    r = []
    for mode in ("2","3"):
        os.environ["B200DCE_GF_FUSED"] = mode
        r.append(timed(lambda: torch.ops.b200dce.guided_filter(x, y, 3, 1e-4)))
    print(f"{B}x{H}x{W}: planes*W = {3*B*W}: chunked {r[0]:.3f} ms, whole columns {r[1]:.3f} ms", flush=True)
