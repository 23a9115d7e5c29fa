import torch, sys, os
sys.path.insert(0, "/root/repo")
from explainn_b200 import ops
import bench_paths
dev = torch.device("cuda", 0)
U, k, L, N = 100, 19, 200, 65536
planes = bench_paths.random_planes(N, L, dev, 1234)
g = torch.Generator().manual_seed(0)
cw = (torch.randn(U, 4, k, generator=g) * 0.3).to(dev)
scale = (torch.rand(U, generator=g) + 0.5).to(dev); shift = (torch.randn(U, generator=g) * 0.3).to(dev)
mx = ops.act_max(planes, L, cw, scale, shift)
thr = (0.05 * mx.half().float()).contiguous()
c = ops.site_counts(planes, L, cw, scale, shift, thr, None)
csum = torch.cumsum(c.to(torch.int64), 1); offs = (csum - c).contiguous()
torch.cuda.synchronize()
print("sites/seq/unit", float(c.float().mean()), flush=True)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
pfm = ops.site_emit(planes, L, cw, scale, shift, thr, offs, 10**9, None, None, None)
e1.record(); torch.cuda.synchronize()
print("pfm ms", e0.elapsed_time(e1))
