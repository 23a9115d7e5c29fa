import torch, sys, os
sys.path.insert(0, "/root/repo")
from explainn_b200 import ops
import bench_paths
dev = torch.device("cuda", 0)
U, k, L, N = 100, 19, 200, 1000000
planes = bench_paths.random_planes(N, L, dev, 1234)
g = torch.Generator().manual_seed(0)
cw = (torch.randn(U, 4, k, generator=g) * 0.3).to(dev)
scale = (torch.rand(U, generator=g) + 0.5).to(dev); shift = (torch.randn(U, generator=g) * 0.3).to(dev)
mx = ops.act_max(planes, L, cw, scale, shift)
thr = (0.5 * mx.half().float()).contiguous()
torch.cuda.synchronize()
print("---- counts", flush=True)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
c = ops.site_counts(planes, L, cw, scale, shift, thr, None)
e1.record(); torch.cuda.synchronize()
print("counts ms", e0.elapsed_time(e1), "sites/seq/unit", float(c.float().mean()))
