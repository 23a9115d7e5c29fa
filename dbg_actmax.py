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
for _ in range(2): ops.act_max(planes, L, cw, scale, shift)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(3): ops.act_max(planes, L, cw, scale, shift)
e1.record(); torch.cuda.synchronize()
print("XNN_DBG", os.environ.get("XNN_DBG"), "act_max ms", e0.elapsed_time(e1) / 3)
