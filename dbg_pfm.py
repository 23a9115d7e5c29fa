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
os.environ.pop("XNN_DBG_SAVE", None)
dbg = os.environ.pop("XNN_DBG", None)
mx = ops.act_max(planes, L, cw, scale, shift)
thr = (0.05 * mx.half().float()).contiguous()
c = ops.site_counts(planes, L, cw, scale, shift, thr, None)
carry = torch.zeros(U, dtype=torch.int64, device=dev)
offs = ops.site_offsets(c, carry)
if dbg: os.environ["XNN_DBG"] = dbg
for _ in range(2): ops.site_emit(planes, L, cw, scale, shift, thr, offs, 10**9, None, None, None)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5): pfm = ops.site_emit(planes, L, cw, scale, shift, thr, offs, 10**9, None, None, None)
e1.record(); torch.cuda.synchronize()
print("XNN_DBG", dbg, "pfm ms", e0.elapsed_time(e1) / 5)
