"""GPU, 2 ranks over NCCL (skipped on a 1-GPU box): the data-parallel training step.

  * a 2-rank step with synchronised BatchNorm on the two halves of a batch equals the 1-rank step on the whole batch
    (loss, every gradient, running statistics) -- the reference's single-process numerics at the global batch;
  * the flat-arena bucket (in-place all-reduce, MLP segment overlapped) and the pack/unpack list bucket agree;
  * the CUDA-graphed StepPipeline with the NCCL all-reduce captured inside reproduces the eager data-parallel loop.
"""
import os
import socket

import numpy as np
import pytest
import torch

import _util as U

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    import torch.distributed as dist
    from explainn_b200 import parallel
    from explainn_b200.architectures import get_loss, get_optimizer
    from explainn_b200.trainer import StepPipeline
    from oracle import explainn_oracle as O
    import faulthandler
    faulthandler.dump_traceback_later(int(os.environ.get("XNN_TEST_WATCHDOG", "240")), exit=True)      # a lost collective must fail with a stack, not hang the box
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    parallel.init_from_env("nccl")
    res = {}
    try:
        Uu, k, L, F, B = 40, 19, 200, 5, 64
        x = O.random_onehot(B, L, seed=99, n_frac=0.01)
        y = torch.randn(B, F, generator=torch.Generator().manual_seed(5))
        lo, hi = parallel.shard_range(B, rank, world)
        crit = get_loss("linear")

        def fresh():
            m = U.random_model(Uu, k, L, F).train().to(dev)
            m.linears[9].p = 0.0
            return m

        # ---- 1 rank, whole batch (no exchange) ----
        m1 = fresh()
        crit(m1(x.to(dev)), y.to(dev)).backward()
        g1 = {n_: p.grad.clone() for n_, p in m1.named_parameters()}
        b1 = {n_: b.clone() for n_, b in m1.named_buffers()}

        # ---- 2 ranks, halves, synchronised BatchNorm, arena bucket ----
        parallel.enable_sync_batchnorm()
        m2 = fresh()
        bucket = parallel.GradBucket(m2)
        loss2 = crit(m2(x[lo:hi].to(dev)), y[lo:hi].to(dev))
        loss2.backward()
        bucket.allreduce_mean()
        torch.cuda.synchronize()
        worst = 0.0
        for n_, p in m2.named_parameters():
            a, b = p.grad.cpu().numpy(), g1[n_].cpu().numpy()
            if n_ in U.ZERO_CLASS:
                assert np.abs(a - b).max() < 2e-6, n_
                continue
            U.assert_close(a, b, rtol=1e-4, atol_scale=1e-5, what="sync-BN 2-rank grad " + n_)
            worst = max(worst, float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30)))
        for n_, b in m2.named_buffers():
            U.assert_close(b.cpu().numpy(), b1[n_].cpu().numpy(), rtol=1e-5, what="sync-BN running " + n_)
        res["sync_bn_worst"] = worst
        bucket.close()

        # ---- per-rank statistics: arena bucket vs list bucket ----
        parallel.disable_sync_batchnorm()
        ma, ml = fresh(), fresh()
        ba, bl = parallel.GradBucket(ma), parallel.GradBucket(list(ml.parameters()))
        for m_, b_ in ((ma, ba), (ml, bl)):
            crit(m_(x[lo:hi].to(dev)), y[lo:hi].to(dev)).backward()
            b_.allreduce_mean()
        for (n_, pa), (_, pl) in zip(ma.named_parameters(), ml.named_parameters()):
            U.assert_close(pa.grad.cpu().numpy(), pl.grad.cpu().numpy(), rtol=1e-6, atol_scale=1e-7, what="bucket " + n_)
        ba.close()

        # ---- graphed pipeline with the all-reduce captured vs the eager data-parallel loop ----
        batches = []
        for i in range(6):
            xi = O.random_onehot(32, L, seed=1000 + 10 * i + rank)
            yi = torch.randn(32, F, generator=torch.Generator().manual_seed(2000 + 10 * i + rank))
            batches.append((xi, yi))
        me, mg = fresh(), fresh()
        oe = get_optimizer(me.parameters(), 0.003)
        be = parallel.GradBucket(me)
        eager = []
        for xi, yi in batches:
            loss = crit(me(xi.to(dev)), yi.to(dev))
            oe.zero_grad()
            loss.backward()
            be.allreduce_sum()
            oe.step(grad_scale=be.scale)
            eager.append(loss.item())
        be.close()
        bg = parallel.GradBucket(mg)
        pipe = StepPipeline(mg, crit, get_optimizer(mg.parameters(), 0.003), bg, dev, graph=True)
        got = [pipe.step(xi.pin_memory(), yi.pin_memory()) for xi, yi in batches][1:] + [pipe.flush()]
        U.assert_close(np.array(got), np.array(eager), rtol=1e-5, what="graphed DP losses")
        for (n_, pe), (_, pg) in zip(me.named_parameters(), mg.named_parameters()):
            if n_ not in U.ZERO_CLASS:
                U.assert_close(pg.detach().cpu().numpy(), pe.detach().cpu().numpy(), rtol=1e-4, atol_scale=1e-4,
                               what="graphed DP " + n_)
        bg.close()
        pipe.close()                       # the graph holds captured NCCL kernels: it goes before the process group
        res["ok"] = True
    except Exception as e:                 # report through the manager dict: the parent prints it
        import traceback
        res["error"] = f"{type(e).__name__}: {e}\n{traceback.format_exc()[-1500:]}"
    finally:
        out[rank] = res
        import gc
        gc.collect()
        torch.cuda.synchronize()
        dist.destroy_process_group()


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_two_rank_data_parallel_step():
    import torch.multiprocessing as mp
    out = mp.Manager().dict()
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    assert out[0].get("ok") and out[1].get("ok"), dict(out)
    print("sync-BN 2-rank vs 1-rank worst gradient difference (relative to max):", out[0]["sync_bn_worst"])
