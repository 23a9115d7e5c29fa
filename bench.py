#!/usr/bin/env python
"""bench.py -- headline benchmark of the ExplaiNN hot path on B200.

    python bench.py [--gpus N --steps K --warmup W] [--workload cfg3|cfg2|cfg1] [--batch B] [--impl reference]
                    [--global-batch G (strong scaling)] [--sync-bn] [--no-graph] [--e2e-input ascii|float|planes]
    python bench.py --workload cfg1_fwd|cfg3_fwd|cfg4|cfg5|metrics   (inference / interpret / PWM scan: bench_paths.py)

Metric (BASELINE.json): sequences/s of the ExplaiNN training step (forward + loss + backward + Adam) on synthetic
one-hot DNA.  Default workload = cfg3, the configuration the metric is quoted on at 1/2/4/8 GPUs:
ExplaiNN(300 units, k=19, 600 bp, 50 outputs), MSE, Adam lr 0.003, batch 4096 per GPU (weak scaling; one rank per GPU,
one all-reduce of the flat gradient arena per step, its MLP segment overlapped with the conv weight gradient).

`value`   : K steps with the float one-hot batches already resident in HBM; the step (pack, forward, loss, backward,
            all-reduce, Adam) is ONE CUDA-graph replay plus the encode kernel of the incoming batch.
`e2e`     : the same step through the host-facing API (explainn_b200/trainer.py StepPipeline, the step of
            selene/__init__.py:283-295) with HOST buffers: every step copies its pinned batch to the device (raw ASCII
            bases by default, the column train.py:266-284 reads; `e2e_float_onehot` = the reference's float one-hot
            tensors) and reads its loss back; the copy of batch i+1 runs on a copy stream under step i.
`roofline`: an eager pass brackets every C-ABI call with CUDA events on the launching stream; the entry point with
            the largest share is reported against its bound.  `traffic` is read from the ncu
            capture summary committed under profiles/ (file named in the line), not typed in here.
`cpu_baseline` / `--impl reference`: the REFERENCE MODULE ITSELF (explainn/architectures/__init__.py, placed under
            oracle/_ref/ by build(); `kind: "reference"`) timed on the host cores on a bounded sample of the same
            workload; the oracle port (`kind: "port"`) only if that file is absent.
"""
import argparse
import importlib.util
import json
import os
import statistics
import subprocess
import sys
import threading
import time
import types

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

import bench_data  # noqa: E402

WORKLOADS = {
    # name: (cnn_units, kernel, length, features, loss kind, per-GPU batch)
    "cfg3": (300, 19, 600, 50, "linear", 4096),
    "cfg2": (300, 19, 60, 66, "linear", 4096),
    "cfg1": (100, 19, 200, 1, "binary", 8192),
}
TRAFFIC_FILE = os.path.join(ROOT, "profiles", "r02_ncu_traffic.json")


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("hbm_gbs", 6650.0), d.get("bf16_tflops_sustained", 1400.0), d.get("bf16_tflops", 1590.0), "measured"
    return 6650.0, 1400.0, 1590.0, "fallback"


def ncu_traffic(entry, workload, B):
    """dram bytes per call of `entry` from the committed ncu --set full summary, or None"""
    if not os.path.exists(TRAFFIC_FILE):
        return None, None
    d = json.load(open(TRAFFIC_FILE))
    if d.get("workload") != workload or d.get("per_gpu_batch") != B:
        return None, None
    return d.get("dram_bytes_per_call", {}).get(entry), os.path.relpath(TRAFFIC_FILE, ROOT)


class ClockSampler:
    """nvidia-smi clocks and throttle reasons during the timed region (B200_PROFILING.md)."""

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.index), "-lms", "100"], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = max((int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()), default=None)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 6 for n, v in zip(names, r[2:6]) if v == "Active"})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": reasons,
                "samples": len(sm)}


def synthetic_batches(n_batches, B, L, F, kind, seed):
    """(float one-hot, targets, ASCII) per batch"""
    xs, ys, asc = [], [], []
    g = torch.Generator().manual_seed(seed)
    for i in range(n_batches):
        codes = bench_data.random_codes(B, L, seed + i)
        xs.append(bench_data.onehot_from_codes(codes))
        asc.append(bench_data.ascii_from_codes(codes))
        ys.append(torch.randn(B, F, generator=g) if kind == "linear" else (torch.rand(B, F, generator=g) < 0.5).float())
    return xs, ys, asc


# ---------------------------------------------------------------------------------------------------------
# CPU arm: the reference module itself
# ---------------------------------------------------------------------------------------------------------
def load_reference_module():
    """oracle/_ref/reference_architectures.py = the reference's explainn/architectures/__init__.py (placed by
    `make -C oracle`, git-ignored); `Bio` is stubbed (imported at module level, never touched by the model)."""
    path = os.path.join(ROOT, "oracle", "_ref", "reference_architectures.py")
    if not os.path.exists(path):
        return None
    for name in ("Bio", "Bio.motifs"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["Bio"].motifs = sys.modules["Bio.motifs"]
    spec = importlib.util.spec_from_file_location("reference_architectures", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def cpu_train_step_times(cfg, B, threads, iters, warmup=1):
    """per-step seconds of the full training step (model(x), loss, zero_grad, backward, Adam step -- the body of
    selene/__init__.py:283-292) on the host; returns (times, kind)"""
    U, k, L, F, kind, _ = cfg
    torch.set_num_threads(threads)
    ref = load_reference_module()
    x = bench_data.random_onehot(B, L, 1234)
    g = torch.Generator().manual_seed(1)
    y = torch.randn(B, F, generator=g) if kind == "linear" else (torch.rand(B, F, generator=g) < 0.5).float()
    times = []
    if ref is not None:
        torch.manual_seed(0)
        model = ref.ExplaiNN(U, k, L, F).train()
        crit, opt = ref.get_loss(kind), ref.get_optimizer(model.parameters(), 0.003)
        for it in range(warmup + iters):
            t0 = time.perf_counter()
            loss = crit(model(x), y)
            opt.zero_grad()
            loss.backward()
            opt.step()
            loss.item()
            if it >= warmup:
                times.append(time.perf_counter() - t0)
        return times, "reference"
    from oracle import explainn_oracle as O        # the one other place bench.py may execute oracle/ (CPU arm only)
    from explainn_b200.architectures import ExplaiNN
    torch.manual_seed(0)
    sd = {k_: v.detach().clone() for k_, v in ExplaiNN(U, k, L, F).state_dict().items()}
    mask = (torch.rand(B, U, 100, generator=g) >= 0.3).to(torch.uint8)
    m = {k_: torch.zeros_like(sd[k_]) for k_ in O.PARAM_KEYS}
    v = {k_: torch.zeros_like(sd[k_]) for k_ in O.PARAM_KEYS}
    for it in range(warmup + iters):
        t0 = time.perf_counter()
        r = O.explainn_train_step(sd, x, y, kind, dropout_mask=mask)
        for k_ in O.PARAM_KEYS:
            sd[k_], m[k_], v[k_] = O.adam_step(sd[k_], r["grads"][k_].reshape(sd[k_].shape), m[k_], v[k_], it + 1, 0.003)
        for k_, val in r["running"].items():
            sd[k_] = val
        if it >= warmup:
            times.append(time.perf_counter() - t0)
    return times, "port"


def cpu_baseline(cfg, sample_B):
    """all host threads at the reference's default batch (100, train.py:36-41) and at `sample_B`, and ONE thread at
    the default batch: median of the timed iterations each"""
    cores = os.cpu_count() or 1
    out = {}
    t, kind = cpu_train_step_times(cfg, sample_B, cores, 5)
    out["value"], out["ms_per_step"] = sample_B / statistics.median(t), statistics.median(t) * 1e3
    t100, _ = cpu_train_step_times(cfg, 100, cores, 5)
    t1, _ = cpu_train_step_times(cfg, 100, 1, 3)
    torch.set_num_threads(cores)
    out.update({"unit": "sequences/s", "cores": cores, "kind": kind,
                "sample": f"median of 5 full training steps of {sample_B} sequences, all {cores} host threads, "
                          + ("the reference module itself (oracle/_ref/reference_architectures.py)" if kind == "reference"
                             else "oracle port (reference module file absent)"),
                "batch100_all_threads": {"value": 100 / statistics.median(t100), "unit": "sequences/s",
                                         "iterations": 5, "note": "reference default batch, train.py:36-41"},
                "batch100_one_thread": {"value": 100 / statistics.median(t1), "unit": "sequences/s", "cores": 1,
                                        "iterations": 3}})
    return out


# ---------------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    import bench_paths
    ap.add_argument("--workload", default="cfg3", choices=list(WORKLOADS) + list(bench_paths.AUX))
    ap.add_argument("--batch", type=int, default=0, help="per-GPU batch (default: workload's)")
    ap.add_argument("--global-batch", type=int, default=0, help="strong scaling: total batch split over the ranks")
    ap.add_argument("--cpu-sample", type=int, default=128, help="sequences per CPU-baseline step")
    ap.add_argument("--sync-bn", action="store_true", help="BatchNorm statistics over the GLOBAL batch (reference numerics)")
    ap.add_argument("--no-graph", action="store_true", help="eager launches instead of one CUDA-graph replay per step")
    ap.add_argument("--e2e-input", default="ascii", choices=["ascii", "float", "planes"])
    ap.add_argument("--no-aux", action="store_true", help="skip the secondary forward fields and the CPU baseline")
    ap.add_argument("--sm-reserve", type=int, default=4,
                    help="SMs the conv weight gradient leaves to the overlapped NCCL all-reduce (N > 1)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.workload in bench_paths.AUX:
        assert args.impl == "ours" and torch.cuda.is_available(), "secondary workloads run on the GPU arm only"
        return bench_paths.run(args.workload, args, rank, world, local, peaks, ClockSampler(local))
    cfg = WORKLOADS[args.workload]
    U, k, L, F, kind, B = cfg
    scaling = "weak"
    if args.global_batch:
        B, scaling = args.global_batch // world, "strong"
    elif args.batch:
        B = args.batch
    workload_name = f"ExplaiNN({U} units,k={k},{L}bp,{F} outputs) train step {kind} loss, Adam lr 0.003"

    if args.impl == "reference":
        if rank != 0:
            return
        steps = max(2, min(args.steps, 8))
        cores = os.cpu_count() or 1
        t, rkind = cpu_train_step_times(cfg, args.cpu_sample, cores, steps, warmup=1)
        ms = sum(t) / len(t) * 1e3
        rate = args.cpu_sample / (sum(t) / len(t))
        print(json.dumps({
            "impl": "reference", "metric": "sequences/sec ExplaiNN fwd+bwd training step", "value": rate,
            "unit": "sequences/s", "n_gpus": args.gpus, "steps": steps, "warmup": 1,
            "ms_per_step": ms, "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": {"workload": workload_name, "per_step_sample": args.cpu_sample},
            "cpu_baseline": {"value": rate, "unit": "sequences/s", "cores": cores, "kind": rkind,
                             "sample": f"{steps} full training steps of {args.cpu_sample} sequences each, all {cores} host "
                                       "threads, " + ("the reference module itself (explainn/architectures/__init__.py "
                                                      "placed under oracle/_ref/ by build())" if rkind == "reference"
                                                      else "oracle port (oracle/_ref/reference_architectures.py absent)")},
            "e2e": {"value": rate, "unit": "sequences/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        }), flush=True)
        return

    assert torch.cuda.is_available(), "bench.py needs a GPU (no CPU fallback)"
    from explainn_b200 import ops, parallel
    from explainn_b200.architectures import ExplaiNN, get_loss, get_optimizer
    from explainn_b200.trainer import StepPipeline
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    parallel.init_from_env("nccl")
    import torch.distributed as dist

    torch.manual_seed(0)
    model = ExplaiNN(U, k, L, F).to(dev).train()
    model.linears.input_check = "lazy"
    parallel.broadcast_parameters(model)
    crit = get_loss(kind)
    opt = get_optimizer(model.parameters(), 0.003)
    bucket = parallel.GradBucket(model, sm_reserve=args.sm_reserve) if world > 1 else None
    if args.sync_bn and world > 1:
        parallel.enable_sync_batchnorm()
    use_graph = not args.no_graph

    # inputs rotate over NB distinct batches so that consecutive steps never re-read an L2-resident input
    NB = max(2, int(140e6 // (B * 4 * L * 4)) + 1)
    xs_h, ys_h, asc_h = synthetic_batches(NB, B, L, F, kind, 1234 + 1000 * rank)
    ys_pin = [y.pin_memory() for y in ys_h]
    xs_d = [x.to(dev) for x in xs_h]
    ys_d = [y.to(dev) for y in ys_h]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            fn(i)
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    def eager_step(x, y):
        loss = crit(model(x), y)
        opt.zero_grad()
        loss.backward()
        if bucket is not None:
            bucket.allreduce_sum()
            opt.step(grad_scale=bucket.scale)
        else:
            opt.step()

    # ---- 1. eager pass with CUDA events around every C-ABI call: which entry point dominates ---------------
    for i in range(args.warmup):
        eager_step(xs_d[i % NB], ys_d[i % NB])
    call_events = {}
    orig_call = ops._call

    def timed_call(name, *a):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        orig_call(name, *a)
        e1.record()
        call_events.setdefault(name, []).append((e0, e1))

    n_break = min(20, args.steps)
    ops._call = timed_call
    k0 = ops.kernel_launches
    ms_eager = timed(lambda i: eager_step(xs_d[i % NB], ys_d[i % NB]), n_break)
    kernels_per_step = (ops.kernel_launches - k0) // n_break
    ops._call = orig_call
    call_ms = {nme: sum(a.elapsed_time(b) for a, b in ev) / n_break for nme, ev in call_events.items()}

    # ---- 2. value: device-resident float one-hot batches, one graph replay per step ---------------------------
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    graph_note = None
    pipe_dev = StepPipeline(model, crit, opt, bucket, dev, graph=use_graph)
    try:
        for i in range(max(3, args.warmup)):
            pipe_dev.step_device(xs_d[i % NB], ys_d[i % NB])
        pipe_dev.flush()
    except Exception as e:                       # a capture failure must not cost the round its number: say so
        graph_note = f"graph capture failed ({type(e).__name__}: {str(e)[:200]}); eager launches timed instead"
        torch.cuda.synchronize()
        use_graph = False
        pipe_dev = StepPipeline(model, crit, opt, bucket, dev, graph=False)
        for i in range(3):
            pipe_dev.step_device(xs_d[i % NB], ys_d[i % NB])
        pipe_dev.flush()
    ms_dev = timed(lambda i: pipe_dev.step_device(xs_d[i % NB], ys_d[i % NB]), args.steps)
    pipe_dev.flush()
    graph_kernels = pipe_dev.graph_kernels if use_graph else 0
    pipe_dev.close()

    # ---- 3. e2e: host batches through StepPipeline.step ---------------------------------------------------------
    def host_inputs(mode):
        if mode == "ascii":
            return [a.pin_memory() for a in asc_h], None
        if mode == "float":
            return [x.pin_memory() for x in xs_h], None
        return [ops.pack_onehot(x).cpu().pin_memory() for x in xs_d], L

    def run_e2e(mode, steps):
        xs_pin, length = host_inputs(mode)
        pipe = StepPipeline(model, crit, opt, bucket, dev, graph=use_graph)
        for i in range(4):
            pipe.step(xs_pin[i % NB], ys_pin[i % NB], length=length)
        pipe.flush()
        h0, d0 = pipe.h2d_bytes, pipe.d2h_bytes

        def all_steps(_):
            for i in range(steps):
                pipe.step(xs_pin[i % NB], ys_pin[i % NB], length=length)
            pipe.flush()

        ms = timed(all_steps, 1)
        pipe.close()
        return ms / steps, (pipe.h2d_bytes - h0) // steps, (pipe.d2h_bytes - d0) // steps

    ms_e2e, h2d_step, d2h_step = run_e2e(args.e2e_input, args.steps)
    other = "float" if args.e2e_input != "float" else "ascii"
    ms_e2e2, h2d2, _ = run_e2e(other, max(10, args.steps // 4))
    clocks = sampler.stop() if rank == 0 else None
    model.check_inputs()

    # ---- 4. secondary: eval forward (the metric is "fwd & fwd+bwd") -------------------------------------------
    fwd = {}
    if not args.no_aux:
        model.eval()
        with torch.no_grad():
            for i in range(3):
                model(xs_d[i % NB])
            n_f = 20
            ms_f = timed(lambda i: model(xs_d[i % NB]), n_f)
        model.train()
        fwd = {"eval_forward_same_config": {"value": world * B * n_f / (ms_f / 1e3), "unit": "sequences/s",
                                            "ms_per_batch": ms_f / n_f, "per_gpu_batch": B,
                                            "input": "float one-hot resident in HBM"}}

    if rank != 0:
        return
    value = world * B * args.steps / (ms_dev / 1e3)
    e2e = world * B / (ms_e2e / 1e3)
    hbm, tf_sus, tf_burst, which = peaks()
    Lc = L - k + 1
    n = Lc // 7
    # algorithmic work per sequence of each entry point
    conv_flops = 2.0 * U * Lc * 4 * k                       # dense-equivalent conv
    pool_bytes = 4.0 * U * n
    ALG = {
        "xnn_frontend_fwd": ("tensor", conv_flops, "conv+BN+exp+maxpool forward, dense-equivalent 2*U*Lc*4k flop"),
        "xnn_frontend_bwd": ("hbm", 9.0 * U * n + 12.0 * ((L + 31) // 32),
                             "conv weight gradient: dy+zmax (fp32) + argmax (u8) + planes read"),
        "xnn_mlp_bwd": ("hbm", 3 * pool_bytes + 48.0 * U,
                        "MLP backward, two passes (BN2-backward reductions; dW1/dy): pool read twice, dy written, "
                        "maskbits+dopre read twice"),
        "xnn_mlp_fwd": ("hbm", pool_bytes + 20.0 * U, "MLP forward: pool read, opre+maskbits written"),
        "xnn_mlp_bn2_train_prepare": ("hbm", pool_bytes, "BN2 batch statistics: pool read"),
        "xnn_pack_onehot_f32": ("hbm", 16.0 * L + 12.0 * ((L + 31) // 32), "float one-hot read, planes written"),
    }
    top = max(call_ms, key=call_ms.get)
    top_ms = call_ms[top]

    def roof(entry, ms):
        bound_, per_seq_, what_ = ALG.get(entry, ("hbm", 0.0, "unlisted"))
        if bound_ == "tensor":
            return bound_, per_seq_, what_, per_seq_ * B / (ms / 1e3) / 1e12, tf_sus, "TFLOP/s", "bf16_tflops_sustained"
        return bound_, per_seq_, what_, per_seq_ * B / (ms / 1e3) / 1e9, hbm, "GB/s", "hbm_gbs"

    bound, per_seq, what, achieved, peak, unit, src = roof(top, top_ms)
    traffic, traffic_src = ncu_traffic(top, args.workload, B)
    per_entry = []
    for nme, ms_ in sorted(call_ms.items(), key=lambda kv: -kv[1])[:6]:
        if nme in ALG:
            b_, ps_, _, ach_, pk_, un_, _ = roof(nme, ms_)
            tr_, _ = ncu_traffic(nme, args.workload, B)
            per_entry.append({"entry": nme, "ms": round(ms_, 4), "bound": b_, "algorithmic_per_seq": ps_,
                              "achieved": ach_, "peak": pk_, "unit": un_, "frac": ach_ / pk_, "traffic": tr_,
                              "traffic_gbs": (tr_ / (ms_ / 1e3) / 1e9) if tr_ else None})
    fe_ms = call_ms.get("xnn_frontend_fwd", 0.0)
    cpu = cpu_baseline(cfg, args.cpu_sample) if not args.no_aux else None
    bn_mode = "synchronised BatchNorm statistics (global batch)" if (args.sync_bn and world > 1) else \
        "per-rank BatchNorm statistics"
    line = {
        "metric": "sequences/sec ExplaiNN fwd+bwd training step", "value": value, "unit": "sequences/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_dev / args.steps,
        "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name, "per_gpu_batch": B, "global_batch": B * world,
                   "parallelism": f"dp{world}, {bn_mode}, one all-reduce of the flat gradient arena per step "
                                  f"(MLP segment issued under the conv weight gradient)",
                   "step_launch": ("one CUDA-graph replay per step (+ the encode kernel of the incoming batch)"
                                   if use_graph else "eager launches") + (f"; {graph_note}" if graph_note else ""),
                   "l2": f"inputs rotate over {NB} distinct batches ({NB * B * 16 * L / 1e6:.0f} MB > 126 MB L2); "
                         f"intermediates exceed L2"},
        "e2e": {"value": e2e, "unit": "sequences/s", "ms_per_step": ms_e2e,
                "h2d_bytes_per_step": h2d_step, "d2h_bytes_per_step": d2h_step, "host_input": args.e2e_input,
                "api": "explainn_b200.trainer.StepPipeline.step(pinned host batch, pinned targets): H2D of batch "
                       "i+1 on a copy stream under step i, loss (+ input-check flag) read back every step one call late"},
        "e2e_" + ("float_onehot" if other == "float" else other):
            {"value": world * B / (ms_e2e2 / 1e3), "unit": "sequences/s", "ms_per_step": ms_e2e2,
             "h2d_bytes_per_step": h2d2, "steps": max(10, args.steps // 4)},
        "gpu_launches": (graph_kernels + 1) * args.steps if use_graph else kernels_per_step * args.steps,
        "launches": {"kernels_per_step": graph_kernels + 1 if use_graph else kernels_per_step,
                     "host_launch_calls_per_step": 2 if use_graph else kernels_per_step,
                     "eager_ms_per_step": ms_eager / n_break},
        "roofline": {"kernel": top, "what": what, "bound": bound, "achieved": achieved, "peak": peak, "unit": unit,
                     "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                     "peak_source": f"{src}, {which}",
                     "kernel_ms": top_ms, "kernel_share_of_step": top_ms / (ms_eager / n_break),
                     "algorithmic_per_seq": per_seq, "entries": per_entry,
                     "frontend_fwd": {"bound": "tensor", "kernel_ms": fe_ms,
                                      "achieved": conv_flops * B / (fe_ms / 1e3) / 1e12 if fe_ms else None,
                                      "peak": tf_sus, "unit": "TFLOP/s"},
                     "per_call_ms": {k_: round(v, 4) for k_, v in sorted(call_ms.items(), key=lambda kv: -kv[1])}},
        "clocks": clocks,
    }
    line.update(fwd)
    if cpu is not None:
        line["cpu_baseline"] = cpu
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    if os.environ.get("XNN_BENCH_WATCHDOG"):        # debugging aid: dump every thread's stack and exit after N seconds
        import faulthandler
        faulthandler.dump_traceback_later(int(os.environ["XNN_BENCH_WATCHDOG"]), exit=True)
    main()
    sys.stdout.flush()
    import torch.distributed as _dist
    if _dist.is_available() and _dist.is_initialized():
        # the line is out; tearing NCCL down must never hold the run: give it 20 s, then leave
        import gc
        gc.collect()
        torch.cuda.synchronize()
        _t = threading.Thread(target=_dist.destroy_process_group, daemon=True)
        _t.start()
        _t.join(20.0)
        if _t.is_alive():
            os._exit(0)
