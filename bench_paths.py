"""Secondary workloads of bench.py (`--workload cfg1_fwd | cfg3_fwd | cfg4 | cfg5 | metrics`): the inference, interpret
and PWM-scan rows of SURVEY.md section 8(d), each printed as one JSON line in bench.py's format.  The default
bench line (cfg3 training step) does not come through here.

Sharding (section 8e): sequences are split across ranks, no collective inside the timed region; the
interpret workload alone exchanges per-unit maxima / site totals (parallel.interpret_sharded).
"""
import json
import os
import time

import numpy as np
import torch

AUX = {
    # name: (kind, cnn_units|G, kernel|w_max, length, features, per-GPU sequences)
    "cfg1_fwd": ("fwd", 100, 19, 200, 1, 65536),
    "cfg3_fwd": ("fwd", 300, 19, 600, 50, 8192),
    "cfg4": ("interpret", 100, 19, 200, 1, 1000000),
    "cfg5": ("scan", 750, 35, 1000, 0, 12500),
    "metrics": ("metrics", 750, 0, 0, 0, 100000),
}


def _timed(fn, steps, world, dev):
    import torch.distributed as dist
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        fn(i)
    e1.record()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item())


def _randomise_bn(model, seed=3):
    """eval-mode runs randomise the BatchNorm buffers/affine so the folding is exercised."""
    g = torch.Generator().manual_seed(seed)
    for m in model.modules():
        if isinstance(m, torch.nn.BatchNorm1d):
            m.running_mean.copy_(torch.randn(m.num_features, generator=g) * 0.1)
            m.running_var.copy_(torch.rand(m.num_features, generator=g) + 0.5)
            m.weight.data.copy_(torch.rand(m.num_features, generator=g) + 0.5)
            m.bias.data.copy_(torch.randn(m.num_features, generator=g) * 0.1)


def synthetic_pwms(G, wmax, seed=7):
    """G log-probability matrices (G,4,wmax): Dirichlet(0.5) columns floored at 1e-4, widths 6..wmax, padded
    with uniform columns to the common width (the effect of pwm-scoring.py:190-213 on shorter matrices)."""
    rng = np.random.default_rng(seed)
    out = np.full((G, 4, wmax), 0.25)
    for g in range(G):
        w = int(rng.integers(6, wmax + 1))
        cols = np.maximum(rng.dirichlet([0.5] * 4, size=w).T, 1e-4)
        o = (wmax - w) // 2
        out[g, :, o:o + w] = cols
    return np.log(out)


def random_planes(N, L, dev, seed):
    """Packed sequences (uint32 planes lo, hi, N mask = 0) of iid uniform bases, made on the device."""
    Wp = (L + 31) // 32
    g = torch.Generator(device=dev).manual_seed(seed)
    pl = torch.randint(-2 ** 31, 2 ** 31 - 1, (N, 3, Wp), dtype=torch.int32, device=dev, generator=g)
    pl[:, 2] = 0
    if L % 32:
        pl[:, :2, Wp - 1] &= (1 << (L % 32)) - 1
    return pl


def _call_timer(ops):
    events = {}
    orig = ops._call

    def timed_call(name, *a):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        orig(name, *a)
        e1.record()
        events.setdefault(name, []).append((e0, e1))

    return events, orig, timed_call


def _cpu_fwd_rate(U, k, L, F, sample):
    from oracle import explainn_oracle as O
    from explainn_b200.architectures import ExplaiNN
    torch.set_num_threads(os.cpu_count() or 1)
    torch.manual_seed(0)
    m = ExplaiNN(U, k, L, F)
    _randomise_bn(m)
    sd = {k_: v.detach().clone() for k_, v in m.state_dict().items()}
    x = O.random_onehot(sample, L, seed=5)
    O.explainn_forward(sd, x, train=False)
    t0 = time.perf_counter()
    reps = 3
    for _ in range(reps):
        O.explainn_forward(sd, x, train=False)
    return reps * sample / (time.perf_counter() - t0)


def run_fwd(name, args, rank, world, local, peaks, sampler):
    """Inference: model(x) on float one-hot batches (predict.py:81-82), sequences sharded across ranks."""
    from explainn_b200 import ops
    from explainn_b200.architectures import ExplaiNN
    from bench_data import random_onehot
    _, U, k, L, F, B = AUX[name]
    B = args.batch or B
    dev = torch.device("cuda", local)
    torch.manual_seed(0)
    model = ExplaiNN(U, k, L, F)
    _randomise_bn(model)
    model = model.to(dev).eval()
    NB = max(2, int(140e6 // (B * 16 * L)) + 1)
    xs_h = [random_onehot(B, L, seed=1234 + 1000 * rank + i).pin_memory() for i in range(NB)]
    xs_d = [x.to(dev) for x in xs_h]

    def step(i):
        with torch.no_grad():
            return model(xs_d[i % NB])

    for i in range(args.warmup):
        step(i)
    if rank == 0:
        sampler.start()
    events, orig, timed_call = _call_timer(ops)
    ops._call = timed_call
    k0 = ops.kernel_launches
    ms_dev = _timed(step, args.steps, world, dev)
    launches = ops.kernel_launches - k0
    ops._call = orig
    call_ms = {n_: sum(a.elapsed_time(b) for a, b in ev) / args.steps for n_, ev in events.items()}

    # end to end: pinned float one-hot -> device (copy stream, one batch ahead) -> logits -> host, every step
    copy_s = torch.cuda.Stream(dev)
    bufs = [torch.empty_like(xs_d[0]) for _ in range(2)]
    ready = [torch.cuda.Event() for _ in range(2)]
    freed = [torch.cuda.Event() for _ in range(2)]
    out_h = torch.empty((B, F), dtype=torch.float32).pin_memory()

    def upload(i):
        s = i & 1
        with torch.cuda.stream(copy_s):
            copy_s.wait_event(freed[s])
            bufs[s].copy_(xs_h[i % NB], non_blocking=True)
            ready[s].record(copy_s)

    def e2e_all(_):
        for s in range(2):
            freed[s].record()
        upload(0)
        for i in range(args.steps):
            if i + 1 < args.steps:
                upload(i + 1)
            torch.cuda.current_stream().wait_event(ready[i & 1])
            with torch.no_grad():
                out = model(bufs[i & 1])
            freed[i & 1].record()
            out_h.copy_(out, non_blocking=True)
        torch.cuda.synchronize()

    e2e_all(0)
    ms_e2e = _timed(e2e_all, 1, world, dev)
    clocks = sampler.stop() if rank == 0 else None
    if rank != 0:
        return
    hbm, tf_sus, _, which = peaks()
    Lc = L - k + 1
    conv_flops = 2.0 * U * Lc * 4 * k
    fe_ms = call_ms.get("xnn_frontend_fwd", 0.0)
    top = max(call_ms, key=call_ms.get)
    if top == "xnn_frontend_fwd":
        roof = {"kernel": top, "bound": "tensor", "achieved": conv_flops * B / (fe_ms / 1e3) / 1e12, "peak": tf_sus,
                "unit": "TFLOP/s", "what": "conv+BN+exp+maxpool, dense-equivalent 2*U*Lc*4k flop per sequence"}
    else:
        per_seq = 16.0 * L + 12.0 * ((L + 31) // 32) if top == "xnn_pack_onehot_f32" else 4.0 * U * (Lc // 7)
        roof = {"kernel": top, "bound": "hbm", "achieved": per_seq * B / (call_ms[top] / 1e3) / 1e9, "peak": hbm,
                "unit": "GB/s", "what": "bytes per sequence: %.0f" % per_seq}
    roof.update({"frac": roof["achieved"] / roof["peak"], "traffic": None, "peak_source": which,
                 "kernel_ms": call_ms[top], "kernel_share_of_step": call_ms[top] / (ms_dev / args.steps),
                 "whole_forward_vs_input_read": {"bytes_per_seq": 16.0 * L + 4.0 * F, "achieved_gbs":
                                                 (16.0 * L + 4 * F) * B * args.steps / (ms_dev / 1e3) / 1e9,
                                                 "peak_gbs": hbm},
                 "per_call_ms": {k_: round(v, 4) for k_, v in sorted(call_ms.items(), key=lambda kv: -kv[1])}})
    cpu = _cpu_fwd_rate(U, k, L, F, args.cpu_sample * 8)
    print(json.dumps({
        "metric": "sequences/sec ExplaiNN forward (inference)", "value": world * B * args.steps / (ms_dev / 1e3),
        "unit": "sequences/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"ExplaiNN({U} units,k={k},{L}bp,{F} outputs) eval forward on float one-hot",
                   "per_gpu_batch": B, "parallelism": f"sequences sharded over {world} rank(s), no collective",
                   "l2": f"inputs rotate over {NB} batches ({NB * B * 16 * L / 1e6:.0f} MB > 126 MB L2)"},
        "e2e": {"value": world * B * args.steps / (ms_e2e / 1e3), "unit": "sequences/s",
                "h2d_bytes_per_step": B * 16 * L, "d2h_bytes_per_step": B * F * 4,
                "api": "ExplaiNN.forward on pinned float one-hot, upload one batch ahead on a copy stream"},
        "gpu_launches": launches, "roofline": roof,
        "cpu_baseline": {"value": cpu, "unit": "sequences/s", "cores": os.cpu_count(), "kind": "port",
                         "sample": f"3 eval forwards of {args.cpu_sample * 8} sequences, oracle port (torch CPU)"},
        "clocks": clocks}))


def _cpu_interpret_rate(U, k, L, sample):
    """Oracle port of test.py:128-166 + interpret.py:363-454 on `sample` sequences, all host threads for the
    torch part (the site loop is the reference's Python loop, one core)."""
    from oracle import explainn_oracle as O
    from explainn_b200.architectures import ExplaiNN
    torch.set_num_threads(os.cpu_count() or 1)
    torch.manual_seed(0)
    m = ExplaiNN(U, k, L, 1)
    _randomise_bn(m)
    sd = {k_: v.detach().clone() for k_, v in m.state_dict().items()}
    x = O.random_onehot(sample, L, seed=5)
    t0 = time.perf_counter()
    acts16 = O.explainn_forward(sd, x, train=False)["acts"].numpy().astype(np.float16)
    idxs = np.arange(sample)
    thr = O.act_thresholds(acts16, idxs)
    codes = O.codes_from_onehot(x.numpy())
    for u in range(U):
        O.pfm_from_sites(codes, O.sites(acts16[:, u, :], idxs, thr[u]), k)
    return sample / (time.perf_counter() - t0)


def run_interpret(name, args, rank, world, local, peaks, sampler):
    """cfg4: thresholds + site counts + PFMs over N packed sequences per rank (interpret.py:363-454)."""
    from explainn_b200 import ops, parallel
    from explainn_b200.architectures import ExplaiNN, PackedDNA
    _, U, k, L, F, N = AUX[name]
    N = args.batch or N
    dev = torch.device("cuda", local)
    torch.manual_seed(0)
    model = ExplaiNN(U, k, L, F)
    _randomise_bn(model)
    model = model.to(dev).eval()
    packed = PackedDNA(random_planes(N, L, dev, 1234 + rank), L)
    res = {}

    def step(_):
        res["r"] = parallel.interpret_sharded(model, packed)

    for i in range(args.warmup):
        step(i)
    if rank == 0:
        sampler.start()
    events, orig, timed_call = _call_timer(ops)
    ops._call = timed_call
    k0 = ops.kernel_launches
    ms_dev = _timed(step, args.steps, world, dev)
    launches = ops.kernel_launches - k0
    ops._call = orig
    call_ms = {n_: sum(a.elapsed_time(b) for a, b in ev) / args.steps for n_, ev in events.items()}

    # end to end: the packed planes start in pinned host memory; thresholds, counts and PFMs end on the host
    planes_h = packed.planes.cpu().pin_memory()

    def e2e(_):
        pk = PackedDNA(planes_h.to(dev, non_blocking=True), L)
        res["r"] = parallel.interpret_sharded(model, pk)

    e2e(0)
    ms_e2e = _timed(e2e, 1, world, dev)
    clocks = sampler.stop() if rank == 0 else None
    if rank != 0:
        return
    hbm, tf_sus, _, which = peaks()
    Lc = L - k + 1
    top = max(call_ms, key=call_ms.get)
    passes = sum(len(ev) for ev in events.values() if ev) / args.steps
    # every pass is the conv front end on the tensor cores: dense-equivalent 2*U*Lc*4k flop per sequence and pass
    # (SURVEY.md 8d); the bf16x3 weight split issues 3x that.  call_ms[top] sums the chunked calls of one step.
    conv_flops = 2.0 * U * Lc * 4 * k
    n_pass = 1.0                                               # each pass runs once (pass 2a counts are reused)
    achieved = conv_flops * N * n_pass / (call_ms[top] / 1e3) / 1e12
    roof = {"kernel": top, "bound": "tensor", "achieved": achieved, "peak": tf_sus, "unit": "TFLOP/s",
            "frac": achieved / tf_sus, "traffic": None, "peak_source": "bf16_tflops_sustained, " + which,
            "what": "dense-equivalent 2*U*Lc*4k flop per sequence and pass (3x issued: bf16x3 weight split)",
            "kernel_ms": call_ms[top], "kernel_share_of_step": call_ms[top] / (ms_dev / args.steps),
            "hbm_input_gbs": 12.0 * ((L + 31) // 32) * N * n_pass / (call_ms[top] / 1e3) / 1e9, "hbm_peak": hbm,
            "per_call_ms": {k_: round(v, 4) for k_, v in sorted(call_ms.items(), key=lambda kv: -kv[1])}}
    sample = max(64, args.cpu_sample * 4)
    cpu = _cpu_interpret_rate(U, k, L, sample)
    r = res["r"]
    print(json.dumps({
        "metric": "sequences/sec interpret (thresholds + sites + PFMs)", "value": world * N * args.steps / (ms_dev / 1e3),
        "unit": "sequences/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32 compute, f16 compare", "data": "synthetic",
        "config": {"workload": f"interpret: ExplaiNN({U} units,k={k},{L}bp) unit activations -> fp16 thresholds -> "
                               f"sites (1e6 cap) -> PFMs over {N} packed sequences per GPU",
                   "parallelism": f"contiguous sequence shards over {world} rank(s); all-reduce(max) of U maxima, "
                                  f"all-gather of site totals, all-reduce(sum) of PFMs",
                   "l2": f"planes {12 * ((L + 31) // 32) * N / 1e6:.0f} MB per pass; activations never stored",
                   "sites_per_unit_median": float(np.median(r["counts"])), "kernel_passes_per_step": passes},
        "e2e": {"value": world * N / (ms_e2e / 1e3), "unit": "sequences/s", "h2d_bytes_per_step": planes_h.numel() * 4,
                "d2h_bytes_per_step": int(r["pfms"].size * 8 + r["counts"].size * 8 + U * 2),
                "api": "parallel.interpret_sharded(model, PackedDNA from pinned host planes)"},
        "gpu_launches": launches, "roofline": roof,
        "cpu_baseline": {"value": cpu, "unit": "sequences/s", "cores": os.cpu_count(), "kind": "port",
                         "sample": f"{sample} sequences: activations fp16 + thresholds + site loop + PFM counts, "
                                   f"oracle port (torch conv on all cores, the reference's Python site loop on one)"},
        "clocks": clocks}))


def _ctool_compare(dev, mbp=4):
    """C-tool semantics (pwm_scoring.c:340-475, sum mode, both strands): one 19-wide matrix over one record,
    xnn_lpm_scan_sum on the device beside the reference's own ELF (oracle/_ref/pwm_scoring) on one host core."""
    import subprocess
    import tempfile
    from explainn_b200 import ops
    root = os.path.dirname(os.path.abspath(__file__))
    tool = os.path.join(root, "oracle", "_ref", "pwm_scoring")
    rng = np.random.default_rng(3)
    n = mbp * 1000000
    codes = rng.integers(0, 4, size=n, dtype=np.int8)
    lpm = rng.dirichlet([0.6] * 4, size=19).T
    lpm5 = np.ascontiguousarray(np.vstack([lpm, np.full((1, 19), 0.25)]))
    bg = np.array([1.0, 1.0, 1.0, 1.0, 0.25])
    cd = torch.tensor(codes, device=dev)
    od = torch.tensor([0, n], dtype=torch.int64, device=dev)
    l5, bgd = torch.tensor(lpm5, device=dev), torch.tensor(bg, device=dev)
    for _ in range(2):
        s = ops.lpm_scan_sum(cd, od, l5, bgd, False)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        s = ops.lpm_scan_sum(cd, od, l5, bgd, False)
    e1.record()
    torch.cuda.synchronize()
    out = {"matrix_width": 19, "record_mbp": mbp, "gpu_mbp_per_s": 5 * mbp / (e0.elapsed_time(e1) / 1e3),
           "gpu_sum": float(s[0])}
    if os.path.exists(tool):
        with tempfile.TemporaryDirectory() as td:
            fa, mat = os.path.join(td, "g.fa"), os.path.join(td, "m.txt")
            with open(fa, "w") as fh:
                fh.write(">chr\n")
                fh.write(np.array(list("ACGT"))[codes].astype("S1").tobytes().decode())
                fh.write("\n")
            np.savetxt(mat, lpm.T, fmt="%.17g")
            t0 = time.perf_counter()
            r = subprocess.run([tool, "-m", mat, fa], capture_output=True, text=True)
            dt = time.perf_counter() - t0
            out.update({"reference_tool_mbp_per_s": mbp / dt, "reference_tool_cores": 1,
                        "reference_tool_stdout": r.stdout.strip()[:80], "gpu_sum_as_%g": "%g" % float(s[0])})
    else:
        out["reference_tool"] = "oracle/_ref/pwm_scoring not present"
    return out


def run_scan(name, args, rank, world, local, peaks, sampler):
    """cfg5: the PWM module (architectures/__init__.py:116-170), G matrices, both strands, max over positions,
    on fixed-length sequences cut from a synthetic genome; sequences sharded across ranks."""
    from explainn_b200 import ops
    from explainn_b200.architectures import PWM, PackedDNA
    from oracle import explainn_oracle as O
    _, G, w, L, _, N = AUX[name]
    N = args.batch or N
    dev = torch.device("cuda", local)
    pw = synthetic_pwms(G, w)
    model = PWM(pw, L, "max").to(dev)
    packed = PackedDNA(random_planes(N, L, dev, 1234 + rank), L)
    res = {}

    def step(_):
        res["s"] = model(packed)

    for i in range(args.warmup):
        step(i)
    if rank == 0:
        sampler.start()
    events, orig, timed_call = _call_timer(ops)
    ops._call = timed_call
    k0 = ops.kernel_launches
    ms_dev = _timed(step, args.steps, world, dev)
    launches = ops.kernel_launches - k0
    ops._call = orig
    call_ms = {n_: sum(a.elapsed_time(b) for a, b in ev) / args.steps for n_, ev in events.items()}
    planes_h = packed.planes.cpu().pin_memory()
    out_h = torch.empty((N, G), dtype=torch.float32).pin_memory()

    def e2e(_):
        pk = PackedDNA(planes_h.to(dev, non_blocking=True), L)
        out_h.copy_(model(pk), non_blocking=True)
        torch.cuda.synchronize()

    e2e(0)
    ms_e2e = _timed(e2e, 1, world, dev)
    clocks = sampler.stop() if rank == 0 else None
    if rank != 0:
        return
    hbm, tf_sus, _, which = peaks()
    top = max(call_ms, key=call_ms.get)
    adds = 2.0 * G * w * (L - w + 1)                  # one-hot-aware gather-adds per sequence, both strands
    dense = 2.0 * 4.0 * adds                           # dense-equivalent flop (2 flop per MAC, 4 channels)
    achieved = dense * N / (call_ms[top] / 1e3) / 1e12
    roof = {"kernel": top, "bound": "tensor", "achieved": achieved, "peak": tf_sus, "unit": "TFLOP/s",
            "frac": achieved / tf_sus, "traffic": None, "peak_source": "bf16_tflops_sustained, " + which,
            "what": "dense-equivalent 2*2*G*4w*(L-w+1) flop per sequence (3x issued: bf16x3 weight split, "
                    "plus padding of 4w to 16 and of G to 128)",
            "kernel_ms": call_ms[top], "kernel_share_of_step": call_ms[top] / (ms_dev / args.steps),
            "gather_adds_per_s": adds * N / (call_ms[top] / 1e3),
            "per_call_ms": {k_: round(v, 4) for k_, v in sorted(call_ms.items(), key=lambda kv: -kv[1])}}
    torch.set_num_threads(os.cpu_count() or 1)
    sample = max(16, args.cpu_sample // 4)
    xs = O.random_onehot(sample, L, seed=9).numpy()
    O.pwm_module_scores(pw, xs[:4], "max")
    t0 = time.perf_counter()
    O.pwm_module_scores(pw, xs, "max")
    cpu = sample / (time.perf_counter() - t0)
    rate = world * N * args.steps / (ms_dev / 1e3)
    print(json.dumps({
        "metric": "sequences/sec PWM log-odds scan (max over both strands)", "value": rate, "unit": "sequences/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_dev / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"pwm_scoring: {G} PWMs (width {w} after resize) over {N} x {L}-bp packed sequences "
                               f"per GPU ({N * L / 1e6:.1f} Mbp; 8 GPUs = 100 Mbp)",
                   "parallelism": f"sequences sharded over {world} rank(s), no collective",
                   "l2": "input 0.4 MB per Mbp: tensor-bound; weights resident in shared memory",
                   "mbp_per_s": rate * L / 1e6, "mbp_pwm_per_s": rate * L * G / 1e6},
        "e2e": {"value": world * N / (ms_e2e / 1e3), "unit": "sequences/s", "h2d_bytes_per_step": planes_h.numel() * 4,
                "d2h_bytes_per_step": N * G * 4, "api": "PWM.forward(PackedDNA from pinned host planes) -> host scores"},
        "gpu_launches": launches, "roofline": roof,
        "cpu_baseline": {"value": cpu, "unit": "sequences/s", "cores": os.cpu_count(), "kind": "port",
                         "sample": f"{sample} sequences of {L} bp x {G} PWMs, oracle port of PWM.forward (torch CPU)"},
        "ctool": _ctool_compare(dev), "clocks": clocks}))


def run_metrics(name, args, rank, world, local, peaks, sampler):
    """The per-PWM AUC loop of pwm-scoring.py:115-120 (get_metrics: aucROC + aucPR of every score column against one
    y_true): G columns x N sequences per GPU, columns sharded across ranks, no collective."""
    from explainn_b200 import metrics, ops
    from oracle import metrics_oracle as MO
    _, G, _, _, _, N = AUX[name]
    N = args.batch or N
    dev = torch.device("cuda", local)
    g = torch.Generator(device=dev).manual_seed(1234 + rank)
    y = (torch.rand(N, device=dev, generator=g) < 0.5).to(torch.uint8)
    cols = torch.randn((G, N), device=dev, generator=g) + 0.5 * y.float()[None, :] * torch.rand((G, 1), device=dev, generator=g)
    res = {}

    def step(_):
        res["m"] = ops.metric_auc(cols, y)

    for i in range(args.warmup):
        step(i)
    if rank == 0:
        sampler.start()
    events, orig, timed_call = _call_timer(ops)
    ops._call = timed_call
    k0 = ops.kernel_launches
    ms_dev = _timed(step, args.steps, world, dev)
    launches = ops.kernel_launches - k0
    ops._call = orig
    call_ms = {n_: sum(a.elapsed_time(b) for a, b in ev) / args.steps for n_, ev in events.items()}
    scores_h = cols.t().contiguous().cpu().pin_memory()          # (N,G) as pwm-scoring.py holds it
    y_h = y.cpu().pin_memory()

    def e2e(_):
        res["e"] = metrics.column_metrics(y_h.to(dev, non_blocking=True), scores_h.to(dev, non_blocking=True), "binary")

    e2e(0)
    ms_e2e = _timed(e2e, 1, world, dev)
    clocks = sampler.stop() if rank == 0 else None
    if rank != 0:
        return
    hbm, _, _, which = peaks()
    top = max(call_ms, key=call_ms.get)
    if top == "xnn_sort_columns":
        # 4 radix passes x (key + index read twice -- histogram, scatter -- and written once) = 4 x 24 B, minus the
        # first pass reading the float column only and plus the 12-byte result of the last: ~ 92 B per element
        alg, what = 92.0 * N * G, ("segmented LSD radix sort, 4 passes of 8 bits: about 92 B per (column, sequence) "
                                   "(keys and indices read twice and written once per pass)")
    else:
        alg, what = 34.0 * N * G, ("34 B per (column, sequence): sorted score, int64 permutation, gathered label and "
                                   "cumulative positives, read in two passes")
    achieved = alg / (call_ms[top] / 1e3) / 1e9
    roof = {"kernel": top, "bound": "hbm", "achieved": achieved, "peak": hbm, "unit": "GB/s", "frac": achieved / hbm,
            "traffic": None, "peak_source": "hbm " + which,
            "what": what,
            "kernel_ms": call_ms[top], "kernel_share_of_step": call_ms[top] / (ms_dev / args.steps),
            "per_call_ms": {k_: round(v, 4) for k_, v in sorted(call_ms.items(), key=lambda kv: -kv[1])}}
    sample = 8
    sc = scores_h[:, :sample].numpy()
    yt = y_h.numpy()
    t0 = time.perf_counter()
    ref = MO.column_metrics(yt, sc, "binary")
    cpu = sample / (time.perf_counter() - t0)
    err = max(float(np.abs(res["e"][k_][:sample] - np.asarray(ref[k_])).max()) for k_ in ref)
    rate = world * G * args.steps / (ms_dev / 1e3)
    print(json.dumps({
        "metric": "score columns/sec aucROC + aucPR (per-PWM metrics)", "value": rate, "unit": "columns/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_dev / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"pwm-scoring metrics: aucROC + aucPR of {G} score columns x {N} sequences per GPU "
                               "against one y_true", "parallelism": f"columns sharded over {world} rank(s), no collective",
                   "l2": f"columns {G * N * 4 / 1e6:.0f} MB + permutation {G * N * 8 / 1e6:.0f} MB per step (> L2)",
                   "max_abs_err_vs_sklearn_on_sample": err},
        "e2e": {"value": world * G / (ms_e2e / 1e3), "unit": "columns/s", "h2d_bytes_per_step": N * G * 4 + N,
                "d2h_bytes_per_step": G * 16, "api": "metrics.column_metrics(pinned host y_true, scores (N,G)) -> host arrays"},
        "gpu_launches": launches, "roofline": roof,
        "cpu_baseline": {"value": cpu, "unit": "columns/s", "cores": 1, "kind": "reference",
                         "sample": f"{sample} columns x {N} sequences through sklearn roc_auc_score + "
                                   "average_precision_score (the functions the reference calls)"},
        "clocks": clocks}))


def run(name, args, rank, world, local, peaks, sampler):
    from explainn_b200 import parallel
    torch.cuda.set_device(local)
    parallel.init_from_env("nccl")
    kind = AUX[name][0]
    {"fwd": run_fwd, "interpret": run_interpret, "scan": run_scan, "metrics": run_metrics}[kind](name, args, rank, world, local, peaks, sampler)
