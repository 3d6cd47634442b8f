"""Benchmark of the hot path: SpMM-style message passing (copy_src + mean, forward + backward)
on the Reddit-shaped synthetic graph of BASELINE.json configs[1]
(232,965 nodes, 114,615,892 edges, 602 fp32 features; GraphSAGE-mean full-graph aggregation).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--scale S]

One "step" = one forward pass (out = mean over in-edges of X[src]) + one backward pass
(grad_X = transposed SpMM of grad_out / deg) over the whole graph through the public API
(dgl_hack_b200.copy_reduce + autograd). metric = GEdges/s = 2*E*N_gpus / step time (2*E edges per step: E forward + E backward).

The same run also measures the PARTITIONED multi-GPU path (bench_partitioned.py): the
products-shaped graph k-way partitioned over the N ranks, copy_src+sum fwd+bwd and the 3-layer GCN
epoch with the halo exchange, asserted against the 1-GPU kernel -> the `partitioned` object of
the JSON line (strong scaling; this is the multi-GPU result, the Reddit headline at N>1 is N
independent replicas).

Prints ONE JSON line (rank 0). Keys follow the driver's contract; see DESIGN.md "Measurement".
--impl reference times the UNMODIFIED reference CPU kernels (oracle/_ref/libdgl.so, OpenMP/minigun,
all host threads) on a bounded sample of the same workload.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
# the barrier of the partitioned leg is a tiny NCCL kernel: on a high-priority stream it is not
# queued behind the CTAs of the aggregation kernel running beside it
os.environ.setdefault("TORCH_NCCL_HIGH_PRIORITY", "1")

N_NODES, N_EDGES, N_FEATS, ALPHA, SEED = 232965, 114615892, 602, 0.6, 2
WORKLOAD = "reddit_shaped_graphsage_mean_copy_src_fwd_bwd"
METRIC = "GEdges/s (SpMM copy_src+mean fwd+bwd; 2*E edges per step)"


def powerlaw_endpoints(n, e, alpha, gen, device):
    """SURVEY 8(d) generator on the device: endpoint rank r drawn with P(r) ~ (r+1)^-alpha by
    inverting the continuous CDF, then mapped through a seeded permutation."""
    import torch
    u = torch.rand(e, generator=gen, device=device, dtype=torch.float64)
    r = ((n ** (1.0 - alpha) - 1.0) * u + 1.0) ** (1.0 / (1.0 - alpha)) - 1.0
    r = r.floor().clamp_(0, n - 1).to(torch.int64)
    perm = torch.randperm(n, generator=gen, device=device)
    return perm[r]


def make_graph_coo(n, e, alpha, seed, device):
    import torch
    gen = torch.Generator(device=device)
    gen.manual_seed(seed)
    src = powerlaw_endpoints(n, e, alpha, gen, device)
    dst = powerlaw_endpoints(n, e, alpha, gen, device)
    return src, dst


class ClockSampler(object):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 6:
                continue
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except ValueError:
                continue
            for nme, v in zip(names, r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nme)
        return {"sm_mhz": statistics.median(sm) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


def algorithmic_bytes(n_rows_out, n_src, e, d, l2_bytes):
    """SpMM fwd bytes = gather + col idx + out write + indptr (fp32 / int32)."""
    s = n_src * d * 4
    gather = e * d * 4 if s > l2_bytes else s
    return gather + e * 4 + n_rows_out * d * 4 + (n_rows_out + 1) * 4


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def reference_sample(n, src, dst, frac):
    """Bounded sample of the workload for the CPU arm: every edge whose destination is one of the
    first n/frac node ids (all sources and the full feature matrix kept, so gathers are as
    cache-unfriendly as in the full problem)."""
    keep = dst < (n // frac)
    return src[keep], dst[keep]


def time_reference_cpu(n, src_np, dst_np, d, steps, warmup):
    """Times kernel._CAPI_DGLKernelCopyReduce('sum') + BackwardCopyReduce of the unmodified
    reference CPU library on host cores (mean's extra degree pass / division is torch-side in the
    reference and not included: this favours the reference). Returns (seconds/step, cores, kind)."""
    import numpy as np
    from oracle import ref_capi as R
    cores = os.cpu_count() or 1
    if R.available():
        # the launcher may have exported OMP_NUM_THREADS=1 (torch.distributed.run does): the CPU
        # arm always runs on all host threads
        cores = R.set_num_threads(cores)
    rs = np.random.RandomState(0)
    x = rs.uniform(-1, 1, (n, d)).astype(np.float32)
    go = rs.uniform(-1, 1, (n, d)).astype(np.float32)
    if R.available():
        kind = "reference"
        g = R.RefGraph(n, src_np, dst_np)
        g.materialize()
        X, GO = R.RefArray(x), R.RefArray(go)
        out = R.RefArray(shape=x.shape, dtype=np.float32)
        gi = R.RefArray(shape=x.shape, dtype=np.float32)

        def step():
            R.copy_reduce("sum", g, 0, X, out)
            R.backward_copy_reduce("sum", g, 0, X, out, GO, gi)
    else:
        kind, cores = "port", 1
        from oracle import oracle as O
        g = O.Graph(n, n, src_np, dst_np)

        def step():
            o = O.copy_reduce("sum", g, 0, x, n)
            O.backward_copy_reduce("sum", g, 0, x, o, go)
    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    return (time.perf_counter() - t0) / max(steps, 1), cores, kind


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU implementation of the path on host cores."""
    if rank != 0:
        return
    import torch
    n, e, d = max(N_NODES // args.scale, 64), max(N_EDGES // args.scale, 64), N_FEATS
    src, dst = make_graph_coo(n, e, ALPHA, SEED, "cpu")
    # size the per-step sample so that (W + K) steps end within a few minutes on this host:
    # probe on a 1/512 slice, then pick the largest power-of-two slice within the time budget
    frac = args.ref_frac
    if args.scale == 1:
        ps, pt = reference_sample(n, src, dst, 512)
        psec, _, _ = time_reference_cpu(n, ps.numpy(), pt.numpy(), d, 1, 0)
        per_edge = psec / max(int(ps.shape[0]), 1)
        budget = 200.0 / max(args.steps + args.warmup, 1)
        while frac < 512 and per_edge * (e / frac) > budget:
            frac *= 2
    s, t = reference_sample(n, src, dst, frac)
    es = int(s.shape[0])
    sec, cores, kind = time_reference_cpu(n, s.numpy(), t.numpy(), d, args.steps, args.warmup)
    value = 2.0 * es / sec / 1e9
    sample = ("edges into dst rows [0, N/%d) of the same graph: %d edges, all %d sources and the "
              "full [N,%d] feature matrix kept; per step 1 fwd + 1 bwd" % (frac, es, n, d))
    line = {
        "impl": "reference", "metric": METRIC, "value": value,
        "unit": "GEdges/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "nodes": n, "edges": e, "feats": d,
                   "sample_edges": es},
        "cpu_baseline": {"value": value, "unit": "GEdges/s", "cores": cores, "kind": kind,
                         "sample": sample, "omp_threads": cores},
        "e2e": {"value": value, "unit": "GEdges/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    import dgl_hack_b200 as B
    from dgl_hack_b200 import kernel as K

    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    n, e, d = max(N_NODES // args.scale, 64), max(N_EDGES // args.scale, 64), N_FEATS
    # weak scaling: every rank owns an independent shard (its own destination rows and their
    # in-edges, sources local) of identical shape; no data-path collective.
    src, dst = make_graph_coo(n, e, ALPHA, SEED + rank, dev)
    cpu_sample = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        s, t = reference_sample(n, src, dst, args.ref_frac)
        cpu_sample = (s.cpu().numpy(), t.cpu().numpy())
    t0 = time.perf_counter()
    g = B.GraphIndex.from_coo(n, n, src, dst, dev)
    torch.cuda.synchronize()
    index_prep_ms = (time.perf_counter() - t0) * 1e3
    del src, dst
    torch.cuda.empty_cache()

    gen = torch.Generator(device=dev)
    gen.manual_seed(1234 + rank)
    X = (torch.rand((n, d), generator=gen, device=dev) * 2 - 1).requires_grad_(True)
    GO = torch.rand((n, d), generator=gen, device=dev) * 2 - 1

    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
    fwd_pairs = []

    def step(record):
        a, b = (ev(), ev()) if record else (None, None)
        if record:
            a.record()
        out = B.copy_reduce("mean", g, 0, X, n)
        if record:
            b.record()
            fwd_pairs.append((a, b))
        out.backward(GO)
        X.grad = None
        return out

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step(False)
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    l0 = K.launch_count()
    start, end = ev(), ev()
    start.record()
    for _ in range(args.steps):
        step(True)
    end.record()
    barrier()
    launches = K.launch_count() - l0
    ms_total = start.elapsed_time(end)
    fwd_ms = statistics.mean(a.elapsed_time(b) for a, b in fwd_pairs)

    # ---- end to end: host buffers in, host buffers out, copies inside the timed region
    Xh = torch.empty((n, d), dtype=torch.float32, pin_memory=True).copy_(X.detach())
    GOh = torch.empty((n, d), dtype=torch.float32, pin_memory=True).copy_(GO)
    outh = torch.empty((n, d), dtype=torch.float32, pin_memory=True)
    gih = torch.empty((n, d), dtype=torch.float32, pin_memory=True)

    # Two-deep software pipeline, as a training loop with a prefetching loader would run it:
    # step k+1's inputs are copied (copy stream) while step k computes, step k's results are
    # copied back (second copy stream) while step k+1 computes, and the host reads the results of
    # step k once its D2H event fires. EVERY step still moves its own 2 inputs in and its 2
    # results out; the timed region ends when the last result is on the host.
    main_s = torch.cuda.current_stream()
    h2d_s, d2h_s = torch.cuda.Stream(), torch.cuda.Stream()
    Xb = [X, torch.empty_like(X).requires_grad_(True)]
    GOb = [GO, torch.empty_like(GO)]
    free = [torch.cuda.Event(), torch.cuda.Event()]
    for f in free:
        f.record(main_s)
    pending = []

    def e2e_issue(k):
        b = k % 2
        ev_x, ev_g, ev_f, ev_b, done = (torch.cuda.Event() for _ in range(5))
        with torch.cuda.stream(h2d_s):
            h2d_s.wait_event(free[b])  # step k-2 no longer reads these buffers
            with torch.no_grad():
                Xb[b].copy_(Xh, non_blocking=True)
                ev_x.record(h2d_s)
                GOb[b].copy_(GOh, non_blocking=True)
                ev_g.record(h2d_s)
        main_s.wait_event(ev_x)
        out = B.copy_reduce("mean", g, 0, Xb[b], n)
        ev_f.record(main_s)
        main_s.wait_event(ev_g)
        out.backward(GOb[b])
        ev_b.record(main_s)
        grad = Xb[b].grad
        Xb[b].grad = None
        free[b] = torch.cuda.Event()
        free[b].record(main_s)
        with torch.cuda.stream(d2h_s):
            d2h_s.wait_event(ev_f)
            outh.copy_(out.detach(), non_blocking=True)
            d2h_s.wait_event(ev_b)
            gih.copy_(grad, non_blocking=True)
            done.record(d2h_s)
        out.record_stream(d2h_s)
        grad.record_stream(d2h_s)
        pending.append((done, out, grad))

    def e2e_run(steps):
        for k in range(steps):
            e2e_issue(k)
            if len(pending) > 1:
                pending.pop(0)[0].synchronize()  # the caller reads step k-1's host result
        while pending:
            pending.pop(0)[0].synchronize()

    e2e_run(min(args.warmup, 3))
    barrier()
    s2, e2 = ev(), ev()
    s2.record()
    e2e_run(args.steps)
    main_s.wait_stream(d2h_s)
    e2.record()
    barrier()
    ms_e2e = s2.elapsed_time(e2)
    clocks = sampler.stop() if rank == 0 else None

    if world > 1:
        t = torch.tensor([ms_total, ms_e2e], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total, ms_e2e = float(t[0]), float(t[1])

    # ---- second precision tier: bf16 feature storage, fp32 accumulation (same workload; the fp32
    # numbers above stay the headline because the reference is fp32-only)
    bf16 = None
    if not args.no_bf16:
        from dgl_hack_b200 import backend as BF
        Xs = BF.to_bf16_storage(X.detach()).requires_grad_(True)
        GOs = BF.to_bf16_storage(GO)
        pairs16 = []

        def step16(record):
            a, b = (ev(), ev()) if record else (None, None)
            if record:
                a.record()
            o = BF.copy_reduce_bf16("mean", g, Xs)
            if record:
                b.record()
                pairs16.append((a, b))
            o.backward(GOs)
            Xs.grad = None
        for _ in range(args.warmup):
            step16(False)
        barrier()
        s16, e16 = ev(), ev()
        s16.record()
        for _ in range(args.steps):
            step16(True)
        e16.record()
        barrier()
        ms16 = s16.elapsed_time(e16) / args.steps
        f16 = statistics.mean(a.elapsed_time(b) for a, b in pairs16)
        if world > 1:
            t16 = torch.tensor([ms16], device=dev, dtype=torch.float64)
            dist.all_reduce(t16, op=dist.ReduceOp.MAX)
            ms16 = float(t16[0])
        dp = int(Xs.shape[1])
        l2b = torch.cuda.get_device_properties(dev).L2_cache_size
        gather16 = e * dp * 2 if n * dp * 2 > l2b else n * dp * 2
        bytes16 = gather16 + e * 4 + n * dp * 2 + (n + 1) * 4
        bf16 = {"dtype": "bf16 storage, f32 accumulate", "value": 2.0 * e * world / (ms16 * 1e-3) / 1e9,
                "unit": "GEdges/s", "ms_per_step": ms16, "kernel_ms_fwd": f16,
                "row_pitch_elements": dp, "algorithmic_bytes_per_launch": bytes16,
                "achieved_gbs": bytes16 / (f16 * 1e-3) / 1e9, "tolerance": "1e-2 relative (north_star)"}
        del Xs, GOs
        torch.cuda.empty_cache()

    # ---- the partitioned multi-GPU path (one graph over the N ranks), same run
    partitioned = None
    if not args.no_partitioned:
        del X, GO, Xb, GOb, Xh, GOh, outh, gih, g
        torch.cuda.empty_cache()
        from bench_partitioned import run_partitioned
        l1 = K.launch_count()
        try:
            partitioned = run_partitioned(rank, world, dev, scale=args.scale, steps=args.steps,
                                          warmup=args.warmup)
            if partitioned is not None:
                partitioned["gpu_launches_rank0"] = K.launch_count() - l1
        except Exception as ex:  # the headline line must still be printed; the failure is reported
            import traceback
            partitioned = {"error": "%s: %s" % (type(ex).__name__, str(ex)[:500]),
                           "traceback_tail": traceback.format_exc()[-1500:], "n_gpus": world}

    if rank == 0:
        ms_step = ms_total / args.steps
        value = 2.0 * e * world / (ms_step * 1e-3) / 1e9
        e2e_value = 2.0 * e * world / (ms_e2e / args.steps * 1e-3) / 1e9
        l2 = torch.cuda.get_device_properties(dev).L2_cache_size
        peak, peak_src = measured_peak()
        bytes_fwd = algorithmic_bytes(n, n, e, d, l2)
        achieved = bytes_fwd / (fwd_ms * 1e-3) / 1e9
        # DRAM bytes per launch of this kernel from the committed `ncu --set full` capture of the
        # same command (profiles/ncu_traffic.json names the kernel and the commit it was taken on)
        traffic, traffic_src = None, None
        tp = os.path.join(ROOT, "profiles", "ncu_traffic.json")
        if os.path.exists(tp) and args.scale == 1:
            try:
                tj = json.load(open(tp))
                traffic = tj.get("spmm_fwd_dram_bytes_per_launch")
                traffic_src = {k: tj.get(k) for k in ("kernel", "commit", "capture", "when")}
            except Exception:
                traffic = None
        line = {
            "metric": METRIC, "value": value,
            "unit": "GEdges/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "nodes": n, "edges": e, "feats": d,
                       "per_gpu": "one independent shard of this shape per GPU",
                       "l2_policy": "inputs larger than L2 (X %.0f MB, CSR %.0f MB vs L2 %.0f MB)"
                       % (n * d * 4 / 1e6, e * 8 / 1e6, l2 / 1e6),
                       "index_prep_ms_once": index_prep_ms},
            "roofline": {"bound": "hbm", "kernel": "spmm_kernel (forward, copy_src+mean)",
                         "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "frac_algorithmic": achieved / peak,
                         "frac_dram": (traffic / (fwd_ms * 1e-3) / 1e9 / peak) if traffic else None,
                         "traffic": traffic, "traffic_source": traffic_src,
                         "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": bytes_fwd, "kernel_ms": fwd_ms,
                         "gedges_per_s_kernel": e / (fwd_ms * 1e-3) / 1e9},
            "e2e": {"value": e2e_value, "unit": "GEdges/s",
                    "h2d_bytes_per_step": 2 * n * d * 4, "d2h_bytes_per_step": 2 * n * d * 4,
                    "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": launches, "clocks": clocks,
            "partitioned": partitioned, "bf16_tier": bf16,
        }
        if cpu_sample is not None:
            sec, cores, kind = time_reference_cpu(n, cpu_sample[0], cpu_sample[1], d, 1, 1)
            es = int(cpu_sample[0].shape[0])
            line["cpu_baseline"] = {
                "value": 2.0 * es / sec / 1e9, "unit": "GEdges/s", "cores": cores, "kind": kind,
                "sample": "edges into dst rows [0, N/%d) of the same graph (%d edges, all sources "
                          "and the full feature matrix kept), 1 warm-up + 1 timed fwd+bwd"
                          % (args.ref_frac, es)}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--scale", type=int, default=1, help="shrink nodes and edges by this factor")
    ap.add_argument("--ref-frac", type=int, default=64,
                    help="CPU arm works on the edges into the first N/ref_frac destination rows")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-bf16", action="store_true", help="skip the bf16-storage tier")
    ap.add_argument("--no-partitioned", action="store_true",
                    help="skip the partitioned products-shaped multi-GPU leg")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        args.warmup = max(args.warmup, 3)
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
