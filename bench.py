#!/usr/bin/env python
"""bench.py — replication task-fits/sec on B200 (BASELINE.json metric), one process per GPU.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--seeds S] [--impl reference]

A *step* is one pass of the hot path over one batch of synthetic input of the named shape
(configs[1] of BASELINE.json: 337 months x 22 factor ETFs / 13 fund indices / rf; the notebook's
sweep L = 1..21, deterministic row order so every fit runs all 1000 epochs; all 13 funds x 145
rolling windows evaluated per fit): kernel (b) fits S x 21 autoencoders, kernel (c) scores them.
S = 1 seed per GPU by default (configs[1]); every rank of a multi-GPU run takes its own seeds
(weak scaling, no data-path collective; one result all-gather outside the hot path).
The 64-seed ensemble of configs[2] is measured in the same run and reported under "ensemble".

`--impl reference` times the CPU restatement of the reference path (oracle/ae_oracle.py; the
reference itself cannot travel to the GPU box and Keras is not installable) on all host cores.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "replication task-fits/sec"
UNIT = "fits/s"
T_ALL, E, F, WINDOW = 337, 22, 13, 24
LATENTS = list(range(1, 22))
EPOCHS = 1000
# dram__bytes_read.sum + dram__bytes_write.sum per fit_kernel launch of the cfg2 sweep, from the committed
# `ncu --set full` capture (profiles/r1_ncu_full_cfg2_summary.txt): the kernel is shared-memory resident.
NCU_DRAM_BYTES_PER_FIT_LAUNCH = 80e3


def synthetic_returns(seed=20240611):
    """Synthetic monthly returns of the reference's shape: K=5 Gaussian factors drive 22 ETFs; 13 funds
    are sparse ETF mixes plus noise; rf constant-ish.  (No network: the real CSVs cannot be assumed.)"""
    rng = np.random.default_rng(seed)
    fac = rng.normal(size=(T_ALL, 5)) * .03
    etf = fac @ rng.normal(size=(5, E)) + rng.normal(size=(T_ALL, E)) * .01
    beta = rng.normal(size=(E, F)) * (rng.uniform(size=(E, F)) < .25) * .3
    hfd = etf @ beta + rng.normal(size=(T_ALL, F)) * .004
    rf = np.abs(rng.normal(size=T_ALL)) * 1e-3 + 1e-3
    return etf, hfd, rf


def flops_per_epoch(L, T_fit=126, T_val=42, batch=48):
    """Algorithmic flop of one epoch of one fit."""
    steps = -(-T_fit // batch)
    mm = 10 * T_fit * E * L + 4 * T_val * E * L
    ew = 2 * T_fit * (L + E) + 3 * T_fit * E + T_val * (L + 4 * E)
    return mm + ew + 28 * E * L * steps


# --------------------------------------------------------------------------------------- CPU arm
def _cpu_task(args):
    os.environ["OMP_NUM_THREADS"] = "1"
    L, seed = args
    from oracle import ae_oracle, fitcore
    etf, hfd, rf = synthetic_returns()
    We0, Wd0 = fitcore.initial_weights(seed, E, L)
    cfg = fitcore.FitConfig(epochs=EPOCHS, early_stop=True)
    t = time.perf_counter()
    ae_oracle.run_task(etf[:168], hfd[:168], etf[168:], hfd[168:], rf, etf, L, We0, Wd0, cfg=cfg)
    return time.perf_counter() - t


def cpu_reference_arm(args):
    """The reference path's CPU restatement on all host cores: each step = the 21-fit sweep."""
    import multiprocessing as mp
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = len(os.sched_getaffinity(0))
    tasks = [(L, 123) for L in LATENTS]
    ctx = mp.get_context("fork")
    times = []
    with ctx.Pool(min(cores, len(tasks))) as pool:
        for i in range(args.warmup + args.steps):
            t = time.perf_counter()
            pool.map(_cpu_task, tasks, chunksize=1)
            dt = time.perf_counter() - t
            if i >= args.warmup:
                times.append(dt)
    total = sum(times)
    value = len(tasks) * len(times) / total
    sample = f"{len(times)} x full sweep L=1..21 (21 fits, 1000 epochs each, all evaluation stages), numpy float32/float64"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "cfg2: 13 indices x 145 windows x L=1..21 x 1 seed (Mode R), synthetic 337x(22+13+1)",
                   "epochs": EPOCHS, "shuffle": False},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": min(cores, len(tasks)), "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def cpu_baseline_single_core(budget_s=20.):
    """Oracle port on ONE core over a bounded sample of the same sweep (about `budget_s` seconds)."""
    sample, spent = [], 0.
    for L in (1, 4, 8, 11, 14, 18, 21, 2, 6, 16, 20, 3, 9, 12, 5, 7, 10, 13, 15, 17, 19):
        dt = _cpu_task((L, 123))
        sample.append(L)
        spent += dt
        if spent > budget_s:
            break
    return {"value": len(sample) / spent, "unit": UNIT, "cores": 1, "kind": "port",
            "sample": f"L in {sorted(sample)} of the cfg2 sweep, full pipeline, {spent:.1f} s on one core"}


def cpu_best_effort():
    """the fit alone as compiled C (oracle/c/fit_oracle.c, gcc -O2 -march=native, OpenMP over the 21
    tasks of the sweep) on all host cores — shown so the speed-up is not only "Python overhead removed".  The
    evaluation stages are NOT included, so this over-states the CPU's task-fits/sec."""
    try:
        from oracle import c_oracle, fitcore, ae_oracle
        etf, _, _ = synthetic_returns()
        x = ae_oracle.minmax_fit_transform(etf[:168]).astype(np.float32)
        W = np.zeros((len(LATENTS), 2 * E * max(LATENTS)), np.float32)
        for i, L in enumerate(LATENTS):
            a, b = fitcore.initial_weights(123, E, L)
            W[i, :E * L], W[i, E * L:2 * E * L] = a.ravel(), b.ravel()
        c_oracle.fit_many(x, LATENTS, W.copy(), epochs=50)                      # warm-up
        t = time.perf_counter()
        _, _, ep = c_oracle.fit_many(x, LATENTS, W, epochs=EPOCHS)
        dt = time.perf_counter() - t
        return {"value": len(LATENTS) / dt, "unit": UNIT, "cores": len(os.sched_getaffinity(0)), "kind": "port-c, fit only",
                "sample": f"21 fits of the sweep, {float(ep.mean()):.0f} epochs on average, no evaluation stages, {dt:.2f} s"}
    except Exception as exc:                                # no gcc on the box: report why, do not fail the bench
        return {"unavailable": str(exc)[:200]}


# --------------------------------------------------------------------------------------- clocks
class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.FIELDS}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [c.strip() for c in line.split(",")]))

    def stop(self, t_from=None):
        """Samples taken since t_from (the start of the timed region).  nvidia-smi needs a few hundred ms to start, so
        the sampler is started before the warm-up steps; if a short timed region caught no sample, the samples of the
        warm-up (the same steps, the same load) are reported and `window` says so."""
        if self.proc:
            self.proc.terminate()
        rows = [r for t, r in self.rows if t_from is None or t >= t_from]
        window = "timed region"
        if not rows:
            rows, window = [r for _, r in self.rows], "warm-up + timed region (the timed region was shorter than one sampling period)"
        sm = [float(r[0]) for r in rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in rows if len(r) >= 6 for n, v in zip(names, r[2:6]) if v.lower() == "active"})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm), "window": window}


# --------------------------------------------------------------------------------------- GPU arm
def gpu_arm(args):
    import torch
    import torch.distributed as dist
    import hfrep
    from hfrep import engine

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    etf, hfd, rf = synthetic_returns()
    # Weak scaling needs the same work on every GPU.  Early stopping makes a sweep's length depend on its seeds (measured:
    # 7.1 to 8.8 ms per 21-fit sweep over four seeds), so every rank runs the sweep of the SAME seeds; the ranks' results are
    # still separate task executions (the job's aggregate counts n_gpus x fits_per_step_per_gpu fits per step).
    seeds = [123 + s for s in range(args.seeds)]
    data = hfrep.ReplicationData(etf, hfd, rf, window=WINDOW, device=dev)
    tasks = sorted(((L, s) for s in seeds for L in LATENTS), key=lambda t: -((t[0] + 3) // 4))   # widest first
    n_tasks = len(tasks)
    w_stride = 2 * E * max(LATENTS)
    W0_host = engine.pack_weights([engine.initial_weights(s, E, L) for L, s in tasks], E, w_stride)
    W0_pinned = torch.from_numpy(W0_host).pin_memory()
    W0d = W0_pinned.to(dev)
    task_L = [L for L, _ in tasks]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)      # > 126 MB L2
    n_buckets = len({(L + 3) // 4 for L in task_L})

    def resident_step(record=None):
        """Inputs already in HBM: kernel (b) and kernel (c) of every width bucket, buckets concurrent."""
        if record:
            record[0].record()
        fit, ev = engine.fit_and_score(data, task_L, W0d, early_stop=1)      # each bucket's kernel (c) right behind its kernel (b)
        if record:
            record[2].record()
        return fit, ev

    def e2e_step():
        """Through the public API with HOST buffers: H2D of the returns + initial kernels, D2H of the results."""
        d = hfrep.ReplicationData(etf, hfd, rf, window=WINDOW, device=dev)
        _, fit, ev = hfrep.run_mode_r(d, LATENTS, seeds=seeds, W0=W0_host)      # same widest-first task order as W0_host
        out = [t.cpu() for t in (ev.ante, ev.post, ev.turnover, ev.metrics, ev.quality, fit.epochs_run)]
        h2d = d.h2d_bytes + W0_host.nbytes
        return h2d, sum(t.numel() * t.element_size() for t in out)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- warm-up (the clock sampler starts here: see ClockSampler.stop)
    sampler = ClockSampler(local)
    sampler.start()
    for _ in range(max(args.warmup, 3)):
        resident_step()
        flush.fill_(1)
    t_wait = time.perf_counter()
    while len(sampler.rows) < 2 and sampler.proc is not None and time.perf_counter() - t_wait < 5.:
        resident_step()                      # more (untimed) warm-up until the sampler has seen the GPU under this load
        flush.fill_(1)
        torch.cuda.synchronize()
    e2e_step()
    barrier()

    # ---- timed: K steps, device time per step from CUDA events (L2 flushed between steps, flush not timed)
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(args.steps)]
    barrier()
    t_wall = t_timed = time.perf_counter()
    last = None
    for k in range(args.steps):
        last = resident_step(evs[k])
        flush.fill_(k & 1)
    barrier()
    wall = time.perf_counter() - t_wall
    step_ms = [e[0].elapsed_time(e[2]) for e in evs]
    total_ms = float(sum(step_ms))
    fit, ev = last
    epochs_run = fit.epochs_run.cpu().numpy()
    assert int(fit.status.abs().sum()) == 0 and int(ev.status.abs().sum()) == 0
    assert bool(torch.isfinite(ev.ante).all())

    # ---- kernel (b) alone (its evaluation is otherwise queued behind it, bucket by bucket): the roofline's duration
    fit_ms = []
    for k in range(args.steps):
        a, b2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        engine.fit_batched(data.Xtr32, task_L, W0d, E=E, T=data.n_train, early_stop=1)
        b2.record()
        torch.cuda.synchronize()
        fit_ms.append(a.elapsed_time(b2))
        flush.fill_(k & 1)

    # ---- end to end through the public API (host buffers in, host results out)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        h2d, d2h = e2e_step()
    barrier()
    e2e_s = time.perf_counter() - t0
    clocks = sampler.stop(t_timed)

    # ---- max over ranks
    tt = torch.tensor([total_ms, e2e_s * 1e3, float(np.mean(fit_ms))], dtype=torch.float64, device=dev)
    per_rank = None
    if world > 1:
        # per-rank step time and SM clock (diagnostic: which rank the max comes from, and whether its clock differs)
        mine = torch.tensor([total_ms / args.steps, clocks["sm_mhz"] or -1.], dtype=torch.float64, device=dev)
        allr = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
        per_rank = {"ms_per_step": [round(float(t[0]), 4) for t in allr], "sm_mhz": [float(t[1]) for t in allr]}
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    total_ms, e2e_ms, fit_ms_mean = (float(v) for v in tt.cpu())
    value = world * n_tasks * args.steps / (total_ms * 1e-3)
    e2e_value = world * n_tasks * args.steps / (e2e_ms * 1e-3)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (fit_kernel): algorithmic flop / measured duration vs measured FFMA peak
    def ffma_peak():
        import ctypes
        sink = torch.zeros(4, dtype=torch.float32, device=dev)
        fl = ctypes.c_double()
        best = 0.
        for _ in range(5):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            hfrep._lib.check(hfrep._lib.lib.hfrep_ffma_peak(1 << 16, hfrep._lib.ptr(sink), ctypes.byref(fl), hfrep._lib.stream_ptr()))
            b.record()
            torch.cuda.synchronize()
            best = max(best, fl.value / (a.elapsed_time(b) * 1e-3))
        return best / 1e12

    peak = ffma_peak()
    flop = float(sum(flops_per_epoch(L) * int(n) for L, n in zip(task_L, epochs_run)))
    achieved = flop / (fit_ms_mean * 1e-3) / 1e12
    roofline = {"bound": "fp32_ffma", "kernel": "fit_kernel", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                "frac": achieved / peak, "traffic": NCU_DRAM_BYTES_PER_FIT_LAUNCH,
                "peak_source": "measured live: hfrep_ffma_peak (8 packed FFMA2 chains/thread, full grid; scalar FFMA reaches 64.4 on the same box); MEASURED_PEAKS.json has no FP32 figure; nominal 148*128*2*1.965 GHz = 74.4",
                "algorithmic_flop_per_launch": flop, "ms_per_launch": fit_ms_mean,
                "note": f"{n_tasks} fits on {n_buckets} concurrent launches: cfg2 exposes only 21 sequential chains, so this is latency-, not FFMA-bound; see 'ensemble'"}

    # ---- the 64-seed ensemble (configs[2]) on this GPU: the throughput regime of the same kernels
    ens = None
    if args.ensemble_seeds > 0:
        es = [1000 + s for s in range(args.ensemble_seeds)]
        et = sorted(((L, s) for s in es for L in LATENTS), key=lambda t: -((t[0] + 3) // 4))
        eW0 = engine.to_device(engine.pack_weights([engine.initial_weights(s, E, L) for L, s in et], E, w_stride), dev)
        eL = [L for L, _ in et]
        times = []
        for i in range(3):
            a, b2, c = (torch.cuda.Event(enable_timing=True) for _ in range(3))
            a.record()
            f2 = engine.fit_batched(data.Xtr32, eL, eW0, E=E, T=data.n_train, early_stop=1)
            b2.record()
            torch.cuda.synchronize()
            flush.fill_(i)
            c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            c0.record()
            engine.fit_and_score(data, eL, eW0, early_stop=1)
            c1.record()
            torch.cuda.synchronize()
            flush.fill_(i)
            times.append((a.elapsed_time(b2), c0.elapsed_time(c1)))
        fms, sms = min(t[0] for t in times[1:]), min(t[1] for t in times[1:])
        eflop = float(sum(flops_per_epoch(L) * int(n) for L, n in zip(eL, f2.epochs_run.cpu().numpy())))
        ens = {"workload": f"cfg3 shape on one GPU: L=1..21 x {args.ensemble_seeds} seeds = {len(et)} fits",
               "value": len(et) / (sms * 1e-3), "unit": UNIT, "ms_per_step": sms, "fit_ms": fms,
               "roofline": {"bound": "fp32_ffma", "achieved": eflop / (fms * 1e-3) / 1e12, "peak": peak,
                            "unit": "TFLOP/s", "frac": eflop / (fms * 1e-3) / 1e12 / peak}}

    # ---- a bounded sample of configs[3] (500 ETFs, 252-day windows, rolling refit) through the wide kernels
    wide = None
    if args.wide_windows > 0 and rank == 0:
        wT, wE, wF, wW, wL, wep = 5000, 500, args.wide_funds, 252, 8, 100
        wr = np.random.default_rng(20240611)                         # SURVEY 8d cfg 4 factor model
        fac = wr.normal(size=(wT, 10)) * .01
        wetf = fac @ wr.normal(size=(10, wE)) + wr.normal(size=(wT, wE)) * .005
        idx = wr.integers(0, wE, size=(wF, 5))
        wts = wr.uniform(.06, .2, size=(wF, 5))
        whfd = sum(wetf[:, idx[:, j]] * wts[:, j] for j in range(5)) + wr.normal(size=(wT, wF)) * .003
        wins = np.linspace(0, wT - wW - 2, args.wide_windows).astype(int)
        best_t = None
        for i in range(3):
            tm = {}
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            wo = hfrep.run_mode_w_wide(wetf, whfd, np.full(wT, 1e-4), train_window=wW, latent_dim=wL, seeds=(1,), windows=wins,
                                       device=dev, max_epochs=wep, early_stop=0, timings=tm)
            torch.cuda.synchronize()
            tm["e2e_ms"] = (time.perf_counter() - t0) * 1e3
            if best_t is None or tm["fit_ms"] < best_t["fit_ms"]:
                best_t = tm
        wflop = float(args.wide_windows) * wep * (10 * 189 * wE * wL + 4 * 63 * wE * wL)
        wide = {"workload": f"cfg4 sample: {args.wide_windows} of 4747 windows x {wF} funds, 500 ETFs, L={wL}, {wep} epochs, rolling refit",
                "fit_ms": best_t["fit_ms"], "eval_ms": best_t["eval_ms"], "e2e_ms_with_h2d": best_t["e2e_ms"],
                "fits_per_s": args.wide_windows / (best_t["fit_ms"] * 1e-3),
                "results_per_s_e2e": args.wide_windows * wF / (best_t["e2e_ms"] * 1e-3),
                "status_ok": bool(int(wo["fit"].status.abs().sum()) == 0 and torch.isfinite(wo["replica"]).all()),
                "roofline": {"bound": "fp32_ffma", "kernel": "fit_wide2_kernel", "achieved": wflop / (best_t["fit_ms"] * 1e-3) / 1e12,
                             "peak": peak, "unit": "TFLOP/s", "frac": wflop / (best_t["fit_ms"] * 1e-3) / 1e12 / peak}}

    cpu = cpu_baseline_single_core() if (world == 1 and not args.no_cpu_baseline) else None
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"cfg2: 13 indices x 145 windows x L=1..21 x {args.seeds} seed(s) per GPU (Mode R), synthetic 337x(22+13+1)",
                   "fits_per_step_per_gpu": n_tasks, "triples_per_step_per_gpu": n_tasks * F * 145,
                   "epochs": EPOCHS, "epochs_run_mean": float(epochs_run.mean()), "shuffle": False,
                   "seeds_per_rank": "the same on every rank (fixed work per GPU; run lengths under early stopping depend on the seed)",
                   "l2": "flushed between steps (256 MiB write, not timed); inputs are KB-sized and smem-resident"},
        "clocks": clocks,
        "per_rank": per_rank,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": e2e_ms / args.steps},
        "gpu_launches": args.steps * 2 * n_buckets,      # fit_kernel + eval_kernel per width bucket
        "roofline": roofline,
        "cpu_baseline": cpu,
        "cpu_best_effort": cpu_best_effort() if cpu is not None else None,
        "ensemble": ens,
        "wide": wide,
        "wall_s_timed_region": wall,
    }
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--seeds", type=int, default=1, help="seeds per GPU (1 = configs[1]; 64 = configs[2] on one GPU)")
    ap.add_argument("--ensemble-seeds", type=int, default=64)
    ap.add_argument("--wide-windows", type=int, default=296, help="windows of the configs[3] sample through the wide kernels (0: skip)")
    ap.add_argument("--wide-funds", type=int, default=2000, help="funds of that sample (configs[3] has 10 000: 400 MB of float64 H2D)")
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        args.steps = min(args.steps, 3)
        args.warmup = min(args.warmup, 1)
        cpu_reference_arm(args)
    else:
        gpu_arm(args)


if __name__ == "__main__":
    main()
