#!/usr/bin/env python
"""bench.py — replication task-fits/sec on B200 (BASELINE.json metric), one process per GPU.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--cpu-procs P]

A *step* is one pass of the hot path over one batch of synthetic input of the named shape
(configs[1] of BASELINE.json: 337 months x 22 factor ETFs / 13 fund indices / rf; the notebook's
sweep L = 1..21, one seed, deterministic row order, EarlyStopping(patience 5) on as in the reference
— run lengths are data dependent, `config.epochs_run_mean` reports them; all 13 funds x 145 rolling
windows evaluated per fit): kernel (b) fits 21 autoencoders, kernel (c) scores them.
Multi-GPU (`--gpus N`, one rank per GPU): every rank runs that sweep for ITS OWN seed (123 + rank):
distinct fits, no data-path collective, `value` = all ranks' fits / the slowest rank's time (weak
scaling).  The same line carries `sharded`: configs[2] (L = 1..21 x 64 seeds = 1 344 fits) dealt
cyclically over the N ranks through the public API with host buffers, finished by the one NCCL
all-gather of result records, timed end to end; `wide`: configs[3] at its stated size (4 747 windows x
10 000 funds x 500 ETFs) and `paths`: configs[4] (1 000 000 generated paths as tasks), their windows /
paths dealt over the N ranks.  At N = 1 the line also carries the throughput regime (`ensemble`), the
HBM-bound kernels (`hbm_kernels`) and the CPU baselines.

`--impl reference` times the CPU restatement of the reference path (oracle/ae_oracle.py — the
reference itself needs Keras, which cannot be installed, and cannot travel to the GPU box) on
`--cpu-procs` host processes (default 16, fewer if the box has fewer cores).
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
WORKLOAD = "cfg2: 13 indices x 145 windows x L=1..21 x 1 seed per GPU (Mode R), synthetic 337x(22+13+1)"
# dram__bytes_read.sum + dram__bytes_write.sum per fit_cluster_kernel launch of the cfg2 sweep, from the committed
# `ncu --set full` capture (profiles/r2_ncu_full_cfg2_summary.txt: 92-140 KB per launch, 0 B written): every CTA of a
# cluster stages its task's 15 KB training block; the run itself is shared-memory resident.
NCU_DRAM_BYTES_PER_FIT_LAUNCH = 110e3


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


def flops_per_epoch(L, T_fit=126, T_val=42, batch=48, E=E):
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
    out = ae_oracle.run_task(etf[:168], hfd[:168], etf[168:], hfd[168:], rf, etf, L, We0, Wd0, cfg=cfg, order="blas")
    return time.perf_counter() - t, int(out["epochs_run"])


CPU_IMPL = ("oracle/ae_oracle.py: the fit as a float32 numpy step loop over BLAS matmuls — the structure of Keras' fit without its "
            "per-step dispatch cost, as in round 1 — and every evaluation stage in numpy float64; one thread per task (the "
            "reference pins TF to one thread, helper.py:38).  cpu_best_effort is the fit alone as compiled C on all cores")


def cpu_reference_arm(args):
    """The reference path's CPU restatement on --cpu-procs host processes: each step = the 21-fit sweep."""
    import multiprocessing as mp
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = len(os.sched_getaffinity(0))
    procs = max(1, min(args.cpu_procs, cores, len(LATENTS)))
    tasks = [(L, 123) for L in LATENTS]
    ctx = mp.get_context("fork")
    times, epochs = [], []
    with ctx.Pool(procs) as pool:
        for i in range(args.warmup + args.steps):
            t = time.perf_counter()
            res = pool.map(_cpu_task, tasks, chunksize=1)
            dt = time.perf_counter() - t
            if i >= args.warmup:
                times.append(dt)
                epochs = [e for _, e in res]
    total = sum(times)
    value = len(tasks) * len(times) / total
    sample = (f"{len(times)} x full sweep L=1..21 (21 fits, up to {EPOCHS} epochs with early stopping: {np.mean(epochs):.0f} run on "
              f"average; all evaluation stages) on {procs} processes of {cores} usable cores; {CPU_IMPL}")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "epochs": EPOCHS, "epochs_run_mean": float(np.mean(epochs)), "shuffle": False,
                   "early_stop": True},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": procs, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def cpu_baseline_single_core(budget_s=20.):
    """Oracle port on ONE core over a bounded sample of the same sweep (about `budget_s` seconds)."""
    sample, spent = [], 0.
    order = (1, 4, 8, 11, 14, 18, 21, 2, 6, 16, 20, 3, 9, 12, 5, 7, 10, 13, 15, 17, 19)
    while spent < budget_s:
        for L in order:
            dt, _ = _cpu_task((L, 123))
            sample.append(L)
            spent += dt
            if spent > budget_s:
                break
    return {"value": len(sample) / spent, "unit": UNIT, "cores": 1, "kind": "port",
            "sample": f"{len(sample)} fits (widths cycling through L = 1..21) of the cfg2 sweep, full pipeline, {spent:.1f} s on one core; {CPU_IMPL}"}


def cpu_best_effort():
    """the fit alone as compiled C in its natural loop order (oracle/c/fit_oracle.c, gcc -O3 -march=native,
    OpenMP over the 21 tasks of the sweep) on all host cores — shown so the speed-up is not only "Python overhead
    removed".  The evaluation stages are NOT included, so this over-states the CPU's task-fits/sec."""
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
def measured_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        return {}


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
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"            # keep NCCL's version banner off stdout: the line below is the output
        dist.init_process_group("nccl", device_id=dev)

    etf, hfd, rf = synthetic_returns()
    seeds = [123 + rank]                    # every rank its own seed: distinct fits, fixed fit count per GPU
    data = hfrep.ReplicationData(etf, hfd, rf, window=WINDOW, device=dev)
    tasks = engine.mode_r_tasks(LATENTS, seeds)
    n_tasks = len(tasks)
    w_stride = 2 * E * max(LATENTS)
    W0_host = engine.pack_weights([engine.initial_weights(s, E, L) for L, s in tasks], E, w_stride)
    W0d = engine.to_device(W0_host, dev)
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

    plan = hfrep.SweepPlan(etf.shape[0], E, hfd.shape[1], LATENTS, seeds, window=WINDOW, device=dev, early_stop=1)

    def e2e_step():
        """Through the public API with HOST buffers — the recorded sweep (`hfrep.SweepPlan.run`: one CUDA graph): H2D of the
        returns + initial kernels, D2H of everything a caller of the reference's AE objects reads — histories, trained
        kernels, quality series, replicas, turnover, metrics."""
        r = plan.run(etf, hfd, rf, W0=W0_host)
        return r.h2d_bytes, r.d2h_bytes

    def one_shot_step():
        """The same job launch by launch (`hfrep.run_mode_r`, no recorded graph: what a first call costs)."""
        d = hfrep.ReplicationData(etf, hfd, rf, window=WINDOW, device=dev)
        _, fit, ev = hfrep.run_mode_r(d, LATENTS, seeds=seeds, W0=W0_host)      # same widest-first task order as W0_host
        engine.to_host([ev.ante, ev.post, ev.turnover, ev.metrics, ev.quality, ev.status, fit.epochs_run, fit.status, fit.loss,
                        fit.val_loss, fit.W])

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
    for _ in range(2):
        e2e_step()
        one_shot_step()
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
    assert int(fit.status.abs().sum()) == 0 and bool(torch.isfinite(ev.ante).all())
    eval_status = sorted(set(int(v) for v in ev.status.cpu().numpy()))       # 2 = a rank-deficient window took the pinv path

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
    t0 = time.perf_counter()
    for _ in range(args.steps):
        one_shot_step()
    barrier()
    one_shot_s = time.perf_counter() - t0
    clocks = sampler.stop(t_timed)

    # ---- the sharded job: configs[2] dealt over the ranks, one all-gather of result records, host buffers in and out
    sharded = sharded_job(args, hfrep, engine, dist, torch, dev, rank, world, (etf, hfd, rf), barrier)

    # ---- configs[3] and configs[4] at their stated sizes, their units dealt over the ranks
    wide_raw = wide_job(args, hfrep, torch, dist, dev, rank, world, barrier) if (args.wide_windows > 0 and not args.no_extras) else None
    paths_raw = paths_job(args, hfrep, engine, torch, dist, dev, rank, world, barrier) if (args.path_tasks > 0 and not args.no_extras) else None
    mode_w = mode_w_record(hfrep, torch, dev, (etf, hfd, rf)) if (rank == 0 and not args.no_extras) else None

    # ---- max over ranks
    tt = torch.tensor([total_ms, e2e_s * 1e3, float(np.mean(fit_ms))], dtype=torch.float64, device=dev)
    per_rank = None
    if world > 1:
        # per-rank step time, run length and SM clock (diagnostic: which rank the max comes from, and why)
        mine = torch.tensor([total_ms / args.steps, clocks["sm_mhz"] or -1., float(epochs_run.mean())], dtype=torch.float64, device=dev)
        allr = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
        per_rank = {"ms_per_step": [round(float(t[0]), 4) for t in allr], "sm_mhz": [float(t[1]) for t in allr],
                    "epochs_run_mean": [round(float(t[2]), 1) for t in allr]}
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
    roofline = {"bound": "fp32_ffma", "kernel": "fit_cluster_kernel", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                "frac": achieved / peak, "traffic": NCU_DRAM_BYTES_PER_FIT_LAUNCH,
                "peak_source": "measured live: hfrep_ffma_peak (8 packed FFMA2 chains/thread, full grid; scalar FFMA reaches 64.4 on the same box); MEASURED_PEAKS.json has no FP32 figure; nominal 148*128*2*1.965 GHz = 74.4",
                "algorithmic_flop_per_launch": flop, "ms_per_launch": fit_ms_mean,
                "note": f"{n_tasks} fits on {n_buckets} concurrent launches: cfg2 exposes only 21 sequential chains, so this is latency-, not FFMA-bound; see 'ensemble'"}

    extras = world == 1 and not args.no_extras           # single-GPU records (ensemble, HBM kernels)
    ens = ensemble_record(args, hfrep, engine, torch, data, dev, flush, peak, w_stride) if extras else None
    hbm = hbm_kernels_record(hfrep, engine, torch, data, dev, flush, task_L, last) if extras else None
    wide = wide_record(wide_raw, peak) if wide_raw else None
    paths = paths_record(paths_raw, peak) if paths_raw else None

    cpu = cpu_baseline_single_core() if (world == 1 and not args.no_cpu_baseline) else None
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "fits_per_step_per_gpu": n_tasks, "triples_per_step_per_gpu": n_tasks * F * 145,
                   "epochs": EPOCHS, "epochs_run_mean": float(epochs_run.mean()), "shuffle": False, "early_stop": True,
                   "seeds_per_rank": "rank r fits seed 123 + r (distinct fits on every GPU; run lengths under early stopping depend on the seed: see per_rank)",
                   "eval_status_values": eval_status,
                   "l2": "flushed between steps (256 MiB write, not timed); inputs are KB-sized and smem-resident"},
        "clocks": clocks,
        "per_rank": per_rank,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": e2e_ms / args.steps, "api": "hfrep.SweepPlan.run (the sweep recorded once as a CUDA graph, replayed per step)",
                "one_shot_ms_per_step": one_shot_s * 1e3 / args.steps, "one_shot_api": "hfrep.ReplicationData + hfrep.run_mode_r + engine.to_host, launch by launch (this rank)"},
        "gpu_launches": args.steps * 2 * n_buckets,      # fit_kernel + eval_kernel per width bucket
        "roofline": roofline,
        "parity": "kernel (b) bit-identical to the CPU oracle (canonical summation order) over whole runs; everything else <= 1e-5 (tests/)",
        "cpu_baseline": cpu,
        "cpu_best_effort": cpu_best_effort() if cpu is not None else None,
        "sharded": sharded,
        "ensemble": ens,
        "hbm_kernels": hbm,
        "wide": wide,
        "paths": paths,
        "mode_w": mode_w,
        "wall_s_timed_region": wall,
    }
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def sharded_job(args, hfrep, engine, dist, torch, dev, rank, world, returns, barrier):
    """configs[2]: the L = 1..21 x S-seed ensemble as ONE job over all ranks, through the public API: host buffers in
    (returns matrices replicated to every GPU, the rank's share of the initial kernels), cyclic task split, fit + score,
    ONE all-gather of packed result records (NCCL over NVLink), host results out on rank 0.  No data-path collective."""
    S = args.sharded_seeds
    if S <= 0:
        return None
    etf, hfd, rf = returns
    es = [1000 + s for s in range(S)]
    tasks = engine.mode_r_tasks(LATENTS, es)
    w_stride = 2 * E * max(LATENTS)
    W0 = engine.pack_weights([engine.initial_weights(s, E, L) for L, s in tasks], E, w_stride)
    recs = []
    out = None
    for i in range(1 + args.sharded_steps):
        out = None                               # the previous job's host results are released (their page-locked buffer is reused)
        barrier()
        t0 = time.perf_counter()
        d = hfrep.ReplicationData(etf, hfd, rf, window=WINDOW, device=dev)
        tm = {}
        _, fit, ev = hfrep.run_mode_r(d, LATENTS, seeds=es, W0=W0, sharded=world > 1, timings=tm)
        d2h = 0
        if rank == 0:
            out = engine.to_host([ev.ante, ev.post, ev.turnover, ev.metrics, ev.quality, fit.epochs_run, fit.status, fit.loss,
                                  fit.val_loss, fit.W])
            d2h = sum(a.nbytes for a in out)
        torch.cuda.synchronize()
        t_local = time.perf_counter() - t0
        barrier()
        t_all = time.perf_counter() - t0
        if i > 0:
            recs.append((t_all, t_local, tm["compute_ms"], tm["gather_ms"], d2h, tm["record_bytes_per_task"]))
    # median over the timed jobs (a job whose 63 MB page-locked result buffer has to be allocated afresh takes ~5 ms longer)
    t_all = float(np.median([r[0] for r in recs]))
    mine = torch.tensor([np.median([r[2] for r in recs]), np.median([r[3] for r in recs]), t_all * 1e3,
                         np.median([r[1] for r in recs]) * 1e3], dtype=torch.float64, device=dev)
    allr = [mine]
    if world > 1:
        allr = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
    t_max = max(float(t[2]) for t in allr) * 1e-3
    ok = bool(int(fit.status.abs().sum()) == 0 and fit.W.shape[0] == len(tasks) and torch.isfinite(ev.ante).all())
    compute = [round(float(t[0]), 3) for t in allr]
    gather = [round(float(t[1]), 3) for t in allr]
    spread = (max(compute) - min(compute)) / max(compute)
    d2h_ms = float(allr[0][3]) - compute[0] - gather[0]      # rank 0: what follows its compute and the gather (result download)
    parts = {"kernel (b)/(c) on each rank's share (a fit is a chain of up to 1000 dependent epochs: with few fits per SM the job "
             "cannot be shorter than the longest fit)": max(compute), "the all-gather": max(gather),
             "rank 0's download of the gathered results to the host": d2h_ms}
    return {"workload": f"cfg3: L=1..21 x {S} seeds = {len(tasks)} fits dealt cyclically over {world} GPU(s), early stopping on, "
                        "host buffers in, gathered results to the host on rank 0",
            "value": len(tasks) / t_max, "unit": UNIT, "ms_per_job": t_max * 1e3, "steps": args.sharded_steps, "statistic": "median over the steps, max over the ranks",
            "compute_ms_per_rank": compute, "gather_ms_per_rank": gather,
            "local_ms_per_rank": [round(float(t[3]), 3) for t in allr],
            "gathered_bytes_per_rank": int(recs[-1][5]) * len(tasks) if world > 1 else 0,
            "d2h_bytes_rank0": int(recs[-1][4]), "h2d_bytes_per_rank": int(d.h2d_bytes + W0.nbytes // world),
            "results_complete_and_finite": ok,
            "fits_per_rank": -(-len(tasks) // world), "result_download_ms_rank0": round(d2h_ms, 3),
            "limiter": max(parts, key=parts.get) + f" — compute {max(compute):.1f} ms (spread over ranks {100 * spread:.1f} %: early-stop "
                       f"raggedness), all-gather {max(gather):.2f} ms, host download {d2h_ms:.1f} ms of {t_max * 1e3:.1f} ms"}


def ensemble_record(args, hfrep, engine, torch, data, dev, flush, peak, w_stride):
    """The 64-seed ensemble (configs[2]) on ONE GPU, inputs resident: the throughput regime of kernels (b) and (c)."""
    if args.ensemble_seeds <= 0:
        return None
    es = [1000 + s for s in range(args.ensemble_seeds)]
    et = engine.mode_r_tasks(LATENTS, es)
    eW0 = engine.to_device(engine.pack_weights([engine.initial_weights(s, E, L) for L, s in et], E, w_stride), dev)
    eL = [L for L, _ in et]
    times = []
    for i in range(3):
        a, b2 = (torch.cuda.Event(enable_timing=True) for _ in range(2))
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
    return {"workload": f"cfg3 shape on one GPU: L=1..21 x {args.ensemble_seeds} seeds = {len(et)} fits",
            "value": len(et) / (sms * 1e-3), "unit": UNIT, "ms_per_step": sms, "fit_ms": fms,
            "roofline": {"bound": "fp32_ffma", "achieved": eflop / (fms * 1e-3) / 1e12, "peak": peak,
                         "unit": "TFLOP/s", "frac": eflop / (fms * 1e-3) / 1e12 / peak}}


def hbm_kernels_record(hfrep, engine, torch, data, dev, flush, task_L, last):
    """The HBM-bound kernels against the measured copy bandwidth (MEASURED_PEAKS.json hbm_gbs): kernel (a) on the
    configs[3] window set (296 windows x 252 rows x 500 ETFs, plain and with the per-window MinMax, float32 and float64
    source) and the output side of kernel (c).  Algorithmic bytes = bytes written + the same read (kernel (a): overlapping
    windows re-read from L2 count as read), CUDA events, L2 flushed before every launch, best of 5."""
    peak = measured_peaks().get("hbm_gbs") or 6450.9

    def timed(fn, reps=5):
        best = float("inf")
        for i in range(reps):
            flush.fill_(i)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            best = min(best, a.elapsed_time(b))
        return best

    lib, ptr, check, sp = hfrep._lib.lib, hfrep._lib.ptr, hfrep._lib.check, hfrep._lib.stream_ptr
    rng = np.random.default_rng(3)
    T, C, W = 5000, 500, 252
    R32 = engine.to_device((rng.normal(size=(T, C)) * .01).astype(np.float32), dev)
    R64 = R32.to(torch.float64)
    out = {"peak_gbs": peak, "peak_source": "MEASURED_PEAKS.json hbm_gbs (copy, read+write bytes)"}
    for n_win in (296, 4747):
        starts = torch.from_numpy(np.linspace(0, T - W - 2, n_win).astype(np.int32)).to(dev)
        dst = torch.empty((n_win, W, C), dtype=torch.float32, device=dev)
        for name, src, mm in (("plain_f32", R32, 0), ("minmax_f32", R32, 1), ("minmax_f64", R64, 1)):
            if n_win == 4747 and name == "plain_f32":
                continue
            fn = lib.hfrep_window_gather if src.dtype == torch.float32 else lib.hfrep_window_gather_f64
            rb = n_win * W * C * (4. + src.element_size()) * (1. + .5 * mm)      # minmax reads every window twice (min/max, scale)
            ms = timed(lambda: check(fn(ptr(src), T, C, ptr(starts), n_win, W, 0, C, mm, ptr(dst), sp())))   # the C ABI call: no host checks in the timed region
            out[f"window_gather_{name}_{n_win}w"] = {"ms": ms, "algorithmic_bytes": rb, "gbs": rb / (ms * 1e-3) / 1e9,
                                                     "frac": rb / (ms * 1e-3) / 1e9 / peak}
        del dst
    # kernel (c), output side: what one launch over many tasks writes (replicas, ex-post, turnover, metrics, quality, masks,
    # beta, norm, status) / its duration — the kernel is compute-bound (float64 QR + quality loops), so this is small
    n = 4096
    fit, _ = last
    W_rep = fit.W[:1].repeat(n, 1).contiguous()
    tL = [task_L[0]] * n
    r, launch = hfrep.predict_metrics(data.Xtr64, data.Xte64, data.Yte64, data.rf64, data.sigma, tL, W_rep, window=WINDOW,
                                      want_oos=False, launch=False)
    wb = sum(t.numel() * t.element_size() for t in (r.ante, r.post, r.turnover, r.metrics, r.mask, r.beta, r.norm, r.status))
    stream = torch.cuda.current_stream(dev).cuda_stream
    ms = timed(lambda: launch(0, n, stream), reps=3)
    out["eval_kernel_output_side_4096_tasks"] = {
        "ms": ms, "bytes_written": wb, "gbs": wb / (ms * 1e-3) / 1e9, "frac": wb / (ms * 1e-3) / 1e9 / peak,
        "note": "kernel (c) writes 37 KB per task and is bound by its float64 QR / statistics, not by HBM"}
    return out


def _max_over_ranks(torch, dist, dev, world, values):
    t = torch.tensor(values, dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [float(v) for v in t.cpu()]


def wide_job(args, hfrep, torch, dist, dev, rank, world, barrier):
    """configs[3] (500 ETFs, 10 000 funds, 5 000-day history, 252-day windows, rolling refit: 4 747 windows, L = 8, 100
    epochs as SURVEY §8d fixes it) through the wide kernels, the windows dealt over the ranks in runs of consecutive windows
    (a rank uploads only its stretch of the histories), every rank's replica rows all-gathered at the end.  Every rank builds the same synthetic history (SURVEY §8d cfg 4 factor model)."""
    from hfrep import distributed
    wT, wE, wF, wW, wL, wep = 5000, 500, args.wide_funds, 252, 8, 100
    wr = np.random.default_rng(20240611)
    fac = wr.normal(size=(wT, 10)) * .01
    wetf = fac @ wr.normal(size=(10, wE)) + wr.normal(size=(wT, wE)) * .005
    idx = wr.integers(0, wE, size=(wF, 5))
    wts = wr.uniform(.06, .2, size=(wF, 5))
    whfd = np.zeros((wT, wF))
    for j in range(5):
        whfd += wetf[:, idx[:, j]] * wts[:, j]
    whfd += wr.normal(size=(wT, wF)) * .003
    n_all = wT - wW - 1                                              # 4 747 windows
    n_win = min(args.wide_windows, n_all)
    wins = np.arange(n_all) if n_win == n_all else np.linspace(0, n_all - 1, n_win).astype(int)
    mine = wins[distributed.shard_blocks(n_win, rank, world)]      # a run of consecutive windows: the rank uploads only its stretch
    W0 = hfrep.engine.pack_weights([hfrep.engine.initial_weights(100003 + int(wi), wE, wL) for wi in mine], wE, 2 * wE * wL)
    best, ok = None, True
    for i in range(3):
        tm = {}
        barrier()
        t0 = time.perf_counter()
        wo = hfrep.run_mode_w_wide(wetf, whfd, np.full(wT, 1e-4), train_window=wW, latent_dim=wL, seeds=(1,), windows=mine,
                                   device=dev, W0=W0, max_epochs=wep, early_stop=0, timings=tm)
        g0 = torch.cuda.Event(enable_timing=True); g1 = torch.cuda.Event(enable_timing=True)
        g0.record()
        replica = distributed.gather_results(wo["replica"][0], n_win, rank, world, order="blocks") if world > 1 else wo["replica"][0]
        g1.record()
        torch.cuda.synchronize()
        tm["gather_ms"] = g0.elapsed_time(g1)
        barrier()
        tm["job_ms"] = (time.perf_counter() - t0) * 1e3
        ok = ok and bool(int(wo["fit"].status.abs().sum()) == 0 and replica.shape[0] == n_win and torch.isfinite(replica).all())
        best = tm if best is None else {k: min(best[k], tm[k]) for k in tm}
    fit_ms, eval_ms, gather_ms, job_ms = _max_over_ranks(torch, dist, dev, world, [best["fit_ms"], best["eval_ms"], best["gather_ms"], best["job_ms"]])
    return dict(n_win=n_win, n_all=n_all, wF=wF, wE=wE, wL=wL, wep=wep, world=world, fit_ms=fit_ms, eval_ms=eval_ms, gather_ms=gather_ms,
                job_ms=job_ms, ok=ok, h2d=int(wo["h2d_bytes"]), gathered=int(n_win * wF * 8))


def wide_record(r, peak):
    wflop = float(r["n_win"]) * r["wep"] * flops_per_epoch(r["wL"], 189, 63, 48, r["wE"])
    per_gpu_flops = wflop / r["world"] / (r["fit_ms"] * 1e-3) / 1e12
    return {"workload": f"cfg4: {r['n_win']} of {r['n_all']} windows x {r['wF']} funds, 500 ETFs, L={r['wL']}, {r['wep']} epochs, rolling refit, "
                        f"windows dealt over {r['world']} GPU(s), replica rows all-gathered",
            "fit_ms": r["fit_ms"], "eval_ms": r["eval_ms"], "gather_ms": r["gather_ms"], "job_ms_with_h2d": r["job_ms"],
            "h2d_bytes_per_rank": r["h2d"], "gathered_bytes": r["gathered"] if r["world"] > 1 else 0,
            "fits_per_s": r["n_win"] / (r["fit_ms"] * 1e-3), "results_per_s_e2e": r["n_win"] * r["wF"] / (r["job_ms"] * 1e-3),
            "status_ok": r["ok"],
            "roofline": {"bound": "fp32_ffma", "kernel": "fit_wide3_kernel", "achieved": per_gpu_flops, "peak": peak, "unit": "TFLOP/s",
                         "frac": per_gpu_flops / peak, "note": "per GPU"}}


def mode_w_record(hfrep, torch, dev, returns):
    """north_star's task read literally — one fit per (window, seed), each scored on every fund: the rolling refit (Mode W,
    our extension: the reference fits once) at the reference's shape, 168 windows x 8 seeds, L = 8, all 1000 epochs, host
    buffers in.  Rank 0 only; best of 3."""
    etf, hfd, rf = returns
    best, n, ok = float("inf"), 0, True
    for _ in range(3):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        out = hfrep.run_mode_w(etf, hfd, rf, train_window=168, latent_dim=8, seeds=range(8), device=dev, max_epochs=EPOCHS, early_stop=0)
        torch.cuda.synchronize()
        best = min(best, time.perf_counter() - t0)
        n = len(out["tasks"])
        ok = ok and int(out["fit"].status.abs().sum()) == 0 and bool(torch.isfinite(out["replica"]).all())
    F = hfd.shape[1]
    return {"workload": f"Mode W: {n} (window, seed) fits (168 windows x 8 seeds, own training block each), L=8, {EPOCHS} epochs, {F} funds scored per fit",
            "ms_per_job_e2e": best * 1e3, "fits_per_s": n / best, "fund_window_seed_triples_per_s": n * F / best, "status_ok": ok}


def paths_job(args, hfrep, engine, torch, dist, dev, rank, world, barrier):
    """configs[4]: generated paths (168 x 36 in MinMax space) as augmented training tasks — every path its own training
    block (`task_x`), L = 8, through the same batched fit (batch-tile mode), the paths dealt over the ranks.  The paths come
    from the reference's own saved generator (weights carried in tests/golden) run on each rank's GPU; without that file,
    from clip(N(.61, .2), -.13, 1.26) as in SURVEY §8d.  Fixed epochs (no early stop) so the flop count is exact."""
    n_all, T, L, ep = args.path_tasks, 168, 8, args.path_epochs
    n = len(range(rank, n_all, world))
    g = torch.Generator(device=dev)
    g.manual_seed(7 + rank)
    X = torch.empty((n, T, E), dtype=torch.float32, device=dev)
    gen_ms = None
    fixture = os.path.join(ROOT, "tests", "golden", "ref_generator.npz")
    if os.path.isfile(fixture):
        gen = hfrep.generator.Generator(hfrep.generator.layers_from_fixture(np.load(fixture)), device=dev)
        sms = torch.cuda.get_device_properties(dev).multi_processor_count
        chunk, gen_ms = 64 * sms * 4, 0.                          # whole waves of the tensor-core kernel's 64-path CTAs
        for lo in range(0, n, chunk):
            m = min(chunk, n - lo)
            noise = torch.empty((m, T, 36), dtype=torch.float32, device=dev).normal_(0., 1., generator=g)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            paths = gen.predict_device(noise)                       # (m, 168, 36); columns 0..21 are a task's ETF block
            b.record()
            torch.cuda.synchronize()
            gen_ms += a.elapsed_time(b)
            X[lo:lo + m] = paths[:, :, :E]
            del noise, paths
    else:
        X.normal_(.61, .2, generator=g).clamp_(-.13, 1.26)
    W0 = engine.to_device(engine.pack_weights([engine.initial_weights(5, E, L)], E, 2 * E * L), dev).repeat(n, 1).contiguous()
    tx = torch.arange(n, dtype=torch.int32, device=dev)
    best, ok = float("inf"), True
    for i in range(2):
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fit = engine.fit_batched(X, [L] * n, W0, E=E, T=T, task_x=tx, max_epochs=ep, early_stop=0)
        b.record()
        torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
        ok = ok and bool(int(fit.status.abs().sum()) == 0 and torch.isfinite(fit.W).all())
        del fit
    fit_ms, gen_max = _max_over_ranks(torch, dist, dev, world, [best, gen_ms or 0.])
    return dict(n_all=n_all, L=L, ep=ep, world=world, fit_ms=fit_ms, gen_ms=gen_max if gen_ms is not None else None, ok=ok,
                blocks_bytes=int(X.numel() * 4))


def paths_record(r, peak):
    n, ep, w = r["n_all"], r["ep"], r["world"]
    flop = float(n) * ep * flops_per_epoch(r["L"])
    gen_rec = None
    if r["gen_ms"] is not None:
        gflop = n * 168 * 2. * ((36 + 100) * 400 + 200 * 400 + 100 * 36) / 1e9
        # tensor-pipe work the tcgen05 kernel issues: 3 TF32 MMAs per product, K padded to 144 / 208, 64-row MMAs
        tc_gflop = n * 168 * 3 * 2. * (400 * (144 + 208) + 40 * 104) / 1e9      # (+ the Dense chain: N = 40, K = 104)
        tf32_peak = (measured_peaks().get("bf16_tflops_sustained") or 1404.3) / 2.
        gen_rec = {"kernel": "generator_tc_kernel (LSTM 36->100->100->36, 168 steps; tcgen05 kind::tf32 in 3xTF32, FP32 accumulators in TMEM)",
                   "ms": r["gen_ms"], "paths_per_s": n / (r["gen_ms"] * 1e-3),
                   "fp32_equivalent_tflops_per_gpu": gflop / w / r["gen_ms"],
                   "roofline": {"bound": "tensor", "achieved": tc_gflop / w / r["gen_ms"], "peak": tf32_peak, "unit": "TFLOP/s",
                                "frac": tc_gflop / w / r["gen_ms"] / tf32_peak,
                                "note": "issued TF32 MMA flops per GPU against half the measured sustained bf16 rate (MEASURED_PEAKS.json has no TF32 figure)"}}
    per_gpu = flop / w / (r["fit_ms"] * 1e-3) / 1e12
    return {"workload": f"cfg5: {n} generated paths (168 x 22 ETF columns each, own training block per task), L={r['L']}, {ep} epochs, "
                        f"paths dealt over {w} GPU(s)",
            "paths_from": "the reference's saved generator run on the GPU(s)" if gen_rec else "clip(N(.61, .2), -.13, 1.26) stand-in",
            "generator": gen_rec, "fit_ms": r["fit_ms"], "fits_per_s": n / (r["fit_ms"] * 1e-3),
            "fits_per_s_at_1000_epochs": n / (r["fit_ms"] * 1e-3) * ep / 1000.,
            "blocks_bytes_per_rank": r["blocks_bytes"], "status_ok": r["ok"],
            "roofline": {"bound": "fp32_ffma", "kernel": "fit_kernel (batch-tile mode)", "achieved": per_gpu, "peak": peak,
                         "unit": "TFLOP/s", "frac": per_gpu / peak, "note": "per GPU"}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--sharded-seeds", type=int, default=64, help="seeds of the sharded configs[2] job (0: skip)")
    ap.add_argument("--sharded-steps", type=int, default=3)
    ap.add_argument("--ensemble-seeds", type=int, default=64)
    ap.add_argument("--wide-windows", type=int, default=4747, help="windows of configs[3] through the wide kernels (4 747 = all; 0: skip)")
    ap.add_argument("--wide-funds", type=int, default=10000, help="funds of that sample (configs[3]: 10 000 = 400 MB of float64 H2D)")
    ap.add_argument("--path-tasks", type=int, default=1000000, help="path tasks of configs[4] (1 000 000 = the stated size; 0: skip)")
    ap.add_argument("--path-epochs", type=int, default=100)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--cpu-procs", type=int, default=16, help="host processes of the --impl reference arm")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the ensemble / hbm_kernels / wide / paths records")
    args = ap.parse_args()
    if args.impl == "reference":
        cpu_reference_arm(args)
    else:
        gpu_arm(args)


if __name__ == "__main__":
    main()
