#!/usr/bin/env python
"""bench.py — MCMC sweeps/sec of the gpirt hot path on B200 (BASELINE.json metric), one JSON line on stdout.

  python bench.py [--gpus N] [--steps K] [--warmup W]            our CUDA path (N>1: launched under torchrun)
  python bench.py --impl reference [--gpus N] --steps K --warmup W   the reference algorithm on the host cores

A step is one full sweep (draw_f, draw_fstar, draw_theta, f<-f* gather, draw_beta, K+Cholesky, draw_threshold,
total log-likelihood: src/gpirtMCMC.cpp:267-331 of the reference) over synthetic data of the named shape.
Workload: BASELINE config C3 — synthetic ordinal n=2000 respondents x 5000 items per GPU, C=5, 5% missing
(weak scaling over items: N GPUs run n=2000 x m=5000*N, item-sharded; the metric counts C3-sized item blocks,
i.e. value = N * sweeps/s, which at N=1 is plain sweeps/s of C3).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

_REAL_STDOUT = sys.stdout

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_RESP, M_PER_GPU, C_CAT = 2000, 5000, 5


def workload(config, gpus):
    """BASELINE.json configs[2] (c3: the configuration the metric is quoted on, weak-scaled by items: 5000 items per GPU) and configs[3]
    (c4: n = 10 000 x m = 50 000, C = 7, one chain item-sharded over the GPUs: strong scaling)."""
    if config == "c4":
        n, m, C = 10000, 50000, 7
        return dict(name="c4", n=n, m=m, m_loc=(m + gpus - 1) // gpus, C=C, seed=20261104 + 4, scaling="strong", blocks=1, metric="mcmc_sweeps_per_sec_c4",
                    text=f"C4 synthetic ordinal n={n} x m={m}, C={C}, 5% missing, one chain item-sharded over {gpus} GPU(s) ({(m + gpus - 1) // gpus} items per GPU)")
    n, m, C = N_RESP, M_PER_GPU * gpus, C_CAT
    return dict(name="c3", n=n, m=m, m_loc=M_PER_GPU, C=C, seed=20261104 + 3, scaling="weak", blocks=gpus, metric="mcmc_sweeps_per_sec_c3_blocks",
                text=f"C3 synthetic ordinal n={n} x m={m} ({M_PER_GPU} items per GPU), C={C}, 5% missing")
FP64_DMMA_PEAK_FALLBACK = 37.1  # profiles/r01_microbench_fp64.txt; used only if the in-run microbenchmark (gpirt_debug_fp64_peak) fails


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return json.load(fh), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0}, "fallback"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i", str(self.index), "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[k] for r in self.rows if len(r) >= 6 for k in range(4) if r[2 + k].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(sm)}


def host_responses(torch, y_i8, pinned=True):
    """The responses as the reference holds them — n x m doubles, NaN = missing, column-major — in pinned (or, for the 4 GB of config 4, pageable)
    host memory, filled in column slabs so that no second full-size temporary exists."""
    n, m = y_i8.shape
    buf = torch.empty((m, n), dtype=torch.float64, pin_memory=True).numpy() if pinned else np.empty((m, n))
    y_host = buf.T
    step = max(1, (1 << 25) // max(n, 1))
    for j0 in range(0, m, step):
        blk = y_i8[:, j0:j0 + step]
        y_host[:, j0:j0 + step] = np.where(blk == 0, np.nan, blk.astype(np.float64))
    return y_host


def synth(n, m, C, seed):
    from tests.problems import make_problem

    return make_problem(n, m, C, seed=seed, missing=0.05, on_grid=False, dtype="i8")


# ------------------------------------------------------------------------------------------------------------------
def cpu_sweeps(n, m, C, seed, *, faithful, warm, max_timed, min_timed, budget_s, state=None, blas=True):
    """Full-size sweeps of the reference algorithm on the host cores (the oracle port; R is absent from this image).

    faithful=1 ("ref-faithful"): the sweep as the reference codes it — second Cholesky per sweep (src/draw-f.cpp:59), both N x m tables copied
    per respondent in draw_theta (src/draw-theta.cpp:144-156), per-item dense gemv / two trsv (src/mvnormal.h:23, src/cholesky-cache.cpp:4-9).
    faithful=0 ("ref-tidy"): the same arithmetic without the redundant Cholesky and the table copies.
    blas: chol / trsm / syrk / gemv / trsv through the OpenBLAS of the scipy wheel (oracle/orc_blas.h), as an R linked against an optimised
    BLAS would run Armadillo's calls; otherwise the oracle's own loops. Threads are set explicitly (torchrun exports OMP_NUM_THREADS=1).
    Every timed sweep is a whole sweep over all m items; nothing is extrapolated. Timed sweeps stop at max_timed, or once min_timed are done
    and the next one would overrun budget_s."""
    from oracle import orc

    threads = os.cpu_count() or 1
    orc.set_threads(threads)
    blas_path = orc.use_blas(True, threads) if blas else (orc.use_blas(False) or None)
    P = synth(n, m, C, seed)
    y = np.where(P["y"] == 0, np.nan, P["y"].astype(np.float64))
    O = orc.Sampler(y, C, P["beta_prior_sds"], P["theta_prior_sds"], faithful=faithful)
    if state is None:
        O.set_state(theta=P["theta0"], beta=P["beta0"], thr=P["thr0"])
        O.compute_mu()
        assert O.update_cholesky() == 0
        f0 = np.asfortranarray(O.L() @ np.random.default_rng(seed).standard_normal((n, m)))  # f0_j = L z_j, src/gpirtMCMC.cpp:158-162
        O.set_state(f=f0)
        start = "the bench's initial state (theta0, beta0, default thresholds, f0 = L z)"
    else:
        O.set_state(theta=state["theta"], beta=state["beta"], thr=state["thresholds"], f=state["f"])
        O.compute_mu()
        assert O.update_cholesky() == 0
        start = "the state the GPU chain reached after its warm-up and timed sweeps"
    t_begin = time.perf_counter()
    it = 0
    for _ in range(warm):
        info, _, _ = O.run_mt(it, 1, store=False)
        assert info == 0
        it += 1
    t_warm = time.perf_counter() - t_begin
    O.times(reset=True)
    times = []
    while len(times) < max_timed:
        t0 = time.perf_counter()
        info, _, _ = O.run_mt(it, 1, store=False)
        assert info == 0
        it += 1
        times.append(time.perf_counter() - t0)
        if len(times) >= min_timed and (time.perf_counter() - t_begin) + float(np.mean(times)) > budget_s:
            break
    fixed, items = (t / len(times) for t in O.times())
    mean = float(np.mean(times))
    name = ("ref-faithful" if faithful else "ref-tidy") + ("+openblas" if blas_path else "")
    return dict(variant=name, value=1.0 / mean, ms_per_step=1e3 * mean, steps=len(times), warmup=warm, threads=O.num_threads(), cores=os.cpu_count(),
                blas=os.path.basename(blas_path) if blas_path else None, sweep_s=[round(t, 3) for t in times], warm_s=round(t_warm, 3),
                shared_s=round(fixed, 3), items_s=round(items, 3),
                sample=f"{len(times)} whole sweep(s) over all {m} items x {n} respondents of the problem given ({name}, {O.num_threads()} threads) after {warm} warm-up sweep(s) from {start}: "
                       f"{mean:.2f} s/sweep = {fixed:.2f} s shared by all items (chol(K), k* solves, chol(K_post)) + {items:.2f} s per-item / per-respondent work; measured, not extrapolated")


def reference_arm(args):
    """--impl reference: the reference's own algorithm on this box's host cores, on the `ours` arm's config. At N = 1 the full C3 workload is
    swept (a faithful sweep takes tens of seconds: the line reports how many sweeps were really timed, never more than --steps). At N > 1 the
    weak-scaled workload (5000 N items) does not fit the time limit on a CPU: the C3 block is timed and the line says extrapolated: true."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    if args.config == "c4":
        return reference_arm_c4(args)
    m_full = M_PER_GPU * args.gpus
    main = cpu_sweeps(N_RESP, M_PER_GPU, C_CAT, 20261104 + 3, faithful=1, warm=1, max_timed=max(1, args.steps), min_timed=min(3, max(1, args.steps)),
                      budget_s=args.ref_budget, blas=True)
    tidy = cpu_sweeps(N_RESP, M_PER_GPU, C_CAT, 20261104 + 3, faithful=0, warm=1, max_timed=min(3, max(1, args.steps)), min_timed=1, budget_s=40.0, blas=True)
    extrap = args.gpus > 1
    # weak-scaling metric: C3-sized item blocks swept per second. A CPU sweeping 5000 N items does N blocks per sweep in ~N times the time.
    value = main["value"]
    workload = (f"C3 synthetic ordinal n={N_RESP} x m={M_PER_GPU}, C={C_CAT}, 5% missing" if not extrap else
                f"C3 block (n={N_RESP} x m={M_PER_GPU}, C={C_CAT}) timed in full; the {args.gpus}-GPU workload n={N_RESP} x m={m_full} is NOT swept on the CPU: "
                f"blocks/s assumed constant in m (per-item work is linear in m), extrapolated")
    line = {"impl": "reference", "metric": "mcmc_sweeps_per_sec_c3_blocks", "value": value, "unit": "sweeps/s", "n_gpus": args.gpus,
            "steps": main["steps"], "warmup": main["warmup"], "steps_requested": args.steps, "warmup_requested": args.warmup,
            "ms_per_step": main["ms_per_step"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "extrapolated": extrap,
            "config": {"workload": workload, "n": N_RESP, "m": M_PER_GPU if extrap else m_full, "C": C_CAT},
            "cpu_baseline": {"value": value, "unit": "sweeps/s", "cores": main["threads"], "kind": "port", "sample": main["sample"]},
            "variants": {main["variant"]: {k: main[k] for k in ("value", "ms_per_step", "steps", "warmup", "threads", "blas", "sweep_s", "shared_s", "items_s")},
                         tidy["variant"]: {k: tidy[k] for k in ("value", "ms_per_step", "steps", "warmup", "threads", "blas", "sweep_s", "shared_s", "items_s")}},
            "variants_note": "value is ref-faithful (the sweep as the reference codes it); ref-tidy drops the reference's per-respondent table copies and its second "
                             "Cholesky (same arithmetic) and is the denominator to use for a speed-up that is not inflated by a reference inefficiency",
            "e2e": {"value": value, "unit": "sweeps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), file=_REAL_STDOUT, flush=True)


def reference_arm_c4(args):
    """Config 4 on the CPU: a whole sweep of 10 000 x 50 000 is out of reach (the faithful sweep copies 8 TB of tables per sweep), so
    the first 64 items are swept with all 10 000 respondents; the oracle clocks the work shared by all items (chol(K) at n = 10 000, the k* solves,
    chol(K_post)) apart from the per-item / per-respondent work, which is linear in m: value = 1 / (shared + items * 50000 / 64). EXTRAPOLATED."""
    W = workload("c4", args.gpus)
    m_sub = 64
    out = {}
    for faithful in (1, 0):
        r = cpu_sweeps(W["n"], m_sub, W["C"], W["seed"], faithful=faithful, warm=1, max_timed=2, min_timed=1, budget_s=120.0, blas=True)
        t_full = r["shared_s"] + r["items_s"] * W["m"] / m_sub
        out[r["variant"]] = dict(value=1.0 / t_full, ms_per_step=1e3 * t_full, steps=r["steps"], warmup=r["warmup"], threads=r["threads"], blas=r["blas"],
                                 sample_sweep_s=r["sweep_s"], shared_s=r["shared_s"], items_s_on_sample=r["items_s"], m_sample=m_sub)
    names = list(out)
    main = out[names[0]]
    sample = (f"first {m_sub} of {W['m']} items, all n={W['n']} respondents, {main['steps']} timed sweep(s) of {names[0]} on {main['threads']} threads: "
              f"{main['shared_s']:.2f} s shared by all items + {main['items_s_on_sample']:.2f} s per-item / per-respondent work on the sample, scaled by {W['m']}/{m_sub}; extrapolated")
    line = {"impl": "reference", "metric": W["metric"], "value": main["value"], "unit": "sweeps/s", "n_gpus": args.gpus, "steps": main["steps"], "warmup": main["warmup"],
            "steps_requested": args.steps, "warmup_requested": args.warmup, "ms_per_step": main["ms_per_step"], "higher_is_better": True, "scaling": W["scaling"],
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "extrapolated": True, "sample_m": m_sub,
            "config": {"workload": f"sub-sample of C4: n={W['n']} x the first {m_sub} of m={W['m']} items, C={W['C']}; per-item work scaled to m={W['m']} (extrapolated, not C4 itself)",
                       "n": W["n"], "m": m_sub, "C": W["C"]},
            "cpu_baseline": {"value": main["value"], "unit": "sweeps/s", "cores": main["threads"], "kind": "port", "sample": sample},
            "variants": out, "e2e": {"value": main["value"], "unit": "sweeps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), file=_REAL_STDOUT, flush=True)


def sharded_equals_single(rank, world, local):
    """Outside the timed region, N > 1: two sweeps of a small ragged problem item-sharded over the N ranks and unsharded on this rank's GPU;
    True iff theta, beta, thresholds and the total log-likelihood agree bit for bit on every rank (Philox streams are keyed by global unit)."""
    import torch
    import torch.distributed as dist

    import gpirt_b200
    from gpirt_b200.distributed import ShardedSampler

    n, m, C = 320, 64 * world + 3, 5
    P = synth(n, m, C, 99)
    out = []
    for sharded in (True, False):
        S = (ShardedSampler(n, m, C, rank=rank, world=world, device=local, seed=7, chain=0) if sharded else
             gpirt_b200.Sampler(n, m, C, device=local, seed=7, chain=0))
        S.set_data(P["y"])
        S.set_priors(P["beta_prior_sds"], P["theta_prior_sds"])
        S.set_state(theta=P["theta0"], beta=P["beta0"], thresholds=P["thr0"])
        S.init_draw(draw_beta=False, draw_f=True)
        S.sweep(2)
        S.sync()
        st = S.get_state(f=False)
        out.append((st["theta"].copy(), st["beta"].copy(), st["thresholds"].copy(), S.total_ll()))
        S.close()
    same = all(np.array_equal(a, b) for a, b in zip(out[0][:3], out[1][:3])) and out[0][3] == out[1][3]
    t = torch.tensor([1 if same else 0], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return bool(t.item())


# ------------------------------------------------------------------------------------------------------------------
def ours(args):
    import torch

    import gpirt_b200
    from gpirt_b200 import _lib
    from gpirt_b200.distributed import ShardedSampler

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("--gpus N > 1 must be launched under torch.distributed.run with N ranks")
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    W = workload(args.config, world)
    n, m, C = W["n"], W["m"], W["C"]
    P = synth(n, m, C, W["seed"])
    common = dict(seed=20261104, chain=0)
    if world == 1:
        S = gpirt_b200.Sampler(n, m, C, device=local, **common)
        S.set_stream(torch.cuda.current_stream().cuda_stream)
        sweep = S.sweep
    else:
        S = ShardedSampler(n, m, C, rank=rank, world=world, device=local, **common)
        sweep = S.sweep
    S.set_data(P["y"])
    S.set_priors(P["beta_prior_sds"], P["theta_prior_sds"])
    S.set_state(theta=P["theta0"], beta=P["beta0"], thresholds=P["thr0"])
    S.init_draw(draw_beta=False, draw_f=True)  # f0_j = L z_j as the reference initialises it (src/gpirtMCMC.cpp:158-162)
    L = _lib.lib()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sweep(args.warmup)
    S.sync()
    barrier()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    launches0 = L.gpirt_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    sweep(args.steps)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    launches = L.gpirt_launch_count() - launches0
    clk = clocks.stop() if rank == 0 else None
    S.sync()
    if world > 1:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ll = S.total_ll()
    ms_per_step = ms / args.steps
    value = W["blocks"] * 1e3 / ms_per_step  # c3: C3-sized item blocks swept per second (weak scaling); c4: sweeps of the whole problem per second

    # ---- per-kernel-group device times (CUDA events inside the library), separate pass ----
    S.profile(True)
    sweep(1)  # not recorded: the first serial sweep still finds the overlapped pass's Cholesky cache and shared f* terms valid and would skip those groups
    S.profile_reset()
    prof_sweeps = 3
    sweep(prof_sweeps)
    sections = S.profile_read()
    S.profile(False)
    per_sweep = {k: v[0] / prof_sweeps for k, v in sections.items()}
    launches_per = {k: v[1] / prof_sweeps for k, v in sections.items()}
    ev = {b: S.diag(b)[0] for b in ("f", "theta", "beta", "thr")}
    inner = S.S if hasattr(S, "S") else S
    ex = {b: inner.exact_evals(b) for b in ("f", "beta", "thr")}
    screened = {b: round(1.0 - float(np.maximum(ex[b][ev[b] > 0] - 1, 0).sum()) / max(1.0, float((ev[b][ev[b] > 0] - 1).sum())), 4) for b in ex}
    items_mode = world > 1 and getattr(S, "theta_shard", "") == "items"
    if items_mode:  # item-sharded theta step: fp64 partial sums of the local items, no fp32 screen
        screened["theta"] = None
    else:
        ext = inner.exact_evals("theta")  # theta step: the current state is screened too; an undecided proposal costs two exact evaluations the first time
        sel = ev["theta"] > 0
        screened["theta"] = round(1.0 - float(np.maximum(ext[sel] - 1, 0).sum()) / max(1.0, float((ev["theta"][sel] - 1).sum())), 4)
    E = {b: float(np.mean(ev[b][ev[b] > 0])) if np.any(ev[b] > 0) else 0.0 for b in ev}
    m_loc, N = W["m_loc"], 1001
    nm = n * m_loc
    algo = {  # algorithmic bytes (HBM-bound groups) or flops (FP64-tensor groups) per sweep, SURVEY.md 8(d)
        "ess_f": ("hbm", 25.0 * nm), "ess_theta": ("hbm", E["theta"] * (n * m_loc if items_mode else (n / world) * m) * 9.0), "ess_beta": ("hbm", 9.0 * nm),
        "ess_thr": ("hbm", 9.0 * nm), "item_ll": ("hbm", 9.0 * nm + 8.0 * (n + 3 * m_loc + (C + 1) * m_loc)), "gather_f": ("hbm", 16.0 * nm),
        "gen_Z": ("hbm", 8.0 * (n + N) * m_loc), "theta_table": ("hbm", 8.0 * N * (m_loc if items_mode else m) * (C + 1)), "add_mustar": ("hbm", 16.0 * N * m_loc), "add_mustar_push": ("hbm", 8.0 * (world + 1) * N * m_loc), "K_build": ("hbm", 8.0 * (n * n + n * N + N * N)),
        "gemm_LZ": ("tensor", 1.0 * n * n * m_loc), "gemm_FtW": ("tensor", 2.0 * N * n * m_loc), "gemm_ZLpost": ("tensor", 1.0 * N * N * m_loc),
        "gemm_Kpost": ("tensor", 1.0 * n * N * N), "chol_K": ("tensor", n ** 3 / 3.0), "chol_Kpost": ("tensor", N ** 3 / 3.0),
        "trsm_kstar": ("tensor", 1.0 * n * n * N), "trsm_W": ("tensor", 1.0 * n * n * N),
        "trtri": ("tensor", n ** 3 / 3.0), "gemm_tmp": ("tensor", 1.0 * n * n * N), "gemm_W": ("tensor", 1.0 * n * n * N),
    }
    peaks, peak_src = measured_peaks()
    try:  # fp64 roofline denominator measured in this run, on this GPU, outside the timed region (MEASURED_PEAKS.json has no fp64 figure)
        from gpirt_b200.api import fp64_peak

        FP64_DMMA_PEAK_TFLOPS, dfma_peak = fp64_peak(local)
        fp64_src = f"gpirt_debug_fp64_peak in this run: DMMA.8x8x4 issue rate {FP64_DMMA_PEAK_TFLOPS:.2f} TFLOP/s, scalar DFMA {dfma_peak:.2f} TFLOP/s"
    except Exception as exc:  # noqa: BLE001
        FP64_DMMA_PEAK_TFLOPS, fp64_src = FP64_DMMA_PEAK_FALLBACK, f"fallback literal (microbenchmark failed: {exc})"
    kernels = []
    for k, t in sorted(per_sweep.items(), key=lambda kv: -kv[1]):
        bound, amount = algo.get(k, ("hbm", 0.0))
        ach = amount / (t * 1e-3) / (1e9 if bound == "hbm" else 1e12) if t > 0 else 0.0
        peak = peaks["hbm_gbs"] if bound == "hbm" else FP64_DMMA_PEAK_TFLOPS
        kernels.append({"group": k, "ms_per_sweep": round(t, 4), "launches_per_sweep": launches_per[k], "bound": bound,
                        "achieved": round(ach, 2), "unit": "GB/s" if bound == "hbm" else "TFLOP/s", "frac": round(ach / peak, 4)})
    # The dominant KERNEL of a sweep is the fp64 DMMA GEMM (one kernel, seven call sites); chol_K is the largest GROUP but it is 64
    # launches of two small kernels and runs on the side stream, off the critical path (profiles/r01_launches_v5.csv).
    gemm_groups = [k for k in per_sweep if k.startswith("gemm_") or k == "trtri"]
    gemm_ms = sum(per_sweep[k] for k in gemm_groups)
    gemm_flops = sum(algo[k][1] for k in gemm_groups)
    gemm_launches = sum(launches_per[k] for k in gemm_groups)
    ach = gemm_flops / (gemm_ms * 1e-3) / 1e12 if gemm_ms > 0 else 0.0
    roofline = {"kernel": "gemm_f64_kernel<64,64,16,32,32,3> (mma.sync m8n8k4 f64 = SASS DMMA.8x8x4; tcgen05 has no fp64 kind)",
                "bound": "tensor", "achieved": round(ach, 2), "peak": FP64_DMMA_PEAK_TFLOPS, "unit": "TFLOP/s", "frac": round(ach / FP64_DMMA_PEAK_TFLOPS, 4),
                "traffic": 167.3e6,
                "traffic_note": "dram read+write of the largest launch (L*Z: algorithmic 176 MB), ncu --set full, profiles/r02_ncu_gemm_band.txt (column tiles in two L2-sized bands; 305.3 MB with every column tile under every row tile, profiles/r01_ncu_gemm64.txt)",
                "peak_source": fp64_src + "; MEASURED_PEAKS.json has no fp64 figure",
                "groups": gemm_groups, "ms_per_sweep": round(gemm_ms, 4), "call_sites_per_sweep": gemm_launches,
                "share_of_sweep": round(gemm_ms / sum(per_sweep.values()), 3),
                "algorithmic_flops_per_sweep": gemm_flops,
                "ncu_share_note": "ncu launch list of the same sweep (profiles/r01_launches_v9_summary.txt; cold-cache, serialised): the kernel's launches outside "
                                  "the Cholesky (templates <...,0,0>, <...,1,0>, <...,1,1>: 70 launches) take 41 % of the summed kernel time, its 184 K = 64 "
                                  "trailing updates inside chol_K (<...,0,1>, ~11 us each, counted in the chol_K group here) another 10 %; the live share above "
                                  "is lower because the latency-bound Cholesky panels (22 % under ncu) are not inflated by serialisation and cold caches there",
                "note": "useful flops (triangular operands counted once) over the summed durations of the kernel's launches; ncu shows the DMMA pipe "
                        "90-91 % busy in these launches. The ESS / likelihood groups in `kernels` are issue-bound on fp64+fp32 arithmetic (2 Phi + 1 log "
                        "per cell and evaluation), not HBM-bound: their algorithmic GB/s stays far below the copy peak by construction."}
    # log-likelihood GB/s (BASELINE metric, second half): the full-matrix ll_bar pass timed alone
    ll_ms = per_sweep.get("item_ll", 0.0)
    loglik_gbs = algo["item_ll"][1] / (ll_ms * 1e-3) / 1e9 if ll_ms > 0 else None

    line = {"metric": W["metric"], "value": value, "unit": "sweeps/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": W["scaling"], "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": W["text"], "n": n, "m": m, "C": C,
                       "l2": "working set per sweep (f, nu, Z, f*, L, W: >350 MB) exceeds the 126 MB L2; no explicit flush",
                       "parallelism": f"items sharded over {world} GPU(s), theta step " + (
                           f"sharded by items too: partial log-likelihoods of 16 proposals per unfinished respondent exchanged per round ({getattr(S, 'theta_rounds', 0)} "
                           "NCCL all-gathers of n x 17 doubles in the last sweep), no f* exchange" if items_mode else "sharded over respondents (f* all-gather)")},
            "gpu_launches": int(launches), "ll_after": ll, "mean_evals": E, "proposals_decided_in_fp32": screened,
            "kernels_note": "per-group device times of a separate pass in the serial order (profiling level 1); the timed pass overlaps the "
                            "theta-only chain (K_build, chol_K, trtri, gemm_tmp, gemm_Kpost, chol_Kpost, gemm_W) with the item kernels on a second stream",
            "kernels": kernels, "roofline": roofline,
            "loglik_gbs": loglik_gbs}
    if rank == 0:
        line["clocks"] = clk
    big = W["name"] == "c4"  # 4 GB of fp64 responses: fewer end-to-end repetitions, pageable host memory, no CPU sweep beside it (--impl reference extrapolates)
    gpu_state = S.get_state() if world == 1 and not args.no_cpu_baseline and not big else None  # where the CPU baseline below starts from
    if world > 1:
        line["sharded_equals_single"] = sharded_equals_single(rank, world, local)

    # ---- e2e: the reference-facing call with HOST buffers (gpirtMCMC -> gpirt_mcmc C-ABI), copies inside the timed region ----
    if world == 1:
        S.close()
        y_host = host_responses(torch, P["y"], pinned=not big)  # column-major n x m
        k = args.steps
        rates = []
        for _ in range(2 if big else 9):  # wall clock on a shared host (set_data / create vary by several ms between calls): the median of nine calls is reported
            t0 = time.perf_counter()
            res = gpirt_b200.gpirtMCMC(y_host, sample_iterations=k, burn_iterations=0, THIN=1, CHAIN=1, beta_prior_sds=P["beta_prior_sds"],
                                       theta_prior_sds=P["theta_prior_sds"], theta_init=P["theta0"], thresholds=P["thr0"], SEED=20261104,
                                       beta_init=P["beta0"], device=local)
            rates.append(k / (time.perf_counter() - t0))
            assert np.all(np.isfinite(res[0]["ll"]))
        h2d = (y_host.nbytes + 8 * (n + 3 * m + m * (C + 1) + 3 * m + 2 * n)) / k
        d2h = 8 * (n + 3 * m + m * (C + 1) + 1)
        line["e2e"] = {"value": float(np.median(rates)), "unit": "sweeps/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "calls": [round(r, 2) for r in rates],
                       "host_memory": "pageable" if big else "pinned",
                       "what": f"gpirtMCMC(sample_iterations={k}) wall clock (median of {2 if big else 9} calls) incl. handle creation, fp64 y upload from pinned memory, initial draws, the kept theta/beta/threshold/ll draws of every sweep copied to the host (buffered on the device, read back in bulk), handle destruction"}
        # ---- batched chains (BASELINE config 5's mode at the C3 shape): the same call with CHAIN = 4, chains side by side on the device ----
        nch = 4
        t0 = time.perf_counter()
        res = [] if big else gpirt_b200.gpirtMCMC(y_host, sample_iterations=k, burn_iterations=0, THIN=1, CHAIN=nch, beta_prior_sds=P["beta_prior_sds"],
                                   theta_prior_sds=P["theta_prior_sds"], theta_init=P["theta0"], thresholds=P["thr0"], SEED=20261104, device=local)
        if not big:
            line["batched_chains"] = {"chains": nch, "value": nch * k / (time.perf_counter() - t0), "unit": "chain-sweeps/s",
                                      "what": f"gpirtMCMC(CHAIN={nch}, sample_iterations={k}) end to end with host buffers (one gpirt_mcmc_chains call: y uploaded "
                                              "once, chains on their own streams, sweeps issued round-robin); each chain is bit-identical to the chain run alone"}
        assert all(np.all(np.isfinite(r["ll"])) for r in res)
        # ---- CPU baseline beside it (bounded sample, rank 0, N=1 only) ----
        if not args.no_cpu_baseline and not big:
            # bounded sample: ONE whole faithful sweep of the full workload (tens of seconds) and two tidy ones, from the GPU chain's own state
            r = cpu_sweeps(n, m, C, 20261104 + 3, faithful=1, warm=0, max_timed=1, min_timed=1, budget_s=0.0, state=gpu_state, blas=True)
            t = cpu_sweeps(n, m, C, 20261104 + 3, faithful=0, warm=0, max_timed=2, min_timed=1, budget_s=30.0, state=gpu_state, blas=True)
            line["cpu_baseline"] = {"value": r["value"], "unit": "sweeps/s", "cores": r["threads"], "kind": "port", "sample": r["sample"],
                                    "variant": r["variant"], "blas": r["blas"],
                                    "tidy": {"value": t["value"], "variant": t["variant"], "sample": t["sample"]}}
    else:
        # the same call sharded: every rank uploads its columns / rows of the pinned fp64 y, runs the chain with the exchanges, and reads
        # the kept draws back; wall clock between barriers, max over ranks
        from gpirt_b200.distributed import gpirtMCMC_sharded

        S.close()
        y_host = host_responses(torch, P["y"], pinned=not big)
        k = args.steps
        rates = []
        for _ in range(2 if big else 5):
            barrier()
            t0 = time.perf_counter()
            res = gpirtMCMC_sharded(y_host, k, 0, THIN=1, rank=rank, world=world, device=local, beta_prior_sds=P["beta_prior_sds"],
                                    theta_prior_sds=P["theta_prior_sds"], theta_init=P["theta0"], thresholds=P["thr0"], SEED=20261104, beta_init=P["beta0"])
            barrier()
            t = torch.tensor([time.perf_counter() - t0], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            rates.append(W["blocks"] * k / float(t.item()))
            assert np.all(np.isfinite(res["ll"]))
        # per rank: its items' columns and its respondents' rows of y (2/world of the matrix), the small state arrays; draws come back on every rank
        h2d = world * (2.0 * y_host.nbytes / world + 8 * (n + 3 * m + m * (C + 1) + 3 * m + 2 * n)) / k
        d2h = world * 8 * (n + 3 * m + m * (C + 1) + 1)
        line["e2e"] = {"value": float(np.median(rates)), "unit": "sweeps/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "calls": [round(r, 2) for r in rates],
                       "host_memory": "pageable" if big else "pinned",
                       "what": f"gpirtMCMC_sharded(sample_iterations={k}) on {world} ranks, wall clock between barriers (max over ranks, median of {2 if big else 5} calls) incl. handle "
                               "creation, each rank's partial fp64 y upload from pinned memory, initial draws, exchanges, kept draws read back on every rank; "
                               + (f"value = {world} C3-sized item blocks x sweeps per second, as `value`" if not big else "value = sweeps of the whole problem per second")}
    if rank == 0:
        print(json.dumps(line), file=_REAL_STDOUT, flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    # The ONE JSON line goes to the real stdout; anything libraries print to fd 1 meanwhile (NCCL's version banner, ...) goes to stderr.
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="c3", choices=["c3", "c4"], help="c3: BASELINE configs[2], the headline (weak scaling over items); c4: configs[3], n=10000 x m=50000, C=7 (strong scaling)")
    ap.add_argument("--ref-budget", type=float, default=150.0, help="--impl reference: wall-clock budget (s) of the faithful variant's warm-up + timed sweeps")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        reference_arm(args)
    else:
        ours(args)


if __name__ == "__main__":
    main()
