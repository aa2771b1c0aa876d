#!/usr/bin/env python
"""bench.py -- GP posterior cells/s per BO step (BASELINE.json metric) on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload C2] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A step = one pass of the hot path (posterior mu/sigma^2 over every cell of this rank's band of grid rows,
CI + safe mask + argmax, and for N > 1 the NCCL gather of the per-rank argmax records) over a seeded
synthetic problem.  Workload at N = 1: BASELINE.json configs[1] = "C2" (1 024 observations x 1 048 576
cells, fp32).  N > 1: every rank owns one C2-sized band of rows of an N-times taller grid (weak scaling,
replicated L^-1/alpha, no data-path collective besides the 48-byte argmax records).

Timing: W warm-up steps, then K steps each bracketed by CUDA events on the stream the kernels run on
(the library is pointed at torch's current stream), L2 flushed between steps (outside the events),
barrier + torch.cuda.synchronize on both sides of the timed region, max over ranks.
`value` is device-resident (inputs already in HBM); `e2e` goes through the host-buffer C-ABI calls a
user makes: sbo_set_observations (H2D + factorisation) + sbo_step (posterior, sets, mask D2H).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="C2", choices=["C1", "C2", "C3", "C4", "C5"])
    ap.add_argument("--kernel", default="auto", choices=["auto", "simt", "tcgen05"])
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="budget of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return d, "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


def credited_flops(m, n):
    """triangular contraction m*n*(n+1) + mu 2*m*n; splits/zeros/K* generation earn nothing."""
    return float(m) * n * (n + 1) + 2.0 * m * n


class ClockSampler:
    """SM clock and clock-event reasons sampled DURING the timed region.  In-process NVML thread (every 2 ms: a C2 step
    lasts 2 ms and the whole timed region tens of ms, which a 200 ms nvidia-smi loop sees once at best); nvidia-smi -lms 200
    as the fallback when pynvml is missing."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    REASON_BITS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown"}

    def __init__(self, device_index):
        self.idx, self.proc, self.file = device_index, None, None
        self.thread, self.stop_flag, self.samples, self.nvml, self.handle = None, False, [], None, None

    def _nvml_open(self):
        import pynvml
        pynvml.nvmlInit()
        try:
            import torch
            uuid = str(torch.cuda.get_device_properties(self.idx).uuid)
            uuid = uuid if uuid.startswith("GPU-") else "GPU-" + uuid
            h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode() if hasattr(uuid, "encode") else uuid)
        except Exception:
            h = pynvml.nvmlDeviceGetHandleByIndex(self.idx)
        self.nvml, self.handle = pynvml, h

    def _loop(self):
        nv, h = self.nvml, self.handle
        reasons_fn = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
        while not self.stop_flag:
            try:
                self.samples.append((nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM), int(reasons_fn(h)),
                                     nv.nvmlDeviceGetPowerUsage(h) / 1000.0))
            except Exception:
                break
            time.sleep(0.002)

    def start(self):
        try:
            import threading
            self._nvml_open()
            self.stop_flag, self.samples = False, []
            self.thread = threading.Thread(target=self._loop, daemon=True)
            self.thread.start()
            return
        except Exception:
            self.thread = None
        try:
            self.file = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.idx)], stdout=self.file, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        if self.thread is not None:
            self.stop_flag = True
            self.thread.join(timeout=2)
            if not self.samples:
                return None
            sm = [x[0] for x in self.samples]
            bits = 0
            for x in self.samples:
                bits |= x[1]
            try:
                mx = float(self.nvml.nvmlDeviceGetMaxClockInfo(self.handle, self.nvml.NVML_CLOCK_SM))
            except Exception:
                mx = float(max(sm))
            return {"sm_mhz": float(np.median(sm)), "sm_min_mhz": float(min(sm)), "sm_max_mhz": mx,
                    "reasons": sorted(n for b, n in self.REASON_BITS.items() if bits & b), "samples": len(sm),
                    "power_w_max": float(max(x[2] for x in self.samples)), "how": "NVML, every 2 ms, during warm-up + timed steps"}
        if not self.proc:
            return None
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.file.flush()
        rows = [r.strip().split(", ") for r in open(self.file.name).read().strip().splitlines() if r.strip()]
        os.unlink(self.file.name)
        sm, mx, reasons = [], [], set()
        for r in rows:
            if len(r) < 9:
                continue
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.strip().lower() == "active":
                    reasons.add(name)
        if not sm:
            return None
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons), "samples": len(sm),
                "how": "nvidia-smi -lms 200"}


def cpu_posterior_sample(p, seconds, prefer=None):
    """Times the CPU restatement (oracle) of the same step on a bounded band of cells of workload p.
    Picks the faster of the C/OpenMP port and the scipy/OpenBLAS port on a calibration band first."""
    from oracle import oracle as O
    hp = O.Hyper(p.sigma_f, p.length, p.noise_var, p.prior_mean)
    g = O.Grid(p.x0, p.y0, p.rx, p.ry, p.W, p.H)
    co = O.COracle()
    t0 = time.perf_counter()
    _, Linv, alpha = O.gp_fit_scipy(p.xy, p.y, hp)
    t_fit = time.perf_counter() - t0

    sets_time = [0.0]

    def run(kind, cells):
        t = time.perf_counter()
        if kind == "c_openmp":
            mu, s2 = co.posterior(p.xy, hp, Linv, alpha, g, 0, cells)
        else:
            mu, s2 = O.gp_posterior_scipy(p.xy, hp, Linv, alpha, g, 0, cells)
        t1 = time.perf_counter()
        q0, q1 = co.confidence_intervals(mu, np.sqrt(np.maximum(s2, 0.0)), p.beta)   # the node's own arithmetic (A4 + A5,
        mask, _ = co.safe_set(q0, p.f_min)                                            # reference :399-416), single thread
        co.argmax_safe(q1, mask)
        sets_time[0] = time.perf_counter() - t1
        return time.perf_counter() - t

    calib = min(p.m, 4096)
    rates = {k: calib / run(k, calib) for k in ("c_openmp", "scipy_openblas")}
    kind = prefer or max(rates, key=rates.get)
    cells = int(min(p.m, max(calib, rates[kind] * seconds)))
    cells = max(p.W, cells // p.W * p.W) if cells >= p.W else cells
    dt = run(kind, cells)
    cores = os.cpu_count() or 1
    return {"value": cells / dt, "unit": "cells/s", "cores": cores, "kind": "port",
            "sample": f"{kind} fp64 oracle, {cells} of {p.m} cells ({cells // max(p.W, 1)} grid rows) of {p.name} "
                      f"n={p.n}: posterior+CI+mask+argmax in {dt:.2f} s on {cores} host threads "
                      f"(factorisation {t_fit:.2f} s not included; calibration {rates})",
            "seconds": dt, "cells": cells,
            "node_arithmetic": {"what": "ComputeSets (CI + mask) + argmax restated in C, 1 thread, on the sample's mu / sigma",
                                "ms": sets_time[0] * 1e3, "cells": cells, "cells_per_s": cells / max(sets_time[0], 1e-9)}}


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU path for this step.  Nothing of /root/reference computes a GP
    posterior or can be compiled here, so the arm is the oracle port of it, with all host threads,
    each step a bounded sample of the same workload.  Rank 0 only."""
    if rank != 0:
        return
    from safe_bayesian_optimization_b200 import synthetic as S
    p = S.make_problem(args.workload)
    per_step = max(2.0, min(args.cpu_seconds, 150.0 / max(1, args.steps + args.warmup)))
    first = cpu_posterior_sample(p, per_step)
    kind = first["sample"].split(" ")[0]
    vals = [first]
    for _ in range(args.warmup + args.steps - 1):
        vals.append(cpu_posterior_sample(p, per_step, prefer=kind))
    timed = vals[args.warmup:]
    cells = sum(v["cells"] for v in timed); secs = sum(v["seconds"] for v in timed)
    value = cells / secs
    line = {"impl": "reference", "metric": "gp_posterior_cells_per_s", "value": value, "unit": "cells/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * secs / max(1, len(timed)),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"{args.workload}: {p.n} observations x {p.m} cells ({p.W}x{p.H}), RBF l={p.length}, "
                                   f"sigma_f={p.sigma_f}, noise_var={p.noise_var}; bounded sample per step"},
            "cpu_baseline": {"value": value, "unit": "cells/s", "cores": first["cores"], "kind": "port",
                             "sample": first["sample"]},
            "e2e": {"value": value, "unit": "cells/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def executed_over_credited(n: int) -> float:
    """Tensor flops the tcgen05 kernel executes per credited flop: 3 split products (lo.hi, hi.lo, hi.hi) times the
    block-triangular padding of its 256-row blocks / 64-wide diagonal chunks (posterior_tc.cu)."""
    n64 = (n + 63) // 64 * 64
    macs = 0
    for jb in range((n + 255) // 256):
        for kc in range(min(n64, (jb + 1) * 256) // 64):
            c = max(kc - 4 * jb, 0)
            macs += 64 * (256 - 64 * c)
    return 3.0 * 2.0 * macs / (n * (n + 1) + 2 * n)


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0")); local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        return run_reference(args, rank, world)

    import torch
    import torch.distributed as dist
    import safe_bayesian_optimization_b200 as pkg
    from safe_bayesian_optimization_b200 import synthetic as S

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    p = S.make_problem(args.workload)
    rows_per_rank = p.H
    H_total = rows_per_rank * world                    # weak scaling: one workload-sized band of rows per rank
    m_local = rows_per_rank * p.W
    m_total = m_local * world
    n_extra = S.CONFIGS[args.workload].get("append", 0)

    ctx = pkg.SboContext(sigma_f=p.sigma_f, length_scale=p.length, noise_var=p.noise_var, prior_mean=p.prior_mean,
                         precision=p.precision, kernel=args.kernel, device=local_rank, max_obs=p.n + max(n_extra, 1) * 8,
                         max_cells=m_local)
    stream = torch.cuda.current_stream()
    ctx.set_stream(stream.cuda_stream)
    r0, r1 = rank * rows_per_rank, (rank + 1) * rows_per_rank
    ctx.set_grid(p.x0, p.y0, p.rx, p.ry, p.W, H_total, r0, r1)
    if world > 1:
        uid = [ctx.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        ctx.comm_init(uid[0], rank, world)
    ctx.set_observations(p.xy, p.y)                    # fit once; device-resident state for `value`

    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")   # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed_loop(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        barrier()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        outs = []
        wall0 = time.perf_counter()
        for a, b in ev:
            flush.fill_(1)                             # L2 flush between timed iterations (outside the events)
            a.record(stream)
            outs.append(fn())
            b.record(stream)
        barrier()
        wall = time.perf_counter() - wall0
        ms = sum(a.elapsed_time(b) for a, b in ev)
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), outs, wall

    # ---- device-resident steps: `value` -------------------------------------------------------------
    sampler = ClockSampler(local_rank)
    l0 = ctx.launch_count
    if rank == 0:
        sampler.start()
    ms_total, outs, wall = timed_loop(lambda: ctx.step_device(p.beta, p.f_min), args.steps, args.warmup)
    clocks = sampler.stop() if rank == 0 else None
    launches = (ctx.launch_count - l0) * args.steps // (args.steps + args.warmup)
    ms_step = ms_total / args.steps
    value = m_total / (ms_step * 1e-3)
    ms_post = float(np.mean([o.ms_posterior for o in outs]))
    ms_sets = float(np.mean([o.ms_sets for o in outs]))
    ms_comm = float(np.mean([o.ms_comm for o in outs]))

    # ---- end to end through the host-buffer API: `e2e` -----------------------------------------------
    mask_host = torch.empty(m_local, dtype=torch.uint8).pin_memory().numpy()
    xy_host = torch.from_numpy(p.xy.copy()).pin_memory().numpy()
    y_host = torch.from_numpy(p.y.copy()).pin_memory().numpy()

    def e2e_step():
        ctx.set_observations(xy_host, y_host)          # H2D of the step's inputs + fp64 factorisation
        res, _ = ctx.step(p.beta, p.f_min, mask_out=mask_host)   # posterior, sets, mask + record D2H
        return res
    e2e_steps = max(2, min(args.steps, 5))
    ms_e2e_total, e2e_outs, _ = timed_loop(e2e_step, e2e_steps, 1)
    ms_e2e = ms_e2e_total / e2e_steps
    e2e_value = m_total / (ms_e2e * 1e-3)

    # ---- the node's steady state, for information: rank-k append (k = the config's, else 1) + incremental step --------
    k_inc = max(n_extra, 1)
    inc_idx = [0]

    def inc_step():
        xy2, y2 = S.extra_observations(p, k_inc, inc_idx[0]); inc_idx[0] += 1
        ctx.append_observations(xy2, y2)                # H2D of the new observations + rank-k update of L, L^-1, alpha
        res, _ = ctx.step(p.beta, p.f_min, mask_out=mask_host, incremental=True)
        return res
    ctx.set_observations(xy_host, y_host)
    ctx.step(p.beta, p.f_min, mask_out=mask_host)
    ms_inc_total, inc_outs, _ = timed_loop(inc_step, 5, 1)
    ms_inc = ms_inc_total / 5

    if rank == 0:
        peaks, peak_src = measured_peaks()
        used_tc = outs[-1].kernel_used == 2
        flops = credited_flops(m_local, p.n)
        if used_tc:
            # long steps run under the power cap: sustained figure; a sub-10 ms kernel timed alone: burst
            key = "bf16_tflops_sustained" if ms_post > 50.0 else "bf16_tflops"
            peak = peaks[key]
            peak_note = f"{key} (kind::f16 tensor rate = the measured cuBLAS bf16 rate) of {peak_src}"
        elif p.precision == "fp64":
            # MEASURED_PEAKS.json has no fp64 entry; profiles/fp64_dmma_r01.md: 37.1 TFLOP/s for DFMA and DMMA alike on this pool's B200s
            peak, peak_note = 37.1, "fp64 pipe (DFMA = DMMA m8n8k4), measured by scripts/probe/fp64_peak.cu (nominal 40)"
        else:
            peak, peak_note = 75.0, "fp32 FMA pipe, nominal (the SIMT cross-check kernel; not the product path)"
        achieved = flops / (ms_post * 1e-3) / 1e12
        kname = {1: "posterior_simt_kernel", 2: "posterior_tc_kernel", 3: "posterior_thin_kernel", 4: "posterior_f64_kernel"}.get(outs[-1].kernel_used, "?")
        if outs[-1].kernel_used == 2 and p.n > 256 and ctx.tc_plan()[0] == 2 and os.environ.get("SBO_TC_PAIR", "1") != "0":
            kname = "posterior_tc_pair_kernel"   # full refresh over >= 2 row blocks of L^-1: the K*-stationary form (posterior_tc_pair.cuh)
        traffic = None
        try:   # per-launch DRAM bytes of the dominant kernel from the committed ncu --set full capture of this workload
            tj = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic_r01.json")))
            traffic = tj[kname][args.workload]["dram_bytes_per_launch"]
        except Exception:
            traffic = None
        roofline = {"bound": "tensor" if used_tc else ("fp64" if p.precision == "fp64" else "fp32"), "achieved": achieved, "peak": peak,
                    "unit": "TFLOP/s", "frac": achieved / peak, "traffic": traffic, "kernel": kname,
                    "peak_source": peak_note, "ms_kernel": ms_post, "credited_flops_per_launch": flops,
                    "executed_over_credited": executed_over_credited(p.n) if used_tc else 1.0}
        if used_tc:
            cg, skipping, kept = ctx.tc_plan()
            roofline["cta_group"] = cg
            if skipping:   # block-sparse K*: the credited (dense triangular) flops include work that is never executed
                roofline["executed_over_credited"] *= kept
                roofline["block_sparse"] = {"kept_fraction_estimate": kept,
                                            "note": "chunks of 64 observations whose K* entries are all < 2^-b sigma_f^2 (b = 36..60, bound 1e-7 sigma_f^2 on sigma^2; DESIGN 3.1) for a "
                                                    "cell tile are skipped; frac counts the dense triangular flops, so it can "
                                                    "exceed 1/3 (SBO_TC_SKIP=0 runs the dense walk)"}
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            cpu = cpu_posterior_sample(p, args.cpu_seconds)
            cpu = {k: cpu[k] for k in ("value", "unit", "cores", "kind", "sample")}
        line = {"metric": "gp_posterior_cells_per_s", "value": value, "unit": "cells/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f16x3" if used_tc else ("f64" if p.precision == "fp64" else "f32"),
                "data": "synthetic",
                "config": {"workload": f"{args.workload}: {p.n} observations x {m_local} cells per GPU ({p.W}x{rows_per_rank} "
                                       f"rows, grid {p.W}x{H_total}), RBF l={p.length} sigma_f={p.sigma_f} noise_var={p.noise_var}, "
                                       f"{p.precision}", "l2": "flushed between timed steps (256 MiB write, outside the events)",
                           "parallelism": f"cells sharded by row bands over {world} GPU(s), replicated L^-1/alpha"},
                "phases_ms": {"posterior": ms_post, "sets": ms_sets, "comm": ms_comm},
                "roofline": roofline, "cpu_baseline": cpu, "clocks": clocks,
                "e2e": {"value": e2e_value, "unit": "cells/s", "ms_per_step": ms_e2e, "h2d_bytes_per_step": int(p.n * 24),
                        "d2h_bytes_per_step": int(m_local + 48), "includes": "H2D observations + fp64 Cholesky/inverse + posterior + sets + mask D2H"},
                "e2e_incremental": {"value": m_total / (ms_inc * 1e-3), "unit": "cells/s", "ms_per_step": ms_inc, "k": k_inc,
                                    "kernel_used": inc_outs[-1].kernel_used,
                                    "includes": "sbo_append_observations (rank-k update) + sbo_step_incremental + mask D2H; "
                                                "not the headline: the full posterior above is"},
                "gpu_launches": int(launches), "wall_s": wall,
                "result": {"safe_count": outs[-1].safe_count, "argmax_ucb": list(outs[-1].argmax_ucb)}}
        print(json.dumps(line), flush=True)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
