#!/usr/bin/env python3
"""bench.py — the hot path's headline benchmark (BASELINE.json config 4: area+diameter-bounded DP, synthetic clustered
n = 200, diameter sweep D in {2,3,4,5,6}, A = 4, K = inf, reconstruct = true — the call shape of src/bench/main.cpp:57-59,100).

One STEP = the whole diameter sweep on one point set: five MT::AntipodalOptimizedAlgorithm-equivalent solves.
Candidate diameter pairs are dealt over the ranks (adk_area_diameter_partial*, shard = rank; cost-sorted round robin), the
only exchanges are a 32-byte record per rank per solve (all-gather) and the witness (one broadcast of n + 8 words from the
rank that owns the winning pair, which reconstructs it from its resident records).  Total work is fixed as N grows: "strong".

  value : DP cell updates/s (one cell = one iteration of AntipodalOptimized.cpp:198,225,271,298; counted in-kernel and
          equal to the instrumented oracle's count), inputs resident in HBM, device-timed over K steps, max over ranks.
  e2e   : the same metric through the reference-facing C-ABI call with HOST buffers (pinned), H2D/D2H inside the timed region.
  --impl reference : the UNMODIFIED reference (oracle/_ref/libadkref.so) on the host cores, on a bounded sample of the workload.

  checks: every answer of the sweep is compared with the unmodified reference's frozen output where one exists (tests/golden: all five
          diameters at sigma = 1.5, D = 2 at sigma = 1.0) and its witness is validated in any case (witness_is_valid).

Usage: python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--sigma S]
                       [--workload ad_n200|area_n1000|cells_batch] [--no-cpu] [--memory-budget-gb G]
"""
import argparse
import json
import math
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "dp_cell_updates_per_sec"
UNIT = "cells/s"
BYTES_PER_CELL = 20      # read-modify-write of the 8-byte destination area + 4-byte packed back-pointer
FLOPS_PER_CELL = 20      # two Orientation tests (2 x 9) + area add + compare
INT_OPS_PER_CELL = 4      # index and count arithmetic per cell update
FP64_NOMINAL_TFLOPS = 37.2   # 148 SM x 64 FP64 lanes x 2 x 1.965 GHz (no FP64 entry in MEASURED_PEAKS.json)
INT32_NOMINAL_TOPS = 18.6    # 148 SM x 64 INT32 lanes x 1.965 GHz
REF_CORES = 16            # replicas of the single-threaded reference timed side by side (fixed, so that every box reports the same arm)


def fp64_peak():
    """FP64 issue peak measured on this pool's B200 by tools/issue_peaks.cu (dependency-free DFMA chains), else nominal."""
    p = os.path.join(ROOT, "profiles", "issue_peaks.json")
    try:
        with open(p) as f:
            d = json.load(f)
        return float(d["fp64_tflops_fma"]), ("measured DFMA issue peak (profiles/issue_peaks.json, tools/issue_peaks.cu: "
                                              f"{d['dfma_per_sm_per_clk_at_max']:.1f} DFMA/SM/clk); 20 flop/cell (SURVEY 8d)")
    except Exception:
        return FP64_NOMINAL_TFLOPS, "nominal 148 SM x 64 lanes x 2 x 1.965 GHz; 20 flop/cell (SURVEY 8d)"


def int32_peak():
    """INT32 issue peak (IMAD per SM per clock x SMs x clock) measured by tools/issue_peaks.cu, else nominal."""
    p = os.path.join(ROOT, "profiles", "issue_peaks.json")
    try:
        with open(p) as f:
            d = json.load(f)
        return float(d["imad_per_s"]) / 1e12, "measured IMAD issue peak (profiles/issue_peaks.json, tools/issue_peaks.cu)"
    except Exception:
        return INT32_NOMINAL_TOPS, "nominal 148 SM x 64 lanes x 1.965 GHz"


def frozen_answers(sigma):
    """The unmodified reference's answers for the sweep (tests/golden/heavy_outputs.json; D = 2 lives in bench_n200.json)."""
    out = {}
    try:
        with open(os.path.join(ROOT, "tests", "golden", "heavy_outputs.json")) as f:
            for r in json.load(f)["runs"]:
                if r["tag"].startswith(f"bench.sigma{sigma}.D"):
                    out[float(r["dmax"])] = r
    except Exception:
        pass
    return out


def witness_is_valid(x, y, sol, amax, dmax):
    """Size-independent check of one answer: the hull is a simple convex cycle, encloses exactly `count` points (boundary
    included), its area is the reported one and within A, and no two of its vertices are farther apart than D."""
    hull = sol["hull"]
    h = len(hull)
    if not sol.get("found", h > 0):
        return sol["count"] == 0
    if h < 3 or len(set(hull)) != h:
        return False
    cr = lambda a, b, px, py: (x[b] - x[a]) * (py - y[a]) - (px - x[a]) * (y[b] - y[a])
    edges = list(zip(hull, hull[1:] + hull[:1]))
    area2 = sum(x[a] * y[b] - x[b] * y[a] for a, b in edges)
    sign = 1.0 if area2 > 0 else -1.0
    if abs(abs(area2) / 2 - sol["area"]) > 1e-7 * max(1.0, sol["area"]) or sol["area"] > amax * (1 + 1e-12):
        return False
    if any(sign * cr(a, b, x[c], y[c]) < -2e-10 for (a, b), c in zip(edges, hull[2:] + hull[:2])):
        return False
    inside = sum(all(sign * cr(a, b, x[p], y[p]) >= -2e-10 for a, b in edges) for p in range(len(x)))
    far = max(math.dist((x[a], y[a]), (x[b], y[b])) for a in hull for b in hull)
    return inside == sol["count"] and far <= dmax * (1 + 1e-12)


def oracle_answers(sigma):
    """Answers of the binary64 ORACLE (not the reference) for solves no frozen reference output exists for, computed with the candidate
    pairs sharded over processes (tests/golden/make_oracle_sharded.py): {dmax: record}.  Only count, area and diameter pair are held."""
    path = os.path.join(ROOT, "tests", "golden", "oracle_sharded.json")
    if not os.path.exists(path):
        return {}
    with open(path) as f:
        return {float(r["dmax"]): r for r in json.load(f)["runs"] if float(r["sigma"]) == float(sigma)}


def load_workload(sigma):
    with open(os.path.join(ROOT, "tests", "golden", "bench_n200.json")) as f:
        fx = json.load(f)
    key = str(float(sigma))
    if key not in fx["sets"]:
        raise SystemExit(f"--sigma {sigma}: not in tests/golden/bench_n200.json (have {sorted(fx['sets'])})")
    st = fx["sets"][key]
    x = [float(p[0]) for p in st["points"]]
    y = [float(p[1]) for p in st["points"]]
    return {"x": x, "y": y, "n": len(x), "amax": fx["amax"], "diameters": fx["diameters"], "sample": st["sample"],
            "sigma": float(sigma), "seed": fx["seed"]}


def workload_config(w, extra=None):
    c = {"workload": f"area+diameter-bounded DP (AntipodalOptimized), synthetic Gaussian n={w['n']} sigma={w['sigma']} seed={w['seed']} "
                     f"in [0,20]^2, A={w['amax']}, K=inf, reconstruct, diameter sweep D={w['diameters']} (BASELINE config 4)",
         "n": w["n"], "sigma": w["sigma"], "amax": w["amax"], "diameters": w["diameters"],
         "step": "one full diameter sweep (5 solves)"}
    if extra:
        c.update(extra)
    return c


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# ---------------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region (B200_PROFILING.md's clocks line).  NVML is queried from a
    thread of this process every 500 ms (nvidia_ml_py); `nvidia-smi -lms` as a fallback.  A light sampler matters here: the
    polling of a separate nvidia-smi process at 5 Hz slowed the timed kernels by several per cent."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index, period_s=0.5):
        self.gpu = gpu_index
        self.period = period_s
        self.proc = None
        self.path = None
        self.thread = None
        self.stop_flag = None
        self.samples = []
        self.reasons = set()
        self.max_mhz = None

    def _nvml_loop(self, nv, h):
        bits = {"hw_slowdown": nv.nvmlClocksEventReasonHwSlowdown if hasattr(nv, "nvmlClocksEventReasonHwSlowdown") else 0x8,
                "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
                "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
                "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4)}
        get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or getattr(nv, "nvmlDeviceGetCurrentClocksThrottleReasons")
        while not self.stop_flag.is_set():
            try:
                self.samples.append(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
                r = int(get_reasons(h))
                for name, bit in bits.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self.stop_flag.wait(self.period)

    def start(self):
        if os.environ.get("ADK_BENCH_NO_CLOCKS"):
            return
        try:
            import threading
            import pynvml as nv
            nv.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = int(vis.split(",")[self.gpu]) if vis and vis.split(",")[0].isdigit() else self.gpu
            h = nv.nvmlDeviceGetHandleByIndex(idx)
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            self.stop_flag = threading.Event()
            self.thread = threading.Thread(target=self._nvml_loop, args=(nv, h), daemon=True)
            self.thread.start()
            return
        except Exception:
            self.thread = None
        try:
            fd, self.path = tempfile.mkstemp(prefix="adk_clocks_", suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", str(int(1000 * self.period))], stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.thread is not None:
            self.stop_flag.set()
            self.thread.join(timeout=5)
            sm = sorted(self.samples)
            if sm:
                out["sm_mhz"] = sm[len(sm) // 2]
                out["samples"] = len(sm)
            out["sm_max_mhz"] = self.max_mhz
            out["reasons"] = sorted(self.reasons)
            out["how"] = f"NVML from a thread of the bench process, every {self.period} s during the timed region"
            return out
        if self.proc is None:
            return out
        try:
            self.proc.terminate()
            self.proc.wait(timeout=5)
        except Exception:
            pass
        try:
            rows = [r.strip().split(", ") for r in open(self.path).read().strip().splitlines() if r.strip()]
            os.unlink(self.path)
        except Exception:
            return out
        sm, reasons = [], set()
        for r in rows:
            if len(r) < 9:
                continue
            try:
                sm.append(float(r[1]))
                out["sm_max_mhz"] = float(r[2])
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.strip().lower() == "active":
                    reasons.add(name)
        if sm:
            sm.sort()
            out["sm_mhz"] = sm[len(sm) // 2]
            out["samples"] = len(sm)
        out["reasons"] = sorted(reasons)
        out["how"] = f"nvidia-smi -lms {int(1000 * self.period)} during the timed region"
        return out


# ---------------------------------------------------------------------------------------------------- reference arm
def _ref_worker(args):
    """One replica: the unmodified reference solving the sample instance (runs in its own process = its own core)."""
    x, y, amax, dmax = args
    from oracle.refpy import RefLib
    lib = RefLib("x87")
    t0 = time.perf_counter()
    r = lib.solve(RefLib.ALGO_ANTIPODAL_OPTIMIZED, x, y, None, amax, dmax, True)
    return time.perf_counter() - t0, r["count"], r["entries_required"]


def _port_worker(args):
    x, y, amax, dmax = args
    from oracle.oraclepy import Oracle
    o = Oracle()
    t0 = time.perf_counter()
    r = o.solve_area_diameter(x, y, None, amax, dmax, True)
    return time.perf_counter() - t0, r["count"], r["states"]


def cpu_reference_run(w, steps, warmup, cores=None, nsolves=1):
    """Times the reference's own CPU implementation on a bounded sample: the first `nsolves` solves of the sweep that the
    fixture holds reference counts for (D = 2 by default), one replica per core on a FIXED number of cores (instance-level
    parallelism is the only kind the single-threaded reference admits).  Returns the cpu_baseline dict."""
    from multiprocessing import get_context
    from oracle.refpy import ref_available
    kind = "reference" if ref_available("x87") else "port"
    worker = _ref_worker if kind == "reference" else _port_worker
    samples = w["sample"][:max(1, nsolves)]
    ncores = cores or max(1, min(os.cpu_count() or 1, REF_CORES))
    ctx = get_context("spawn")
    times, single = [], None
    with ctx.Pool(ncores) as pool:
        for it in range(warmup + steps):
            t0 = time.perf_counter()
            rep = 0.0
            for smp in samples:
                res = pool.map(worker, [(w["x"], w["y"], w["amax"], smp["dmax"])] * ncores)
                for _, cnt, _ in res:
                    assert cnt == smp["count"], (cnt, smp["count"])
                rep += min(r[0] for r in res)
            dt = time.perf_counter() - t0
            if it >= warmup:
                times.append(dt)
                single = rep if single is None else min(single, rep)
    total = sum(times)
    cells = sum(smp["cells"] for smp in samples)
    states = sum(smp["states"] for smp in samples)
    what = " + ".join(f"D={smp['dmax']}" for smp in samples)
    # The whole sweep on this host, ESTIMATED: the unmodified reference's single-core times of all five solves are frozen from the
    # build container (tests/golden/bench_n200.json, heavy_outputs.json); the D = 2 solve timed live here gives the speed factor
    # between the two hosts.  A record beside the measured sample, not a measurement: the reference's cost is per hash entry and
    # the later solves have more cell updates per entry, so cells/s taken on the sample alone understates its whole-sweep rate.
    whole = None
    try:
        frozen = frozen_answers(w["sigma"])
        first = w["sample"][0]
        rest = [d for d in w["diameters"] if float(d) != float(first["dmax"])]
        if kind == "reference" and single and first.get("ref_solve_ms_build_container") and all(float(d) in frozen for d in rest):
            frozen_s = first["ref_solve_ms_build_container"] / 1e3 + sum(frozen[float(d)]["solve_ms"] / 1e3 for d in rest)
            factor = single / (first["ref_solve_ms_build_container"] / 1e3) if len(samples) == 1 else None
            if factor:
                whole = {"core_seconds_frozen_build_container": frozen_s, "host_speed_factor_from_live_sample": factor,
                         "core_seconds_this_host_estimated": frozen_s * factor, "cores": ncores,
                         "sweeps_per_sec_estimated": ncores / (frozen_s * factor),
                         "note": "estimate, not a measurement: frozen single-core times of the five solves x (live D=2 time / frozen D=2 time)"}
    except Exception:
        whole = None
    return {"value": cells * ncores * len(times) / total, "unit": UNIT, "cores": ncores, "kind": kind,
            "sample": f"{what} solve(s) of the sweep ({cells} cells, {states} states = {cells / states:.1f} cells per state; the whole sweep has "
                      f"27 cells per state), one replica per core on {ncores} cores, {len(times)} timed passes, {total / len(times):.2f} s per pass",
            "states_per_sec": states * ncores * len(times) / total, "cells_per_sec_per_core": cells * len(times) / total,
            "states_per_sec_per_core": states * len(times) / total,
            "ms_per_step": 1e3 * total / len(times), "single_replica_s": single, "whole_sweep_estimate": whole}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    w = load_workload(args.sigma)
    cb = cpu_reference_run(w, args.steps, args.warmup, nsolves=args.ref_solves)
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": cb["ms_per_step"], "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f80 (x87 long double)" if cb["kind"] == "reference" else "f64", "data": "synthetic",
            "config": workload_config(w, {"reference_step": "bounded sample: " + cb["sample"]}),
            "states_per_sec": cb["states_per_sec"], "single_replica_s": cb["single_replica_s"],
            "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample", "states_per_sec", "cells_per_sec_per_core",
                                                "states_per_sec_per_core", "single_replica_s", "whole_sweep_estimate")},
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------- B200 arm
def run_b200_arm(args):
    import torch
    import torch.distributed as dist
    import __graft_entry__ as ge

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback; use --impl reference for the host baseline)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    pkg = ge.load_package()
    import importlib
    adist = importlib.import_module("area_diameter_algorithm_b200.dist")

    w = load_workload(args.sigma)
    n, amax, diams = w["n"], w["amax"], w["diameters"]
    solver = pkg.Solver(local)
    if args.memory_budget_gb > 0:
        solver.set_memory_budget(int(args.memory_budget_gb * 2**30))
    stream = torch.cuda.Stream(device=dev)
    solver.set_stream(stream.cuda_stream)
    solver.set_reuse(False)      # every solve recomputes the count tables and orders, as the reference does on every call

    d_x = torch.tensor(w["x"], dtype=torch.float64, device=dev)
    d_y = torch.tensor(w["y"], dtype=torch.float64, device=dev)
    h_x = torch.tensor(w["x"], dtype=torch.float64).pin_memory()
    h_y = torch.tensor(w["y"], dtype=torch.float64).pin_memory()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()

    def step(resident):
        """One diameter sweep.  Returns (cells, states, dp_ms, launches, results) of THIS rank."""
        cells = states = 0
        dp_ms = 0.0
        results = []
        flush.zero_()                                   # L2 flush between steps (on the timed stream)
        for d in diams:
            if resident:
                part, st = solver.area_diameter_partial_dev(d_x.data_ptr(), d_y.data_ptr(), n, None, amax, d, rank, world)
            elif world == 1:
                r = solver.solve_area_diameter_host(h_x.data_ptr(), h_y.data_ptr(), n, None, amax, d, True)
                cells += r["cells_executed"]; states += r["states"]; dp_ms += r["dp_ms"]
                results.append(r)
                continue
            else:
                part, st = solver.area_diameter_partial_host(h_x.data_ptr(), h_y.data_ptr(), n, None, amax, d, rank, world)
            cells += st["cells_executed"]; states += st["states"]; dp_ms += st["dp_ms"]
            parts = adist.gather_partials(part, device=dev)
            best = adist.merge_partials(parts)
            owner = adist.owner_of(best, parts)          # its records are still resident: no re-run of the winning pair
            r = None
            if rank == owner:
                if resident:
                    r = solver.area_diameter_finish_dev(d_x.data_ptr(), d_y.data_ptr(), n, None, amax, d, best, True)
                else:
                    r = solver.area_diameter_finish_host(h_x.data_ptr(), h_y.data_ptr(), n, None, amax, d, best, True)
            results.append(adist.share_result(r, owner, n, device=dev))
        return cells, states, dp_ms, results

    def timed(resident, steps, warmup):
        with torch.cuda.stream(stream):
            for _ in range(warmup):
                step(resident)
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            barrier(); torch.cuda.synchronize()
            l0 = solver.launch_count()
            t0 = time.perf_counter()
            ev0.record(stream)
            cells = states = 0
            dp_ms = 0.0
            last = None
            for _ in range(steps):
                c, s, d, last = step(resident)
                cells += c; states += s; dp_ms += d
            ev1.record(stream)
            torch.cuda.synchronize(); barrier()
            wall_ms = 1e3 * (time.perf_counter() - t0)
            dev_ms = ev0.elapsed_time(ev1)
            launches = solver.launch_count() - l0 + steps     # + the flush memset per step
        agg = torch.tensor([cells, states, launches], dtype=torch.float64, device=dev)
        mx = torch.tensor([dev_ms, wall_ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(agg, op=dist.ReduceOp.SUM)
            dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        return {"cells": int(agg[0].item()), "states": int(agg[1].item()), "launches": int(agg[2].item()),
                "dev_ms": float(mx[0].item()), "wall_ms": float(mx[1].item()), "my_cells": cells, "my_dp_ms": dp_ms, "last": last}

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    res = timed(True, args.steps, args.warmup)
    clocks = sampler.stop() if rank == 0 else None
    e2e = timed(False, args.steps, max(1, args.warmup // 3))

    # Several GPUs behind the reference-facing call: ONE process (rank 0 here) hands host buffers to adk_multi_solve_area_diameter,
    # the library shards the candidate pairs over the N devices with one host thread each and exchanges the local bests with NCCL
    # (all-gather + witness broadcast) inside the call.  This is what the reference's C++ callers get through libadk_host.so
    # with ADK_DEVICES set.  The other ranks free their devices' workspaces and wait.
    multi = None
    if world > 1:
        solver.close()
        del flush
        torch.cuda.empty_cache()
        torch.cuda.synchronize()
        # the waiting ranks must not spin in a kernel on the devices rank 0 is about to use (an NCCL barrier does): wait on the host
        host_group = dist.new_group(backend="gloo")
        dist.barrier(group=host_group)
        if rank == 0:
            m = pkg.MultiSolver(list(range(world)))
            m.set_reuse(False)
            fl = [torch.empty(256 << 20, dtype=torch.uint8, device=torch.device("cuda", d)) for d in range(world)]

            def msweep():
                for f_ in fl:
                    f_.zero_()
                for d_ in range(world):
                    torch.cuda.synchronize(d_)
                t0_ = time.perf_counter()
                cells_ = 0; states_ = 0; exch_ = 0.0; dp_ = 0.0; outs_ = []
                for d in diams:
                    r_ = m.solve_area_diameter_host(h_x.data_ptr(), h_y.data_ptr(), n, None, amax, d, True)
                    st_ = m.last_stats()
                    cells_ += r_["cells_executed"]; states_ += r_["states"]; exch_ += st_["max_exchange_ms"]; dp_ += st_["max_dp_ms"]
                    outs_.append(r_)
                return time.perf_counter() - t0_, cells_, states_, exch_, dp_, outs_
            for _ in range(2):
                msweep()
            tt = cc = ss = ee = dd_ = 0.0
            for _ in range(args.steps):
                t_, c_, s_, e_, d_, mouts = msweep()
                tt += t_; cc += c_; ss += s_; ee += e_; dd_ += d_
            multi = {"value": cc / tt, "unit": UNIT, "ms_per_step": 1e3 * tt / args.steps, "states_per_sec": ss / tt,
                     "exchange_ms_per_step": ee / args.steps, "slowest_rank_dp_ms_per_step": dd_ / args.steps,
                     "api": "adk_multi_solve_area_diameter (host buffers; shards, NCCL all-gather and witness broadcast inside libadk.so)",
                     "answers": [{"D": d, "count": r_["count"], "area": r_["area"], "diam": r_["diam"], "hull": r_["hull"]} for d, r_ in zip(diams, mouts)]}
            m.close()
            del fl
        dist.barrier(group=host_group)

    if rank == 0:
        secs = res["dev_ms"] / 1e3
        value = res["cells"] / secs
        peak, peak_src = measured_peaks()
        # dominant kernel: k_ad_cells, timed by the library's own CUDA events on the launching stream, live in this run
        dp_s = res["my_dp_ms"] / 1e3
        launches_dp = args.steps * len(diams)
        achieved = res["my_cells"] * BYTES_PER_CELL / dp_s / 1e9 if dp_s > 0 else 0.0
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "roofline_traffic.json")
        if os.path.exists(tpath):
            try:
                traffic = json.load(open(tpath)).get("k_ad_cells_dram_bytes_per_launch")
            except Exception:
                traffic = None
        fp64_tflops = res["my_cells"] * FLOPS_PER_CELL / dp_s / 1e12 if dp_s > 0 else 0.0
        fp64_pk, fp64_src = fp64_peak()
        int_tops = res["my_cells"] * INT_OPS_PER_CELL / dp_s / 1e12 if dp_s > 0 else 0.0
        int_pk, int_src = int32_peak()
        # The roofline the north_star names for this config is FP64 / INT issue: the 20 B/cell read-modify-write of the
        # reference happens in registers and shared memory here, so HBM is not the bound (measured DRAM traffic is a small
        # fraction of the algorithmic bytes); the HBM figure is kept as `hbm_nominal`.
        roof = {"bound": "fp64_issue", "kernel": "k_ad_cells", "achieved": fp64_tflops, "peak": fp64_pk, "unit": "TFLOP/s",
                "frac": fp64_tflops / fp64_pk, "peak_source": fp64_src,
                "traffic": traffic, "traffic_note": "dram__bytes_read.sum + dram__bytes_write.sum per k_ad_cells launch (ncu, profiles/roofline_traffic.json)",
                "algorithmic_flops_per_launch": res["my_cells"] * FLOPS_PER_CELL / max(launches_dp, 1),
                "algorithmic_bytes_per_launch": res["my_cells"] * BYTES_PER_CELL / max(launches_dp, 1),
                "avg_launch_ms": res["my_dp_ms"] / max(launches_dp, 1), "share_of_step": res["my_dp_ms"] / res["dev_ms"],
                "int32_issue": {"achieved": int_tops, "peak": int_pk, "unit": "Tinst/s", "frac": int_tops / int_pk, "peak_source": int_src},
                "hbm_nominal": {"achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "peak_source": peak_src,
                                "note": "20 B/cell (SURVEY 8d) x cells / kernel time; nominal: the bytes never reach HBM"}}
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": res["dev_ms"] / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic",
                "config": workload_config(w, {"parallelism": f"candidate pairs sharded over {world} rank(s)",
                                              "l2": "256 MiB memset on the timed stream before every step (L2 flush)",
                                              "states_per_step": res["states"] // args.steps, "cells_per_step": res["cells"] // args.steps}),
                "solve_ms_per_instance": res["dev_ms"] / args.steps / len(diams),
                "states_per_sec": res["states"] / secs,
                "e2e": ({"value": e2e["cells"] / (e2e["wall_ms"] / 1e3), "unit": UNIT, "ms_per_step": e2e["wall_ms"] / args.steps,
                         "h2d_bytes_per_step": len(diams) * 2 * n * 8,
                         "d2h_bytes_per_step": len(diams) * (64 + 4 * (n + 4) + 112), "timer": "host wall clock around the C-ABI calls",
                         "api": "adk_solve_area_diameter (host buffers)"} if multi is None else
                        {"value": multi["value"], "unit": UNIT, "ms_per_step": multi["ms_per_step"],
                         "h2d_bytes_per_step": len(diams) * world * 2 * n * 8,
                         "d2h_bytes_per_step": len(diams) * world * (64 + 32 * world + 4 * (n + 16)), "timer": "host wall clock around the C-ABI calls",
                         "api": multi["api"], "exchange_ms_per_step": multi["exchange_ms_per_step"],
                         "slowest_rank_dp_ms_per_step": multi["slowest_rank_dp_ms_per_step"]}),
                "e2e_one_process_per_gpu": None if multi is None else
                        {"value": e2e["cells"] / (e2e["wall_ms"] / 1e3), "unit": UNIT, "ms_per_step": e2e["wall_ms"] / args.steps,
                         "api": "adk_area_diameter_partial / _finish per rank + torch.distributed all-gather and broadcast"},
                "gpu_launches": res["launches"], "roofline": roof, "clocks": clocks,
                "answers": [{"D": d, "count": r["count"], "area": r["area"], "diam": r["diam"]} for d, r in zip(diams, res["last"])]}
        # every solve of the sweep must reproduce the unmodified reference's frozen answer (tests/golden/bench_n200.json for the
        # sample, tests/golden/heavy_outputs.json for the rest)
        want = {float(r_["dmax"]): r_ for r_ in w["sample"]}
        want.update(frozen_answers(w["sigma"]))
        checked = []
        for d, got in zip(diams, res["last"]):
            ref = want.get(float(d))
            if ref is None:
                continue
            assert got["count"] == ref["count"] and got["hull"] == ref["hull"] and sorted(got["diam"]) == sorted(ref["diam"]), (d, got, ref)
            assert abs(got["area"] - float(ref["area"])) <= 1e-9 * float(ref["area"]), (d, got["area"], ref["area"])
            checked.append(d)
        if multi is not None:
            for d, got in zip(diams, multi["answers"]):
                ref = want.get(float(d))
                if ref is not None:
                    assert got["count"] == ref["count"] and got["hull"] == ref["hull"] and sorted(got["diam"]) == sorted(ref["diam"]), ("multi", d, got, ref)
        line["answers_checked_against_frozen_reference"] = checked
        # where the reference is out of a CPU's reach, the binary64 oracle (pinned to the reference elsewhere) is the second opinion
        by_oracle = []
        for d, got in zip(diams, res["last"]):
            o = oracle_answers(w["sigma"]).get(float(d))
            if o is None or d in checked:
                continue
            assert got["count"] == o["count"] and sorted(got["diam"]) == sorted(o["diam"]), ("oracle", d, got, o)
            assert abs(got["area"] - float(o["area"])) <= 1e-9 * float(o["area"]), ("oracle", d, got["area"], o["area"])
            by_oracle.append(d)
        line["answers_checked_against_oracle"] = by_oracle
        # the diameters with no frozen answer (sigma != 1.5 beyond the D = 2 sample): the witness itself is checked — simple convex
        # cycle, enclosed-point count = the reported count, area <= A, farthest pair <= D — and the counts must not fall as D grows
        line["witness_valid"] = [d for d, got in zip(diams, res["last"]) if witness_is_valid(w["x"], w["y"], got, w["amax"], d)]
        assert len(line["witness_valid"]) == len(diams), ("invalid witness", line["witness_valid"])
        counts = [r["count"] for r in res["last"]]
        assert all(a <= b for a, b in zip(counts, counts[1:])) or diams != sorted(diams), counts
        if world == 1 and not args.no_cpu:
            cb = cpu_reference_run(w, steps=2, warmup=0)
            line["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample", "states_per_sec", "cells_per_sec_per_core",
                                                       "states_per_sec_per_core", "single_replica_s", "whole_sweep_estimate")}
            if line["cpu_baseline"]["whole_sweep_estimate"]:
                wse = dict(line["cpu_baseline"]["whole_sweep_estimate"])
                wse["cells_per_sec_estimated"] = (res["cells"] // args.steps) * wse["sweeps_per_sec_estimated"]
                line["cpu_baseline"]["whole_sweep_estimate"] = wse
        print(json.dumps(line), flush=True)
    if world == 1:
        solver.close()
    if world > 1:
        dist.destroy_process_group()


# ---------------------------------------------------------------------------------------------------- other configs
def _aux_ref_worker(args):
    algo, x, y, kmax, amax, dmax = args
    from oracle.refpy import RefLib
    lib = RefLib("x87")
    t0 = time.perf_counter()
    r = lib.solve(algo, x, y, kmax, amax, dmax, True)
    return time.perf_counter() - t0, r["count"]


def run_aux_workload(args):
    """BASELINE configs 3 and 5 as secondary bench lines (same JSON shape; the driver's headline stays config 4).
      area_n1000  : area-bounded DP (Eppstein path), synthetic clustered n = 1000 (sigma 2, seed 1000), A = 4, K = inf; anchors
                    sharded over the ranks.  value = NOMINAL cells/s, (K-2) n (n-1) (n-2) / 3, executed cells beside it.
      cells_batch : the ten cell samples (n = 362..699), A = 4, D = 4, K = inf, reconstruct; instances sharded over the ranks
                    (adk_solve_batch).  value = cells/s over the whole batch."""
    import numpy as np
    import torch
    import torch.distributed as dist
    import importlib
    import __graft_entry__ as ge
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    pkg = ge.load_package()
    adist = importlib.import_module("area_diameter_algorithm_b200.dist")
    capi = importlib.import_module("area_diameter_algorithm_b200.capi")
    solver = pkg.Solver(local)
    if args.memory_budget_gb > 0:
        solver.set_memory_budget(int(args.memory_budget_gb * 2**30))
    stream = torch.cuda.Stream(device=dev)
    solver.set_stream(stream.cuda_stream)
    solver.set_reuse(False)      # every solve recomputes the count tables and orders, as the reference does on every call
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    amax = 4.0
    if args.workload == "area_n1000":
        synth = importlib.import_module("area_diameter_algorithm_b200.synth")
        pts = synth.gaussian_points(1000, 2.0, 1000)
        n = len(pts)
        x = [p[0] for p in pts]; y = [p[1] for p in pts]
        d_x = torch.tensor(x, dtype=torch.float64, device=dev); d_y = torch.tensor(y, dtype=torch.float64, device=dev)
        h_x = torch.tensor(x, dtype=torch.float64).pin_memory(); h_y = torch.tensor(y, dtype=torch.float64).pin_memory()
        nominal = (n - 2) * n * (n - 1) * (n - 2) // 3
        desc = f"area-bounded DP (Eppstein), synthetic Gaussian n={n} sigma=2.0 seed=1000 in [0,20]^2, A={amax}, K=inf, reconstruct (BASELINE config 3)"
        kernel, bytes_per_cell = "k_area_rows", 16
        try:
            area_traffic = json.load(open(os.path.join(ROOT, "profiles", "roofline_traffic.json"))).get("k_area_rows_dram_bytes_per_launch")
        except Exception:
            area_traffic = None

        def step(resident):
            flush.zero_()
            if resident or world > 1:
                part, st = (solver.area_partial_dev(d_x.data_ptr(), d_y.data_ptr(), n, None, amax, rank, world) if resident else
                            solver.area_partial(x, y, None, amax, rank, world))
                parts = adist.gather_partials(part, device=dev)
                best = adist.merge_partials(parts)
                # the rank whose anchor won finishes: its cluster still holds the polygon it walked back during the sweep (no re-run)
                owner = adist.owner_of(best, parts)
                r = solver.area_finish_dev(d_x.data_ptr(), d_y.data_ptr(), n, None, amax, best, True) if rank == owner else None
                r = adist.share_result(r, owner, n, device=dev)
                return st["cells_executed"], st["dp_ms"], r
            r = solver.solve_area_host(h_x.data_ptr(), h_y.data_ptr(), n, None, amax, True)
            return r["cells_executed"], r["dp_ms"], r
        h2d, d2h = 2 * n * 8, 64 + 4 * (n + 4) + 112
    else:
        with open(os.path.join(ROOT, "tests", "golden", "points.json")) as f:
            allp = json.load(f)
        names = [f"real/{i}/0" for i in range(10)]
        sets = [np.array(allp[k], dtype=np.float64) for k in names]
        dmax = 4.0
        desc = (f"area+diameter-bounded DP on the ten cell samples (n={[len(s_) for s_ in sets]}), A={amax}, D={dmax}, K=inf, reconstruct, "
                "one adk_solve_batch call (BASELINE config 5)")
        kernel, bytes_per_cell, nominal = "k_ad_cells", BYTES_PER_CELL, None
        area_traffic = None

        def step(resident):
            flush.zero_()
            if world > 1 and not args.instance_sharding:
                # one slide dominates the batch: shard every instance's candidate pairs instead of whole instances
                allr, cells, dp = adist.solve_batch_pair_sharded(solver, sets, None, amax, dmax, True, device=dev)
                return cells, dp, allr
            mine = solver.solve_batch(sets, capi.ALGO_AREA_DIAMETER, None, amax, dmax, True, shard=rank, nshards=world)
            cells = sum(r["cells_executed"] for r in mine if r is not None)
            dp = sum(r["dp_ms"] for r in mine if r is not None)
            return cells, dp, adist.gather_batch_results(mine)
        h2d = sum(len(s_) * 16 for s_ in sets); d2h = sum(64 + 4 * (len(s_) + 4) + 112 for s_ in sets)

    def barrier():
        if world > 1:
            dist.barrier()

    def timed(resident, steps, warmup):
        with torch.cuda.stream(stream):
            for _ in range(warmup):
                step(resident)
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            barrier(); torch.cuda.synchronize()
            l0 = solver.launch_count(); t0 = time.perf_counter(); ev0.record(stream)
            cells = 0; dp = 0.0; last = None
            for _ in range(steps):
                c, d, last = step(resident)
                cells += c; dp += d
            ev1.record(stream); torch.cuda.synchronize(); barrier()
            wall = 1e3 * (time.perf_counter() - t0); devms = ev0.elapsed_time(ev1)
            launches = solver.launch_count() - l0 + steps
        agg = torch.tensor([cells, launches], dtype=torch.float64, device=dev)
        mx = torch.tensor([devms, wall], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(agg, op=dist.ReduceOp.SUM); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        return {"cells": int(agg[0].item()), "launches": int(agg[1].item()), "dev_ms": float(mx[0].item()), "wall_ms": float(mx[1].item()),
                "my_cells": cells, "my_dp_ms": dp, "last": last}

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    res = timed(True, args.steps, args.warmup)
    clocks = sampler.stop() if rank == 0 else None
    e2e = timed(False, args.steps, 1)
    if rank == 0:
        secs = res["dev_ms"] / 1e3
        units = res["cells"]          # EXECUTED cell updates (the area kernels skip the levels that cannot be finite; the nominal count is beside it)
        peak, peak_src = measured_peaks()
        dp_s = res["my_dp_ms"] / 1e3
        achieved = res["my_cells"] * bytes_per_cell / dp_s / 1e9 if dp_s > 0 else 0.0     # EXECUTED cells of this rank (levels that cannot be finite are skipped)
        line = {"metric": METRIC, "value": units / secs, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": res["dev_ms"] / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic" if nominal else "the reference's cell samples (tests/golden/points.json)",
                "config": {"workload": desc, "l2": "256 MiB memset on the timed stream before every step (L2 flush)",
                           "cells": "executed cell updates, counted in-kernel",
                           "cells_executed_per_step": res["cells"] // args.steps},
                "executed_cells_per_sec": res["cells"] / secs,
                "e2e": {"value": e2e["cells"] / (e2e["wall_ms"] / 1e3), "unit": UNIT,
                        "ms_per_step": e2e["wall_ms"] / args.steps, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "timer": "host wall clock around the C-ABI calls, max over ranks"},
                "gpu_launches": res["launches"],
                "roofline": {"bound": "hbm", "kernel": kernel, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "traffic": area_traffic if world == 1 else None, "peak_source": peak_src,
                             "algorithmic_bytes_per_launch": res["my_cells"] * bytes_per_cell / args.steps,
                             "avg_launch_ms": res["my_dp_ms"] / args.steps, "share_of_step": res["my_dp_ms"] / res["dev_ms"]},
                "clocks": clocks}
        last = res["last"]
        if nominal:
            line["answers"] = [{"count": last["count"], "area": last["area"], "hull": last["hull"]}]
            line["solve_ms_per_instance"] = res["dev_ms"] / args.steps
            line["nominal_cells_per_sec"] = nominal * args.steps / secs
            line["config"]["nominal_cells_per_step"] = nominal
            # the binary64 oracle's answer for this very solve (anchors sharded over processes in the build container,
            # tests/golden/make_oracle_sharded.py --area): count and hull cycle exact, area to 1e-9
            opath = os.path.join(ROOT, "tests", "golden", "oracle_sharded_area.json")
            if os.path.exists(opath):
                with open(opath) as f:
                    o = next((r_ for r_ in json.load(f)["runs"] if r_["tag"].startswith("area.n1000.K")), None)
                if o is not None:
                    hg, ho = list(last["hull"]), list(o["hull"])
                    rot = hg.index(ho[0]) if ho and ho[0] in hg else -1
                    assert last["count"] == o["count"] and rot >= 0 and hg[rot:] + hg[:rot] == ho, ("oracle", last["count"], hg, o["count"], ho)
                    assert abs(last["area"] - float(o["area"])) <= 1e-9 * float(o["area"]), ("oracle", last["area"], o["area"])
                    line["answer_checked_against_oracle"] = o["tag"]
            line["config"]["nominal_note"] = ("(K-2) n (n-1) (n-2) / 3 with K = n: what the reference's loops would execute (Eppstein.cpp:192-235); the kernels "
                                              "stop at the deepest reachable level, so `value` counts the executed updates only")
        else:
            line["answers"] = [{"n": len(s_), "count": r["count"], "area": r["area"], "diam": r["diam"], "ms": r["kernel_ms"]} for s_, r in zip(sets, last)]
            line["config"]["parallelism"] = ("one GPU" if world == 1 else "whole instances per rank (adk_solve_batch)" if args.instance_sharding
                                             else "candidate pairs of every instance sharded over the ranks")
            line["instances_per_sec"] = len(sets) * args.steps / secs
        if world == 1 and not args.no_cpu:
            from multiprocessing import get_context
            from oracle.refpy import RefLib, ref_available
            if ref_available("x87"):
                ncores = max(1, min(os.cpu_count() or 1, REF_CORES))
                if nominal:
                    sub = 100        # same K policy as the GPU arm (K = inf); the reference executes every nominal cell
                    job = (RefLib.ALGO_EPPSTEIN, x[:sub], y[:sub], None, amax, float("inf"))
                    cpu_cells = (sub - 2) * sub * (sub - 1) * (sub - 2) // 3
                    what = (f"first {sub} points of the same set, K=inf as on the GPU ({cpu_cells} cells, all executed by the reference; "
                            "the un-capped n=1000 table is 32 GB for the reference)")
                else:
                    k9 = names.index("real/9/0")
                    job = (RefLib.ALGO_ANTIPODAL_OPTIMIZED, sets[k9][:, 0].tolist(), sets[k9][:, 1].tolist(), None, amax, dmax)
                    cpu_cells = last[k9]["cells_executed"]
                    what = f"the lightest slide real/9 (n={len(sets[k9])}, {cpu_cells} cells)"
                with get_context("spawn").Pool(ncores) as pool:
                    t0 = time.perf_counter(); out = pool.map(_aux_ref_worker, [job] * ncores); dt = time.perf_counter() - t0
                line["cpu_baseline"] = {"value": cpu_cells * ncores / dt, "unit": UNIT, "cores": ncores, "kind": "reference",
                                        "sample": f"{what}, one replica per core, {dt:.2f} s"}
        print(json.dumps(line), flush=True)
    solver.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--sigma", type=float, default=1.5, help="cluster width of the synthetic n=200 set (fixture: 3.0, 2.0, 1.5, 1.0, 0.5)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--memory-budget-gb", type=float, default=0.0, help="development: cap the DP workspace (adk_set_memory_budget); 0 = the library's default")
    ap.add_argument("--ref-solves", type=int, default=1, help="--impl reference: solves of the sweep per step (1 = D=2, 2 = D=2 and D=3; "
                    "D=3 alone is ~10 core-minutes per replica)")
    ap.add_argument("--instance-sharding", action="store_true", help="cells_batch on N > 1 GPUs: whole instances per rank (adk_solve_batch) "
                    "instead of every instance's candidate pairs")
    ap.add_argument("--workload", default="ad_n200", choices=["ad_n200", "area_n1000", "cells_batch"],
                    help="ad_n200 = the headline (BASELINE config 4); area_n1000 = config 3; cells_batch = config 5")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    elif args.workload != "ad_n200":
        run_aux_workload(args)
    else:
        run_b200_arm(args)


if __name__ == "__main__":
    main()
