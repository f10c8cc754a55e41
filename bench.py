#!/usr/bin/env python
"""bench.py -- NDsort + crowding dominance-pairs/s (BASELINE.json metric).

One "step" = one NSGA-II environmental selection over one synthetic population:
non_dominated_sort(fitness) -> worst_rank mask -> crowding_distance(fitness, mask) -> survivor
order, i.e. exactly reference nsga2.py:90-96, on N rows x M objectives of uniform-random float32.
value = N^2 ordered pairs per step / device time per step (inputs resident in HBM).
e2e   = the same through the public Python API with HOST buffers (H2D of the fitness matrix and
        D2H of the survivor order inside the timed region).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c4|c2|c5|c3|NxM] [--impl reference]

Launched under torchrun for N > 1 (one rank per GPU, NCCL).  Rank 0 prints ONE JSON line.
`--impl reference` times the CPU restatement of the reference (oracle/, C + OpenMP, all host
cores) on a bounded sample of the same workload; JAX is not installable in this image, so the
reference's own JAX-CPU path cannot be run.
"""

from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np

METRIC = "NDsort+crowding dominance-pairs/s"
UNIT = "pairs/s"
WORKLOADS = {
    # name: (N, M, description)  -- SURVEY.md section 8: config -> (N, M)
    "c4": (262144, 5, "BASELINE configs[3]: N=262144 M=5 uniform-random f32 (the size the metric is quoted on)"),
    "c2": (65536, 3, "BASELINE configs[1]: N=65536 M=3 uniform-random f32"),
    "c5": (65536, 10, "BASELINE configs[4] shape: N=65536 M=10 uniform-random f32"),
    "c3": (32768, 3, "BASELINE configs[2] shape: N=2*16384 M=3 uniform-random f32"),
}


def parse_workload(name: str):
    if name in WORKLOADS:
        return WORKLOADS[name]
    n, m = name.lower().split("x")
    return int(n), int(m), f"custom N={n} M={m} uniform-random f32"


def make_fitness(n: int, m: int, seed: int = 0) -> np.ndarray:
    return np.random.default_rng(seed).random((n, m), dtype=np.float32)


def library_source_hash() -> str:
    """sha256 (16 hex digits) over the CUDA sources of libevoxb.so -- ties a committed ncu capture to the code."""
    import glob
    import hashlib

    h = hashlib.sha256()
    for path in sorted(glob.glob(os.path.join(ROOT, "evox-myc_b200", "csrc", "*.cu*"))):
        h.update(open(path, "rb").read())
    return h.hexdigest()[:16]


def build_uses_levels(n, m):
    """Whether the library builds the dominance matrix of this shape from 15-bit levels (csrc/levels.cuh)."""
    import ctypes

    from evox_myc_b200 import _lib

    path = ctypes.c_int(0)
    _lib.check(_lib.lib().evoxb_nds_build_path(None, n, m, None, ctypes.byref(path)))
    return path.value == 1


def ncu_traffic_per_launch(kernel_substr: str, n: int, m: int, world: int = 1):
    """DRAM bytes (read + write) per launch of a kernel from the committed ncu launch list of the same
    workload (profiles/).  The capture's side file (<csv>.meta.json, written by the capture script) holds
    the source hash it was taken from; a capture of older kernels is refused (traffic null, reason given)."""
    import csv
    import glob

    # multi-GPU: the per-rank kernel is captured on a one-rank communicator (never ncu a multi-rank command);
    # a rank of `world` streams 1 / world of those bytes
    tag = f"N{n}_M{m}" + ("_w1" if world > 1 else "")
    best = None
    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "r*", f"launches_*_{tag}_v*.csv"))):
        best = path  # the latest version of the latest round
    if best is None:
        return None, "no committed ncu capture of this workload"
    meta = best + ".meta.json"
    rel = os.path.relpath(best, ROOT)
    try:
        want = json.load(open(meta))["src_hash"]
    except Exception:
        return None, f"{rel} has no source hash beside it (taken before the kernels changed?)"
    if want != library_source_hash():
        return None, f"{rel} is stale: taken from sources {want}, library is {library_source_hash()}"
    tot, launches = 0.0, 0
    try:
        rows = list(csv.reader(l for l in open(best) if l.startswith('"')))
        h = rows[0]
        ki, vi, mi = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Name")
        for r in rows[1:]:
            if kernel_substr in r[ki]:
                if r[mi] in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                    tot += float(r[vi].replace(",", ""))
                elif r[mi] == "gpu__time_duration.sum":
                    launches += 1
    except Exception as ex:
        return None, f"{rel}: unreadable ({ex!r})"
    if not (launches and tot):
        return None, f"{rel} holds no launch of {kernel_substr}"
    note = f"ncu dram__bytes_read.sum + dram__bytes_write.sum per launch, {rel} (sources {want})"
    if world > 1:
        return tot / launches / world, note + f", one-rank capture divided by {world} ranks"
    return tot / launches, note


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)", float(p.get("sm_max_mhz", 1965.0))
    return 6650.0, "fallback (B200_PROFILING.md)", 1965.0


# --------------------------------------------------------------------------------------------
# CPU arm: the oracle port on the host cores (bounded sample)
# --------------------------------------------------------------------------------------------
def host_cores() -> int:
    """Host threads the CPU arm uses: every core this process may run on.  Set explicitly on each oracle
    call, because torchrun exports OMP_NUM_THREADS=1 to its workers and OpenMP would obey it."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:  # pragma: no cover
        return max(1, os.cpu_count() or 1)


def cpu_sample_rows(n: int, m: int, budget_s: float) -> int:
    """Largest power-of-two sample size <= n whose estimated oracle time fits the budget."""
    from oracle import c_oracle

    probe = min(n, 8192)
    x = make_fitness(probe, m, 123)
    t0 = time.perf_counter()
    c_oracle.nsga2_select(x, probe // 2, nthreads=host_cores())
    dt = max(time.perf_counter() - t0, 1e-4)
    per_pair = dt / (probe * probe)
    ns = probe
    while ns * 2 <= n and per_pair * (ns * 2) ** 2 <= budget_s:
        ns *= 2
    return ns


def cpu_step(x: np.ndarray, want_result: bool = False):
    from oracle import c_oracle

    t0 = time.perf_counter()
    res = c_oracle.nsga2_select(x, x.shape[0] // 2, nthreads=host_cores())
    dt = time.perf_counter() - t0
    return (dt, res) if want_result else dt


def run_reference_arm(args, n, m, desc):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return  # the CPU arm runs on rank 0 alone
    from oracle import c_oracle

    cores = host_cores()
    # the same sample at every --gpus value: the largest power of two whose step stays within a fixed
    # per-step budget on this host (threads pinned above, so the choice does not depend on the launcher)
    ns = cpu_sample_rows(n, m, budget_s=max(2.0, 120.0 / (args.steps + args.warmup)))
    x = make_fitness(ns, m, 0)
    for _ in range(args.warmup):
        cpu_step(x)
    times = [cpu_step(x) for _ in range(args.steps)]
    t = sum(times) / len(times)
    value = ns * ns / t
    sample = f"{ns} of {n} rows (same distribution, M={m}); {ns}^2 ordered pairs per step"
    line = {
        "impl": "reference",
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": desc, "N": n, "M": m, "sample_rows": ns, "cores": cores,
                   "step": "non_dominated_sort + worst-rank mask + crowding_distance + survivor order (nsga2.py:90-96)"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "CPU restatement of the reference algorithm (oracle/nds_oracle.c, OpenMP); the reference's JAX-CPU path is unavailable on this box (no jax)",
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------
# clocks
# --------------------------------------------------------------------------------------------
class ClockSampler:
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
              "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, smax, reasons = [], [], set()
        try:
            for ln in open(self.path):
                f = [v.strip() for v in ln.split(",")]
                if len(f) < 9:
                    continue
                try:
                    sm.append(float(f[1])); smax.append(float(f[2]))
                except ValueError:
                    continue
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            busy = [v for v in sm if v >= 0.5 * max(sm)] or sm
            out.update(sm_mhz=statistics.median(busy), sm_max_mhz=max(smax), reasons=sorted(reasons), samples=len(sm))
        return out


# --------------------------------------------------------------------------------------------
# GPU arm
# --------------------------------------------------------------------------------------------
def run_gpu_arm(args, n, m, desc):
    import torch

    import evox_myc_b200 as e
    from evox_myc_b200 import _lib

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=dev)
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world}"

    pop = n // 2
    x_host = make_fitness(n, m, 0)
    x_dev = torch.from_numpy(x_host).to(dev)

    if world > 1:
        from evox_myc_b200 import sharded

        ctx = sharded.ShardedSelector(n, m, dev)
        step_dev = lambda: ctx.nsga2_select(x_dev, pop)  # noqa: E731
        step_host = lambda: ctx.nsga2_select(x_host, pop)  # noqa: E731
    else:
        step_dev = lambda: e.nsga2_select(x_dev, pop)  # noqa: E731
        step_host = lambda: e.nsga2_select(x_host, pop)  # noqa: E731

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # ---- device-resident timing -------------------------------------------------------------
    # K steps enqueued back to back (no host round trip inside the timed region); one event pair around the
    # whole run gives `value`, one pair per step gives best / median.  Stage times come from a separate,
    # untimed pass with the library's stage events enabled.
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()  # nvidia-smi needs a moment to start; samples are filtered to the busy ones
    for _ in range(max(args.warmup, 3)):
        step_dev()
    barrier()
    launches0 = _lib.launch_count()
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    evs[0].record()
    for i in range(args.steps):
        step_dev()
        evs[i + 1].record()
    barrier()
    launches = _lib.launch_count() - launches0
    clocks = sampler.stop() if rank == 0 else None
    t_ms = evs[0].elapsed_time(evs[-1]) / args.steps
    per_step = sorted(evs[i].elapsed_time(evs[i + 1]) for i in range(args.steps))
    if dist is not None:
        tt = torch.tensor([t_ms, per_step[0], per_step[len(per_step) // 2]], device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        t_ms, best_ms, median_ms = (float(v) for v in tt.tolist())
    else:
        best_ms, median_ms = per_step[0], per_step[len(per_step) // 2]
    _lib.profile_enable(True)
    stage_acc: dict[str, float] = {}
    prof_steps = min(args.steps, 10)
    for _ in range(prof_steps):
        step_dev()
        for k, v in _lib.profile_read().items():  # synchronises on the step's last stage event
            stage_acc[k] = stage_acc.get(k, 0.0) + v
    _lib.profile_enable(False)
    barrier()

    # ---- end-to-end through the public API with host buffers -----------------------------------
    for _ in range(2):
        step_host()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        out = step_host()
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3 / args.steps
    if dist is not None:
        tt = torch.tensor([e2e_ms], device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        e2e_ms = float(tt.item())
    h2d = x_host.nbytes
    d2h = int(np.asarray(out).nbytes)

    # ---- what ran, and is it right (outside every timed region) ------------------------------
    parity = None
    exchange = None
    if world > 1:
        order_s, rank_s, crowd_s = ctx.nsga2_select(x_dev, pop, return_details=True)
        p2p, peer_err, fronts = ctx.status()
        exchange = {"path": ctx.last_path(), "peer_memory": p2p, "error": peer_err, "fronts": fronts}
        if rank == 0 and not args.no_parity:
            from oracle import c_oracle

            w_order, w_rank, w_crowd = c_oracle.nsga2_select(x_host, pop, nthreads=host_cores())
            fin = np.isfinite(w_crowd)
            crowd_np = crowd_s.cpu().numpy()
            parity = {"ranks_equal": bool(np.array_equal(rank_s.cpu().numpy(), w_rank)),
                      "order_equal": bool(np.array_equal(order_s.cpu().numpy(), w_order)),
                      "crowding_rtol_1e-6": bool(np.array_equal(np.isfinite(crowd_np), fin)
                                                 and np.allclose(crowd_np[fin], w_crowd[fin], rtol=1e-6, atol=0)),
                      "checker": "oracle/nds_oracle.c on rank 0, full size, after the timed regions"}

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    hbm_gbs, peak_src, sm_max = measured_peaks()
    stages = {k: v / prof_steps for k, v in stage_acc.items()}
    pairs = float(n) * float(n)
    peel_ms = stages.get("peel")
    cnt_ms = stages.get("dominator_count")
    build_ms = stages.get("dominance_build")
    # Algorithmic work is that of the reference formulation: every packed row
    # streamed once (N^2/8 bytes over all ranks) and 2*M compares per ordered pair.  The presorted,
    # upper-triangular layout really touches about half of those bytes and executes (M-1)+1 compares on
    # half of the pairs; both figures are reported so a fraction above 1 is not mistaken for a timing error.
    tri = m <= 16
    alg_bytes = float(n) * n / 8.0 / world
    real_bytes = alg_bytes * (0.5 + 0.5 * 4096.0 / n) if tri else alg_bytes  # chunks c >= row/4096 only
    roofline = None
    if world == 1:
        flow_possible = n <= 131072 and m <= 16
        peel_kernel = ("peel_flow_kernel or peel_persistent_kernel (both enqueued, front-0 size decides on the device)"
                       if flow_possible else "peel_persistent_kernel")
    else:
        peel_kernel = {"data-flow peer-memory kernel": "peel_flow_kernel (tagged words stored to the peers over NVLink)",
                       "grid-wide peer-memory kernel": "peel_p2p_kernel (tagged words stored to the peers over NVLink)",
                       }.get(exchange["path"], "front_subtract_kernel (+ detect / ncclAllGather / expand per front)")
    if world == 1:
        # both peel kernels may have been enqueued; the one that returned at once moved (almost) no bytes
        cands = [ncu_traffic_per_launch(k, n, m, 1) for k in (("peel_flow_kernel", "peel_persistent_kernel") if flow_possible
                                                                else ("peel_persistent_kernel",))]
        ncu_bytes, ncu_src = max(cands, key=lambda c: c[0] or 0.0)
    else:
        ncu_bytes, ncu_src = ncu_traffic_per_launch("peel_flow_kernel" if exchange["path"].startswith("data-flow") else "peel_p2p_kernel",
                                                    n, m, world)
    if peel_ms:
        ach = alg_bytes / (peel_ms * 1e-3) / 1e9
        roofline = {"kernel": peel_kernel,
                    "bound": "hbm", "achieved": ach, "peak": hbm_gbs, "unit": "GB/s",
                    "frac": ach / hbm_gbs, "traffic": ncu_bytes, "traffic_source": ncu_src,
                    "triangle_bytes": real_bytes,
                    "executed_gbs": real_bytes / (peel_ms * 1e-3) / 1e9, "executed_frac": real_bytes / (peel_ms * 1e-3) / 1e9 / hbm_gbs,
                    "peak_source": peak_src, "algorithmic_bytes_per_launch": alg_bytes, "ms_per_launch": peel_ms}
    extra = []
    if cnt_ms and build_uses_levels(n, m):
        extra.append({"kernel": "dominator_count_kernel", "ms_per_launch": cnt_ms,
                      "note": "returns at once: the level build counts the dominators while it builds the matrix; the "
                              "counting pass over the matrix only runs when that build falls back to the float kernel"})
    elif cnt_ms:
        ach = alg_bytes / (cnt_ms * 1e-3) / 1e9
        extra.append({"kernel": "dominator_count_kernel", "bound": "hbm", "achieved": ach, "peak": hbm_gbs, "unit": "GB/s",
                      "frac": ach / hbm_gbs, "executed_gbs": real_bytes / (cnt_ms * 1e-3) / 1e9,
                      "executed_frac": real_bytes / (cnt_ms * 1e-3) / 1e9 / hbm_gbs, "ms_per_launch": cnt_ms})
    if build_ms:
        # compare-pipe roofline: 64 lanes/clk/SM.  algorithmic = 2*M compares per ordered pair (common.py:98).
        # The presorted kernel executes far fewer (upper triangle only, M-1 float compares + 1 uid compare at
        # most, whole blocks discarded by their objective bands): how many is data dependent, so only the
        # algorithmic rate is computed here; ncu's ALU-pipe utilisation is in profiles/.
        clk = (clocks or {}).get("sm_mhz") or sm_max
        peak_ops = 148 * 64 * clk * 1e6
        pairs_rank = float(n) * n / world
        ach_ops = 2.0 * m * pairs_rank / (build_ms * 1e-3)
        levels = build_uses_levels(n, m)
        extra.append({"kernel": ("dominance_build_lvl_kernel (15-bit levels; the stage also holds the level_* set-up and fix-up kernels)"
                                 if levels else "dominance_build_tri_kernel" if tri else "dominance_build_kernel"),
                      "bound": ("integer pipes: ALU (LOP3 + VOTE) and FMA (IMAD), 64 lanes/clk/SM each; peak below = the float "
                                "kernel's compare pipe" if levels else "alu (fp32 compare pipe, 64 lanes/clk/SM)"),
                      "achieved": ach_ops / 1e12, "peak": peak_ops / 1e12, "unit": "Tcompare/s", "frac": ach_ops / peak_ops,
                      "executed": None, "executed_frac": None,
                      "ms_per_launch": build_ms, "pairs_per_s": pairs_rank / (build_ms * 1e-3)})

    nsga2_info = None
    nsga3_info = None
    if world == 1 and not args.no_nsga2:
        nsga2_info = bench_nsga2(args, dev)
        nsga3_info = bench_nsga3(args, dev)

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        from oracle import c_oracle

        ns = cpu_sample_rows(n, m, budget_s=20.0)
        xs = make_fitness(ns, m, 0)
        dt, (w_order, w_rank, w_crowd) = cpu_step(xs, want_result=True)
        cpu = {"value": ns * ns / dt, "unit": UNIT, "cores": host_cores(), "kind": "port",
               "sample": f"{ns} of {n} rows (same distribution, M={m}), one step = {dt:.2f} s"}
        if not args.no_parity:
            # the baseline's step doubles as the checker: same rows through the GPU path, outside every timed region
            order_s, rank_s, crowd_s = e.nsga2_select(torch.from_numpy(xs).to(dev), ns // 2, return_details=True)
            fin = np.isfinite(w_crowd)
            crowd_np = crowd_s.cpu().numpy()
            parity = {"ranks_equal": bool(np.array_equal(rank_s.cpu().numpy(), w_rank)),
                      "order_equal": bool(np.array_equal(order_s.cpu().numpy(), w_order)),
                      "crowding_rtol_1e-6": bool(np.array_equal(np.isfinite(crowd_np), fin)
                                                 and np.allclose(crowd_np[fin], w_crowd[fin], rtol=1e-6, atol=0)),
                      "checker": f"oracle/nds_oracle.c, {ns} of {n} rows (the cpu_baseline step's own result), after the timed regions"}

    line = {
        "metric": METRIC, "value": pairs / (t_ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": t_ms, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": desc, "N": n, "M": m, "pop_size": pop,
                   "step": "non_dominated_sort + worst-rank mask + crowding_distance + survivor order (nsga2.py:90-96)",
                   "l2": "working set (bit-packed dominance matrix, N^2/8 bytes per step) exceeds the 126 MB L2"
                         if n * n / 8 / world > 126e6 else "L2 flushed by the matrix rewrite of each step",
                   "cores": host_cores(),
                   "matrix_build": ("15-bit integer levels of the float objectives + exact fix-up (csrc/levels.cuh); ranks bit-exact "
                                    "with the float compares" if build_uses_levels(n, m) else "float compares"),
                   "parallelism": "single GPU" if world == 1 else
                   f"dominance matrix sharded by column chunk over {world} GPUs; front bitmask (N/8 bytes per front) exchanged by: {exchange['path']}"},
        "e2e": {"value": pairs / (e2e_ms * 1e-3), "unit": UNIT, "ms_per_step": e2e_ms,
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "gpu_launches": launches,
        "ms_per_step_best": best_ms, "ms_per_step_median": median_ms,
        "timing": "CUDA events on the launching stream; K steps enqueued back to back, max over ranks",
        "exchange": exchange,
        "parity": parity,
        "clocks": clocks,
        "roofline": roofline,
        "kernels": extra,
        "stages_ms": {k: round(v, 4) for k, v in stages.items()},
        "cpu_baseline": cpu,
        "nsga2": nsga2_info,
        "nsga3": nsga3_info,
    }
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def bench_nsga2(args, dev):
    """The other half of BASELINE.json's metric: NSGA-II generations/s.

    GPU: BASELINE configs[2] -- DTLZ2, pop = 16384, M = 3, D = 12; one generation = ask (SBX + polynomial
    mutation + clip) + evaluate + tell (ranking, crowding, survivor gather), all native kernels.
    CPU: BASELINE configs[0] shape -- pop = 100, 100 generations -- with the NumPy restatement
    (oracle/), one core; trajectories are not comparable with JAX's (different PRNG).
    """
    import torch

    from evox_myc_b200 import nsga2

    pop, m, d = 16384, 3, 12
    algo = nsga2.NSGA2(np.zeros(d), np.ones(d), m, pop, device=dev)
    prob = nsga2.DTLZ2(d=d, m=m)
    state = algo.setup(0)
    x, state = algo.init_ask(state)
    f, _ = prob.evaluate(None, x)
    state = algo.init_tell(state, f)

    def gen(state):
        x, state = algo.ask(state)
        f, _ = prob.evaluate(None, x)
        return algo.tell(state, f)

    for _ in range(5):
        state = gen(state)
    gens = max(args.steps, 20)
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(gens):
        state = gen(state)
    e1.record()
    torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1) / gens
    out = {"gpu": {"config": "BASELINE configs[2]: NSGA-II on DTLZ2, pop=16384, M=3, D=12", "generations": gens,
                   "ms_per_generation": ms, "gens_per_s": 1e3 / ms,
                   "note": "early generations (deep fronts); ask/evaluate/tell all native kernels"}}
    try:
        from oracle import nds_oracle, variation_oracle as vo

        P = np.random.default_rng(0).random((100, d), dtype=np.float32)
        fP = vo.dtlz2(P, m)
        t0 = time.perf_counter()
        for g in range(100):
            Q = vo.nsga2_ask(P, np.zeros(d, np.float32), np.ones(d, np.float32), 0, g)
            fQ = vo.dtlz2(Q, m)
            P, fP = nds_oracle.nsga2_tell(P, Q, fP, fQ, 100)
        dt = time.perf_counter() - t0
        out["cpu_port"] = {"config": "BASELINE configs[0] shape: NSGA-II on DTLZ2, pop=100, M=3, D=12, 100 generations",
                           "gens_per_s": 100 / dt, "cores": 1, "kind": "port (NumPy restatement)"}
    except Exception as ex:  # pragma: no cover
        out["cpu_port"] = {"error": repr(ex)}
    return out


def bench_nsga3(args, dev):
    """BASELINE configs[4]: NSGA-III on LSMOP1, M = 10, pop = 32768, d = 1000 -- generations/s with ask (SBX +
    polynomial mutation + clip), LSMOP1 evaluate and the whole tell (ranking, normalisation, association with
    29315 reference directions, niche selection) as native kernels; stage times of one tell beside it."""
    import torch

    from evox_myc_b200 import _lib, nsga2, nsga3

    pop, m, d = 32768, 10, 1000
    lb = np.zeros(d, np.float32)
    ub = np.concatenate([np.ones(m - 1), 10 * np.ones(d - m + 1)]).astype(np.float32)  # the box of tests/test_lsmop.py:10-14
    algo = nsga3.NSGA3(lb, ub, m, pop, device=dev)
    prob = nsga2.LSMOP1(d=d, m=m)
    state = algo.setup(0)
    x, state = algo.init_ask(state)
    f, _ = prob.evaluate(None, x)
    state = algo.init_tell(state, f)

    def gen(state):
        x, state = algo.ask(state)
        f, _ = prob.evaluate(None, x)
        return algo.tell(state, f)

    for _ in range(3):
        state = gen(state)
    gens = max(min(args.steps, 20), 5)
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(gens):
        state = gen(state)
    e1.record()
    torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1) / gens
    _lib.profile_enable(True)
    state = gen(state)
    stages = _lib.profile_read()
    _lib.profile_enable(False)
    return {"config": "BASELINE configs[4]: NSGA-III on LSMOP1, M=10, pop=32768, d=1000, 29315 reference points",
            "generations": gens, "ms_per_generation": ms, "gens_per_s": 1e3 / ms,
            "tell_stages_ms": {k: round(v, 4) for k, v in stages.items()},
            "note": "ask / evaluate / tell all native kernels; tell consumes the JAX-layout threefry key like the reference"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="evoxb", choices=["evoxb", "reference"])
    ap.add_argument("--workload", default="c4")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-nsga2", action="store_true", help="skip the NSGA-II generations/s measurement")
    ap.add_argument("--no-parity", action="store_true", help="multi-GPU: skip the full-size oracle comparison on rank 0")
    args = ap.parse_args()
    n, m, desc = parse_workload(args.workload)
    if args.impl == "reference":
        run_reference_arm(args, n, m, desc)
    else:
        run_gpu_arm(args, n, m, desc)


if __name__ == "__main__":
    main()
