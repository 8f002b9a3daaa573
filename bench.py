#!/usr/bin/env python3
"""bench.py -- headline benchmark of the B200 LP engine ("Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--batch B] [--dist D1|D2|D3] [--n 100] [--m 100]
                    [--no-single] [--no-gj] [--no-e2e] [--no-cpu-baseline] [--scaling weak|strong]

ONE JSON line covers both halves of BASELINE.json's metric and the three bench-size configs:

  configs[3] (headline, `value`): B = 65536 random feasible 100x100 LPs per GPU (distribution D1 of
      linear_programming_b200/workloads.py, integer data stored as fp64), sharded by problem index across
      ranks with NO data-path collective.  One "step" = one launch of simplex_batched_kernel over the batch.
        value        LPs/s, whole job, inputs already resident in HBM when the timed region starts
        e2e          LPs/s through the reference-facing C ABI with HOST (pinned) buffers: host->device copies,
                     solve and device->host copies all inside the timed region
        roofline     the dominant kernel against what binds it: the tableau is shared-memory resident, so the
                     bound is the shared-memory pipe ("smem"); the SURVEY 8(d) figure (16*(n+1)*(m+1) bytes per
                     pivot against the HBM peak) is kept as effective_vs_hbm
        strong_scaling  (N > 1) the same 65536 LPs split over the N ranks, as configs[3] words it
        cpu_baseline the C oracle port timed on this box's host cores on a bounded sample
  configs[4] (`roofline_single`, `parity_single`): ONE dense LP with a 16384 x 32768 condensed tableau on the
      grid-wide kernel -- on one GPU, or COLUMN-SHARDED over all N ranks under torchrun (strong scaling; the only
      multi-GPU path with an exchange: in-kernel entering-key all-to-all + pivot-column broadcast).  ms per pivot,
      achieved HBM GB/s = 16*(n+1)*(m+1) / ms_per_pivot against N x the measured HBM peak, the pivot-sequence CRC
      (must not depend on N), and full-size parity: the first pivots, the objective row and the constant column
      after them against the CPU oracle (rows swept on all host threads).  N > 1 adds `sharded_soak`: many small
      LPs back to back on one communicator (>= 10^5 exchange-dominated pivots), every answer against the oracle.
  configs[2] (`gauss_jordan`): 4096 systems of 512 x 512 per GPU, ALL of rank 0's checked bit for bit.

`--impl reference` times the CPU implementation of the path (the reference solver is Python and cannot travel to
the GPU box, so this is the oracle port -- pinned bit-exact to the reference -- on all host threads), same
metric/config, bounded sample per step.
"""
from __future__ import annotations

import argparse
import hashlib
import json
import os
import subprocess
import sys
import threading
import time
import zlib

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "LPs/sec (batched 100x100 simplex)"
UNIT = "LPs/s"


def parse_args(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=65536)
    ap.add_argument("--dist", default="D1", choices=["D1", "D2", "D3"])
    ap.add_argument("--n", type=int, default=100)
    ap.add_argument("--m", type=int, default=100)
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: every rank solves --batch LPs (default); strong: --batch LPs in total, split by "
                         "problem index (the headline `value` is then the strong-scaling figure)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    # configs[4]: the single large LP
    ap.add_argument("--no-single", action="store_true")
    ap.add_argument("--single-n", type=int, default=16383)
    ap.add_argument("--single-m", type=int, default=32767)
    ap.add_argument("--single-pivots", type=int, default=200)
    ap.add_argument("--single-parity-pivots", type=int, default=32,
                    help="first pivots of the large LP replayed on the CPU oracle (0 = skip)")
    ap.add_argument("--soak-lps", type=int, default=64, help="N > 1: small LPs solved back to back, sharded")
    # configs[2]: batched Gauss-Jordan
    ap.add_argument("--no-gj", action="store_true")
    ap.add_argument("--gj-batch", type=int, default=4096)
    ap.add_argument("--gj-n", type=int, default=512)
    ap.add_argument("--gj-check", type=int, default=-1, help="systems of rank 0 checked against the oracle (-1: all)")
    return ap.parse_args(argv)


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


KERNEL_SOURCES = {  # the files a kernel's device code comes from (host-side edits elsewhere do not stale its ncu facts)
    "simplex_batched_kernel": ["simplex_batched.cu", "simplex_batched_impl.cuh", "lp_common.cuh"],
    "simplex_single_kernel": ["simplex_single.cu", "lp_common.cuh"],
    "gauss_jordan_batched_kernel": ["gauss_jordan_batched.cu", "lp_common.cuh"],
    "verify_bounds_kernel": ["verify_bounds.cu", "lp_common.cuh"],
}


def source_sha16(kernel=None):
    """Hash of the CUDA/C++ sources the library is built from (all of csrc/, or the files of one kernel): ties
    committed ncu summaries to a build."""
    h = hashlib.sha256()
    csrc = os.path.join(ROOT, "linear_programming_b200", "csrc")
    names = sorted(KERNEL_SOURCES[kernel]) if kernel in KERNEL_SOURCES else \
        sorted(n for n in os.listdir(csrc) if n.endswith((".cu", ".cuh", ".h", ".cpp")))
    for name in names:
        with open(os.path.join(csrc, name), "rb") as f:
            h.update(name.encode())
            h.update(f.read())
    return h.hexdigest()[:16]


def ncu_facts(kernel: str):
    """Per-launch ncu facts of `kernel` from profiles/ncu_facts.json (written by tools/ncu_facts.py from a
    --set full capture of this very command).  Each entry records the source hash of the build it was taken
    from; facts of another build are NOT reported (traffic: null)."""
    path = os.path.join(ROOT, "profiles", "ncu_facts.json")
    try:
        with open(path) as f:
            rec = json.load(f).get(kernel)
    except Exception:
        return None
    if not rec:
        return None
    if rec.get("src_sha16") != source_sha16(kernel):
        return {"stale": True, "captured_src_sha16": rec.get("src_sha16"), "file": rec.get("file")}
    return rec


def smem_peak(sm_mhz):
    """Shared-memory bandwidth of the whole GPU: the committed micro-benchmark result when there is one
    (tools/gpu/micro/smem_bw.cu -> profiles/r2_smem_peak.json), else 148 SMs x 128 B/clk x the SM clock."""
    try:
        with open(os.path.join(ROOT, "profiles", "r2_smem_peak.json")) as f:
            d = json.load(f)
            return float(d["smem_GBps"]), "measured (profiles/r2_smem_peak.json, LDS/STS.128 streaming)"
    except Exception:
        mhz = sm_mhz or 1965.0
        return 148 * 128 * mhz * 1e6 / 1e9, "nominal (148 SMs x 128 B/clk x %.0f MHz)" % mhz


def config_dict(args, batch, world=1):
    strong = args.scaling == "strong"
    return {
        "workload": "configs[3]: batched simplex, %d random feasible %dx%d LPs %s (%s), "
                    "sharded by problem index" % (batch, args.n, args.m, "in total" if strong else "per GPU", args.dist),
        "batch_per_gpu": batch // world if strong else batch, "n": args.n, "m": args.m, "distribution": args.dist,
        "tolerance": 1e-9, "parallelism": "problem-sharded, no collective",
        "l2": "inputs (%.1f GB per GPU) larger than L2; every step re-reads them from HBM"
              % ((batch // world if strong else batch) * (args.n * args.m + args.n + args.m) * 8 / 1e9),
        "also_measured": "configs[4] single 16384x32768 LP (roofline_single, column-sharded over the ranks), "
                         "configs[2] 4096 x 512x512 Gauss-Jordan (gauss_jordan)",
    }


# ------------------------------------------------------------------------------------------
# clocks sampler (nvidia-smi during the timed regions)
# ------------------------------------------------------------------------------------------
class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.lines = []
        self.proc = None
        self.thread = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.FIELDS,
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None
            return
        self.thread = threading.Thread(target=self._read, daemon=True)
        self.thread.start()

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def mark(self):
        """Index of the next sample line: brackets a timed region inside one long recording."""
        return len(self.lines)

    def window(self, lo=0, hi=None):
        """Clock statistics of the samples [lo, hi) (the recording keeps running)."""
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        window = self.lines[lo:hi] if hi is not None and hi > lo else self.lines[lo:]
        if not window:
            window = self.lines[max(0, lo - 2):]
        for ln in window:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
                power.append(float(parts[2]))
            except ValueError:
                continue
            for nm, val in zip(names, parts[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "power_w_max": float(max(power)) if power else None, "samples": len(sm)}

    def close(self):
        if self.proc is None:
            return
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()

    def stop(self, lo=0, hi=None):
        time.sleep(0.15)
        out = self.window(lo, hi)
        self.close()
        return out


# ------------------------------------------------------------------------------------------
# CPU arm: the oracle port on the host cores (bounded sample of the same workload)
# ------------------------------------------------------------------------------------------
def cpu_sample_size(args, oracle, workloads, target_s=12.0):
    """Pick a sample of the first LPs of the shard worth ~target_s of wall time on all cores."""
    A, b, c = workloads.simplex_batch(args.dist, 64, args.n, args.m, seed=2)
    t0 = time.perf_counter()
    oracle.simplex_batch(A, b, c, max_pivots=50000, threads=0)
    dt = max(time.perf_counter() - t0, 1e-4)
    rate = 64 / dt
    cnt = int(min(args.batch, max(64, rate * target_s)))
    return (cnt + 63) // 64 * 64


def run_cpu_baseline(args, sample=None, repeats=1):
    import oracle
    from linear_programming_b200 import workloads
    oracle.build()
    if sample is None:
        sample = cpu_sample_size(args, oracle, workloads)
    A, b, c = workloads.simplex_batch(args.dist, sample, args.n, args.m, seed=2)
    best = None
    piv = 0
    used = 1
    for _ in range(repeats):
        t0 = time.perf_counter()
        st, x, obj, npiv, used = oracle.simplex_batch(A, b, c, max_pivots=50000, threads=0)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
        piv = int(npiv.sum())
    return {"value": sample / best, "unit": UNIT, "cores": used, "kind": "port",
            "sample": "first %d LPs of the same %s batch, C oracle (bit-exact restatement of "
                      "week-2/diet/simplex.py), one LP per host thread, %.2f s" % (sample, args.dist, best),
            "pivots_per_s": piv / best, "seconds": best, "lps": sample}


def reference_arm(args):
    """--impl reference: the CPU implementation of the path on all host threads.  The reference
    solver itself is pure Python under /root/reference (absent on the GPU box, cannot be
    copied), so this times the oracle port; each step is a bounded sample of the workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import oracle
    from linear_programming_b200 import workloads
    oracle.build()
    sample = cpu_sample_size(args, oracle, workloads, target_s=6.0)
    A, b, c = workloads.simplex_batch(args.dist, sample, args.n, args.m, seed=2)
    used = 1
    for _ in range(args.warmup):
        oracle.simplex_batch(A[:64], b[:64], c[:64], max_pivots=50000, threads=0)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        _, _, _, npiv, used = oracle.simplex_batch(A, b, c, max_pivots=50000, threads=0)
    dt = time.perf_counter() - t0
    value = sample * args.steps / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": config_dict(args, args.batch, max(1, args.gpus)),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": used, "kind": "port",
                         "sample": "each step = first %d LPs of the %s batch on %d host threads"
                                   % (sample, args.dist, used)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "pivots_per_s": float(npiv.sum()) * args.steps / dt,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------
class Ctx:
    """Process-wide state of the GPU arm: rank, device, library, collectives for the plumbing."""

    def __init__(self):
        import torch
        import torch.distributed as dist
        from linear_programming_b200 import _lib
        self.torch, self.dist = torch, dist
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise RuntimeError("bench.py needs a CUDA device: the engine has no CPU fallback")
        torch.cuda.set_device(self.local_rank)
        self.dev = torch.device("cuda", self.local_rank)
        if self.world > 1:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group("nccl", device_id=self.dev)
            # the ranks of one box share its host cores: each rank's narrowing pool (second lane of the
            # host-pointer path) gets its share instead of every rank spawning one thread per core
            os.environ.setdefault("LPB_HOST_THREADS", str(max(1, (os.cpu_count() or 1) // self.world)))
        self._lib = _lib
        self.L = _lib.lib()
        self.stream = torch.cuda.current_stream(self.dev).cuda_stream
        self.sampler = ClockSampler(self.local_rank)

    def check(self, rc):
        self._lib.check(rc)

    def barrier(self):
        if self.world > 1:
            self.dist.barrier(device_ids=[self.local_rank])

    def _reduce(self, v, op):
        if self.world == 1:
            return v
        t = self.torch.tensor([v], dtype=self.torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=op)
        return float(t.item())

    def max_over_ranks(self, v: float) -> float:
        return self._reduce(v, self.dist.ReduceOp.MAX)

    def min_over_ranks(self, v: float) -> float:
        return self._reduce(v, self.dist.ReduceOp.MIN)

    def sum_over_ranks(self, v: float) -> float:
        return self._reduce(v, self.dist.ReduceOp.SUM)

    def timed(self, fn, reps=1):
        """fn() `reps` times between two events on the launching stream, barrier + synchronize on both sides;
        returns (max-over-ranks milliseconds, clocks of the window)."""
        torch = self.torch
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        self.barrier()
        torch.cuda.synchronize()
        lo = self.sampler.mark()
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        hi = self.sampler.mark()
        ms = self.max_over_ranks(e0.elapsed_time(e1))
        self.barrier()
        return ms, (lo, hi)


def merge_clocks(sampler, windows):
    """Clock record over several (lo, hi) windows of one recording."""
    keep = []
    for lo, hi in windows:
        keep.extend(sampler.lines[lo:max(hi, lo + 1)])
    tmp = ClockSampler(sampler.index)
    tmp.proc = sampler.proc
    tmp.lines = keep if keep else sampler.lines[-3:]
    return tmp.window(0, None)


# ---- configs[3]: the batched 100x100 simplex (headline) ----
def bench_batch(ctx, args):
    import ctypes
    torch = ctx.torch
    from linear_programming_b200 import workloads
    from linear_programming_b200.sharded import problem_shard
    L, dev, rank, world = ctx.L, ctx.dev, ctx.rank, ctx.world
    n, m = args.n, args.m
    tol, cap = 1e-9, 50000
    if args.scaling == "strong":
        lo_p, hi_p = problem_shard(args.batch, world, rank)
        B, start = hi_p - lo_p, lo_p
    else:
        B, start = args.batch, rank * args.batch  # problem index = rank*B + p
    A, b, c = workloads.simplex_batch(args.dist, B, n, m, seed=2, start=start)
    hA, hb, hc = (torch.from_numpy(v).pin_memory() for v in (A, b, c))
    del A, b, c
    dA, db, dc = (v.to(dev, non_blocking=True) for v in (hA, hb, hc))
    d_st = torch.empty(B, dtype=torch.int32, device=dev)
    d_x = torch.empty((B, m), dtype=torch.float64, device=dev)
    d_obj = torch.empty(B, dtype=torch.float64, device=dev)
    d_np = torch.empty(B, dtype=torch.int32, device=dev)
    stream = ctx.stream

    def launch(cnt):
        ctx.check(L.lpb200_simplex_batch_device(cnt, n, m, dA.data_ptr(), db.data_ptr(), dc.data_ptr(), tol, cap,
                                                d_st.data_ptr(), d_x.data_ptr(), d_obj.data_ptr(),
                                                d_np.data_ptr(), None, 0, stream))

    def step():
        launch(B)

    for _ in range(max(args.warmup, 3)):
        step()
    torch.cuda.synchronize()
    K = args.steps
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(K + 1)]
    ctx.barrier()
    torch.cuda.synchronize()
    lo = ctx.sampler.mark()
    evs[0].record()
    for k in range(K):
        step()
        evs[k + 1].record()
    torch.cuda.synchronize()
    time.sleep(0.12)
    hi = ctx.sampler.mark()
    ctx.barrier()
    clocks = ctx.sampler.window(lo, hi)
    total_ms = ctx.max_over_ranks(evs[0].elapsed_time(evs[K]))
    kern_ms = [evs[k].elapsed_time(evs[k + 1]) for k in range(K)]
    pivots_step = int(d_np.sum().item())
    total_lps = ctx.sum_over_ranks(float(B))
    value = total_lps * K / (total_ms * 1e-3)
    peak, peak_src = measured_peak()
    avg_kern_s = float(np.mean(kern_ms)) * 1e-3
    bytes_per_pivot = 16.0 * (n + 1) * (m + 1)
    alg_bytes = bytes_per_pivot * pivots_step
    achieved = alg_bytes / avg_kern_s / 1e9
    input_bytes = B * (n * m + n + m) * 8 + B * (m * 8 + 16)
    speak, speak_src = smem_peak(clocks.get("sm_mhz"))
    facts = ncu_facts("simplex_batched_kernel")
    same_cfg = (B == 65536 and args.dist == "D1" and n == 100 and m == 100)
    traffic = wavefronts = None
    traffic_src = "no ncu capture of this build in profiles/ncu_facts.json"
    if facts and not facts.get("stale") and same_cfg:
        traffic = facts.get("dram_bytes")
        if facts.get("smem_wavefronts") and facts.get("pivots"):
            wavefronts = facts["smem_wavefronts"] / facts["pivots"]
        traffic_src = facts.get("file")
    elif facts and facts.get("stale"):
        traffic_src = "profiles/ncu_facts.json was captured from another build (%s): not reported" \
                      % facts.get("captured_src_sha16")
    roofline = {
        "bound": "smem", "achieved": achieved, "peak": speak, "unit": "GB/s", "frac": achieved / speak,
        "peak_source": speak_src,
        "traffic": traffic, "traffic_source": traffic_src,
        "effective_vs_hbm": achieved / peak, "hbm_peak": peak, "hbm_peak_source": peak_src,
        "compulsory_hbm_bytes_per_launch": input_bytes,
        "dram_traffic_ratio": (traffic / input_bytes) if traffic else None,
        "hbm_frac_physical": (traffic / avg_kern_s / 1e9 / peak) if traffic else None,
        "wavefronts_per_pivot": wavefronts,
        "ideal_wavefronts_per_pivot": bytes_per_pivot / 128.0,
        "kernel": "simplex_batched_kernel", "launch_ms": avg_kern_s * 1e3,
        "launch_ms_each": [round(v, 3) for v in kern_ms],
        "algorithmic_bytes_per_launch": alg_bytes, "algorithmic_bytes_per_pivot": bytes_per_pivot,
        "pivots_per_launch": pivots_step, "src_sha16": source_sha16("simplex_batched_kernel"),
        "note": "the tableau of an LP stays in shared memory for the whole solve, so the pipe that binds the kernel "
                "is shared memory (128 B/clk/SM): frac = algorithmic bytes (16*(n+1)*(m+1) per pivot, SURVEY 8(d)) / "
                "time / shared-memory peak.  The same bytes against the HBM peak (effective_vs_hbm) exceed 1: HBM "
                "only carries the inputs once and the results once (compulsory_hbm_bytes_per_launch).",
    }
    all_pivots = ctx.sum_over_ranks(float(pivots_step))

    # ---- parity + CPU baseline (rank 0, outside every timed region) ----
    cpu = None
    parity = None
    if rank == 0 and not args.no_cpu_baseline:
        import oracle
        oracle.build()
        cpu = run_cpu_baseline(args)
        s = min(cpu["lps"], B)
        A0, b0, c0 = workloads.simplex_batch(args.dist, s, n, m, seed=2, start=start)
        st, x, obj, npiv, _ = oracle.simplex_batch(A0, b0, c0, max_pivots=cap)
        ok = (np.array_equal(d_st[:s].cpu().numpy(), st) and np.array_equal(d_np[:s].cpu().numpy(), npiv)
              and np.array_equal(d_x[:s].cpu().numpy().view(np.int64), x.view(np.int64))
              and np.array_equal(d_obj[:s].cpu().numpy().view(np.int64), obj.view(np.int64)))
        parity = {"checked_lps": int(s), "bit_exact_vs_oracle": bool(ok)}
        cpu = {k: v for k, v in cpu.items() if k not in ("seconds", "lps")}

    # ---- strong scaling (N > 1, weak mode): the SAME `--batch` problems split over the ranks ----
    strong = None
    if world > 1 and args.scaling == "weak":
        # rank r takes the first B/N LPs it already holds: the per-rank workload of a 65536-LP job cut N ways
        # (every shard is B/N D1 problems; which indices they are does not change the work)
        Bs = problem_shard(args.batch, world, rank)
        cnt = Bs[1] - Bs[0]
        launch(cnt)
        ms, win = ctx.timed(lambda: launch(cnt), reps=K)
        strong = {"total_lps": args.batch, "lps_per_rank": cnt, "value": args.batch * K / (ms * 1e-3), "unit": UNIT,
                  "ms_per_step": ms / K, "clocks": ctx.sampler.window(*win),
                  "note": "configs[3] as worded: one 65536-LP batch sharded by problem over the N ranks; "
                          "efficiency = value / (N x the N=1 `value`)"}
        launch(B)  # leave the full-batch answers in place for the checks below
        torch.cuda.synchronize()

    # ---- end to end through the C ABI with host (pinned) buffers ----
    e2e = None
    e2e_compact = None
    if not args.no_e2e:
        h_st = torch.empty(B, dtype=torch.int32).pin_memory()
        h_x = torch.empty((B, m), dtype=torch.float64).pin_memory()
        h_obj = torch.empty(B, dtype=torch.float64).pin_memory()
        h_np = torch.empty(B, dtype=torch.int32).pin_memory()

        def e2e_step():
            ctx.check(L.lpb200_simplex_batch(B, n, m, hA.data_ptr(), hb.data_ptr(), hc.data_ptr(), tol, cap,
                                             h_st.data_ptr(), h_x.data_ptr(), h_obj.data_ptr(),
                                             h_np.data_ptr(), None, 0))

        def last_lanes():
            out = (ctypes.c_int64 * 4)()
            L.lpb200_simplex_batch_last_lanes(out)
            return [int(v) for v in out]
        Ke = max(1, min(K, 3))
        e2e_step()  # warm-up
        e2e_step()  # second warm-up: the lane gate measures on the first call of a shape
        ctx.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(Ke):
            e2e_step()   # synchronous: returns after the last device->host copy
        torch.cuda.synchronize()
        dt = ctx.max_over_ranks(time.perf_counter() - t0)
        ctx.barrier()
        same = bool(torch.equal(h_st, d_st.cpu()) and torch.equal(h_np, d_np.cpu())
                    and torch.equal(h_x.view(torch.int64), d_x.cpu().view(torch.int64)))
        lanes = last_lanes()  # of the last timed step on this rank
        e2e = {"value": total_lps * Ke / dt, "unit": UNIT,
               "h2d_bytes_per_step": int(lanes[3] + B * (n + m) * 8),
               "host_fp64_bytes_per_step": int(B * (n * m + n + m) * 8),
               "d2h_bytes_per_step": int(B * (m * 8 + 8 + 4 + 4)), "steps": Ke,
               "ms_per_step": dt / Ke * 1e3, "matches_device_path": same,
               "lanes": {"copy_engine_chunks": lanes[0], "host_narrowing_chunks": lanes[1],
                         "narrowed_to_int8": lanes[2], "host_threads": int(L.lpb200_host_threads())},
               "api": "lpb200_simplex_batch (fp64 host pointers, pinned; chunked copy/solve pipeline, two lanes: "
                      "the copy engine ships fp64 chunks while the host cores narrow other chunks to int8 "
                      "exactly -- answers bit-identical, see include/lp_b200.h)"}

        # ---- the same call with the constraint matrix held as int8 on the host (the reference parses
        #      integers, simplex.py:712-717; Config 4's coefficients are in [-100, 100]): 1/8 of the A bytes
        #      over PCIe, answers bit-identical (checked).  Reported beside e2e, not instead of it:
        #      Config 4 is defined with fp64 storage. ----
        hA8 = hA.to(torch.int8).pin_memory()
        lossless = bool(torch.equal(hA8.to(torch.float64), hA))
        c_st = torch.empty(B, dtype=torch.int32).pin_memory()
        c_x = torch.empty((B, m), dtype=torch.float64).pin_memory()
        c_obj = torch.empty(B, dtype=torch.float64).pin_memory()
        c_np = torch.empty(B, dtype=torch.int32).pin_memory()

        def compact_step():
            ctx.check(L.lpb200_simplex_batch_ex(B, n, m, hA8.data_ptr(), 4, hb.data_ptr(), hc.data_ptr(), tol,
                                                cap, 0, c_st.data_ptr(), c_x.data_ptr(), c_obj.data_ptr(),
                                                c_np.data_ptr(), None, 0))
        compact_step()
        ctx.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(Ke):
            compact_step()
        torch.cuda.synchronize()
        dtc = ctx.max_over_ranks(time.perf_counter() - t0)
        ctx.barrier()
        same_c = bool(lossless and torch.equal(c_st, h_st) and torch.equal(c_np, h_np)
                      and torch.equal(c_x.view(torch.int64), h_x.view(torch.int64)))
        e2e_compact = {"value": total_lps * Ke / dtc, "unit": UNIT, "a_dtype": "int8",
                       "h2d_bytes_per_step": int(B * (n * m + (n + m) * 8)),
                       "d2h_bytes_per_step": int(B * (m * 8 + 8 + 4 + 4)), "steps": Ke,
                       "ms_per_step": dtc / Ke * 1e3, "matches_fp64_call": same_c,
                       "api": "lpb200_simplex_batch_ex (LPB200_DTYPE_I8 A; b, c fp64; host pointers, pinned)"}
        del hA8

    # ---- extra: pivot-heavy distribution D2 on a smaller batch (pivots/s), rank 0 only, untimed region ----
    extra = {}
    if rank == 0 and args.dist == "D1":
        B2 = min(B, 4096)
        A2, b2, c2 = (torch.from_numpy(v).to(dev) for v in workloads.simplex_batch("D2", B2, n, m, seed=2))
        st2 = torch.empty(B2, dtype=torch.int32, device=dev)
        np2 = torch.empty(B2, dtype=torch.int32, device=dev)
        x2 = torch.empty((B2, m), dtype=torch.float64, device=dev)
        o2 = torch.empty(B2, dtype=torch.float64, device=dev)

        def step2():
            ctx.check(L.lpb200_simplex_batch_device(B2, n, m, A2.data_ptr(), b2.data_ptr(), c2.data_ptr(), tol,
                                                    cap, st2.data_ptr(), x2.data_ptr(), o2.data_ptr(),
                                                    np2.data_ptr(), None, 0, stream))
        step2()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        step2()
        e1.record()
        torch.cuda.synchronize()
        ms2 = e0.elapsed_time(e1)
        piv2 = int(np2.sum().item())
        extra = {"D2_batch": B2, "D2_ms": ms2, "D2_pivots": piv2, "D2_pivots_per_s": piv2 / (ms2 * 1e-3),
                 "D2_lps_per_s": B2 / (ms2 * 1e-3),
                 "D2_effective_GBps": bytes_per_pivot * piv2 / (ms2 * 1e-3) / 1e9,
                 "D2_frac_of_smem_peak": bytes_per_pivot * piv2 / (ms2 * 1e-3) / 1e9 / speak}
    if rank == 0:
        # the unmodified Python reference cannot run on this box; its timing on samples of these very
        # inputs was recorded where it exists (tools/time_reference_batch.py) and is quoted, labelled
        try:
            with open(os.path.join(ROOT, "tests", "golden", "reference_batch_times.json")) as f:
                rt = json.load(f)
            key = "config4_" + args.dist
            if key in rt and n == 100 and m == 100:
                extra["python_reference_recorded"] = {
                    "lps_per_s_per_core": rt[key]["lps_per_s_per_core"], "ms_per_pivot": rt[key]["ms_per_pivot"],
                    "sample_lps": rt[key]["lps"], "where": rt["note"]}
        except Exception:
            pass
    return {"value": value, "ms_per_step": total_ms / K, "pivots_per_s": all_pivots * K / (total_ms * 1e-3),
            "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu, "parity": parity, "e2e": e2e,
            "e2e_compact": e2e_compact, "strong_scaling": strong, "extra": extra, "B": B}


# ---- configs[4]: one large LP, grid-wide kernel, column-sharded over the ranks ----
def bench_single(ctx, args):
    torch = ctx.torch
    from linear_programming_b200 import workloads
    from linear_programming_b200.sharded import ShardedComm, column_shard
    L, dev, rank, world = ctx.L, ctx.dev, ctx.rank, ctx.world
    n, m, P = args.single_n, args.single_m, args.single_pivots
    Kp = min(args.single_parity_pivots, P)
    nv = n + m + 1
    comm = None
    if world > 1:
        comm = ShardedComm(max_n=max(n, 128))
        lo, cnt = column_shard(m, world, rank)
    else:
        lo, cnt = 0, m
    A_full = None
    t_gen = time.perf_counter()
    if rank == 0 and Kp > 0 and not args.no_cpu_baseline:
        A_full, b, c = workloads.single_lp(n, m, seed=5)  # the oracle needs the whole matrix
        A_mine = A_full if world == 1 else np.ascontiguousarray(A_full[:, lo:lo + cnt])
    else:
        A_mine, b, c = workloads.single_lp(n, m, seed=5, col_lo=lo, col_cnt=cnt)
    t_gen = time.perf_counter() - t_gen
    dA = torch.from_numpy(A_mine).to(dev)
    db = torch.from_numpy(b).to(dev)
    dc = torch.from_numpy(c).to(dev)
    del A_mine
    cap = max(P, 1)
    if world == 1:
        wsb = int(L.lpb200_simplex_single_workspace_bytes(n, m))
    else:
        wsb = int(L.lpb200_simplex_sharded_workspace_bytes(n, m, world))
    ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
    st = torch.zeros(1, dtype=torch.int32, device=dev)
    npv = torch.zeros(1, dtype=torch.int32, device=dev)
    x = torch.zeros(m, dtype=torch.float64, device=dev)
    obj = torch.zeros(1, dtype=torch.float64, device=dev)
    log = torch.full((cap, 2), -1, dtype=torch.int32, device=dev)
    launches = [0]

    def launch(maxpiv):
        launches[0] += 1
        if world == 1:
            ctx.check(L.lpb200_simplex_single_ex_device(n, m, dA.data_ptr(), db.data_ptr(), dc.data_ptr(), 1e-9,
                                                        maxpiv, 0, st.data_ptr(), x.data_ptr(), obj.data_ptr(),
                                                        npv.data_ptr(), log.data_ptr(), cap, ws.data_ptr(), wsb,
                                                        ctx.stream))
        else:
            ctx.check(L.lpb200_simplex_sharded_ex_device(comm._comm, n, m, dA.data_ptr() if cnt else None,
                                                         dA.stride(0) if cnt else 0, db.data_ptr(), dc.data_ptr(),
                                                         1e-9, maxpiv, 0, st.data_ptr(), x.data_ptr(),
                                                         obj.data_ptr(), npv.data_ptr(), log.data_ptr(), cap,
                                                         ws.data_ptr(), wsb, 30.0, ctx.stream))

    def run(maxpiv):
        ms, win = ctx.timed(lambda: launch(maxpiv))
        return ms, int(npv.item()), int(st.item()), win

    run(2)  # warm-up
    run(2)
    run(2)
    setup_ms = min(run(0)[0] for _ in range(2))
    best = None
    windows = []
    for _ in range(2):
        r = run(P)
        windows.append(r[3])
        if best is None or r[0] < best[0]:
            best = r
    ms, piv, status, _ = best
    crc = zlib.crc32(log[:piv].cpu().numpy().tobytes())
    log_full = log[:piv].cpu().numpy().copy()
    per_pivot_ms = (ms - setup_ms) / max(piv, 1)
    alg = 16.0 * (n + 1) * (m + 1)
    peak, peak_src = measured_peak()
    gbs = alg / (per_pivot_ms * 1e-3) / 1e9
    out = {
        "workload": "configs[4]: single dense LP, condensed tableau %d x %d (n=%d, m=%d), D1, "
                    "column-sharded over %d GPU(s)" % (n + 1, m + 1, n, m, world),
        "kernel": "simplex_single_kernel", "bound": "hbm", "n_gpus": world, "pivots": piv, "status": status,
        "pivot_crc": crc, "total_ms": ms, "setup_ms": setup_ms, "ms_per_pivot": per_pivot_ms,
        "algorithmic_bytes_per_pivot": alg, "achieved": gbs, "unit": "GB/s", "peak": peak * world,
        "peak_per_gpu": peak, "peak_source": peak_src, "frac": gbs / (peak * world), "pivots_per_s": 1e3 / per_pivot_ms,
        "scaling": "strong (one LP, columns split over the ranks)",
        "clocks": merge_clocks(ctx.sampler, windows), "gpu_launches": None, "host_generate_s": t_gen,
        "exchange": "none" if world == 1 else
                    "in-kernel LL packets over peer-mapped (CUDA IPC) buffers: 8-byte entering-key all-to-all + "
                    "pivot-column broadcast per pivot, no NCCL call on the pivot path",
    }
    facts = ncu_facts("simplex_single_kernel")
    if facts and not facts.get("stale") and world == 1:
        out["traffic"] = facts.get("dram_bytes_per_pivot")
        out["traffic_source"] = facts.get("file")
    else:
        out["traffic"] = None

    # ---- full-size parity: first Kp pivots, objective row and constant column vs the CPU oracle ----
    parity = None
    if Kp > 0 and not args.no_cpu_baseline:
        launch(Kp)
        torch.cuda.synchronize()
        ov = torch.zeros(nv + 1, dtype=torch.float64, device=dev)
        cv = torch.zeros(nv, dtype=torch.float64, device=dev)
        ctx.check(L.lpb200_simplex_single_state_device(n, m, ws.data_ptr(), ov.data_ptr(), cv.data_ptr(), ctx.stream))
        torch.cuda.synchronize()
        obj_const = float(ov[nv].item())
        ovv = ov[:nv].clone()
        cv_all_same = True
        if world > 1:
            ctx.dist.all_reduce(ovv, op=ctx.dist.ReduceOp.SUM)  # every column is owned by exactly one rank
            cmax, cmin = cv.clone(), cv.clone()
            ctx.dist.all_reduce(cmax, op=ctx.dist.ReduceOp.MAX)
            ctx.dist.all_reduce(cmin, op=ctx.dist.ReduceOp.MIN)
            cv_all_same = bool(torch.equal(cmax, cmin))  # the constant column is replicated
        glog = log[:Kp].cpu().numpy().copy()
        gpiv, gst = int(npv.item()), int(st.item())
        if rank == 0:
            import oracle
            oracle.build()
            t0 = time.perf_counter()
            o0 = oracle.simplex_state(A_full, b, c, max_pivots=0, threads=0)
            t_setup = time.perf_counter() - t0
            t0 = time.perf_counter()
            o = oracle.simplex_state(A_full, b, c, max_pivots=Kp, threads=0)
            t_run = time.perf_counter() - t0
            del o0
            want_log = np.array(o["pivots"], dtype=np.int32).reshape(-1, 2)
            k = o["npivots"]
            g_ov, g_cv = ovv.cpu().numpy(), cv.cpu().numpy()
            parity = {
                "pivots_checked": int(k),
                "pivot_sequence_equal": bool(gpiv == k and gst == o["status"] and np.array_equal(glog[:k], want_log)),
                "timed_run_prefix_equal": bool(np.array_equal(log_full[:k], want_log[:min(k, len(log_full))])),
                "objective_row_equal": bool(np.array_equal(g_ov, o["obj_by_var"])),
                "objective_row_nonzeros": int(np.count_nonzero(g_ov)),
                "objective_constant_equal": bool(np.float64(obj_const).view(np.int64) ==
                                                 np.float64(o["obj_const"]).view(np.int64)),
                "constant_column_equal": bool(np.array_equal(g_cv, o["const_by_var"])) and cv_all_same,
                "constant_column_crc": zlib.crc32(np.ascontiguousarray(g_cv + 0.0).tobytes()),
                "oracle_constant_column_crc": zlib.crc32(np.ascontiguousarray(o["const_by_var"] + 0.0).tobytes()),
                "oracle": "oracle/simplex_oracle.c lp_oracle_simplex_state, rows of each pivot swept on all host "
                          "threads (bit-identical for any thread count), %d x %d tableau" % (n + 1, m + 1),
            }
            parity["bit_exact_vs_oracle"] = bool(parity["pivot_sequence_equal"] and parity["objective_row_equal"]
                                                 and parity["objective_constant_equal"]
                                                 and parity["constant_column_equal"])
            cpu_ms = (t_run - t_setup) / max(k, 1) * 1e3
            out["cpu_baseline"] = {"value": cpu_ms, "unit": "ms/pivot", "cores": os.cpu_count(), "kind": "port",
                                   "sample": "first %d pivots of the same LP on the C oracle, all host threads "
                                             "(set-up %.1f s subtracted)" % (k, t_setup)}
    out["gpu_launches"] = launches[0]
    del A_full, dA, ws
    return out, parity, comm


# ---- N > 1: soak of the in-kernel exchange -- many small LPs back to back on one communicator ----
def bench_soak(ctx, args, comm):
    torch = ctx.torch
    from linear_programming_b200 import workloads
    from linear_programming_b200.sharded import column_shard
    import oracle
    L, dev, rank, world = ctx.L, ctx.dev, ctx.rank, ctx.world
    cntlp, n, m, cap = args.soak_lps, 100, 100, 20000
    A, b, c = workloads.simplex_batch("D2", cntlp, n, m, seed=9)
    lo, cnt = column_shard(m, world, rank)
    dA = torch.from_numpy(np.ascontiguousarray(A[:, :, lo:lo + cnt])).to(dev) if cnt else None
    db, dc = torch.from_numpy(b).to(dev), torch.from_numpy(c).to(dev)
    wsb = int(L.lpb200_simplex_sharded_workspace_bytes(n, m, world))
    ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
    st = torch.zeros(cntlp, dtype=torch.int32, device=dev)
    npv = torch.zeros(cntlp, dtype=torch.int32, device=dev)
    x = torch.zeros((cntlp, m), dtype=torch.float64, device=dev)
    obj = torch.zeros(cntlp, dtype=torch.float64, device=dev)

    def all_lps():
        for i in range(cntlp):
            ctx.check(L.lpb200_simplex_sharded_ex_device(
                comm._comm, n, m, dA[i].data_ptr() if cnt else None, cnt if cnt else 0, db[i].data_ptr(),
                dc[i].data_ptr(), 1e-9, cap, 0, st[i:].data_ptr(), x[i].data_ptr(), obj[i:].data_ptr(),
                npv[i:].data_ptr(), None, 0, ws.data_ptr(), wsb, 30.0, ctx.stream))

    ms, win = ctx.timed(all_lps)
    ost, ox, oobj, onp, _ = oracle.simplex_batch(A, b, c, max_pivots=cap, threads=0)
    ok = (np.array_equal(st.cpu().numpy(), ost) and np.array_equal(npv.cpu().numpy(), onp)
          and np.array_equal(x.cpu().numpy().view(np.int64), ox.view(np.int64))
          and np.array_equal(obj.cpu().numpy().view(np.int64), oobj.view(np.int64)))
    all_ok = ctx.min_over_ranks(1.0 if ok else 0.0) == 1.0
    piv = int(onp.sum())
    return {"lps": cntlp, "shape": "%dx%d D2" % (n, m), "pivots": piv, "ms": ms,
            "us_per_pivot": ms * 1e3 / max(piv, 1), "every_rank_bit_exact_vs_oracle": bool(all_ok),
            "back_to_back_solves_on_one_communicator": cntlp, "clocks": ctx.sampler.window(*win),
            "note": "each pivot = key all-to-all + pivot-column broadcast between %d processes; the tableau slice of "
                    "a rank is a few KB, so this is the latency of the exchange, not bandwidth" % world}


# ---- configs[2]: batched Gauss-Jordan ----
def bench_gj(ctx, args):
    torch = ctx.torch
    L, dev, rank = ctx.L, ctx.dev, ctx.rank
    B, n = args.gj_batch, args.gj_n
    g = torch.Generator(device=dev)
    g.manual_seed(20260 + rank)
    a = torch.rand((B, n, n), generator=g, device=dev, dtype=torch.float64)
    a.mul_(2.0).sub_(1.0)
    a.diagonal(dim1=1, dim2=2).add_(float(n))
    bb = torch.rand((B, n), generator=g, device=dev, dtype=torch.float64)
    bb.mul_(2.0).sub_(1.0)
    planted = 0
    if n >= 2:  # exact zeros on the diagonal of ~1 % of the systems: the reference's row swap (energy_values.cc:100-110)
        plant = (torch.rand(B, generator=g, device=dev) < 0.01).nonzero().flatten().tolist()
        ks = torch.randint(0, n - 1, (B,), generator=g, device=dev).tolist()
        for s in plant:
            k = ks[s]
            rows = a[s, [k + 1, k]].clone()
            a[s, k] = rows[0]
            a[s, k + 1] = rows[1]
            a[s, k, k] = 0.0
        planted = len(plant)
    x = torch.empty_like(bb)
    wsb = int(L.lpb200_gauss_jordan_workspace_bytes(B, n))
    ws = torch.empty(max(wsb, 1), dtype=torch.uint8, device=dev)

    def launch():
        ctx.check(L.lpb200_gauss_jordan_batch_device(B, n, a.data_ptr(), bb.data_ptr(), x.data_ptr(),
                                                     ws.data_ptr(), wsb, ctx.stream))
    launch()
    launch()
    launch()
    best, windows = None, []
    for _ in range(3):
        ms, win = ctx.timed(launch)
        windows.append(win)
        best = ms if best is None else min(best, ms)
    ms = best
    total = ctx.sum_over_ranks(float(B))
    peak, peak_src = measured_peak()
    out = {"workload": "configs[2]: batched Gauss-Jordan, %d systems %dx%d fp64 per GPU (a = U(-1,1) + n*I, "
                       "%d systems with a planted zero pivot)" % (B, n, n, planted),
           "kernel": "gauss_jordan_batched_kernel", "ms": ms, "systems_per_s": total / (ms * 1e-3),
           "n_gpus": ctx.world, "scaling": "weak (independent systems, no collective)",
           "effective_GBps_right_looking": 8.0 * n * n * (n + 3) * B / (ms * 1e-3) / 1e9,
           "effective_GBps_full_sweep": 16.0 * n * n * (n + 1) * B / (ms * 1e-3) / 1e9,
           "fp64_ops_per_s": 2.0 * n * n * (n + 1) * B / (ms * 1e-3),
           "compulsory_hbm_bytes": (2.0 * n * n + 3.0 * n) * 8 * B,
           "clocks": merge_clocks(ctx.sampler, windows), "gpu_launches": 6, "hbm_peak": peak}
    facts = ncu_facts("gauss_jordan_batched_kernel")
    if facts and not facts.get("stale"):
        out["dram_bytes_per_system"] = facts.get("dram_bytes_per_system")
        out["fp64_pipe_pct"] = facts.get("fp64_pipe_pct")
        out["ncu_source"] = facts.get("file")
        if facts.get("dram_bytes_per_system"):
            out["dram_GBps"] = facts["dram_bytes_per_system"] * B / (ms * 1e-3) / 1e9
            out["dram_frac_of_hbm_peak"] = out["dram_GBps"] / peak
    else:
        out["dram_bytes_per_system"] = None
    if rank == 0 and not args.no_cpu_baseline:
        import oracle
        oracle.build()
        k = B if args.gj_check < 0 else min(B, args.gj_check)
        if k > 0:
            ha, hb = a[:k].cpu().numpy(), bb[:k].cpu().numpy()
            t0 = time.perf_counter()
            want, used = oracle.gauss_jordan_batch(ha, hb, threads=0)
            dt = time.perf_counter() - t0
            got = x[:k].cpu().numpy()
            out["bit_exact_vs_oracle"] = bool(np.array_equal(got.view(np.int64), want.view(np.int64)))
            out["checked_systems"] = int(k)
            out["finite_solutions"] = int(np.isfinite(got).all(axis=1).sum())
            ks = min(k, 64)
            t0 = time.perf_counter()
            oracle.gauss_jordan_batch(ha[:ks], hb[:ks], full_sweep=True, threads=0)
            dts = time.perf_counter() - t0
            out["cpu_baseline"] = {"value": ks / dts, "unit": "systems/s", "cores": used, "kind": "port",
                                   "sample": "%d systems, full sweep like energy_values.cc:146-158, all host "
                                             "threads; the right-looking checker did %d systems in %.1f s"
                                             % (ks, k, dt)}
            del ha, hb
    del a, ws
    return out


def ours_arm(args):
    ctx = Ctx()
    torch = ctx.torch
    # nvidia-smi is started BEFORE the warm-up (its NVML start-up can stall a launch by a few ms)
    # and keeps sampling through every timed region; only the samples inside them are reported
    ctx.sampler.start()
    res = bench_batch(ctx, args)
    torch.cuda.empty_cache()
    single = parity_single = soak = gj = None
    comm = None
    if not args.no_single:
        single, parity_single, comm = bench_single(ctx, args)
        torch.cuda.empty_cache()
        if comm is not None and args.soak_lps > 0:
            soak = bench_soak(ctx, args, comm)
        if comm is not None:
            ctx.barrier()
            comm.close()
    if not args.no_gj:
        gj = bench_gj(ctx, args)
    ctx.sampler.close()
    if ctx.rank == 0:
        K = args.steps
        launches = K
        line = {
            "metric": METRIC, "value": res["value"], "unit": UNIT, "n_gpus": ctx.world, "steps": K,
            "warmup": max(args.warmup, 3), "ms_per_step": res["ms_per_step"], "higher_is_better": True,
            "scaling": args.scaling, "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config_dict(args, args.batch, ctx.world),
            "pivots_per_s": res["pivots_per_s"], "clocks": res["clocks"],
            "roofline": res["roofline"], "cpu_baseline": res["cpu_baseline"], "e2e": res["e2e"],
            "gpu_launches": launches, "e2e_compact": res["e2e_compact"], "parity": res["parity"],
            "strong_scaling": res["strong_scaling"],
            "roofline_single": single, "parity_single": parity_single, "sharded_soak": soak,
            "sharded_parity": (bool(parity_single["bit_exact_vs_oracle"]) and
                               (soak is None or soak["every_rank_bit_exact_vs_oracle"]))
                              if (parity_single and ctx.world > 1) else None,
            "gauss_jordan": gj, "extra": res["extra"],
        }
        print(json.dumps(line), flush=True)
    if ctx.world > 1:
        ctx.dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        reference_arm(args)
    else:
        ours_arm(args)


if __name__ == "__main__":
    main()
