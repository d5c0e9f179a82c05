#!/usr/bin/env python
"""bench.py — image pairs matched/s and G descriptor-comparisons/s of the pairwise matching stage.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--workload c3|c5|c1|c4|c4cap|c1real] [--mode knn2_ratio|crosscheck]
                    [--variant auto|popc|popc_csa|tensor] [--no-sub]

One step = one pass of the hot path over the whole pair list of the workload (default C3:
synthetic 21 images x 50 000 ORB descriptors, all 210 pairs, kNN k=2 + ratio test — the
configuration BASELINE.json quotes its metric and its >=100x target on).  Prints ONE JSON line.

  value     pairs/s with the banks already resident in HBM (CUDA events, max over ranks); a 256 MiB
            device memset between steps flushes the L2
  e2e       the same through the public API with HOST buffers: pinned host bank -> device (+ NCCL
            broadcast at N>1), match, results gathered to rank 0 and copied back to the host, every
            step, over the SAME number of steps and with the same L2 flush as `value`
  roofline  dominant kernel against the measured int8 tensor-pipe ceiling (profiles/int8_ceiling_*.json)
  cpu_baseline  cv2.BFMatcher (the reference's own CPU call) on a bounded sample, rank 0, N=1; its
            matches are compared with the GPU's for the same pairs ("parity_in_run")
  sub       (default invocation only) the other configurations the path is judged on, each with its
            own value / e2e / kernel_ms / roofline: C3 cross-check (the reference's LIVE call), C5
            (19 900 pairs x 20 k, the pair-sharding configuration), C4 and C4 at the 15 000 cap
            (star polygons, kernels 2 + 3)

--impl reference times the reference's CPU implementation of the path (cv2.BFMatcher with all
host threads, custom_bruteforce_matcher.py:160-172 / :193-199) on a bounded sample per step.
"""
import argparse
import glob
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    "c3": dict(n_images=21, n_desc=50_000, seed=0,
               name="synthetic 21 images x 50k ORB 256-bit descriptors, all 210 pairs"),
    "c5": dict(n_images=200, n_desc=20_000, seed=1,
               name="synthetic 200-image sweep x 20k descriptors, 19900 pairs"),
    "c1": dict(n_images=21, n_desc=1_000, seed=0,
               name="synthetic stand-in for example_01b: 21 images x 1000 descriptors, 210 pairs"),
    # whole matcher stage of example_01b through the reference-facing class; see run_stage
    "c1real": dict(stage=True, name="example_01b: ORB(1000) descriptors of the 21 forest JPEGs (frozen fixture), "
                                    "CustomBruteForceMatcher.match_n_features, all 210 pairs incl. RANSAC + MatchesInfo"),
    # star-polygon path (kernels 2 + 3); see polygon_record
    "c4": dict(n_images=21, n_stars=2_000, seed=2, polygon=True,
               name="synthetic star fields 21 images x 2k stars, f=1135 px, 894x1342, 5-gon polygon matching, all 210 pairs"),
    "c4cap": dict(n_images=21, n_polygons=15_000, seed=3, polygon=True,
                  name="polygon banks at the reference's cap: 21 images x 15000 polygons, all 210 pairs (kernel 3 only)"),
}
FLOP_PER_EVALUATION = 45         # 15 sub + 15 mul + 15 add per polygon-pair evaluation
IMG_W, IMG_H, FOCAL = 1342, 894, 1135.0
POPC_PER_COMPARISON = 8          # algorithmic: 8 x (32-bit XOR + POPC) per 256-bit distance
OPS_PER_COMPARISON = 512         # tensor variant: 256 int8 multiply-adds = 512 ops per distance
I8_OPS_PER_CLK_PER_SM = 16384    # tcgen05 kind::i8: 8192 MAC/clk/SM (M128 N256 K32 issues every 128 clk)
BYTES_PER_ROW = 32
FLUSH_BYTES = 256 << 20          # > 126 MB L2
L2_NOTE = "256 MiB device memset between steps (inside the timed region) flushes the 126 MB L2"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c3", choices=list(WORKLOADS))
    ap.add_argument("--mode", default="knn2_ratio", choices=["knn2_ratio", "crosscheck"])
    ap.add_argument("--variant", default="auto")
    ap.add_argument("--cpu-sample-pairs", type=int, default=10)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-sub", action="store_true", help="default invocation: skip the sub-records")
    ap.add_argument("--sub", default="crosscheck,c5,c4,c4cap", help="which sub-records the default invocation adds")
    ap.add_argument("--sub-steps", type=int, default=3)
    ap.add_argument("--fp32", action="store_true", help="polygon workloads: non-parity fp32 compare")
    ap.add_argument("--ransac", default="opencv", choices=["opencv", "device"],
                    help="c1real: cv.findHomography per pair (reference, parity) or the batched device RANSAC (row N3)")
    return ap.parse_args()


# ---- clocks during the timed region -------------------------------------------------------------
class ClockSampler:
    """Samples SM clock / throttle reasons of one GPU every 100 ms through NVML."""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown",
               0x4: "sw_power_cap", 0x80: "hw_power_brake_slowdown"}

    def __init__(self, index=0):
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _loop(self):
        while not self._stop.is_set():
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                get = getattr(self.nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
                    self.nv.nvmlDeviceGetCurrentClocksThrottleReasons
                mask = get(self.h)
                for bit, name in self.REASONS.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.1)

    def __enter__(self):
        if self.nv is not None:
            self._thread = threading.Thread(target=self._loop, daemon=True)
            self._thread.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._thread:
            self._thread.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                    "note": "NVML unavailable"}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ---- one process per GPU ----------------------------------------------------------------------------
class Runtime:
    """Rank / device / process-group state shared by every record of one bench invocation."""

    def __init__(self):
        import torch
        self.torch = torch
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        self.dist = None
        torch.cuda.set_device(self.local_rank)
        self.dev = torch.device("cuda", self.local_rank)
        if self.world > 1:
            import torch.distributed as dist
            import datetime
            # a rank that dies inside a sub-record must not leave the others waiting for ten minutes
            dist.init_process_group("nccl", device_id=self.dev, timeout=datetime.timedelta(seconds=240))
            self.dist = dist
        # ONE stream for everything: torch ops, NCCL's stream waits and the engine's kernels (a dedicated
        # stream has a non-zero handle; torch's default stream reads as 0 = "engine's own stream" in the C ABI)
        self.stream = torch.cuda.Stream(self.dev)
        torch.cuda.set_stream(self.stream)
        self.flush = torch.empty(FLUSH_BYTES, dtype=torch.uint8, device=self.dev)

    def engine(self):
        import ssp_b200
        eng = ssp_b200.Engine(self.local_rank)
        eng.set_stream(self.stream.cuda_stream)
        return eng

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize(self.dev)

    def max_over_ranks(self, x):
        if self.dist is None:
            return float(x)
        t = self.torch.tensor([x], dtype=self.torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(self, xs):
        if self.dist is None:
            return [int(x) for x in xs]
        t = self.torch.tensor(list(xs), dtype=self.torch.int64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return [int(v) for v in t.tolist()]

    def close(self):
        if self.dist is not None:
            self.dist.barrier()
            self.dist.destroy_process_group()


def timed_steps(rt, step, steps, warmup):
    """W untimed steps, then EXACTLY `steps` steps between CUDA events on the engine's stream,
    barrier + synchronize on both sides, max over ranks.  Returns (ms per step, clocks, last result)."""
    out = None
    for _ in range(warmup):
        out = step()
    rt.barrier()
    with ClockSampler(rt.local_rank) as clocks:
        ev0, ev1 = rt.torch.cuda.Event(enable_timing=True), rt.torch.cuda.Event(enable_timing=True)
        ev0.record(rt.stream)
        for _ in range(steps):
            out = step()
        ev1.record(rt.stream)
        rt.barrier()
    return rt.max_over_ranks(ev0.elapsed_time(ev1)) / steps, clocks.summary(), out


def timed_e2e(rt, step, steps, warmup):
    """The end-to-end leg: host buffers in, host buffers out, wall clock between barriers (host
    work is part of it), max over ranks, same step count and L2 treatment as the resident leg."""
    out = None
    for _ in range(max(1, min(warmup, 2))):
        out = step()
    rt.barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        out = step()
    rt.barrier()
    return rt.max_over_ranks(time.perf_counter() - t0) / steps, out


# ---- CPU reference arm ----------------------------------------------------------------------------
def cv_match_pair(cv, q, t, mode, keep=False):
    """The reference's own CPU call for one pair; returns the number of surviving matches (and, with
    keep=True, the survivors as an int array (n, 3) {queryIdx, trainIdx, distance})."""
    if mode == "crosscheck":
        ms = cv.BFMatcher(cv.NORM_HAMMING, crossCheck=True).match(cv.UMat(q), cv.UMat(t))
    else:
        ms = [m for m, n in cv.BFMatcher(cv.NORM_HAMMING).knnMatch(q, t, k=2) if m.distance < 0.75 * n.distance]
    if not keep:
        return len(ms)
    return np.array([[m.queryIdx, m.trainIdx, int(m.distance)] for m in ms], dtype=np.int32).reshape(-1, 3)


def cpu_sample(sets, pairs, mode, n_sample):
    """cv2.BFMatcher on `n_sample` evenly spaced pairs of the workload; keeps the matches so that the
    caller can compare them with the GPU's for the same pairs."""
    import cv2 as cv
    cv.setNumThreads(os.cpu_count() or 1)
    n_sample = max(1, min(n_sample, len(pairs)))
    stride = max(1, len(pairs) // n_sample)
    sample = [pairs[i * stride] for i in range(n_sample)]
    matches, cmps, dt = [], 0, 0.0
    for a, b in sample:
        t0 = time.perf_counter()
        n = cv_match_pair(cv, sets[a], sets[b], mode)           # timed: the call + the ratio loop only
        dt += time.perf_counter() - t0
        matches.append(cv_match_pair(cv, sets[a], sets[b], mode, keep=True))   # untimed: for the parity check
        assert len(matches[-1]) == n
        cmps += sets[a].shape[0] * sets[b].shape[0]
    return {"pairs": sample, "matches": matches, "seconds": dt, "pairs_per_s": len(sample) / dt,
            "gcmp_per_s": cmps / dt / 1e9, "cores": cv.getNumThreads(), "cv2": cv.__version__}


def run_reference(args, wl, sets, pairs):
    """--impl reference: cv2.BFMatcher on a bounded sample per step (one pair per step)."""
    import cv2 as cv
    cv.setNumThreads(os.cpu_count() or 1)
    rng = np.random.default_rng(0)
    order = rng.permutation(len(pairs))
    per_step = 1
    k = 0

    def step():
        nonlocal k
        cmps = 0
        for _ in range(per_step):
            a, b = pairs[order[k % len(pairs)]]
            k += 1
            cv_match_pair(cv, sets[a], sets[b], args.mode)
            cmps += sets[a].shape[0] * sets[b].shape[0]
        return cmps

    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    cmps = sum(step() for _ in range(args.steps))
    dt = time.perf_counter() - t0
    value = args.steps * per_step / dt
    sample = (f"cv2 {cv.__version__} BFMatcher {'crossCheck match' if args.mode == 'crosscheck' else 'knnMatch(k=2)+ratio 0.75'}, "
              f"{per_step} of {len(pairs)} pairs per step, {cv.getNumThreads()} threads")
    return {
        "impl": "reference", "metric": "image_pairs_matched_per_sec", "value": value, "unit": "pairs/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u8 (256-bit Hamming, OpenCV SIMD popcount)", "data": "synthetic",
        "config": {"workload": wl["name"], "mode": args.mode, "pairs": len(pairs)},
        "gcmp_per_s": cmps / dt / 1e9,
        "cpu_baseline": {"value": value, "unit": "pairs/s", "cores": cv.getNumThreads(), "kind": "reference",
                         "sample": sample},
        "e2e": {"value": value, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }


# ---- peaks ------------------------------------------------------------------------------------------
def _latest_profile(pattern):
    found = sorted(glob.glob(os.path.join(ROOT, "profiles", pattern)))
    return found[-1] if found else None


def popc_peak_per_clk_per_sm():
    """POPC issue rate measured on this pool's B200 by tools/pipe_microbench (committed summary),
    else the CUDA programming-guide planning figure."""
    try:
        with open(_latest_profile("pipe_microbench_r*.json")) as f:
            return float(json.load(f)["popc"]), "measured (profiles/pipe_microbench_r01.json)"
    except Exception:
        return 16.0, "planning figure (CUDA programming guide, 16 POPC/clk/SM)"


def measured_int8_ceiling(sm_count, sm_max_mhz):
    """Dense int8 ceiling of THIS kernel's tensor path in TOP/s: the MMA-only probe (the product
    kernel with the epilogue compiled out: TMA feed + tcgen05.mma issue, nothing else) timed on this
    pool's B200 with its clock record, committed under profiles/ (tools/int8_ceiling.sh).  Fallback:
    the physical issue rate of the pipe at the maximum SM clock."""
    path = _latest_profile("int8_ceiling_r*.json")
    try:
        with open(path) as f:
            j = json.load(f)
        return float(j["tops"]), (f"measured: MMA-only probe of the same kernel, {j['kernel_ms']:.2f} ms per C3 launch at "
                                  f"{j.get('sm_mhz')} MHz ({os.path.relpath(path, ROOT)})"), "measured"
    except Exception:
        peak = sm_count * I8_OPS_PER_CLK_PER_SM * sm_max_mhz * 1e6 / 1e12
        return peak, f"fallback: {sm_count} SMs x {I8_OPS_PER_CLK_PER_SM} int8 ops/clk/SM x {sm_max_mhz:g} MHz (no committed probe)", "fallback"


def bf16_proxy_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            j = json.load(f)
        return {"two_x_cublas_bf16_burst": 2.0 * float(j["bf16_tflops"]),
                "two_x_cublas_bf16_sustained": 2.0 * float(j.get("bf16_tflops_sustained", j["bf16_tflops"])),
                "source": "MEASURED_PEAKS.json (cuBLAS bf16 is power-capped near 1.3 GHz on this pool: a proxy, not the ceiling)"}
    except Exception:
        return None


def ncu_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of the dominant kernel on this
    workload, from the committed `ncu --set full` capture (profiles/traffic_r*.json)."""
    try:
        path = _latest_profile("traffic_r*.json")
        with open(path) as f:
            t = json.load(f)[kernel]
        return float(t["dram_bytes"]), t["source"]
    except Exception:
        return None, None


def measured_bf16_burst():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["bf16_tflops"])
    except Exception:
        return 1650.0


def measured_hbm_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


# ---- kernel 1 workloads ---------------------------------------------------------------------------
def hamming_record(rt, wl, sets, pairs, mode, variant, steps, warmup, cpu_pairs):
    """value / e2e / roofline / cpu_baseline of one Hamming workload in one mode.  Returns the record
    on rank 0 (None elsewhere)."""
    import ssp_b200
    from ssp_b200.distributed import (shard_pairs_contiguous, shard_bounds, broadcast_descriptor_bank,
                                      gather_matches_resident)
    torch, world, rank = rt.torch, rt.world, rt.rank
    eng = rt.engine()

    # host-side bank in pinned memory (rank 0 owns the input, like the reference's single process)
    bank = offs = pinned = None
    offs_known = np.zeros(len(sets) + 1, dtype=np.int64)
    np.cumsum([s.shape[0] for s in sets], out=offs_known[1:])
    if rank == 0:
        bank_np, offs = ssp_b200.Engine.pack_descriptors(sets)
        pinned = torch.from_numpy(bank_np).pin_memory()
        bank = pinned.numpy()
    bank_bytes = int(offs_known[-1]) * BYTES_PER_ROW
    cost = [int(sets[a].shape[0]) * int(sets[b].shape[0]) for a, b in pairs]
    local_pairs, lo, hi = shard_pairs_contiguous(pairs, rank, world, cost)
    max_block = int(np.diff(shard_bounds(len(pairs), world, cost)).max())   # every rank can compute it

    # ---- resident bank for the `value` leg ----------------------------------------------------
    if world > 1:
        bank_view, offs_all = broadcast_descriptor_bank(bank, offs, 0)
        bank_t = bank_view.clone()          # the receive buffer is reused by the e2e leg
    else:
        bank_t, offs_all = pinned.to(rt.dev), offs
    eng.bind_descriptor_bank(bank_t.data_ptr(), offs_all, keepalive=bank_t)
    kernel_ms, launches = [], [0]

    def step_resident():
        rt.flush.zero_()
        res = eng.hamming_match_pairs(local_pairs, mode=mode, variant=variant, fetch=False)
        st = eng.hamming_last_stats()
        kernel_ms.append(st["main_kernel_ms"])
        launches[0] += st["launches"]
        return res

    ms_per_step, clk, res = timed_steps(rt, step_resident, steps, warmup)
    kernel_ms = kernel_ms[warmup:]
    st = eng.hamming_last_stats()
    launches_timed = st["launches"] * steps      # the same launches every step (bank stays expanded)
    comparisons_local = st["comparisons"]
    comparisons, n_matches, launches_all = rt.sum_over_ranks([comparisons_local, int(res.offsets[-1]), launches_timed])
    value = len(pairs) / (ms_per_step / 1e3)

    # ---- e2e leg: host buffers in, host buffers out, every step -------------------------------
    def step_e2e():
        rt.flush.zero_()
        if world > 1:
            b_t, o_all = broadcast_descriptor_bank(bank, offs, 0)       # H2D on rank 0 + ONE NCCL broadcast
            eng.bind_descriptor_bank(b_t.data_ptr(), o_all, keepalive=b_t)
            local = eng.hamming_match_pairs(local_pairs, mode=mode, variant=variant, fetch=False)
            # survivors: device -> exact-length NCCL transfers into rank 0's arrays -> pinned host memory
            return gather_matches_resident(eng, local.offsets, dst=0, max_block=max_block)
        eng.upload_descriptor_bank(bank, offs)                           # pinned host -> device
        return eng.hamming_match_pairs(local_pairs, mode=mode, variant=variant, pinned=True)  # D2H of survivors

    e2e_s, out = timed_e2e(rt, step_e2e, steps, warmup)
    e2e_value = len(pairs) / e2e_s
    d2h = (int(out.offsets[-1]) * 12 + (len(pairs) + 1) * 8) if rank == 0 else 0
    if rank == 0:
        assert int(out.offsets[-1]) == n_matches, "e2e and resident legs disagree on the survivor count"

    # one more end-to-end step with a device synchronisation after every piece: where its time goes
    # (max over ranks per piece; the pieces overlap a little less than in the timed steps)
    timeline = {}

    def piece(name, fn):
        rt.barrier()
        t0 = time.perf_counter()
        r = fn()
        torch.cuda.synchronize(rt.dev)
        timeline[name] = rt.max_over_ranks((time.perf_counter() - t0) * 1e3)
        return r

    if world > 1:
        b_t, o_all = piece("h2d_on_rank0+broadcast", lambda: broadcast_descriptor_bank(bank, offs, 0))
        piece("bind", lambda: eng.bind_descriptor_bank(b_t.data_ptr(), o_all, keepalive=b_t))
        local = piece("expand+kernel1+finalize", lambda: eng.hamming_match_pairs(local_pairs, mode=mode, variant=variant, fetch=False))
        piece("gather_to_rank0+d2h", lambda: gather_matches_resident(eng, local.offsets, dst=0, max_block=max_block))
    else:
        piece("h2d", lambda: eng.upload_descriptor_bank(bank, offs))
        local = piece("expand+kernel1+finalize", lambda: eng.hamming_match_pairs(local_pairs, mode=mode, variant=variant, fetch=False))
        n_loc = int(local.offsets[-1])
        bufs = [eng.pinned_array(k, n_loc, np.int32) for k in ("hq", "ht", "hd")]
        piece("d2h", lambda: eng.hamming_fetch_into(*(b.ctypes.data for b in bufs), n_loc, device=False))

    # ---- roofline of the dominant kernel -------------------------------------------------------
    info = eng.device_info()
    main_ms = float(np.mean(kernel_ms)) if kernel_ms else float("nan")
    gcmp = comparisons_local / (main_ms / 1e3) / 1e9 if main_ms > 0 else 0.0   # Gcmp/s, this rank
    sm_max_mhz = clk.get("sm_max_mhz") or info["sm_clock_khz"] / 1e3
    held = (clk["sm_mhz"] / sm_max_mhz) if clk.get("sm_mhz") else None
    hbm_peak, hbm_src = measured_hbm_peak()
    rows_q = sum(sets[a].shape[0] for a, _ in local_pairs)
    rows_t = sum(sets[b].shape[0] for _, b in local_pairs)
    tensor = variant in ("auto", "tensor")
    if tensor:
        # algorithmic work: one 256-term int8 dot product per comparison = 512 integer ops, counted
        # ONCE per image pair also in cross-check mode (one distance serves both directions)
        achieved = gcmp * OPS_PER_COMPARISON / 1e3                      # TOP/s
        peak, peak_src, peak_kind = measured_int8_ceiling(info["sm_count"], sm_max_mhz)
        pipe_peak = info["sm_count"] * I8_OPS_PER_CLK_PER_SM * sm_max_mhz * 1e6 / 1e12
        algo_bytes = (rows_q + rows_t) * 256 + rows_q * 16              # expanded operands once + best-2 records
        roofline = {
            "bound": "tensor", "kernel": "hamming_tcgen05_kernel", "achieved": achieved, "peak": peak,
            "unit": "TFLOP/s", "frac": achieved / peak if peak else None, "of": peak_kind,
            "ops": "int8 multiply-add counted as 2 ops; s8 x s8 -> s32, exact",
            "peak_definition": peak_src,
            "pipe_peak_at_max_clock": pipe_peak,
            "pipe_peak_definition": f"{info['sm_count']} SMs x {I8_OPS_PER_CLK_PER_SM} int8 ops/clk/SM (tcgen05 kind::i8 "
                                    f"M128 N256 K32 = 128 clk) x {sm_max_mhz:g} MHz",
            "frac_of_pipe_at_max_clock": achieved / pipe_peak,
            "frac_of_pipe_at_held_clock": achieved / (pipe_peak * held) if held else None,
            "other_denominators": bf16_proxy_peak(),
            "gcmp_per_s": gcmp,
        }
    else:
        popc_rate, popc_src = popc_peak_per_clk_per_sm()
        popc_per_cmp = 4 if variant == "popc_csa" else POPC_PER_COMPARISON
        peak = info["sm_count"] * popc_rate * sm_max_mhz * 1e6 / POPC_PER_COMPARISON / 1e9
        algo_bytes = (rows_q + rows_t) * BYTES_PER_ROW + rows_q * 16    # each operand once + best-2 records
        roofline = {
            "bound": "int-pipe (POPC)", "kernel": "hamming_popc_kernel", "achieved": gcmp, "peak": peak,
            "unit": "Gcmp/s", "frac": gcmp / peak if peak else None,
            "peak_definition": f"{info['sm_count']} SMs x {popc_rate:g} POPC/clk/SM [{popc_src}] x {sm_max_mhz:g} MHz / "
                               f"{POPC_PER_COMPARISON} POPC per 256-bit comparison (the CSA variant issues {popc_per_cmp})",
            "frac_at_held_clock": gcmp / (peak * held) if held else None,
        }
    traffic, traffic_src = ncu_traffic(roofline["kernel"] + ("_crosscheck" if mode == "crosscheck" else ""))
    roofline.update({
        "kernel_ms_per_launch": main_ms, "comparisons_per_launch": comparisons_local,
        "traffic": traffic, "traffic_source": traffic_src,
        "hbm": {"achieved": algo_bytes / (main_ms / 1e3) / 1e9 if main_ms > 0 else None, "peak": hbm_peak,
                "unit": "GB/s", "algorithmic_bytes_per_launch": algo_bytes, "peak_source": hbm_src,
                "frac": (algo_bytes / (main_ms / 1e3) / 1e9) / hbm_peak if main_ms > 0 else None},
    })

    rec = None
    if rank == 0:
        rec = {
            "value": value, "unit": "pairs/s", "n_gpus": world, "steps": steps, "warmup": warmup,
            "ms_per_step": ms_per_step, "kernel_ms": main_ms,
            "dtype": "s8 x s8 -> s32 (exact)" if tensor else "u32 popcount",
            "config": {"workload": wl["name"], "mode": mode, "variant": variant, "pairs": len(pairs),
                       "descriptors_per_image": wl["n_desc"],
                       "sharding": f"contiguous cost-balanced blocks of the pair list over {world} rank(s)", "l2": L2_NOTE},
            "gcmp_per_s": comparisons / (ms_per_step / 1e3) / 1e9,
            "matches_per_step": n_matches,
            "e2e": {"value": e2e_value, "unit": "pairs/s", "h2d_bytes_per_step": bank_bytes + len(offs_known) * 8,
                    "d2h_bytes_per_step": d2h, "steps": steps, "ms_per_step": e2e_s * 1e3,
                    "gcmp_per_s": comparisons / e2e_s / 1e9, "timeline_ms": timeline},
            "gpu_launches": launches_all,
            "clocks": clk,
            "roofline": roofline,
        }
        if world == 1 and cpu_pairs > 0:
            cpu = cpu_sample(sets, pairs, mode, cpu_pairs)
            got = eng.hamming_match_pairs(cpu["pairs"], mode=mode, variant=variant)
            same = all(np.array_equal(np.stack(got.pair(k), axis=1), cpu["matches"][k]) for k in range(len(cpu["pairs"])))
            rec["parity_in_run"] = bool(same)
            rec["parity_in_run_detail"] = (f"survivors (queryIdx, trainIdx, distance) of {len(cpu['pairs'])} pairs: GPU step vs "
                                           f"cv2 {cpu['cv2']} BFMatcher, {sum(len(m) for m in cpu['matches'])} matches compared")
            rec["cpu_baseline"] = {
                "value": cpu["pairs_per_s"], "unit": "pairs/s", "cores": cpu["cores"], "kind": "reference",
                "gcmp_per_s": cpu["gcmp_per_s"],
                "sample": f"cv2 {cpu['cv2']} BFMatcher ({mode}) on {len(cpu['pairs'])} of {len(pairs)} pairs "
                          f"({cpu['seconds']:.1f} s), {cpu['cores']} threads; the reference's own CPU call"}
    eng.close()
    del bank_t
    torch.cuda.empty_cache()
    return rec


# ---- star-polygon path (kernels 2 + 3) ------------------------------------------------------------
def polygon_inputs(wl):
    """Host-side inputs of the polygon workloads.  c4: star fields + the reference's own candidate
    reduction (cbpm:261-443, Ward clustering on the host exactly as the reference does it; brightness
    supplied directly).  c4cap: ready-made normalised 15-dim polygon rows at the cap."""
    from ssp_b200 import synthetic as synth
    if "n_polygons" in wl:
        return {"feat": synth.polygon_feature_bank(wl["n_images"], wl["n_polygons"], seed=wl["seed"])}
    import cv2
    import ssp_b200
    xy, br, _ = synth.star_fields(wl["n_images"], wl["n_stars"], seed=wl["seed"], img_w=IMG_W, img_h=IMG_H, focal=FOCAL)
    frame = np.zeros((IMG_H, IMG_W, 3), dtype=np.uint8)       # only .shape is read
    bf = ssp_b200.BFPolygonMatcher("/tmp", "a", "b", 0, 1, frame.shape, FOCAL, write_debug_images=False)
    cands = []
    for i in range(wl["n_images"]):
        kps = tuple(cv2.KeyPoint(float(p[0]), float(p[1]), 4.0) for p in xy[i].astype(np.float32))
        c = bf.reduce_amount_of_keypoints_and_create_polygon_spanning_tuples_of_keypoints(
            kps, f"img{i}", frame, 5, 15000, _brightness=br[i])
        cands.append(np.array(sorted(c), dtype=np.int32).reshape(-1, 5))
    return {"xy": xy, "br": br, "cands": cands}


def polygon_cpu_pair(inp, oracle, k, pairs, state):
    """CPU port of ONE pair's share of the path, as the reference amortises it: every image is
    measured once for all its pairs (polygon_data_store cache, cbpm:744-748), i.e. n_images/n_pairs
    of an image per pair, then the (P1, P2) numpy compare + row argmin + threshold of that pair."""
    a, b = pairs[k % len(pairs)]
    evals = 0
    if "cands" in inp:
        share = len(inp["xy"]) / state["n_pairs_total"]
        c = inp["cands"][a]
        n = max(1, int(round(len(c) * share)))
        lo = (k * n) % max(1, len(c) - n)
        oracle.characteristics(inp["xy"][a], inp["br"][a], c[lo:lo + n])
        f1, f2 = state["feat"][a], state["feat"][b]
    else:
        f1, f2 = inp["feat"][a][:state["rows"]], inp["feat"][b]
    m = oracle.compare(f1, f2)
    oracle.select(m)
    evals += f1.shape[0] * f2.shape[0]
    return evals


def polygon_cpu_state(inp, oracle, pairs, steps, n_pairs_total):
    """Untimed preparation of the CPU leg: the polygon rows the sampled pairs compare."""
    if "cands" not in inp:
        return {"rows": 1500, "n_pairs_total": n_pairs_total}     # 1500 x 15000 rows per step: numpy needs ~1.6 GB of temporaries
    need = sorted({i for k in range(steps) for i in pairs[k % len(pairs)]})
    feat = {}
    for i in need:
        feat[i] = np.ascontiguousarray(oracle.characteristics(inp["xy"][i], inp["br"][i], inp["cands"][i])["feat"])
    return {"feat": feat, "n_pairs_total": n_pairs_total}


def run_polygon_reference(args, wl, inp, pairs):
    from oracle.polygon_oracle import PolygonOracle
    oracle = PolygonOracle((IMG_H, IMG_W, 3), FOCAL)
    sample_pairs = [pairs[0], pairs[len(pairs) // 2]]
    state = polygon_cpu_state(inp, oracle, sample_pairs, args.steps + args.warmup, len(pairs))
    k = 0
    for _ in range(args.warmup):
        polygon_cpu_pair(inp, oracle, k, sample_pairs, state)
        k += 1
    t0 = time.perf_counter()
    evals = 0
    for _ in range(args.steps):
        evals += polygon_cpu_pair(inp, oracle, k, sample_pairs, state)
        k += 1
    dt = time.perf_counter() - t0
    scale = 1.0
    if "cands" not in inp:      # a step compared only state["rows"] of the first image's rows
        scale = state["rows"] / inp["feat"][0].shape[0]
    value = args.steps * scale / dt
    sample = ("CPU port (oracle/polygon_oracle.py: Python measure loop + numpy (P1,P2) compare, the reference's "
              "algorithm, 1 thread) on one pair per step incl. its 1/10-image share of polygon measuring"
              if "cands" in inp else
              f"CPU port numpy compare of {state['rows']} x {inp['feat'][1].shape[0]} polygon rows per step, scaled to a full pair, 1 thread")
    return {
        "impl": "reference", "metric": "image_pairs_matched_per_sec", "value": value, "unit": "pairs/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": wl["name"], "pairs": len(pairs)},
        "polygon_evals_per_s": evals / dt,
        "cpu_baseline": {"value": value, "unit": "pairs/s", "cores": 1, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }


def fp_pipe_peaks():
    """FP64 / FP32 issue rates measured on this pool's B200 by tools/pipe_microbench (committed
    summary, thread-ops per clock per SM); None when the summary has no such entry."""
    try:
        with open(_latest_profile("pipe_microbench_r*.json")) as f:
            j = json.load(f)
        return {k: float(j[k]) for k in ("dfma", "dsub_dmul_dadd_total", "ffma") if k in j}
    except Exception:
        return {}


def polygon_parity_in_run(eng, inp, pairs, measured, max_rows=1500):
    """Kernel 3's answer for one sampled pair against numpy's on the SAME polygon banks (the banks the GPU
    holds, fetched back): argmin and FP64 distance of every row (the first `max_rows` at the cap, where
    numpy's (P1, P2, 5) temporaries would not fit) and the thresholded matches, bit for bit.  Where the
    step measures its polygons (kernel 2), the first 300 candidates of the first image are also measured by
    the CPU port and must agree (same polygons, same vertex order, features within 1e-9)."""
    from oracle.polygon_oracle import PolygonOracle
    a, b = pairs[len(pairs) // 2]
    fa = eng.polygon_fetch(a, raw=False)
    fb = eng.polygon_fetch(b, raw=False)
    n = len(fa["feat"]) if len(fb["feat"]) <= 8000 else min(len(fa["feat"]), max_rows)
    chk = eng.polygon_match_pairs([(a, b)], return_rowmin=True)
    m = PolygonOracle.compare(fa["feat"][:n], fb["feat"])
    same = bool(np.array_equal(chk.argmin[:n], np.argmin(m, axis=1)) and np.array_equal(chk.mindist[:n], m.min(axis=1)))
    rows, cols, d = PolygonOracle.select(m, 0.1566)
    r, c, dd = chk.pair(0)
    keep = r < n
    same = same and bool(np.array_equal(r[keep], rows) and np.array_equal(c[keep], cols) and np.array_equal(dd[keep], d))
    detail = (f"pair ({a}, {b}): argmin, FP64 distance and thresholded matches of {n} rows x {len(fb['feat'])} columns, "
              f"GPU (kernel {eng.polygon_last_stats()['tensor_screen']['used'] and 'poly_screen_tcgen05 + FP64 candidates' or 'CUDA-core'}) "
              f"vs numpy on the same banks: exact")
    if measured:
        oracle = PolygonOracle((IMG_H, IMG_W, 3), FOCAL)
        want = oracle.characteristics(inp["xy"][a], inp["br"][a], inp["cands"][a][:300])
        nv = int(want["valid"].sum())
        same = same and bool(np.array_equal(fa["star_idx"][:nv], want["star_idx"]) and
                             np.allclose(fa["feat"][:nv], want["feat"], rtol=0, atol=1e-9))
        detail += f"; kernel 2: the first {nv} polygons of image {a} vs the CPU port (vertex order equal, features within 1e-9)"
    return same, detail


def polygon_record(rt, wl, inp, pairs, steps, warmup, fp32, parity=False):
    from ssp_b200.distributed import (shard_pairs_contiguous, broadcast_star_bank, measure_polygons_sharded,
                                      gather_polygon_matches_resident, _broadcaster)
    torch, world, rank, dev, dist = rt.torch, rt.world, rt.rank, rt.dev, rt.dist
    eng = rt.engine()
    n_img = wl["n_images"]
    local_pairs, lo, hi = shard_pairs_contiguous(pairs, rank, world)
    measured = "cands" in inp

    # ---- resident inputs ------------------------------------------------------------------------
    if measured:
        if world > 1:
            src = (inp["xy"], inp["br"], inp["cands"]) if rank == 0 else (None, None, None)
            so, co, xy_v, br_v, cand_v = broadcast_star_bank(*src, src=0)
            xy_t, br_t, cand_t = xy_v.clone(), br_v.clone(), cand_v.clone()   # the receive buffer is reused by the e2e leg
        else:
            eng.upload_star_bank(inp["xy"], inp["br"], IMG_W, IMG_H, FOCAL)
            co = np.zeros(n_img + 1, dtype=np.int64)
            np.cumsum([len(c) for c in inp["cands"]], out=co[1:])
            cand_host = torch.from_numpy(np.concatenate(inp["cands"])).pin_memory()
            cand_t = cand_host.to(dev)
        h2d = sum(a.nbytes for a in inp["xy"]) + sum(b.nbytes for b in inp["br"]) + sum(c.nbytes for c in inp["cands"])
    else:
        if rank == 0 or world == 1:
            feat_host = torch.from_numpy(np.concatenate(inp["feat"])).pin_memory()
        fo = np.zeros(n_img + 1, dtype=np.int64)
        np.cumsum([len(f) for f in inp["feat"]], out=fo[1:])
        for i in range(n_img):
            eng.polygon_upload_features(i, inp["feat"][i])
        h2d = sum(f.nbytes for f in inp["feat"])
    h2d += len(pairs) * 8

    def measure_all_local():
        eng.polygon_measure_batch_resident(list(range(n_img)), cand_t.data_ptr(), co)   # ONE launch pair for all images
        return 2

    kernel_ms = []

    def step_resident():
        n = 0
        if measured:
            if world > 1:
                measure_polygons_sharded(eng, so, co, xy_t, br_t, cand_t, IMG_W, IMG_H, FOCAL)
                n += 2 + 1   # star rays + measure batch + compaction
            else:
                n += measure_all_local()
        res = eng.polygon_match_pairs(local_pairs, fp32=fp32, fetch=False)
        st = eng.polygon_last_stats()
        kernel_ms.append(st["main_kernel_ms"])
        return res, n + st["launches"]

    def step_e2e():
        if measured:
            if world > 1:
                src = (inp["xy"], inp["br"], inp["cands"]) if rank == 0 else (None, None, None)
                so2, co2, xy2, br2, cand2 = broadcast_star_bank(*src, src=0)       # H2D on rank 0 + ONE broadcast
                measure_polygons_sharded(eng, so2, co2, xy2, br2, cand2, IMG_W, IMG_H, FOCAL)
            else:
                eng.upload_star_bank(inp["xy"], inp["br"], IMG_W, IMG_H, FOCAL)
                cand_t.copy_(cand_host, non_blocking=True)                           # one H2D for all candidate tables
                measure_all_local()
        else:
            if world > 1:
                arrays = [fo, feat_host.numpy()] if rank == 0 else None
                _, feat_v = _broadcaster("polyfeat", 0).broadcast(arrays, to_host=(0,))
                for i in range(n_img):
                    eng.polygon_upload_features_resident(i, feat_v[int(fo[i]):].data_ptr(), 0, int(fo[i + 1] - fo[i]))
            else:
                for i in range(n_img):
                    eng.polygon_upload_features(i, inp["feat"][i])
        if world > 1:
            local = eng.polygon_match_pairs(local_pairs, fp32=fp32, fetch=False)
            return gather_polygon_matches_resident(eng, local.offsets, dst=0)
        return eng.polygon_match_pairs(local_pairs, fp32=fp32)

    ms_per_step, clk, (res, launches_step) = timed_steps(rt, step_resident, steps, warmup)
    kernel_ms = kernel_ms[warmup:]
    st = eng.polygon_last_stats()
    evals_local = st["evaluations"]
    executed_local = st.get("executed_flops")
    n_polys = [eng.polygon_count(i) for i in range(n_img)]
    evals, n_matches, launches = rt.sum_over_ranks([evals_local, int(res.offsets[-1]), launches_step * steps])
    e2e_s, out = timed_e2e(rt, step_e2e, steps, warmup)
    d2h = (int(out.offsets[-1]) * 16 + (len(pairs) + 1) * 8) if rank == 0 else 0
    if rank == 0:
        assert int(out.offsets[-1]) == n_matches, "e2e and resident legs disagree on the polygon match count"

    info = eng.device_info()
    sm_max_mhz = clk.get("sm_max_mhz") or info["sm_clock_khz"] / 1e3
    main_ms = float(np.mean(kernel_ms))
    rates = fp_pipe_peaks()
    sm_hz = sm_max_mhz * 1e6
    n_sm = info["sm_count"]
    # the kernel prunes exactly (a column whose partial sum already reaches the row's best is dropped):
    # the roofline counts the flops it EXECUTED (counted on the device), never the 45 per evaluation it skipped
    algorithmic = evals_local * FLOP_PER_EVALUATION / (main_ms / 1e3) / 1e12 if main_ms > 0 else None
    achieved = executed_local / (main_ms / 1e3) / 1e12 if (executed_local and main_ms > 0) else None
    f64_rate = rates.get("dsub_dmul_dadd_total")                     # thread-ops per clock per SM, one flop each
    f32_rate = max(rates.get("ffma", 0.0), 128.0)                    # nominal 4 x 32 lanes; the r01 microbench reached less
    peak, peak_def, split = None, "not measured", None
    kernel_name = "poly_compare_f32_kernel" if fp32 else "poly_compare_f64_kernel"
    if fp32:
        # the fp32 mode needs 15 FSUB + 15 FFMA = 45 flop in 30 issue slots
        peak = n_sm * f32_rate * 1.5 * sm_hz / 1e12
        peak_def = f"{n_sm} SMs x {f32_rate:.0f} thread-ops/clk/SM x 1.5 flop per op (15 FSUB + 15 FFMA) x {sm_max_mhz:g} MHz"
    elif f64_rate and executed_local:
        # parity mode: the first-group FILTER of the compacted-pruning kernel runs in single precision
        # (conservative, 10 ops for 14 flop per column), every result-bearing operation in unfused FP64.
        # peak = the rate at which the two pipes, one after the other, would get through exactly this mix
        f32_flops = st.get("executed_flops_fp32") or 0
        f64_flops = executed_local - f32_flops
        screen = f32_flops == 44 * evals_local      # every cell screened on all 15 properties: 15 FADD + FMUL + 14 FFMA
        flop_per_op = 44.0 / 30.0 if screen else 14.0 / 10.0
        kernel_name = "poly_compare_f64_screen_kernel" if screen else "poly_compare_f64_queue_kernel"
        t64 = f64_flops / (n_sm * f64_rate * sm_hz)
        t32 = f32_flops / (n_sm * f32_rate * flop_per_op * sm_hz)
        peak = executed_local / (t64 + t32) / 1e12
        split = {"fp64_flops": f64_flops, "fp32_flops": f32_flops, "fp64_pipe_ms": t64 * 1e3, "fp32_pipe_ms": t32 * 1e3}
        peak_def = (f"executed flops / (FP64 flops / ({n_sm} SMs x {f64_rate:.1f} unfused DSUB+DMUL+DADD per clk per SM, measured: "
                    f"profiles/pipe_microbench_r01.json) + single-precision screen flops / ({n_sm} SMs x {f32_rate:.0f} ops/clk/SM x {flop_per_op:.3f} flop "
                    f"per op)) at {sm_max_mhz:g} MHz: the two arithmetic pipes one after the other")
    ts = st.get("tensor_screen") or {}
    if ts.get("used"):
        # default parity compare: every cell goes through a K = 64 TF32 contraction on the tensor cores
        # (operands split in two TF32 halves, |b|^2 folded in), FP64 only for the candidates
        flop_per_cell = 128.0
        tc_flops = evals_local * flop_per_cell
        achieved = tc_flops / (main_ms / 1e3) / 1e12 if main_ms > 0 else None
        peak = n_sm * 4096.0 * sm_hz / 1e12
        kernel_name = "poly_screen_tcgen05_kernel"
        split = {"tf32_tensor_flops": tc_flops, "fp64_flops": executed_local, "candidates": ts.get("candidates"),
                 "screen_worst_error_coef": ts.get("worst_error_coef"), "screen_bound_coef": ts.get("bound_coef")}
        peak_def = (f"{n_sm} SMs x 4096 flop/clk/SM (tcgen05 kind::tf32 M128 N256 K8 = 128 clk) x {sm_max_mhz:g} MHz; achieved counts the "
                    f"K = 64 contraction the screen executes per cell (128 flop), over the whole compare (operand banks + screen + FP64 "
                    f"evaluation of the candidates); cuBLAS bf16 burst / 2 on this pool = {measured_bf16_burst() / 2:.0f} TFLOP/s (power-capped proxy)")
    roofline = {
        "bound": "fp32-pipe" if fp32 else ("tensor" if ts.get("used") else "fp64-pipe"), "kernel": kernel_name,
        "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak if peak and achieved else None,
        "achieved_counts": "flops executed by the kernel (device counter); pruned terms are not counted",
        "executed_fraction_of_algorithmic_flops": (executed_local / (evals_local * FLOP_PER_EVALUATION)) if executed_local and evals_local else None,
        "algorithmic_tflops_equivalent": algorithmic, "executed_by_precision": split,
        "peak_definition": peak_def,
        "binding_unit": None if fp32 else ("tensor pipe + epilogue issue (profiles/INDEX.md, kernel 3)" if ts.get("used") else
                                           "ncu (profiles/INDEX.md, kernel 3): the screen kernel is bound by the FP32 pipe, the queue kernel by the shared-memory data pipe and instruction issue"),
        "tensor_screen": ts or None,
        "flop_per_evaluation": FLOP_PER_EVALUATION, "kernel_ms_per_launch": main_ms,
        "evaluations_per_launch": evals_local, "evaluations_per_s": evals_local / (main_ms / 1e3) if main_ms > 0 else None,
        "traffic": None,
    }
    if not measured and not fp32 and world == 1:   # the committed ncu capture is of exactly this launch
        roofline["traffic"], roofline["traffic_source"] = ncu_traffic(
            "poly_screen_tcgen05_kernel" if ts.get("used") else "poly_compare_f64_kernel")
    rec = None
    if rank == 0:
        rec = {
            "value": len(pairs) / (ms_per_step / 1e3), "unit": "pairs/s", "n_gpus": world, "steps": steps,
            "warmup": warmup, "ms_per_step": ms_per_step, "kernel_ms": main_ms,
            "dtype": "f32" if fp32 else "f64",
            "config": {"workload": wl["name"], "pairs": len(pairs), "polygons_per_image": [min(n_polys), max(n_polys)],
                       "candidates_per_image": [min(map(len, inp["cands"])), max(map(len, inp["cands"]))] if measured else None,
                       "threshold": 0.1566,
                       "sharding": f"kernel 2 by image, kernel 3 by contiguous blocks of the pair list over {world} rank(s)",
                       "l2": "inputs are far below L2 size by nature of the path (<= 38 MB polygon bank); no flush"},
            "polygon_evals_per_s": evals / (ms_per_step / 1e3), "matches_per_step": n_matches,
            "e2e": {"value": len(pairs) / e2e_s, "unit": "pairs/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "steps": steps, "ms_per_step": e2e_s * 1e3},
            "gpu_launches": launches, "clocks": clk, "roofline": roofline,
        }
    if rec is not None and parity and world == 1 and not fp32:
        rec["parity_in_run"], rec["parity_in_run_detail"] = polygon_parity_in_run(eng, inp, pairs, measured)
    eng.close()
    return rec


# ---- whole matcher stage at BASELINE config 1 (real ORB descriptors) -------------------------------
class CvBFMatcherBackend:
    """The reference's own arithmetic call (cv.BFMatcher(NORM_HAMMING, crossCheck=True).match per
    pair, custom_bruteforce_matcher.py:193-199) behind the engine's two methods, so that the CPU arm
    runs the SAME host logic (RANSAC, confidence, mirrored MatchesInfo) as ours."""

    def __init__(self):
        self.sets = None

    def upload_descriptor_bank(self, rows):
        self.sets = rows

    def hamming_match_pairs(self, pairs, mode="crosscheck", ratio=0.75, variant="auto"):
        import cv2 as cv
        from ssp_b200.engine import HammingResult
        qs, ts, ds, counts = [], [], [], []
        for a, b in pairs:
            ms = cv.BFMatcher(cv.NORM_HAMMING, crossCheck=True).match(cv.UMat(self.sets[a]), cv.UMat(self.sets[b]))
            qs += [m.queryIdx for m in ms]
            ts += [m.trainIdx for m in ms]
            ds += [int(m.distance) for m in ms]
            counts.append(len(ms))
        o = np.zeros(len(pairs) + 1, dtype=np.int64)
        np.cumsum(counts, out=o[1:])
        return HammingResult(o, np.array(qs, np.int32), np.array(ts, np.int32), np.array(ds, np.int32))


def run_stage(args, wl):
    """One step = CustomBruteForceMatcher(...).match_n_features() from cv2.detail.ImageFeatures to
    the tuple of 441 cv2.detail.MatchesInfo (the call stitching_detailed_enhanced.py:971-987
    makes).  Host-side by nature: value == e2e, wall clock."""
    import tempfile
    import cv2 as cv
    import ssp_b200
    g = np.load(os.path.join(ROOT, "tests", "golden", "orb_forest_c1.npz"), allow_pickle=False)
    offs = g["offsets"]
    n = len(offs) - 1
    h, w = int(g["img_shape"][0]), int(g["img_shape"][1])
    orb = cv.ORB_create()
    feats = []
    for i in range(n):
        ft = cv.detail.computeImageFeatures2(orb, np.zeros((h, w, 3), np.uint8))
        ft.descriptors = cv.UMat(np.ascontiguousarray(g["descriptors"][offs[i]:offs[i + 1]]))
        ft.keypoints = tuple(cv.KeyPoint(float(p[0]), float(p[1]), 30.0) for p in g["points"][offs[i]:offs[i + 1]])
        ft.img_size = (w, h)
        ft.img_idx = -1
        feats.append(ft)
    imgs = [np.zeros((h, w, 3), np.uint8) for _ in range(n)]
    names = [str(x) for x in g["names"][:n]]
    tmp = tempfile.mkdtemp(prefix="ssp_bench_stage_")
    reference = args.impl == "reference"
    engine = CvBFMatcherBackend() if reference else ssp_b200.Engine(0)
    n_pairs = n * (n - 1) // 2

    def step():
        m = ssp_b200.CustomBruteForceMatcher(
            imgs_features_orb_and_stars=feats, imgs_features_len_orb=[len(f.keypoints) for f in feats],
            conf_thresh=0.2, img_names=names, work_scale_imgs=imgs, pairwise_matches_output_dir=tmp,
            disable_prompts=True, output_dir_polygon_matching=tmp, predefined_overlaps=[],
            focal_length_pinhole=1135, engine=engine, write_debug_images=False, verbose=False,
            ransac="opencv" if reference else args.ransac)
        return m.match_n_features()

    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        infos = step()
    dt = (time.perf_counter() - t0) / args.steps
    kept = sum(1 for mi in infos if mi.src_img_idx >= 0)
    bytes_in = int(g["descriptors"].nbytes)
    # what comes back per step: (query, train, distance) int32 per cross-check match + the offsets table;
    # with the device RANSAC also one 3x3 f64 model, an inlier count per pair and a mask byte per match
    n_matches = sum(len(mi.matches) for mi in infos if 0 <= mi.src_img_idx < mi.dst_img_idx)
    bytes_out = 0 if reference else n_matches * 12 + (n_pairs + 1) * 8
    if not reference and args.ransac == "device":
        bytes_in += int(g["points"].nbytes)
        bytes_out += n_pairs * (72 + 4) + n_matches
    line = {
        "metric": "image_pairs_matched_per_sec", "value": n_pairs / dt, "unit": "pairs/s", "n_gpus": 1,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "u8 Hamming + f64 RANSAC (OpenCV)", "data": "real (frozen ORB fixture)",
        "config": {"workload": wl["name"], "pairs": n_pairs, "mode": "crosscheck", "kept_directed_entries": kept,
                   "ransac": "opencv" if reference else args.ransac,
                   "note": "whole stage incl. cv.findHomography per pair and MatchesInfo construction on the host; "
                           "the matching itself is one engine call in front of the pair loop"},
        "e2e": {"value": n_pairs / dt, "unit": "pairs/s", "h2d_bytes_per_step": 0 if reference else bytes_in,
                "d2h_bytes_per_step": bytes_out,
                "bytes_note": "d2h counts the matches of the pairs the stage KEEPS (a lower bound of what the engine returns)"},
    }
    if reference:
        line["impl"] = "reference"
        line["cpu_baseline"] = {"value": n_pairs / dt, "unit": "pairs/s", "cores": cv.getNumThreads(), "kind": "reference",
                                "sample": f"cv2 {cv.__version__} BFMatcher crossCheck per pair + the same host logic, all {n_pairs} pairs per step"}
    else:
        # the engine's own count of the last match call (+ the one RANSAC launch), per step
        line["gpu_launches"] = args.steps * (engine.hamming_last_stats()["launches"] + (args.ransac == "device"))
        engine.close()
    return line


_JSON_FD = None


def emit(line):
    """The ONE JSON line goes to the process's original stdout; everything else a library prints
    there (NCCL's version banner at N > 1) was re-routed to stderr in main()."""
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def top_level(rec, args):
    """A record as the ONE line of the bench contract."""
    line = {"metric": "image_pairs_matched_per_sec", "value": rec["value"], "unit": rec["unit"], "n_gpus": rec["n_gpus"],
            "steps": rec["steps"], "warmup": rec["warmup"], "ms_per_step": rec["ms_per_step"], "higher_is_better": True,
            # the pair list of a workload is fixed and divided over the ranks at every N
            "scaling": "strong", "vs_baseline": None, "dtype": rec["dtype"], "data": "synthetic"}
    line.update({k: v for k, v in rec.items() if k not in line})
    return line


def main():
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    args = parse()
    wl = WORKLOADS[args.workload]
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference" and rank != 0:
        return 0  # rank 0 alone runs the CPU arm
    from ssp_b200 import synthetic as synth
    if wl.get("stage"):
        if rank == 0:
            emit(run_stage(args, wl))
        return 0
    cpu_pairs = 0 if (args.no_cpu_baseline or world > 1) else args.cpu_sample_pairs
    if wl.get("polygon"):
        inp = polygon_inputs(wl)
        pairs = synth.all_pairs(wl["n_images"])
        if args.impl == "reference":
            emit(run_polygon_reference(args, wl, inp, pairs))
            return 0
        rt = Runtime()
        rec = polygon_record(rt, wl, inp, pairs, args.steps, args.warmup, args.fp32, parity=bool(cpu_pairs))
        if rank == 0:
            if cpu_pairs:
                ref_args = argparse.Namespace(**{**vars(args), "steps": 2, "warmup": 0})
                rec["cpu_baseline"] = run_polygon_reference(ref_args, wl, inp, pairs)["cpu_baseline"]
            emit(top_level(rec, args))
        rt.close()
        return 0
    # every rank generates the same seeded input (cheap next to shipping it); rank 0's copy is
    # the "host input" of the e2e leg
    sets = synth.descriptor_bank(wl["n_images"], wl["n_desc"], seed=wl["seed"])
    pairs = synth.all_pairs(wl["n_images"])
    if args.impl == "reference":
        emit(run_reference(args, wl, sets, pairs))
        return 0
    rt = Runtime()
    rec = hamming_record(rt, wl, sets, pairs, args.mode, args.variant, args.steps, args.warmup, cpu_pairs)
    default_invocation = (args.workload == "c3" and args.mode == "knn2_ratio" and args.variant == "auto" and not args.no_sub)
    sub = {}
    if default_invocation:
        want = [s for s in args.sub.split(",") if s]
        k, w = max(1, min(args.steps, args.sub_steps)), max(3, min(args.warmup, 3))

        broken = []

        def guarded(name, fn):
            # a sub-record must never take the headline line down with it; after a failure the process
            # group may be out of step, so the remaining sub-records are skipped rather than risked
            if broken:
                r = {"error": f"skipped after the failure of sub-record {broken[0]!r}"} if rank == 0 else None
            else:
                try:
                    r = fn()
                except Exception as e:                                     # noqa: BLE001
                    broken.append(name)
                    r = {"error": f"{type(e).__name__}: {e}"[:400]} if rank == 0 else None
            if rank == 0:
                sub[name] = r

        if "crosscheck" in want:
            # the call the reference's pipelines actually make (custom_bruteforce_matcher.py:193-199)
            guarded("crosscheck", lambda: hamming_record(rt, wl, sets, pairs, "crosscheck", "auto", k, w, min(cpu_pairs, 6)))
        del sets
        if "c5" in want:
            w5 = WORKLOADS["c5"]
            sets5 = synth.descriptor_bank(w5["n_images"], w5["n_desc"], seed=w5["seed"])
            pairs5 = synth.all_pairs(w5["n_images"])
            guarded("c5", lambda: hamming_record(rt, w5, sets5, pairs5, "knn2_ratio", "auto", k, w, min(cpu_pairs * 3, 32)))
            del sets5
        for name in ("c4", "c4cap"):
            if name in want:
                wp = WORKLOADS[name]
                inp = polygon_inputs(wp)
                pp = synth.all_pairs(wp["n_images"])

                def poly(wp=wp, inp=inp, pp=pp):
                    r = polygon_record(rt, wp, inp, pp, k, w, False, parity=bool(cpu_pairs))
                    if rank == 0 and cpu_pairs:
                        ref_args = argparse.Namespace(**{**vars(args), "steps": 2, "warmup": 0})
                        r["cpu_baseline"] = run_polygon_reference(ref_args, wp, inp, pp)["cpu_baseline"]
                    return r
                guarded(name, poly)
    if rank == 0:
        line = top_level(rec, args)
        if sub:
            line["sub"] = sub
        emit(line)
    rt.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
