#!/usr/bin/env python
"""bench.py — candidate pairs scored+selected per second on N B200s (BASELINE.json metric).

One step = one pass of the hot path (score every pair with the K-learner siamese head -> committee
reduction -> utility -> top-B, + the cross-shard key merge when N > 1) over one batch of synthetic
input. Workload at N = 1: BASELINE.json configs[1] (1 M pairs, 512-d ArcFace-shaped table, K = 5,
top-1000, fp32 bit-exact mode); weak scaling: every rank owns a 1 M-pair shard of an N x 1 M pool.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--mode refine|fp32|bf16] [--impl native|reference]
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

from alink_b200 import synth  # noqa: E402

METRIC = "pairs_scored_selected_per_s"
UNIT = "pairs/s"
FLOPS_PER_PAIR = lambda K, D, H1, H2, C: 2.0 * K * (D * H1 + H1 * H2 + H2 * C)


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm_gbs=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sustained=d["bf16_tflops_sustained"],
                    source="measured (MEASURED_PEAKS.json)")
    return dict(hbm_gbs=6650.0, tf_burst=1590.0, tf_sustained=1400.0, source="fallback (B200_PROFILING.md)")


class ClockSampler:
    """SM clock / throttle reasons DURING the timed region (B200_PROFILING.md recipe). The timed region of
    this step is tens of milliseconds, too short for `nvidia-smi -lms`, so the same NVML counters are polled
    in-process every 2 ms (pynvml); `nvidia-smi` is the fallback when pynvml cannot be loaded."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    BITS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, gpu_index=0, uuid=None):
        self.rows, self.proc, self.gpu = [], None, gpu_index
        self.sm, self.mask, self.h, self.nv, self._stop = [], 0, None, None, False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            try:
                self.h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + str(uuid)) if uuid and not str(uuid).startswith("GPU-") else str(uuid))
            except Exception:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
            self.max_sm = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.h = None

    def _poll(self):
        nv = self.nv
        reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
        while not self._stop:
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                self.mask |= int(reasons(self.h))
            except Exception:
                pass
            time.sleep(0.002)

    def start(self):
        if self.h is not None:
            self.t = threading.Thread(target=self._poll, daemon=True)
            self.t.start()
            return
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.h is not None:
            self._stop = True
            self.t.join(timeout=1)
            return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.max_sm,
                    "reasons": sorted(n for b, n in self.BITS.items() if self.mask & b), "samples": len(self.sm),
                    "source": "NVML polled in-process every 2 ms"}
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            pass
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 9 for n, v in zip(names, r[5:9]) if v.lower() == "active"})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm), "source": "nvidia-smi -lms 100"}


def cpu_oracle_throughput(pool, sample_pairs, threads=0, repeats=1):
    """Oracle (C restatement, oracle/) on the first `sample_pairs` pairs + top-B; returns pairs/s."""
    from oracle import oracle_ffi
    n = min(sample_pairs, pool.N)
    ia, ib = pool.ia[:n], pool.ib[:n]
    best = None
    for _ in range(repeats):
        t0 = time.perf_counter()
        out = oracle_ffi.score(pool.table, pool.table, ia, ib, pool.heads, measure=pool.measure,
                               want_per_learner=False, want_consensus=False, threads=threads)
        oracle_ffi.topk(out["util"], min(pool.B, n))
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return n / best, n, best, oracle_ffi.lib().alink_oracle_get_threads()


def workload_desc(pool, n_gpus, mode, cfg=2, table_rows=None):
    h = pool.heads
    rows = table_rows or pool.table.shape[0]
    shape = "ArcFace-shaped fp32" if cfg == 2 else "ResNet-50-shaped bf16"
    tbytes = rows * pool.D * (4 if cfg == 2 else 2)
    return {"workload": f"BASELINE configs[{cfg - 1}]: {pool.N:,} pairs/GPU x {n_gpus} GPU, {pool.D}-d {shape} table "
                        f"(M={rows:,}), K={h.K} committee {h.D_in}->{h.H1}->{h.H2}->{h.C} {h.act}, "
                        f"measure={pool.measure}, top-{pool.B}",
            "pairs_per_gpu": pool.N, "D": pool.D, "K": h.K, "B": pool.B, "compute": mode,
            "sharding": f"dp{n_gpus} by pair index; alink_query_sharded: ncclAllGather of the top-B keys + merge inside the library" if n_gpus > 1 else "single GPU",
            "l2": f"inputs larger than L2 (table {tbytes / 1e6:.0f} MB + {pool.N * 8 / 1e6:.0f} MB indices, random row "
                  "gather; no explicit flush)"}


def run_reference(args, rank, emit):
    """--impl reference: the reference's CPU path for this step. Keras/TF is not installable here, so
    this is the oracle port (oracle/alink_oracle.c) with all host threads, on a bounded sample."""
    if rank != 0:
        return
    if args.config == 3:
        pool = synth.make_pool(3, N=args.ref_sample, M=20_000, measure=args.measure)
    else:
        pool = synth.make_pool(2, N=args.ref_sample, measure=args.measure)
    times = []
    threads = 0
    for i in range(args.warmup + args.steps):
        v, n, dt, threads = cpu_oracle_throughput(pool, args.ref_sample)
        if i >= args.warmup:
            times.append(dt)
    ms = 1e3 * float(np.mean(times))
    value = args.ref_sample / (ms / 1e3)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": dict(workload_desc(pool, 1, "cpu-oracle", args.config),
                           note="Keras/TF reference unavailable on this box; timed path = C oracle port, "
                                f"bounded sample of {args.ref_sample:,} pairs per step"),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": f"first {args.ref_sample:,} pairs of configs[1], all {threads} host threads"},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


def run_config45(args, rank, local_rank, world, dev, peaks, emit, torch, dist):
    """BASELINE configs[3] (A2-LINK noise-augmented pool) and configs[4] (DFW-style all-pairs) through the product's
    device-resident entry points; same timing protocol as the headline (W warm-up steps, K timed steps between barriers,
    CUDA events + wall clock, max over ranks). These are parity-test shapes measured for the record (profiles/), not
    the driver's bench line."""
    from alink_b200 import _ffi, alink
    from alink_b200.dist import decode_keys, merge_shards, sharded_disparity_select
    torch.cuda.set_stream(torch.cuda.Stream(device=dev))
    ctx = _ffi.Context(local_rank)
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    if world > 1:
        uid = [_ffi.Context.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        ctx.comm_init(uid[0], rank, world)
    mode = "bf16" if args.mode == "refine" and args.config == 4 else args.mode  # config 4 needs scores, not a selection
    cid = _ffi.COMPUTE_ID[mode]
    B = 1000

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        sampler = None if args.no_clock_sampling else ClockSampler(local_rank, getattr(torch.cuda.get_device_properties(local_rank), "uuid", None))
        barrier()
        if sampler:
            sampler.start()
        l0 = ctx.launches
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for _ in range(steps):
            res = fn()
        e1.record()
        barrier()
        wall = 1e3 * (time.perf_counter() - t0)
        ms = torch.tensor([max(wall, e0.elapsed_time(e1))], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()) / steps, res, ctx.launches - l0, (sampler.stop() if sampler else None)

    if args.config == 4:
        T = args.noise_types
        n_total = args.pairs or 1_000_000
        pool = synth.make_pool(2, N=n_total)
        lo, hi = (n_total * rank) // world, (n_total * (rank + 1)) // world
        n = hi - lo
        m2 = synth.make_pool(2, N=4096, K=1).heads
        table = torch.from_numpy(pool.table).to(dev)
        ia, ib = torch.from_numpy(pool.ia[lo:hi]).to(dev), torch.from_numpy(pool.ib[lo:hi]).to(dev)
        y = torch.from_numpy(pool.genuine[lo:hi].astype(np.float32)).to(dev)
        ctx.set_table(0, table)
        ctx.set_heads(0, pool.heads, compute=cid)
        ctx.set_heads(1, m2, compute=cid)
        seed = synth.BASE_SEED + 4
        K1 = pool.heads.K

        def step():
            idx, active, top_idx, top_util, (p1, p2) = alink.a2link_pool_select(
                ctx, 0, 1, 0, ia, ib, n, K1, T=T, seed=seed, y=y if world == 1 else None, B=min(B, n), rule=(world == 1), pair_base=lo)
            if world > 1:  # the per-noise prefix is a GLOBAL order statistic: radix select across the ranks on the device
                idx, active = ctx.disparity_select_comm(p2, p1, lo, n_total, 0.25, 0.05, False, y)
                cnt = torch.tensor([len(idx)], dtype=torch.int64, device=dev)
                dist.all_reduce(cnt)
                idx = np.empty((int(cnt.item()),), np.int64)  # (only the count is reported; the indices stay sharded)
                keys = torch.zeros((B,), dtype=torch.int64, device=dev)
                _, _, k = ctx.select_topk(min(B, n), base=lo)
                keys[:len(k)] = torch.from_numpy(k.view(np.int64)).to(dev)
                merged = merge_shards(keys, B, ctx.topk_keys, dist=dist, device=dev)
                top_idx, top_util = decode_keys(merged)
            return idx, active, top_idx, top_util

        ms, (idx, active, top_idx, top_util), launches, clocks = timed(step, args.steps, args.warmup)
        scorings = n_total * (1 + T)
        h = pool.heads
        flops = (FLOPS_PER_PAIR(K1, h.D_in, h.H1, h.H2, h.C) + T * FLOPS_PER_PAIR(1, h.D_in, h.H1, h.H2, h.C)) * n
        gen_bytes = T * 2 * n * pool.D * 4 * 2  # perturbed rows: written once, read once by the M2 head
        line = {"metric": METRIC, "value": scorings / (ms / 1e3), "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "bf16" if mode == "bf16" else "f32", "data": "synthetic (perturbed copies generated on the device)",
                "config": {"workload": f"BASELINE configs[3]: A2-LINK pool, {n_total:,} base pairs x T={T} perturbed copies of both sides "
                                       f"(alink_perturb_rows, 40 coordinates, U(-0.25, 0.25), renormalised), M1 K={K1} on the base pairs + M2 K=1 "
                                       f"on all {n_total * T:,} perturbed pairs, 512-d fp32 table, disparity rule (ratio 0.25, eps 0.05) + top-{B} by "
                                       "min-disparity; a unit = one pair SCORING", "compute": mode, "pair_scorings_per_step": scorings,
                           "sharding": f"base pairs split over {world} GPU(s)", "l2": "perturbed tables of 2 x 2 GB per noise type stream through HBM"},
                "roofline": {"bound": "hbm", "achieved": gen_bytes / (ms / 1e3) / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                             "frac": gen_bytes / (ms / 1e3) / 1e9 / peaks["hbm_gbs"], "traffic": None,
                             "note": "whole step against the bytes of the generated pool alone (K = 1 scoring of fp32 rows is the HBM-bound case of SURVEY 8d)",
                             "tensor_TFLOPs_whole_step": flops / (ms / 1e3) / 1e12},
                "result": {"rule_selected": int(len(idx)), "active_count": int(active), "top1_pair": int(top_idx[0]), "top1_min_disparity": float(top_util[0])},
                "gpu_launches": launches, "clocks": clocks}
        if rank == 0:
            emit(line)
        return

    # ---- config 5: P x G all-pairs, probe rows sharded over the ranks
    P = G = args.probe_rows
    pool = synth.make_pool(2, N=4096, M=P)
    ctx.set_table(0, torch.from_numpy(pool.table).to(dev))
    ctx.set_heads(0, pool.heads, compute=cid)
    r0, r1 = (P * rank) // world, (P * (rank + 1)) // world

    def step5():
        _, _, keys = ctx.query_allpairs_keys(0, 0, 0, r0, r1 - r0, args.measure, B)
        if world == 1:
            return keys
        kd = torch.from_numpy(keys.view(np.int64)).to(dev)
        return merge_shards(kd, B, ctx.topk_keys, dist=dist, device=dev)

    ms, merged, launches, clocks = timed(step5, args.steps, args.warmup)
    n_pairs = P * G
    h = pool.heads
    tf_gpu = FLOPS_PER_PAIR(h.K, h.D_in, h.H1, h.H2, h.C) * (r1 - r0) * G / (ms / 1e3) / 1e12
    idx, util = decode_keys(merged)
    line = {"metric": METRIC, "value": n_pairs / (ms / 1e3), "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "bf16" if mode == "bf16" else "f32",
            "data": "synthetic", "config": {"workload": f"BASELINE configs[4]: DFW-style all-pairs {P:,} x {G:,} x 512-d, K={h.K}, {args.measure} top-{B}, "
                                                        f"probe rows sharded over {world} GPU(s) (pairs enumerated arithmetically, no index arrays)",
                                            "compute": {"fp32": "FP32_EXACT", "bf16": "BF16_TC", "refine": "FP32_REFINE"}[mode], "pairs": n_pairs,
                                            "l2": "2 x 102 MB tables: L2-resident gather"},
            "roofline": {"bound": "tensor", "achieved": tf_gpu, "peak": peaks["tf_sustained"], "unit": "TFLOP/s", "frac": tf_gpu / peaks["tf_sustained"],
                         "frac_burst": tf_gpu / peaks["tf_burst"], "traffic": None, "note": "whole step per GPU (filter + refine + select)"},
            "result": {"top1_pair": [int(idx[0]) // G, int(idx[0]) % G], "top1_utility": float(util[0]),
                       "refine": ctx.refine_report() if mode == "refine" else None},
            "gpu_launches": launches, "clocks": clocks}
    if rank == 0:
        emit(line)


def main():
    # stdout carries exactly ONE JSON line: libraries that print banners there (NCCL) go to stderr instead
    json_fd = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(os.dup(2), "w")

    def emit(obj):
        os.write(json_fd, (json.dumps(obj) + "\n").encode())

    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--mode", default="refine", choices=["refine", "fp32", "bf16"],
                    help="refine = fp32 bit-exact selection via tensor-core filter + exact re-score (default: "
                         "BASELINE configs[1] is the fp32 bit-exact mode); fp32 = CUDA-core exact; bf16 = tensor cores only")
    ap.add_argument("--measure", default="uncertainty")
    ap.add_argument("--pairs", type=int, default=None, help="pairs per GPU (default 1,000,000; config 3: 1,250,000)")
    ap.add_argument("--config", type=int, default=2, choices=[2, 3, 4, 5],
                    help="BASELINE.json configs ordinal: 2 = 1 M pairs x 512-d (default, the N=1 workload); "
                         "3 = 10 M pairs x 2048-d bf16 over 8 GPUs (1.25 M pairs per GPU, bf16 table, BF16_TC); "
                         "4 = A2-LINK noise-augmented pool (1 M base pairs x 16 perturbed copies generated on the device, "
                         "M1 K=5 + M2 K=1, disparity rule + top-1000 by min-disparity; base pairs sharded over the GPUs); "
                         "5 = DFW-style 50k x 50k all-pairs, sharded by probe-row block over the GPUs")
    ap.add_argument("--noise-types", type=int, default=16, help="config 4: T perturbed copies per base pair")
    ap.add_argument("--probe-rows", type=int, default=50_000, help="config 5: P = G rows")
    ap.add_argument("--table-rows", type=int, default=None)
    ap.add_argument("--ref-sample", type=int, default=100_000)
    ap.add_argument("--cpu-sample", type=int, default=100_000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-mode-compare", action="store_true")
    ap.add_argument("--no-target", action="store_true", help="skip the north-star target block at --gpus 8")
    ap.add_argument("--no-clock-sampling", action="store_true", help="for runs under ncu: no NVML polling thread")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank, emit)
        return

    import torch
    import torch.distributed as dist
    from alink_b200 import _ffi
    from alink_b200.dist import decode_keys, merge_shards

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    n_gpus = world
    peaks = load_peaks()
    if args.config in (4, 5):
        run_config45(args, rank, local_rank, world, dev, peaks, emit, torch, dist)
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    # ---- synthetic workload (identical table/heads on every rank; each rank draws its own shard of pairs)
    if args.pairs is None:
        args.pairs = 1_250_000 if args.config == 3 else 1_000_000
    table_dev = None
    if args.config == 3:
        # 2048-d ResNet-50-shaped features: generated ON THE DEVICE (500 k x 2048), stored bf16 (2 GB); the host
        # only holds a small fp32 slice for head calibration and the CPU sample
        args.mode = "bf16"
        M, D, I = args.table_rows or 500_000, 2048, 125_000
        gen = torch.Generator(device=dev).manual_seed(synth.BASE_SEED + 3)
        cen = torch.randn((I, D), generator=gen, device=dev, dtype=torch.float32)
        table_dev = torch.empty((M, D), device=dev, dtype=torch.bfloat16)
        per = max(M // I, 1)
        for lo in range(0, M, 50_000):
            hi = min(lo + 50_000, M)
            ident = torch.clamp(torch.arange(lo, hi, device=dev) // per, max=I - 1)
            table_dev[lo:hi] = torch.relu(cen[ident] + 0.5 * torch.randn((hi - lo, D), generator=gen, device=dev)).to(torch.bfloat16)
        del cen
        small = table_dev[:20_000].float().cpu().numpy()
        ia, ib, genuine = synth.make_pairs(args.pairs, M, I, seed=synth.BASE_SEED + 103 + 1000 * rank)
        cal_a, cal_b, _ = synth.make_pairs(4096, 20_000, 5_000, seed=7)
        heads = synth.make_heads(5, D, synth.pair_features(small, small, cal_a, cal_b))
        pool = synth.Pool("cfg3", small, ia, ib, heads, 1000, args.measure, genuine)
    else:
        pool = synth.make_pool(2, N=args.pairs, measure=args.measure)
        if world > 1:
            ia, ib, _ = synth.make_pairs(args.pairs, pool.table.shape[0], 50_000, seed=synth.BASE_SEED + 102 + 1000 * rank)
            pool.ia, pool.ib = ia, ib
    N, B, K = pool.N, pool.B, pool.heads.K
    base = rank * N  # global pair index of this shard's first pair
    mode = _ffi.COMPUTE_ID[args.mode]

    # a non-default stream for everything timed: the legacy default stream synchronises implicitly with every blocking
    # stream, which would serialise the e2e pipeline (table feed on a side stream) behind each kernel of the query
    torch.cuda.set_stream(torch.cuda.Stream(device=dev))
    ctx = _ffi.Context(local_rank)
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    ctx.set_heads(0, pool.heads, compute=mode)
    if world > 1:
        # the library's own communicator (alink_comm_init): the id travels through the process group that torchrun set up
        uid = [_ffi.Context.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        ctx.comm_init(uid[0], rank, world)
    if table_dev is None:
        table_dev = torch.from_numpy(pool.table).to(dev)
    ia_dev = torch.from_numpy(pool.ia).to(dev)
    ib_dev = torch.from_numpy(pool.ib).to(dev)
    table_pin = table_dev.cpu().pin_memory() if args.config == 3 else torch.from_numpy(pool.table).pin_memory()
    ia_pin = torch.from_numpy(pool.ia).pin_memory()
    ib_pin = torch.from_numpy(pool.ib).pin_memory()
    out = (np.empty((B,), np.int64), np.empty((B,), np.float32), np.empty((B,), np.uint64))

    keys_dev = torch.zeros((B,), dtype=torch.int64, device=dev)  # u64 key bit patterns, left on the device
    out_dev = (None, None, keys_dev)

    def step_resident():
        """inputs already in HBM; per-shard keys stay on the device, one all-gather + merge when N > 1, and
        the merged B keys are read back."""
        if world == 1:
            _, _, keys, t = ctx.query(0, 0, 0, ia_dev, ib_dev, N, args.measure, B, base=base, timing=True, out=out)
            return keys, t
        _, _, keys, t = ctx.query_sharded(0, 0, 0, ia_dev, ib_dev, N, args.measure, B, base=base, timing=True)
        return keys, t  # shard top-B -> ncclAllGather -> merge, all inside the library (alink_query_sharded)

    e2e_slot = {"cur": 1, "primed": False}
    feed = None
    if world > 1:
        # N > 1: every rank uploads 1/N of the (replicated) table over its own PCIe link and one NCCL all-gather over
        # NVLink completes the replica on every GPU (alink_b200.dist.ShardedTableFeed); round 1 uploaded the whole table
        # N times from one NUMA node (e2e efficiency 0.42 at N = 8)
        from alink_b200.dist import ShardedTableFeed
        feed = ShardedTableFeed(ctx, table_pin.shape[0], table_pin.shape[1], dist=dist, device=dev, dtype=table_pin.dtype,
                                gather=os.environ.get("ALINK_FEED_GATHER", "nccl"))
        shard_pin = feed.shard_of(table_pin)

    def step_e2e():
        """host buffers in, host result out: table + indices cross PCIe inside EVERY step. Two table buffers: while
        step i is scored from one, the table of step i + 1 is uploaded into the other on a side stream, so a step
        costs max(upload, query) instead of their sum. Each step uploads exactly one table (the first call of a run
        uploads its own synchronously). N = 1: alink_set_table_async (copy stream inside the library)."""
        cur, nxt = e2e_slot["cur"], 3 - e2e_slot["cur"]
        if feed is not None:
            if not e2e_slot["primed"]:
                feed.feed_async(cur - 1, shard_pin, (ia_pin, ib_pin))
                e2e_slot["primed"] = True
            feed.feed_async(nxt - 1, shard_pin, (ia_pin, ib_pin))  # the next step's table shard AND pair indices
            ia_e, ib_e = feed.wait(cur - 1, cur)
        else:
            if not e2e_slot["primed"]:
                ctx.set_table(cur, table_pin)
                e2e_slot["primed"] = True
            ctx.set_table_async(nxt, table_pin)
        e2e_slot["cur"] = nxt
        if world == 1:
            _, _, keys, t = ctx.query(0, cur, cur, ia_pin.numpy(), ib_pin.numpy(), N, args.measure, B, base=base,
                                      timing=True, out=out)
            return keys, t
        _, _, keys, t = ctx.query_sharded(0, cur, cur, ia_e, ib_e, N, args.measure, B, base=base, timing=True)
        return keys, t

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(step_fn, steps, warmup, sample_clocks=False):
        for _ in range(warmup):
            step_fn()
        sampler = ClockSampler(local_rank, getattr(torch.cuda.get_device_properties(local_rank), "uuid", None)) if sample_clocks else None
        barrier()
        if sampler:
            sampler.start()
        l0 = ctx.launches
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        stage = {"ms_score": 0.0, "ms_utility": 0.0, "ms_select": 0.0, "ms_h2d": 0.0, "ms_d2h": 0.0}
        t0 = time.perf_counter()
        e0.record()
        for _ in range(steps):
            merged, t = step_fn()
            for k_ in stage:
                stage[k_] += getattr(t, k_)
        e1.record()
        barrier()
        wall_ms = 1e3 * (time.perf_counter() - t0)
        dev_ms = e0.elapsed_time(e1)
        clocks = sampler.stop() if sampler else None
        launches = ctx.launches - l0
        ms = torch.tensor([max(wall_ms, dev_ms)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()) / steps, {k_: v / steps for k_, v in stage.items()}, launches, clocks, merged, dev_ms / steps

    ctx.set_table(0, table_dev)
    ms_step, stage, launches, clocks, merged, dev_ms = timed(step_resident, args.steps, args.warmup, sample_clocks=not args.no_clock_sampling)
    refine_stats = ctx.refine_stats() if args.mode == "refine" else None
    refine_detail = ctx.refine_detail() if args.mode == "refine" else None
    refine_report = ctx.refine_report() if args.mode == "refine" else None
    reserve = int(os.environ.get("ALINK_RESERVE_SMS", "0"))
    if feed is not None and reserve:
        ctx.set_reserved_sms(reserve)  # the feed's communication kernels run beside the persistent head kernel
    ms_e2e, stage_e2e, _, _, merged_e2e, _ = timed(step_e2e, max(4, min(args.steps, 8)), 2)
    if feed is not None and reserve:
        ctx.set_reserved_sms(0)
    gather_ms = None
    if feed is not None:  # the table all-gather alone (device time, max over ranks), for its bus bandwidth
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        best = None
        for _ in range(3):
            barrier()
            g0.record()
            lo = feed.rank * feed.rows
            dist.all_gather_into_tensor(feed.bufs[0], feed.bufs[0][lo:lo + feed.rows])
            g1.record()
            torch.cuda.synchronize()
            tms = g0.elapsed_time(g1)
            best = tms if best is None else min(best, tms)
        tt = torch.tensor([best], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        gather_ms = float(tt.item())
    assert np.array_equal(merged, merged_e2e), "resident and e2e paths selected different pairs"

    value = n_gpus * N / (ms_step / 1e3)
    e2e_value = n_gpus * N / (ms_e2e / 1e3)
    h = pool.heads
    flops = FLOPS_PER_PAIR(K, h.D_in, h.H1, h.H2, h.C) * N
    head_ms = stage["ms_score"]  # fp32 / bf16: the head kernel(s); refine: the tensor-core filter over all N pairs
    achieved_tf = flops / (head_ms / 1e3) / 1e12
    step_tf = flops / (dev_ms / 1e3) / 1e12  # the whole step (what `value` measures), device time
    roofline = {"bound": "tensor", "achieved": achieved_tf, "peak": peaks["tf_sustained"], "unit": "TFLOP/s",
                "frac": achieved_tf / peaks["tf_sustained"], "traffic": None,
                "peak_burst": peaks["tf_burst"], "frac_burst": achieved_tf / peaks["tf_burst"],
                "step_achieved": step_tf, "step_frac": step_tf / peaks["tf_sustained"],
                "step_frac_burst": step_tf / peaks["tf_burst"],
                "denominators": "frac / step_frac: measured sustained cuBLAS bf16 peak (a seconds-long loop under the power "
                                "cap); frac_burst / step_frac_burst: measured burst peak — the timed region here is tens of "
                                "milliseconds at full clocks, so the burst figure is the stricter, matching denominator",
                "kernel": "siamese head (gather+|a-b| fused layer-1 GEMM, layer-2/3 + softmax)",
                "flops_per_pair": FLOPS_PER_PAIR(K, h.D_in, h.H1, h.H2, h.C), "launch_ms": head_ms,
                "peak_source": peaks["source"]}
    mode_name = {"fp32": "FP32_EXACT", "bf16": "BF16_TC", "refine": "FP32_REFINE"}[args.mode]
    if args.config == 2 and args.mode != "fp32" and N == 1_000_000:
        # DRAM bytes of one launch of this kernel on this workload, from the committed ncu --set full capture
        try:
            if args.mode == "refine":  # ncu --set full of this build's filter launch (tools/r2_ncu.sh)
                with open(os.path.join(ROOT, "profiles", "r2_ncu_head_full_summary.json")) as f:
                    roofline["traffic"] = json.load(f)["traffic_bytes_per_launch"]
                roofline["traffic_source"] = "profiles/r2_ncu_head_full_summary.json (dram__bytes_read.sum + dram__bytes_write.sum, one launch)"
            else:  # the bf16 K = 5 launch of the HBM-stage capture
                with open(os.path.join(ROOT, "profiles", "r2_ncu_hbm_stages.json")) as f:
                    hk = [l for l in json.load(f)["launches"] if "head_kernel<0, 2, 0, 1>" in l["kernel"] and l["us"] > 1500]
                roofline["traffic"] = hk[0]["dram_read_B"] + hk[0]["dram_write_B"]
                roofline["traffic_source"] = "profiles/r2_ncu_hbm_stages.json (dram__bytes_read.sum + dram__bytes_write.sum of the K = 5 launch)"
        except Exception:
            pass
    if args.mode == "refine":
        roofline["kernel"] = ("tc::head_kernel (tcgen05 kind::f16 filter over every pair, "
                              f"{refine_detail['filter']} operands / fp32 accumulate; the exact re-score of the "
                              "survivors is extra, non-algorithmic work)")
        roofline["share_of_step"] = head_ms / ms_step
    if args.mode == "fp32":
        clk = (clocks or {}).get("sm_mhz") or 1965.0
        ffma_peak = 2 * 128 * 148 * clk * 1e6 / 1e12
        roofline["note"] = ("FP32_EXACT is the bit-exact correctness mode: CUDA-core FFMA, not tensor cores; "
                            f"FFMA peak at the sampled {clk:.0f} MHz = {ffma_peak:.1f} TFLOP/s")
        roofline["frac_of_ffma_peak"] = achieved_tf / ffma_peak
    h2d = int((feed.bytes_h2d if feed is not None else table_pin.numel() * table_pin.element_size()) + pool.ia.nbytes + pool.ib.nbytes)
    d2h = int(B * 8)
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": n_gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16" if args.mode == "bf16" else "f32", "data": "synthetic",
            "config": workload_desc(pool, n_gpus, mode_name, args.config, int(table_dev.shape[0])),
            "stage_ms": stage, "device_ms_per_step": dev_ms,
            "roofline": roofline,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms_e2e, "stage_ms": stage_e2e,
                    "call": ("alink_set_table_async(HOST table of the next step, second slot) + alink_query(HOST indices) -> host top-B; every step uploads one table, overlapped with the previous step's query"
                             if feed is None else
                             "ShardedTableFeed: each rank uploads the next step's pair indices and its 1/N of the next step's table from pinned host memory (h2d_bytes_per_step is PER RANK) + in-place NCCL all-gather of the table over NVLink, all on a side stream overlapped with the current step's alink_query -> device keys -> all-gather + merge -> host top-B"),
                    **({"table_allgather": {"mode": ("peer-to-peer copies out of symmetric memory (copy engines, no SMs)" if feed.mode == "copy" else "ncclAllGather"),
                                             "bytes_received_per_rank": feed.bytes_gather, "ranks": world, "ms_alone_nccl": gather_ms,
                                             "busbw_GBps": feed.bytes_gather / (gather_ms / 1e3) / 1e9}} if feed is not None else {})},
            "gpu_launches": launches, "clocks": clocks}

    if world == 8 and args.config == 2 and not args.no_target:
        # north_star target shape, measured by the same run: 10 M pairs x 512-d, K = 5, top-1000 over 8 GPUs (1.25 M pairs per
        # GPU), fp32-identical selection (FP32_REFINE); device-resident, same timing protocol, outside the headline value
        NT = 10_000_000 // world
        ta, tb, _ = synth.make_pairs(NT, pool.table.shape[0], 50_000, seed=synth.BASE_SEED + 202 + 1000 * rank)
        ta_dev, tb_dev = torch.from_numpy(ta).to(dev), torch.from_numpy(tb).to(dev)
        tbase = rank * NT

        def step_target():
            _, _, keys, t = ctx.query_sharded(0, 0, 0, ta_dev, tb_dev, NT, args.measure, B, base=tbase, timing=True)
            return keys, t

        ms_t, stage_t, _, _, merged_t, dev_t = timed(step_target, max(3, args.steps), 3)
        tflops = FLOPS_PER_PAIR(K, h.D_in, h.H1, h.H2, h.C) * NT
        tf_gpu = tflops / (ms_t / 1e3) / 1e12
        line["target"] = {"shape": "north_star: 10,000,000 pairs x 512-d, K=5, top-1000 over 8 GPUs, FP32_REFINE (selection identical to fp32)",
                          "pairs_total": NT * world, "ms_per_step": ms_t, "pairs_per_s": NT * world / (ms_t / 1e3), "stage_ms": stage_t,
                          "tflops_per_gpu_whole_step": tf_gpu, "step_frac": tf_gpu / peaks["tf_sustained"],
                          "step_frac_burst": tf_gpu / peaks["tf_burst"],
                          "head_frac": tflops / (stage_t["ms_score"] / 1e3) / 1e12 / peaks["tf_sustained"],
                          "head_frac_burst": tflops / (stage_t["ms_score"] / 1e3) / 1e12 / peaks["tf_burst"],
                          "refine": ctx.refine_report(), "top1": [int(v) for v in decode_keys(merged_t[:1])[0]]}
    if refine_stats is not None:
        line["refine"] = {"survivors_rescored_exactly": refine_stats[0], "max_abs_utility_error_seen": refine_stats[1],
                          **refine_detail, "report": refine_report, "stage_ms": {"tc_filter": stage["ms_score"], "mark_and_exact_rescore": stage["ms_utility"],
                                                    "select": stage["ms_select"]}}
    if n_gpus == 1 and not args.no_mode_compare and args.config == 2:
        # outside the timed region: the same step in the other compute modes, and how the selections compare
        cmp_ = {}
        sel = {}
        for name, cid in (("fp32_exact", _ffi.FP32_EXACT), ("bf16_tc", _ffi.BF16_TC), ("fp32_refine", _ffi.FP32_REFINE)):
            ctx.set_heads(1, pool.heads, compute=cid)
            best = None
            for _ in range(3):
                idx_m, _, keys_m, t = ctx.query(1, 0, 0, ia_dev, ib_dev, N, args.measure, B, base=base, timing=True)
                best = t.ms_total if best is None else min(best, t.ms_total)
            sel[name] = set(idx_m.tolist())
            cmp_[name] = {"ms_per_step": best, "pairs_per_s": N / (best / 1e3)}
        cmp_["fp32_refine"]["selection_identical_to_fp32_exact"] = sel["fp32_refine"] == sel["fp32_exact"]
        cmp_["bf16_tc"]["top_B_overlap_with_fp32_exact"] = len(sel["bf16_tc"] & sel["fp32_exact"]) / max(B, 1)
        line["modes"] = cmp_
    if rank == 0 and n_gpus == 1 and not args.no_cpu_baseline:
        # bounded CPU sample: calibrate on 10k pairs, then size the sample for ~10-20 s
        if args.config == 3:  # the host only holds a 20 k-row fp32 slice: same per-pair work on folded indices
            pool = synth.Pool(pool.name, pool.table, pool.ia % 20_000, pool.ib % 20_000, pool.heads, pool.B, pool.measure,
                              pool.genuine)
        v0, _, dt0, threads = cpu_oracle_throughput(pool, 10_000)
        n_s = int(min(pool.N, max(20_000, min(args.cpu_sample * 4, v0 * 12.0))))
        v, n_s, dt, threads = cpu_oracle_throughput(pool, n_s)
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                                "sample": f"first {n_s:,} pairs of the same workload, {dt:.1f} s, C oracle "
                                          f"(oracle/alink_oracle.c) on {threads} host threads; Keras/TF reference "
                                          "unavailable on this box"}
    if rank == 0:
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
