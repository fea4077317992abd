#!/usr/bin/env python
"""bench.py -- chain_dp throughput on B200 (BASELINE.json metric), one JSON line on stdout.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --steps K --warmup W    # the reference's CPU mm_chain_dp on the host cores

A step = one pass of the hot path (mm_chain_dp over every read of the workload: prep, fill, compaction,
backtrack, gather) over one batch of synthetic reads.  Workload at N=1: BASELINE.json configs[1]
("map-ont: 100k synthetic 10 kb Nanopore-like reads").  For N>1 every rank owns its own 100k reads (weak scaling,
read-batch sharding, no collective on the data path).

* value : visited anchor pairs / s, whole job, anchors already resident in HBM, CUDA events, max over ranks
* e2e   : the same through the C ABI call a reference-side batcher makes (mm2b200_chain_batch) with pinned HOST
          buffers: H2D of the anchors, all kernels, D2H of the chains inside the timed region
* roofline / roofline_issue : the fill kernel against HBM and against the measured integer-issue peak
* cpu_baseline : the reference's own chain.c (oracle/_ref) under its kt_for on all host cores, bounded sample

"visited pair" = one execution of the reference's inner loop body (chain.c:59-88); the count comes from the
device (mm2b200_result_t.n_visited) and is cross-checked against the oracle on the CPU-baseline sample.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")   # before CUDA initialises (see chain_api.cu: sub-batch streams)

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "chain_dp_anchor_pair_evals_per_sec"
UNIT = "pairs/s"
INT_OPS_PER_PAIR = 28          # integer ops the algorithm needs per visited pair
BYTES_PER_ANCHOR = 40          # 16 B anchor in + 16 B anchor out + 8 B f,p
CPU_SAMPLE_READS = 40_000      # bounded sample for the CPU arm (~10-30 s of CPU work on the map-ont workload)


def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as fh:
            d = json.load(fh)
        return float(d.get("hbm_gbs", 6650.0)), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed regions run."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        try:
            self.proc.terminate()
            self.proc.wait(timeout=5)
        except Exception:
            pass
        sm, smax, reasons = [], [], set()
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 9:
                    continue
                try:
                    sm.append(float(f[1])); smax.append(float(f[2]))
                except ValueError:
                    continue
                for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                    if val.lower().startswith("active"):
                        reasons.add(name)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            # median over the samples taken under load (clock above idle)
            busy = [x for x in sm if x > 0.5 * max(sm)] or sm
            out.update(sm_mhz=float(np.median(busy)), sm_max_mhz=float(max(smax)), reasons=sorted(reasons), samples=len(sm))
        return out


def cpu_reference_run(spec, offs, a, threads, want_outputs=False):
    """The reference's mm_chain_dp over the sample under its kt_for (oracle/_ref), else the oracle port."""
    import oracle
    prm = oracle.params(spec.preset)
    if oracle.have_ref():
        res = oracle.ref_chain_batch(prm, offs, a, n_threads=threads, use_kalloc=True, keep=want_outputs)
        return res, "reference", threads
    t0 = time.perf_counter()
    res = oracle.chain_batch(prm, offs, a)
    res.seconds = time.perf_counter() - t0
    return res, "port", 1


def count_pairs_cpu(spec, offs, a, threads):
    """Metric numerator for the CPU arm: reference-visited pairs, counted by the instrumented oracle (not timed)."""
    import oracle
    from concurrent.futures import ThreadPoolExecutor
    prm = oracle.params(spec.preset)
    nr = offs.shape[0] - 1
    cuts = np.linspace(0, nr, min(threads, max(nr, 1)) + 1).astype(np.int64)

    def part(k):
        lo, hi = int(cuts[k]), int(cuts[k + 1])
        if hi <= lo:
            return 0
        o = offs[lo:hi + 1] - offs[lo]
        return oracle.count_visited(prm, o, a[offs[lo]:offs[hi]])

    with ThreadPoolExecutor(max_workers=threads) as ex:
        return int(sum(ex.map(part, range(len(cuts) - 1))))


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on this box's host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from minimap2_chaindp_b200 import workload
    import __graft_entry__ as ge
    ge.build()
    spec = workload.CONFIGS[args.workload]
    n_sample = min(args.cpu_reads, spec.n_reads)
    offs, a = workload.generate(spec, 0, n_sample)
    threads = host_cores()
    pairs = count_pairs_cpu(spec, offs, a, threads)
    kind = "reference"
    for _ in range(args.warmup):
        _, kind, used = cpu_reference_run(spec, offs, a, threads)
    secs = []
    for _ in range(args.steps):
        res, kind, used = cpu_reference_run(spec, offs, a, threads)
        secs.append(res.seconds)
    total = float(sum(secs))
    value = pairs * args.steps / total
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int32", "data": "synthetic",
        "config": {"workload": f"{spec.name}: G-anc synthetic anchors, preset {spec.preset}", "reads_per_step": int(n_sample),
                   "anchors_per_step": int(offs[-1]), "note": "CPU arm: bounded sample of the same read stream"},
        "reads_per_s": n_sample * args.steps / total,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": used, "kind": kind,
                         "sample": f"first {n_sample} reads of {spec.name} ({int(offs[-1])} anchors, {pairs} visited pairs) per step; "
                                   f"reference mm_chain_dp under its kt_for, one kalloc arena per thread"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="map-ont-100k")
    ap.add_argument("--reads", type=int, default=0, help="override reads per GPU (default: the config's)")
    ap.add_argument("--cpu-reads", type=int, default=CPU_SAMPLE_READS)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--slot-anchors", type=int, default=0)
    ap.add_argument("--slots", type=int, default=0)
    ap.add_argument("--wire", action="store_true", help="e2e: upload through the host batcher's 8-byte wire format (default: raw mm128_t)")
    ap.add_argument("--wire-threads", type=int, default=0, help="e2e: host encoder threads (0 = library default)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    from minimap2_chaindp_b200 import api, workload
    import __graft_entry__ as ge

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if rank == 0:
        ge.build()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the chaining path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ["NCCL_DEBUG"] = "WARN"      # keep stdout to the one JSON line (NCCL is only used for the barrier / reductions)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        dist.barrier()
    dev = torch.device("cuda", local)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce_max(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def reduce_sum(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    # ---- this rank's shard of the read stream: reads [rank*n, (rank+1)*n), generated straight into pinned memory
    spec = workload.CONFIGS[args.workload]
    n_reads = args.reads or spec.n_reads
    first = rank * n_reads
    offs = workload.count(spec, first, n_reads)
    n_anchors = int(offs[-1])
    pin = api.PinnedArray((n_anchors, 2), np.uint64)
    workload.fill(spec, offs, pin.array, first)
    prm = api.params(spec.preset)

    ctx = api.ChainContext(device=local, n_slots=args.slots, slot_anchors=args.slot_anchors, profile=True)
    if args.wire:
        ctx.set_upload(True, args.wire_threads)
    stream = torch.cuda.current_stream()
    d_a = torch.from_numpy(pin.array.view(np.int64)).to(dev)
    d_offs = torch.from_numpy(offs).to(dev)
    torch.cuda.synchronize()

    def device_step():
        return ctx.chain_device(prm, n_reads, n_anchors, d_offs.data_ptr(), d_a.data_ptr(), stream.cuda_stream)

    def e2e_step():
        return ctx.chain_batch_raw(prm, n_reads, offs.ctypes.data, pin.ptr)

    # ---- warm-up (allocates workspaces and pinned result arenas outside the timed regions)
    for _ in range(max(args.warmup, 1)):
        out = device_step()
    torch.cuda.synchronize()
    for _ in range(max(args.warmup, 1)):
        res = e2e_step()
    import ctypes as C
    # metric numerator: reference-visited pairs, counted once by the instrumented kernel variant (untimed)
    with api.ChainContext(device=local, count_visited=True) as cctx:
        cres = cctx.chain_batch_raw(prm, n_reads, offs.ctypes.data, pin.ptr)
        pairs_step = int(cres.n_visited)
        lanes_step = int(cres.n_lanes)
    chains_step = int(np.frombuffer((C.c_char * 8).from_address(res.u_off + 8 * n_reads), dtype=np.int64)[0])
    canch_step = int(np.frombuffer((C.c_char * 8).from_address(res.b_off + 8 * n_reads), dtype=np.int64)[0])
    n_units = int(res.n_units)
    n_sub = int(res.n_sub_batches)

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()

    # ---- timed region 1: device-resident inputs, CUDA events on the launching stream
    ctx.stage_reset()
    launches0 = ctx.launch_count()
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    for _ in range(args.steps):
        device_step()
    ev1.record(stream)
    barrier()
    dev_ms = reduce_max(ev0.elapsed_time(ev1))
    launches = ctx.launch_count() - launches0
    stage = {s: ctx.stage_ms(s) for s in ("prep", "fill", "compact", "tail", "gather")}

    # ---- timed region 2: end to end through the C ABI with host buffers (H2D + kernels + D2H inside)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        res = e2e_step()
    torch.cuda.synchronize()
    e2e_s = reduce_max(time.perf_counter() - t0)
    barrier()
    clocks = sampler.stop() if rank == 0 else None

    total_pairs = reduce_sum(float(pairs_step))
    total_reads = reduce_sum(float(n_reads))
    total_anchors = reduce_sum(float(n_anchors))
    value = total_pairs * args.steps / (dev_ms * 1e-3)
    e2e_value = total_pairs * args.steps / e2e_s
    h2d, wire_sub = ctx.upload_stats()      # bytes the last e2e step copied host -> device (counted by the library)
    d2h = n_reads * (4 + 8 + 8) + chains_step * 8 + canch_step * 16 + n_sub * 64

    # ---- rooflines of the dominant kernel (k_fill): per-launch averages from the events around it
    fill_ms, fill_launches = stage["fill"]
    fill_ms_per_step = fill_ms / max(args.steps, 1)
    hbm_peak, peak_src = peaks()
    hbm_achieved = n_anchors * BYTES_PER_ANCHOR / (fill_ms_per_step * 1e-3) / 1e9 if fill_ms_per_step > 0 else 0.0
    int_peak_mix = ctx.int_peak(1)
    int_peak_alu = ctx.int_peak(0)
    int_achieved = pairs_step * INT_OPS_PER_PAIR / (fill_ms_per_step * 1e-3) if fill_ms_per_step > 0 else 0.0

    traffic = None   # dram bytes of one k_fill launch from the committed ncu capture of this exact workload, else null
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "r01_fill_traffic.json")))
        if tj.get("workload") == spec.name and int(tj.get("reads", 0)) == n_reads:
            traffic = int(tj["dram_bytes_per_launch"])
    except Exception:
        traffic = None

    line = None
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 1),
            "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "int32", "data": "synthetic",
            "config": {"workload": f"{spec.name}: G-anc synthetic anchors, preset {spec.preset} "
                                   f"(max_dist {prm.max_dist_x}, bw {prm.bw}, max_skip {prm.max_skip}, max_iter unbounded = reference)",
                       "reads_per_gpu": n_reads, "anchors_per_gpu": n_anchors, "visited_pairs_per_gpu": pairs_step,
                       "runs_per_gpu": n_units, "sharding": f"reads x{world} (one process per GPU, no collective)",
                       "l2": f"inputs larger than L2 ({n_anchors * 16 / 2**20:.0f} MiB of anchors + workspace vs 126 MB)",
                       "sub_batches_e2e": n_sub},
            "reads_per_s": total_reads * args.steps / (dev_ms * 1e-3),
            "anchors_per_s": total_anchors * args.steps / (dev_ms * 1e-3),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "reads_per_s": total_reads * args.steps / e2e_s, "ms_per_step": 1e3 * e2e_s / args.steps,
                    "api": "mm2b200_chain_batch (pinned host anchors in, pinned host chains out)",
                    "upload": (f"host batcher wire format ({wire_sub}/{n_sub} sub-batches, {h2d / max(n_anchors, 1):.2f} B/anchor, "
                               f"encoded by host threads inside the timed region)") if wire_sub else "raw mm128_t (16 B/anchor)"},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"kernel": "k_fill", "bound": "hbm", "achieved": hbm_achieved, "peak": hbm_peak, "unit": "GB/s",
                         "frac": hbm_achieved / hbm_peak, "traffic": traffic, "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": n_anchors * BYTES_PER_ANCHOR, "ms_per_launch": fill_ms_per_step,
                         "note": "k_fill is integer-issue bound, not HBM bound (see roofline_issue); 40 B/anchor algorithmic"},
            "roofline_issue": {"kernel": "k_fill", "bound": "int-issue", "achieved": int_achieved / 1e12, "peak": int_peak_mix / 1e12,
                               "unit": "T lane-op/s", "frac": int_achieved / int_peak_mix if int_peak_mix else None,
                               "peak_alu_pipe_only": int_peak_alu / 1e12, "ops_per_pair": INT_OPS_PER_PAIR,
                               "lane_efficiency": pairs_step / lanes_step if lanes_step else None,
                               "peak_source": "measured live (mm2b200_int_peak: add/xor/mad.lo streams on all SMs)"},
            "stage_ms_per_step": {k: v[0] / max(args.steps, 1) for k, v in stage.items()},
        }
    # ---- CPU baseline beside it (rank 0, N=1 only): the reference on a bounded sample, and a parity spot check
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        n_sample = min(args.cpu_reads, n_reads)
        so = offs[:n_sample + 1].copy()
        sa = pin.array[:int(so[-1])]
        threads = host_cores()
        ref, kind, used = cpu_reference_run(spec, so, sa, threads, want_outputs=True)
        s_pairs = count_pairs_cpu(spec, so, sa, threads)
        u_off = np.frombuffer((C.c_char * (8 * (n_sample + 1))).from_address(res.u_off), dtype=np.int64)
        b_off = np.frombuffer((C.c_char * (8 * (n_sample + 1))).from_address(res.b_off), dtype=np.int64)
        gu = np.frombuffer((C.c_char * (8 * int(u_off[-1]))).from_address(res.u), dtype=np.uint64) if u_off[-1] else np.zeros(0, np.uint64)
        gb = np.frombuffer((C.c_char * (16 * int(b_off[-1]))).from_address(res.b), dtype=np.uint64).reshape(-1, 2) if b_off[-1] else np.zeros((0, 2), np.uint64)
        parity = bool(np.array_equal(u_off, ref.u_off) and np.array_equal(b_off, ref.b_off)
                      and np.array_equal(gu, ref.u) and np.array_equal(gb, ref.b))
        line["cpu_baseline"] = {"value": s_pairs / ref.seconds, "unit": UNIT, "cores": used, "kind": kind,
                                "reads_per_s": n_sample / ref.seconds,
                                "sample": f"first {n_sample} reads of {spec.name} ({int(so[-1])} anchors, {s_pairs} visited pairs), "
                                          f"reference mm_chain_dp under its kt_for, one kalloc arena per thread, 1 pass",
                                "parity_with_gpu_on_sample": parity}
        if not parity:
            line["parity_error"] = "GPU chains differ from the reference on the CPU-baseline sample"
    if rank == 0:
        print(json.dumps(line))
    ctx.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
