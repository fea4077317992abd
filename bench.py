#!/usr/bin/env python
"""bench.py -- chain_dp throughput on B200 (BASELINE.json metric), one JSON line on stdout.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --steps K --warmup W    # the reference's CPU mm_chain_dp on the host cores

A step = one pass of the hot path (mm_chain_dp over every read of the workload: prep, fill, compaction,
backtrack, gather) over one batch of synthetic reads.

Headline (the line's top-level keys): BASELINE.json configs[1] "map-ont: 100k synthetic 10 kb Nanopore-like reads".
For N>1 every rank owns its own 100k reads (weak scaling, read-batch sharding, no collective on the data path).

* value : visited anchor pairs / s, whole job, anchors already resident in HBM, CUDA events, max over ranks
* e2e   : the same through the C ABI call a reference-side batcher makes (mm2b200_chain_batch) with pinned HOST
          buffers: H2D of the anchors, all kernels, D2H of the chains inside the timed region; plus the pageable-memory
          variant, the PCIe ceiling of the same byte volume, rank 0 alone (for an e2e scaling efficiency) and -- at
          N>1 -- `e2e_multi`: ONE host process driving all N GPUs through mm2b200_chain_batch_multi
* roofline : the fill kernel against the measured integer-issue peak (the binding roofline); roofline_hbm: the
          same kernel against the measured copy bandwidth (secondary: the kernel is not HBM bound)
* cpu_baseline : the reference's own chain.c (oracle/_ref) under its kt_for on all host cores, SAME 100k reads
* configs : the other BASELINE.json configs (ava-ont-50k, map-pb-ul-20k, asm5-50mb), each a FIXED workload split over
          the N ranks (strong scaling), with value, e2e, fill roofline, stage times and parity against the compiled
          reference on every read of a sample

"visited pair" = one execution of the reference's inner loop body (chain.c:59-88); the count comes from the
device (mm2b200_result_t.n_visited) and is cross-checked against the oracle on the CPU-baseline sample.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import struct
import sys
import tempfile
import time

import numpy as np

os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")   # before CUDA initialises (see chain_api.cu: sub-batch streams)
if os.environ.get("OMP_NUM_THREADS") == "1" and "LOCAL_WORLD_SIZE" in os.environ:
    # torchrun's default of one OpenMP thread per rank would serialise the (untimed) workload generator
    try:
        _cores = len(os.sched_getaffinity(0))
    except Exception:
        _cores = os.cpu_count() or 1
    os.environ["OMP_NUM_THREADS"] = str(max(1, _cores // max(int(os.environ.get("LOCAL_WORLD_SIZE", "1")), 1)))

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "chain_dp_anchor_pair_evals_per_sec"
UNIT = "pairs/s"
INT_OPS_PER_PAIR = 28          # integer ops the algorithm needs per visited pair
BYTES_PER_ANCHOR = 40          # 16 B anchor in + 16 B anchor out + 8 B f,p

# the other BASELINE.json configs: reads per chunk (bounds pinned + device memory) and reads in the parity sample
CONFIG_BLOCK = {
    "ava-ont-50k": dict(chunk_reads=8000, parity_reads=2000),
    "map-pb-ul-20k": dict(chunk_reads=6000, parity_reads=2000),
    "asm5-50mb": dict(chunk_reads=8, parity_reads=8),
}


def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def numa_nodes() -> int:
    try:
        return len([d for d in os.listdir("/sys/devices/system/node") if d.startswith("node") and d[4:].isdigit()])
    except Exception:
        return 1


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as fh:
            d = json.load(fh)
        return float(d.get("hbm_gbs", 6650.0)), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 50 ms while the timed regions run."""
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
            busy = [x for x in sm if x > 0.5 * max(sm)] or sm     # samples taken under load (clock above idle)
            out.update(sm_mhz=float(np.median(busy)), sm_max_mhz=float(max(smax)), reasons=sorted(reasons), samples=len(sm))
        return out


# ---------------------------------------------------------------------------------------------------------
# CPU arm: the reference's own mm_chain_dp under its kt_for (oracle/_ref), else the oracle port
# ---------------------------------------------------------------------------------------------------------
def cpu_reference_timed(spec, offs, a, threads, passes=3, warmup=1, want_outputs=False):
    """Best of `passes` after `warmup` untimed passes, for both of the reference's memory modes:
    "malloc" (km = NULL, the authors' --no-kalloc run, run.sh:3; inputs copied per call OUTSIDE the timed region, as
    BASELINE.md section 3 states) and "kalloc" (one km_init() arena per thread; the copy into the arena is inside the
    timed region because the arena belongs to whichever worker steals the read).  Returns the faster one."""
    import oracle
    prm = oracle.params(spec.preset)
    if not oracle.have_ref():
        t0 = time.perf_counter()
        res = oracle.chain_batch(prm, offs, a)
        res.seconds = time.perf_counter() - t0
        return res, dict(kind="port", cores=1, mode="oracle restatement, 1 thread", seconds=res.seconds, all={})
    best, best_mode, allm, keep = None, None, {}, None
    for mode, use_kalloc in (("malloc", False), ("kalloc", True)):
        secs = []
        for k in range(warmup + passes):
            res = oracle.ref_chain_batch(prm, offs, a, n_threads=threads, use_kalloc=use_kalloc,
                                         keep=want_outputs and k == warmup + passes - 1)
            if k >= warmup:
                secs.append(res.seconds)
        allm[mode] = min(secs)
        if best is None or min(secs) < best:
            best, best_mode, keep = min(secs), mode, res
    keep.seconds = best
    return keep, dict(kind="reference", cores=threads, mode=best_mode, seconds=best, all=allm)


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
    """--impl reference: the reference's CPU implementation of the path on this box's host cores, on the SAME reads as
    the b200 arm's headline (rank 0's 100k map-ont reads).  Loads only the workload generator and the oracle."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from minimap2_chaindp_b200 import build as pkg_build, workload
    import oracle
    pkg_build.build_workload()
    oracle.build()
    spec = workload.CONFIGS[args.workload]
    n_reads = args.reads or spec.n_reads
    offs, a = workload.generate(spec, 0, n_reads)
    threads = host_cores()
    pairs = count_pairs_cpu(spec, offs, a, threads)
    _, info = cpu_reference_timed(spec, offs, a, threads, passes=max(args.steps, 1), warmup=max(args.warmup, 1))
    secs = info["seconds"]
    value = pairs / secs
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * secs, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int32", "data": "synthetic",
        "config": {"workload": f"{spec.name}: G-anc synthetic anchors, preset {spec.preset}", "reads_per_step": int(n_reads),
                   "anchors_per_step": int(offs[-1]), "visited_pairs_per_step": int(pairs),
                   "note": "CPU arm: the same reads as the b200 arm's rank 0; a rate, so comparable at every N"},
        "reads_per_s": n_reads / secs,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": info["cores"], "kind": info["kind"],
                         "sample": f"all {n_reads} reads of {spec.name} ({int(offs[-1])} anchors, {pairs} visited pairs) per step; "
                                   f"reference mm_chain_dp under its kt_for; best step of {max(args.steps, 1)} after {max(args.warmup, 1)} warm-up; "
                                   f"memory mode '{info['mode']}' (faster of malloc/pre-copied inputs and kalloc arenas: "
                                   + ", ".join(f"{k} {v * 1e3:.0f} ms" for k, v in info["all"].items()) + ")"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))
    return 0


# ---------------------------------------------------------------------------------------------------------
# b200 arm
# ---------------------------------------------------------------------------------------------------------
def result_arrays(res, n_reads):
    """numpy views of a mm2b200_result_t (context-owned pinned memory; valid until the next call on the context)."""
    u_off = np.frombuffer((C.c_char * (8 * (n_reads + 1))).from_address(res.u_off), dtype=np.int64)
    b_off = np.frombuffer((C.c_char * (8 * (n_reads + 1))).from_address(res.b_off), dtype=np.int64)
    gu = np.frombuffer((C.c_char * (8 * int(u_off[-1]))).from_address(res.u), dtype=np.uint64) if u_off[-1] else np.zeros(0, np.uint64)
    gb = (np.frombuffer((C.c_char * (16 * int(b_off[-1]))).from_address(res.b), dtype=np.uint64).reshape(-1, 2)
          if b_off[-1] else np.zeros((0, 2), np.uint64))
    return u_off, b_off, gu, gb


def parity_against_reference(spec, offs, a, res, n_sample, threads):
    """Every read of the first n_sample reads: GPU chains (u, b, offsets) == compiled reference under kt_for."""
    so = offs[:n_sample + 1].copy()
    sa = a[:int(so[-1])]
    ref, info = cpu_reference_timed(spec, so, sa, threads, passes=1, warmup=0, want_outputs=True)
    u_off, b_off, gu, gb = result_arrays(res, offs.shape[0] - 1)
    nu, nb = int(u_off[n_sample]), int(b_off[n_sample])
    ok = bool(np.array_equal(u_off[:n_sample + 1], ref.u_off) and np.array_equal(b_off[:n_sample + 1], ref.b_off)
              and np.array_equal(gu[:nu], ref.u) and np.array_equal(gb[:nb], ref.b))
    return ok, ref, info, so, sa


class Bench:
    def __init__(self, args):
        import torch
        import torch.distributed as dist
        from minimap2_chaindp_b200 import api, workload
        self.torch, self.dist, self.api, self.workload = torch, dist, api, workload
        self.args = args
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device -- the chaining path has no CPU fallback (use --impl reference for the CPU arm)")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        self.cpu_group = None
        if self.world > 1:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            # NCCL_DEBUG is left to the driver (it reads NCCL's rank lines); unset, NCCL prints nothing to stdout
            dist.init_process_group("nccl", device_id=self.dev)
            # a CPU-side group: ranks that only WAIT (while rank 0 drives their GPUs in the single-process legs) must not
            # sit in an NCCL kernel on the GPU they are lending
            self.cpu_group = dist.new_group(backend="gloo")
            dist.barrier()
        self.stream = torch.cuda.current_stream()

    # -- collectives (NCCL carries nothing but these scalars and the barrier) --
    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def cpu_barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier(group=self.cpu_group)

    def _reduce(self, x, op):
        if self.world == 1:
            return float(x)
        t = self.torch.tensor([x], dtype=self.torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=op)
        return float(t.item())

    def rmax(self, x):
        return self._reduce(x, self.dist.ReduceOp.MAX if self.world > 1 else None)

    def rsum(self, x):
        return self._reduce(x, self.dist.ReduceOp.SUM if self.world > 1 else None)

    # -- one chunk of reads, resident in pinned memory --------------------------------------------------------
    def measure_chunk(self, ctx, cctx, prm, offs, pin, steps, warmup, e2e_steps):
        """Device-resident passes (CUDA events) and end-to-end passes (host clock) over one chunk.  Returns a dict of
        totals over the timed steps plus the last e2e result struct."""
        torch, api = self.torch, self.api
        n_reads, n_anchors = offs.shape[0] - 1, int(offs[-1])
        d_a = torch.from_numpy(pin.array[:n_anchors].view(np.int64)).to(self.dev)
        d_offs = torch.from_numpy(offs).to(self.dev)
        torch.cuda.synchronize()
        st = self.stream.cuda_stream
        for _ in range(max(warmup, 1)):
            ctx.chain_device(prm, n_reads, n_anchors, d_offs.data_ptr(), d_a.data_ptr(), st)
        torch.cuda.synchronize()
        res = None
        for _ in range(max(min(warmup, 1), 1)):
            res = ctx.chain_batch_raw(prm, n_reads, offs.ctypes.data, pin.ptr)
        # metric numerator: reference-visited pairs, counted once by the instrumented kernel variant (untimed)
        cres = cctx.chain_batch_raw(prm, n_reads, offs.ctypes.data, pin.ptr)
        out = dict(reads=n_reads, anchors=n_anchors, pairs=int(cres.n_visited), lanes=int(cres.n_lanes),
                   units=int(res.n_units), sub_batches=int(res.n_sub_batches))
        ctx.stage_reset()
        l0 = ctx.launch_count()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        ev0.record(self.stream)
        for _ in range(steps):
            ctx.chain_device(prm, n_reads, n_anchors, d_offs.data_ptr(), d_a.data_ptr(), st)
        ev1.record(self.stream)
        torch.cuda.synchronize()
        out["dev_ms"] = float(ev0.elapsed_time(ev1))
        out["launches"] = ctx.launch_count() - l0
        # per-kernel durations: one more pass per step with every kernel alone on the stream (see run())
        ctx.set_split(False)
        ctx.stage_reset()
        for _ in range(steps):
            ctx.chain_device(prm, n_reads, n_anchors, d_offs.data_ptr(), d_a.data_ptr(), st)
        torch.cuda.synchronize()
        out["stage"] = {s: ctx.stage_ms(s)[0] for s in ("prep", "fill", "compact", "tail", "gather")}
        ctx.set_split(True)
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            res = ctx.chain_batch_raw(prm, n_reads, offs.ctypes.data, pin.ptr)
        torch.cuda.synchronize()
        out["e2e_s"] = time.perf_counter() - t0
        h2d, wire_sub = ctx.upload_stats()
        u_tot = int(np.frombuffer((C.c_char * 8).from_address(res.u_off + 8 * n_reads), dtype=np.int64)[0])
        b_tot = int(np.frombuffer((C.c_char * 8).from_address(res.b_off + 8 * n_reads), dtype=np.int64)[0])
        out.update(h2d=int(h2d), wire_sub=int(wire_sub), chains=u_tot, chain_anchors=b_tot,
                   d2h=n_reads * (4 + 8 + 8) + u_tot * 8 + b_tot * 16 + int(res.n_sub_batches) * 64, res=res)
        del d_a, d_offs
        return out

    # -- a fixed workload split over the ranks (strong scaling), processed in chunks ----------------------------
    def run_config(self, name, steps, int_peak):
        api, workload = self.api, self.workload
        spec = workload.CONFIGS[name]
        cb = CONFIG_BLOCK[name]
        prm = api.params(spec.preset)
        total_reads = spec.n_reads
        lo = total_reads * self.rank // self.world
        hi = total_reads * (self.rank + 1) // self.world
        acc = dict(reads=0, anchors=0, pairs=0, lanes=0, dev_ms=0.0, e2e_s=0.0, launches=0, h2d=0, d2h=0, chunks=0,
                   stage={s: 0.0 for s in ("prep", "fill", "compact", "tail", "gather")})
        parity = cpu = None
        ctx = api.ChainContext(device=self.local, profile=True)
        cctx = api.ChainContext(device=self.local, count_visited=True)
        try:
            first = lo
            while first < hi:
                n = min(cb["chunk_reads"], hi - first)
                offs = workload.count(spec, first, n)
                pin = api.PinnedArray((max(int(offs[-1]), 1), 2), np.uint64)
                workload.fill(spec, offs, pin.array, first)
                m = self.measure_chunk(ctx, cctx, prm, offs, pin, steps, 1, 1)
                for k in ("reads", "anchors", "pairs", "lanes", "launches", "h2d", "d2h"):
                    acc[k] += m[k]
                acc["dev_ms"] += m["dev_ms"] / steps
                acc["e2e_s"] += m["e2e_s"]
                for s in acc["stage"]:
                    acc["stage"][s] += m["stage"][s] / steps
                acc["chunks"] += 1
                if parity is None and self.rank == 0:
                    # every read of the sample against the compiled reference under kt_for (same pinned anchors)
                    ns = min(cb["parity_reads"], n)
                    ok, ref, info, so, sa = parity_against_reference(spec, offs, pin.array, m["res"], ns, host_cores())
                    s_pairs = count_pairs_cpu(spec, so, sa, host_cores())
                    parity = {"ok": ok, "reads": int(ns), "anchors": int(so[-1])}
                    cpu = {"value": s_pairs / ref.seconds, "unit": UNIT, "cores": info["cores"], "kind": info["kind"],
                           "reads_per_s": ns / ref.seconds, "sample": f"first {ns} reads ({int(so[-1])} anchors), 1 pass, mode {info['mode']}"}
                pin.free()
                first += n
        finally:
            ctx.close()
            cctx.close()
        # whole-job figures: every rank's share of the fixed workload, time = max over ranks
        dev_ms = self.rmax(acc["dev_ms"])
        e2e_s = self.rmax(acc["e2e_s"])
        pairs, reads, anchors = self.rsum(acc["pairs"]), self.rsum(acc["reads"]), self.rsum(acc["anchors"])
        fill_ms = acc["stage"]["fill"]
        frac = (acc["pairs"] * INT_OPS_PER_PAIR / (fill_ms * 1e-3) / int_peak) if fill_ms > 0 and int_peak else None
        if self.rank != 0:
            return None
        return {
            "workload": f"{spec.name}: G-anc synthetic anchors, preset {spec.preset}; the FIXED {total_reads} reads split over {self.world} rank(s)",
            "scaling": "strong", "reads": int(reads), "anchors": int(anchors), "visited_pairs": int(pairs),
            "value": pairs / (dev_ms * 1e-3), "unit": UNIT, "ms_per_pass": dev_ms, "reads_per_s": reads / (dev_ms * 1e-3),
            "e2e": {"value": pairs / e2e_s, "unit": UNIT, "ms_per_pass": 1e3 * e2e_s, "reads_per_s": reads / e2e_s,
                    "h2d_bytes_rank0": acc["h2d"], "d2h_bytes_rank0": acc["d2h"]},
            "roofline": {"kernel": "k_fill", "bound": "int-issue", "frac": frac, "fill_ms_rank0": fill_ms,
                         "lane_efficiency": acc["pairs"] / acc["lanes"] if acc["lanes"] else None},
            "stage_ms_rank0": acc["stage"], "chunks_rank0": acc["chunks"], "steps_per_chunk": steps,
            "parity_with_reference": parity, "cpu_baseline": cpu,
        }

    # -- PCIe ceiling of the headline's byte volume: plain copies, every rank at once -----------------------------
    def pcie_ceiling_ms(self, pin, n_anchors, d2h_bytes, reps=3):
        torch = self.torch
        d_in = torch.empty(n_anchors * 2, dtype=torch.int64, device=self.dev)
        d_out = torch.empty(max(d2h_bytes // 8, 1), dtype=torch.int64, device=self.dev)
        h_in = torch.from_numpy(pin.array[:n_anchors].view(np.int64).reshape(-1))
        h_out = torch.empty(max(d2h_bytes // 8, 1), dtype=torch.int64).pin_memory()
        s_up, s_dn = torch.cuda.Stream(), torch.cuda.Stream()
        best = None
        for _ in range(reps + 1):
            self.barrier()
            t0 = time.perf_counter()
            with torch.cuda.stream(s_up):
                d_in.copy_(h_in, non_blocking=True)
            with torch.cuda.stream(s_dn):
                h_out.copy_(d_out, non_blocking=True)
            torch.cuda.synchronize()
            dt = self.rmax(time.perf_counter() - t0)
            best = dt if best is None or dt < best else best
        return 1e3 * best

    # -- the headline ------------------------------------------------------------------------------------------
    def run(self):
        args, api, workload, torch = self.args, self.api, self.workload, self.torch
        rank, world, local = self.rank, self.world, self.local
        spec = workload.CONFIGS[args.workload]
        n_reads = args.reads or spec.n_reads
        first = rank * n_reads
        offs = workload.count(spec, first, n_reads)
        n_anchors = int(offs[-1])
        pin = api.PinnedArray((n_anchors, 2), np.uint64)       # generated straight into pinned memory
        workload.fill(spec, offs, pin.array, first)
        prm = api.params(spec.preset)
        ctx = api.ChainContext(device=local, n_slots=args.slots, slot_anchors=args.slot_anchors, profile=True)
        cctx = api.ChainContext(device=local, count_visited=True)
        if args.wire:
            ctx.set_upload(True, args.wire_threads)
        steps, warm = args.steps, max(args.warmup, 1)

        d_a = torch.from_numpy(pin.array.view(np.int64)).to(self.dev)
        d_offs = torch.from_numpy(offs).to(self.dev)
        torch.cuda.synchronize()
        st = self.stream.cuda_stream

        def device_step():
            return ctx.chain_device(prm, n_reads, n_anchors, d_offs.data_ptr(), d_a.data_ptr(), st)

        def e2e_step():
            return ctx.chain_batch_raw(prm, n_reads, offs.ctypes.data, pin.ptr)

        # warm-up (allocates workspaces and pinned result arenas outside the timed regions)
        for _ in range(warm):
            device_step()
        torch.cuda.synchronize()
        for _ in range(warm):
            res = e2e_step()
        cres = cctx.chain_batch_raw(prm, n_reads, offs.ctypes.data, pin.ptr)
        pairs_step, lanes_step = int(cres.n_visited), int(cres.n_lanes)
        cctx.close()
        chains_step = int(np.frombuffer((C.c_char * 8).from_address(res.u_off + 8 * n_reads), dtype=np.int64)[0])
        canch_step = int(np.frombuffer((C.c_char * 8).from_address(res.b_off + 8 * n_reads), dtype=np.int64)[0])
        n_units, n_sub = int(res.n_units), int(res.n_sub_batches)
        d2h = n_reads * (4 + 8 + 8) + chains_step * 8 + canch_step * 16 + n_sub * 64

        sampler = ClockSampler(local)
        if rank == 0:
            sampler.start()

        # ---- timed region 1: device-resident inputs, CUDA events on the launching stream
        ctx.stage_reset()
        launches0 = ctx.launch_count()
        self.barrier()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record(self.stream)
        for _ in range(steps):
            device_step()
        ev1.record(self.stream)
        self.barrier()
        dev_ms = self.rmax(ev0.elapsed_time(ev1))
        launches = ctx.launch_count() - launches0
        stage_overlapped = {s: ctx.stage_ms(s) for s in ("prep", "fill", "compact", "tail", "gather")}
        # per-kernel durations for the roofline: the same passes with every kernel ALONE on the stream (the timed region
        # above runs the two halves of the batch on two streams, where the durations of concurrent kernels overlap)
        ctx.set_split(False)
        device_step()
        torch.cuda.synchronize()
        ctx.stage_reset()
        ev2, ev3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev2.record(self.stream)
        for _ in range(steps):
            device_step()
        ev3.record(self.stream)
        torch.cuda.synchronize()
        unsplit_ms = ev2.elapsed_time(ev3) / steps
        stage = {s: ctx.stage_ms(s) for s in ("prep", "fill", "compact", "tail", "gather")}
        ctx.set_split(True)

        # ---- timed region 2: end to end through the C ABI with pinned host buffers (H2D + kernels + D2H inside), all ranks at once
        # The W warm-up steps of this region run again right in front of it: the device-resident passes above leave the host link
        # idle for a while, and on a fresh box the first transfers after that were seen to run at half speed (one cold run: 40 ms
        # per step where every later one gave 27-28).  Then up to 8 more untimed steps until one is within 5 % of the best seen.
        e2e_warm = 0
        best_w = None
        for k in range(warm + 8):
            tw = time.perf_counter()
            res = e2e_step()
            tw = time.perf_counter() - tw
            e2e_warm += 1
            settled = best_w is not None and tw <= 1.05 * best_w
            best_w = tw if best_w is None else min(best_w, tw)
            if k + 1 >= warm and self.rmax(0.0 if settled else 1.0) == 0.0:
                break
        torch.cuda.synchronize()
        self.barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            res = e2e_step()
        torch.cuda.synchronize()
        e2e_s = self.rmax(time.perf_counter() - t0)
        self.barrier()
        clocks = sampler.stop() if rank == 0 else None
        h2d, wire_sub = ctx.upload_stats()      # bytes the last e2e step copied host -> device (counted by the library)

        # ---- e2e side legs (untimed for the headline): rank 0 alone, pageable input, the PCIe ceiling
        solo_s = page_s = regs_s = None
        n_regs_out = 0
        self.cpu_barrier()
        if rank == 0:
            t0 = time.perf_counter()
            for _ in range(steps):
                e2e_step()
            torch.cuda.synchronize()
            solo_s = (time.perf_counter() - t0) / steps
            pageable = np.array(pin.array, copy=True)          # ordinary (pageable) host memory, what kalloc hands a caller
            ctx.chain_batch_raw(prm, n_reads, offs.ctypes.data, pageable.ctypes.data)
            t0 = time.perf_counter()
            res = ctx.chain_batch_raw(prm, n_reads, offs.ctypes.data, pageable.ctypes.data)
            torch.cuda.synchronize()
            page_s = time.perf_counter() - t0
            del pageable
            res = e2e_step()
            # regions instead of chains (row f3): mm2b200_chain_regs_batch, same pinned anchors in, mm_reg1_t out
            from minimap2_chaindp_b200 import post as mpost
            hashes = ((np.arange(n_reads, dtype=np.uint64) * np.uint64(2654435761)) & np.uint64(0xffffffff)).astype(np.uint32)
            qlens = np.full(n_reads, 12000, np.int32)
            pp = mpost.post_params(spec.preset if spec.preset in mpost.POST_PRESETS else "map-ont")
            rres = mpost._Regs()

            def regs_step():
                rc = mpost._lib().mm2b200_chain_regs_batch(ctx._h, C.byref(prm), C.byref(pp), n_reads, offs.ctypes.data, pin.ptr, hashes.ctypes.data,
                                                           qlens.ctypes.data, 0, C.byref(rres))
                if rc:
                    raise api.ChainError(rc, api.last_error())
            regs_step(); regs_step()
            t0 = time.perf_counter()
            for _ in range(steps):
                regs_step()
            regs_s = (time.perf_counter() - t0) / steps
            n_regs_out = int(np.frombuffer((C.c_char * 8).from_address(rres.reg_off + 8 * n_reads), np.int64)[0])
            res = e2e_step()
        self.cpu_barrier()
        pcie_ms = self.pcie_ceiling_ms(pin, n_anchors, d2h)

        total_pairs = self.rsum(float(pairs_step))
        total_reads = self.rsum(float(n_reads))
        total_anchors = self.rsum(float(n_anchors))
        value = total_pairs * steps / (dev_ms * 1e-3)
        e2e_value = total_pairs * steps / e2e_s

        # ---- rooflines of the dominant kernel (k_fill): per-launch averages from the events around it
        fill_ms, _ = stage["fill"]
        fill_ms_per_step = fill_ms / max(steps, 1)
        hbm_peak, peak_src = peaks()
        hbm_achieved = n_anchors * BYTES_PER_ANCHOR / (fill_ms_per_step * 1e-3) / 1e9 if fill_ms_per_step > 0 else 0.0
        int_peak_mix = ctx.int_peak(1)
        int_peak_alu = ctx.int_peak(0)
        int_achieved = pairs_step * INT_OPS_PER_PAIR / (fill_ms_per_step * 1e-3) if fill_ms_per_step > 0 else 0.0
        traffic, traffic_src = None, "not captured for this workload size"
        try:
            tj = json.load(open(os.path.join(ROOT, "profiles", "r02_fill_traffic.json")))
            if tj.get("workload") == spec.name and int(tj.get("reads", 0)) == n_reads:
                traffic, traffic_src = int(tj["dram_bytes_per_launch"]), tj.get("source", "profiles/r02_fill_traffic.json")
        except Exception:
            pass

        # ---- e2e_multi (N>1): ONE host process (rank 0) drives all N GPUs through mm2b200_chain_batch_multi
        e2e_multi = None
        if world > 1:
            e2e_multi = self.run_multi(spec, prm, n_reads, steps, solo_s)

        line = None
        if rank == 0:
            solo_value = pairs_step / solo_s
            line = {
                "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warm,
                "ms_per_step": dev_ms / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "int32", "data": "synthetic",
                "config": {"workload": f"{spec.name}: G-anc synthetic anchors, preset {spec.preset} "
                                       f"(max_dist {prm.max_dist_x}, bw {prm.bw}, max_skip {prm.max_skip}, max_iter unbounded = reference)",
                           "reads_per_gpu": n_reads, "anchors_per_gpu": n_anchors, "visited_pairs_per_gpu": pairs_step,
                           "runs_per_gpu": n_units, "sharding": f"reads x{world} (one process per GPU, no collective)",
                           "l2": f"inputs larger than L2 ({n_anchors * 16 / 2**20:.0f} MiB of anchors + workspace vs 126 MB)",
                           "sub_batches_e2e": n_sub,
                           "host": {"cores": host_cores(), "numa_nodes": numa_nodes(),
                                    "affinity": "no-op: single NUMA node" if numa_nodes() <= 1 else "not pinned"}},
                "reads_per_s": total_reads * steps / (dev_ms * 1e-3),
                "anchors_per_s": total_anchors * steps / (dev_ms * 1e-3),
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                        "reads_per_s": total_reads * steps / e2e_s, "ms_per_step": 1e3 * e2e_s / steps,
                        "api": "mm2b200_chain_batch (pinned host anchors in, pinned host chains out), one process per GPU",
                        "upload": (f"host batcher wire format ({wire_sub}/{n_sub} sub-batches, {h2d / max(n_anchors, 1):.2f} B/anchor, "
                                   f"encoded by host threads inside the timed region)") if wire_sub else "raw mm128_t (16 B/anchor)",
                        "warmup_steps": e2e_warm,
                        "warmup_note": "untimed e2e steps run directly in front of the timed ones (W, then until a step is within 5 % of the best seen, at most W + 8)",
                        "pcie_ceiling_ms": pcie_ms,
                        "pcie_ceiling_note": "plain H2D of the anchors + D2H of the result bytes on two streams, all ranks at once, max over ranks",
                        "rank0_alone_ms_per_step": 1e3 * solo_s,
                        "scaling_efficiency": e2e_value / (world * solo_value),
                        "pageable_ms_per_step": 1e3 * page_s,
                        "pageable_note": "anchors in ordinary host memory: the library stages every sub-batch through its own pinned buffer",
                        "regs": {"api": "mm2b200_chain_regs_batch (pinned host anchors in, mm_reg1_t out: mm_chain_dp + mm_gen_regs + chain_post on the device), rank 0 alone",
                                 "ms_per_step": 1e3 * regs_s, "reads_per_s": n_reads / regs_s, "value": pairs_step / regs_s, "unit": UNIT,
                                 "d2h_bytes_per_step": n_regs_out * 80 + (n_reads + 1) * 8, "regions": n_regs_out}},
                "gpu_launches": int(launches),
                "clocks": clocks,
                "roofline": {"kernel": "k_fill", "bound": "int-issue", "achieved": int_achieved / 1e12, "peak": int_peak_mix / 1e12,
                             "unit": "T lane-op/s", "frac": int_achieved / int_peak_mix if int_peak_mix else None,
                             "traffic": traffic, "traffic_source": traffic_src,
                             "peak_alu_pipe_only": int_peak_alu / 1e12, "ops_per_pair": INT_OPS_PER_PAIR,
                             "lane_efficiency": pairs_step / lanes_step if lanes_step else None, "ms_per_launch": fill_ms_per_step,
                             "peak_source": "measured live (mm2b200_int_peak: add/xor/mad.lo streams on all SMs)",
                             "note": "the binding roofline of this path: 28 integer ops per reference-visited pair (SURVEY 8d) against the issue port"},
                "roofline_hbm": {"kernel": "k_fill", "bound": "hbm", "achieved": hbm_achieved, "peak": hbm_peak, "unit": "GB/s",
                                 "frac": hbm_achieved / hbm_peak, "peak_source": peak_src,
                                 "algorithmic_bytes_per_launch": n_anchors * BYTES_PER_ANCHOR,
                                 "note": "secondary: 40 B/anchor algorithmic; the kernel is not HBM bound"},
                "stage_ms_per_step": {k: v[0] / max(steps, 1) for k, v in stage.items()},
                "stage_note": ("stage_ms_per_step, roofline.ms_per_launch: CUDA events around each kernel in an extra pass with every "
                               "kernel alone on the stream (ms_per_step_unsplit); the timed region of `value` runs the two halves of the "
                               "batch on two streams (stage_ms_overlapped: the same events there, durations of concurrent kernels overlap)"),
                "ms_per_step_unsplit": unsplit_ms,
                "stage_ms_overlapped": {k: v[0] / max(steps, 1) for k, v in stage_overlapped.items()},
            }
            if e2e_multi is not None:
                line["e2e_multi"] = e2e_multi
        # ---- CPU baseline beside it (rank 0, N=1 only): the reference on the SAME reads, and parity on every read
        if rank == 0 and world == 1 and not args.no_cpu_baseline:
            threads = host_cores()
            ok, ref, info, so, sa = parity_against_reference(spec, offs, pin.array, res, n_reads, threads)
            ref_t, info_t = cpu_reference_timed(spec, so, sa, threads, passes=3, warmup=1)
            s_pairs = count_pairs_cpu(spec, so, sa, threads)
            line["cpu_baseline"] = {"value": s_pairs / info_t["seconds"], "unit": UNIT, "cores": info_t["cores"], "kind": info_t["kind"],
                                    "reads_per_s": n_reads / info_t["seconds"],
                                    "sample": f"all {n_reads} reads of {spec.name} ({int(so[-1])} anchors, {s_pairs} visited pairs), "
                                              f"reference mm_chain_dp under its kt_for, best of 3 after 1 warm-up, memory mode "
                                              f"'{info_t['mode']}' ({', '.join(f'{k} {v * 1e3:.0f} ms' for k, v in info_t['all'].items())})",
                                    "parity_with_gpu_on_sample": ok, "pairs_match_device_count": bool(s_pairs == pairs_step)}
            if not ok:
                line["parity_error"] = "GPU chains differ from the reference on the CPU-baseline sample"
        ctx.close()
        del d_a, d_offs
        pin.free()
        torch.cuda.empty_cache()

        # ---- the other BASELINE.json configs, each a fixed workload split over the ranks
        if args.configs != "none" and args.workload == "map-ont-100k":
            names = list(CONFIG_BLOCK) if args.configs == "all" else [n for n in args.configs.split(",") if n in CONFIG_BLOCK]
            block = {}
            for name in names:
                self.barrier()
                t0 = time.perf_counter()
                c = self.run_config(name, args.config_steps, int_peak_mix)
                if rank == 0:
                    c["wall_s_incl_generation"] = time.perf_counter() - t0
                    block[name] = c
            if rank == 0:
                line["configs"] = block
                bad = [k for k, v in block.items() if v["parity_with_reference"] and not v["parity_with_reference"]["ok"]]
                if bad:
                    line["parity_error"] = f"GPU chains differ from the reference on {bad}"
        if rank == 0 and world == 1 and args.seeded_reads > 0 and args.workload == "map-ont-100k":
            line["seeded"] = self.run_seeded(args.seeded_reads, max(1, min(steps, 3)))
            if not line["seeded"]["parity_with_oracle"]["ok"]:
                line["parity_error"] = "seeded path differs from the oracle"
        if rank == 0 and world == 1 and args.mapped_reads > 0 and args.workload == "map-ont-100k":
            line["mapped"] = self.run_mapped(args.mapped_reads, max(1, min(steps, 3)))
            if not line["mapped"]["parity_with_oracle"]["ok"]:
                line["parity_error"] = "composed path (bases in, regions out) differs from the oracle composition"
        if rank == 0:
            print(json.dumps(line))
        if world > 1:
            self.dist.barrier()
            self.dist.destroy_process_group()
        return 0

    # ---- SURVEY 8(f) rows composed: bases in, regions out (f4 sketch -> f1 seeding -> DP -> f3 regions), and the driver boundary (f2)
    def run_mapped(self, n_reads, steps):
        """One GPU, rank 0.  G-seq workload (synthetic genome + reads as BASES).  Timed: mm2b200_map_regs_batch end to end from
        pinned host bases; the sketching stage alone (its own roofline); the same reads as request packets through the fork's
        driver symbols.  Beside it on the host cores: the same composition from the reference's own sketch.c / chain.c / hit.c
        (compiled unmodified) + the oracle's collect_seed_hits port, on a bounded sample, which is also the parity check."""
        from concurrent.futures import ThreadPoolExecutor
        from minimap2_chaindp_b200 import fpgadrv, post as mpost, seed as mseed, sketch as msketch
        api, workload, torch = self.api, self.workload, self.torch
        spec = workload.SEQ_CONFIGS["map-ont-100k-seq"]
        w, k, hpc = msketch.SKETCH_PRESETS[spec.preset]
        t0 = time.perf_counter()
        W = workload.SeqWorkload(spec)
        ctx = api.ChainContext(device=self.local, profile=True)
        g_off, g_mv = msketch.sketch_batch(ctx, (W.contig_offsets(), W.genome), w, k, bool(hpc), rid=np.arange(spec.n_contigs))
        streams = W.build_index(g_off, g_mv, k)
        build_s = time.perf_counter() - t0
        ix = mseed.SeedIndex(ctx).load_streams(*streams)
        off, bid = W.reads_count(0, n_reads)
        n_bases = int(off[-1])
        pin = api.PinnedArray((n_bases,), np.uint8)
        W.reads_fill(0, off, pin.array)
        hashes = ((np.arange(n_reads, dtype=np.uint64) * np.uint64(2654435761)) & np.uint64(0xffffffff)).astype(np.uint32)
        chain = api.params(spec.preset)
        sk, sp, pp = msketch.SketchParams(w, k, hpc), mseed.seed_params(chain, spec.flag, spec.max_occ), mpost.post_params(spec.preset)
        L = mseed._lib()
        res = mseed._MapResult()

        def call():
            rc = L.mm2b200_map_regs_batch(ix._h, C.byref(sk), C.byref(sp), C.byref(pp), n_reads, off.ctypes.data, pin.ptr, bid.ctypes.data,
                                          hashes.ctypes.data, C.byref(res))
            if rc:
                raise api.ChainError(rc, api.last_error())
        call(); call()                           # warm-up: allocations, first touch
        ctx.stage_reset()
        s0, k0, p0 = ix.stage_ms(), msketch.sketch_stage_ms(ctx), mpost.post_stage_ms(ctx)
        l0 = ix.launch_count() + ctx.launch_count()
        t0 = time.perf_counter()
        for _ in range(steps):
            call()
        dt = (time.perf_counter() - t0) / steps
        launches = (ix.launch_count() + ctx.launch_count() - l0) // steps
        s1, k1, p1 = ix.stage_ms(), msketch.sketch_stage_ms(ctx), mpost.post_stage_ms(ctx)
        stage = {"sketch": (k1[0] - k0[0]) / steps}
        stage.update({kk: (s1[kk] - s0[kk]) / steps for kk in s1})
        stage.update({kk: ctx.stage_ms(kk)[0] / steps for kk in ("prep", "fill", "compact", "tail", "gather")})
        stage["post"] = (p1[0] - p0[0]) / steps
        n_regs = int(np.frombuffer((C.c_char * 8).from_address(res.reg_off + 8 * n_reads), np.int64)[0])
        n_mv, n_anchors, n_chains = int(res.n_minimizers), int(res.n_anchors), int(res.n_chains)
        h2d = n_bases + n_reads * (8 + 4 + 4 + 4) + 8
        d2h = n_regs * 80 + n_reads * (8 + 4) + 8 + 3 * 64
        # ---- the sketching stage alone, device resident (CUDA events): bases/s and its HBM roofline
        d_seq = torch.from_numpy(pin.array).to(self.dev)
        d_off = torch.from_numpy(off).to(self.dev)
        dsk, n_out = msketch._DevSketch(), C.c_int64()
        st = self.stream.cuda_stream

        def sketch_dev():
            rc = msketch._lib().mm2b200_sketch_device(ctx._h, C.byref(sk), n_reads, n_bases, d_off.data_ptr(), d_seq.data_ptr(), None,
                                                      C.c_void_p(st), C.byref(dsk), C.byref(n_out))
            if rc:
                raise api.ChainError(rc, api.last_error())
        sketch_dev(); torch.cuda.synchronize()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record(self.stream)
        for _ in range(steps):
            sketch_dev()
        ev1.record(self.stream)
        torch.cuda.synchronize()
        sk_ms = ev0.elapsed_time(ev1) / steps
        hbm_peak, peak_src = peaks()
        sk_bytes = n_bases + 2 * 2 * 16 * n_mv + 16 * n_reads          # bases read; minimizers written to their slot, read back and packed
        del d_seq, d_off
        # ---- parity + CPU beside it: the same composition on the host, sample of reads, all cores
        import oracle
        from oracle import post as opost, seed as oseed
        oix = oseed.SeedIndex(*[bytes(x) for x in streams])
        ns = min(2000, n_reads)
        got = ix.map_regs_batch(sk, sp, pp, off[:ns + 1], pin.array[:int(off[ns])], bid[:ns], hashes[:ns])
        oprm, opp = oracle.params(spec.preset), opost.post_params(spec.preset)
        use_ref = opost.have_ref() and oracle.have_ref()
        f_sketch = opost.ref_sketch if use_ref else opost.sketch
        f_chain = oracle.ref_chain_dp if use_ref else oracle.chain_dp
        f_post = opost.ref_chain_post if use_ref else opost.chain_post
        threads = host_cores()
        times = np.zeros((threads, 4))

        def part(t):
            ok = True
            for r in range(t, ns, threads):
                sq = pin.array[off[r]:off[r + 1]].tobytes()
                c0 = time.perf_counter()
                mv = f_sketch(sq, w, k, 0, bool(hpc))
                c1 = time.perf_counter()
                a, rep, _ = oseed.collect_seed_hits(oix, spec.flag, spec.max_occ, mv, int(bid[r]), len(sq))
                c2 = time.perf_counter()
                u, b = f_chain(oprm, a)
                c3 = time.perf_counter()
                want, _, _ = f_post(opp, int(hashes[r]), len(sq), u, b)
                c4 = time.perf_counter()
                times[t] += (c1 - c0, c2 - c1, c3 - c2, c4 - c3)
                g = got.regs[got.reg_off[r]:got.reg_off[r + 1]]
                ok = ok and g.tobytes() == want.tobytes() and int(got.rep_len[r]) == rep
            return ok
        with ThreadPoolExecutor(threads) as ex:
            t0 = time.perf_counter()
            oks = list(ex.map(part, range(threads)))
            cpu_s = time.perf_counter() - t0
        share = times.sum(0) / max(times.sum(), 1e-9)
        # ---- the driver boundary (row f2): the same reads as request packets through fpga_* (map.c:286-324 in, map.c:533-565 out)
        shim = None
        try:
            nd = min(int(os.environ.get("MM2B200_BENCH_SHIM_READS", "20000")), n_reads)
            m_off, m_mv = msketch.sketch_batch(ctx, (off[:nd + 1], pin.array[:int(off[nd])]), w, k, bool(hpc))
            packets = []
            for r0 in range(0, nd, 8):
                body = b""
                for r in range(r0, min(nd, r0 + 8)):
                    m = m_mv[m_off[r]:m_off[r + 1]]
                    body += struct.pack("<iiiiIIhbxQ28x", chain.max_dist_y, chain.max_dist_x, m.shape[0], int(off[r + 1] - off[r]), r, int(bid[r]), 1, spec.b, 0)
                    raw = m.tobytes()
                    body += raw + b"\0" * (-len(raw) % 64)
                packets.append(struct.pack("<IIHHBB50x", r0, 64 + len(body), 0, min(nd, r0 + 8) - r0, 3, 0) + body)
            fpgadrv.init()
            for sname, t in zip(streams, (fpgadrv.TYPE_INDEX_B, fpgadrv.TYPE_INDEX_H, fpgadrv.TYPE_INDEX_V, fpgadrv.TYPE_INDEX_P)):
                fpgadrv.load_index(sname, t)
            fpgadrv.set_params(chain.bw, chain.is_cdna, chain.max_skip, chain.min_sc, spec.flag, spec.max_occ)
            # the fork's side of the boundary in C (libmm2workload.so: wl_shim_pump): 8 submitting threads + 1 receiver
            WL = workload.lib()
            WL.wl_shim_pump.restype = C.c_double
            WL.wl_shim_pump.argtypes = [C.c_int, C.c_int64, C.c_void_p, C.c_void_p] + [C.c_void_p] * 4 + [C.POINTER(C.c_int64)]
            FL = fpgadrv._lib()
            fptr = [C.cast(getattr(FL, nm), C.c_void_p) for nm in ("fpga_get_writebuf_thread", "fpga_writebuf_submit", "fpga_get_retbuf", "fpga_release_retbuf")]
            pk_off = np.concatenate([[0], np.cumsum([len(x) for x in packets])]).astype(np.int64)
            pk_raw = np.frombuffer(b"".join(packets), np.uint8)
            best = None
            all_ms = []
            for rep_i in range(6):
                b0 = fpgadrv.stats()
                rb = C.c_int64()
                el = WL.wl_shim_pump(8, len(packets), pk_off.ctypes.data, pk_raw.ctypes.data, *fptr, C.byref(rb))
                b1 = fpgadrv.stats()
                assert b1["packets"] - b0["packets"] == len(packets)
                if rep_i:
                    all_ms.append(round(1e3 * el, 2))
                    if best is None or el < best[0]:
                        best = (el, b1["batches"] - b0["batches"], b1["err_reads"] - b0["err_reads"], int(rb.value))
            fpgadrv.finalize()
            shim = {"api": "fpga_get_writebuf_thread / fpga_writebuf_submit / fpga_get_retbuf (include/mm2fpgadrv_b200.h), 8 sender threads + 1 receiver, packets of 8 reads",
                    "reads": nd, "packets": len(packets), "reads_per_s": nd / best[0], "ms": 1e3 * best[0], "device_batches": best[1], "err_reads": best[2],
                    "bytes_in": int(sum(len(x) for x in packets)), "bytes_out": best[3], "ms_all": all_ms, "reads_per_s_median": nd / (1e-3 * float(np.median(all_ms))),
                    "note": "best of 5 after 1 warm-up (ms_all: every pass); the fork's side (get buffer, memcpy, submit; receive, release) runs as C threads (csrc/workload.c: wl_shim_pump)"}
        except Exception as e:  # noqa: BLE001
            shim = {"error": repr(e)}
        blk = {
            "workload": f"{spec.name}: G-seq synthetic genome ({spec.n_contigs} x {spec.contig_len} bases, {spec.p_repeat:.0%} repeats), index from its "
                        f"(w={w}, k={k}) minimizers ({g_mv.shape[0]} sites, {W.n_keys} keys, streams {sum(x.size for x in streams) / 2**20:.0f} MiB, max occurrence "
                        f"{W.max_occ_seen}, mid_occ {spec.max_occ}), {n_reads} reads x {spec.read_len} bases at {spec.err:.0%} error given as BASES",
            "api": "mm2b200_map_regs_batch: pinned host bases in, pinned host mm_reg1_t + rep_len out (one call; minimizers, anchors, chains stay in HBM)",
            "reads": n_reads, "bases": n_bases, "minimizers": n_mv, "anchors": n_anchors, "chains": n_chains, "regions": n_regs,
            "ms_per_step": 1e3 * dt, "reads_per_s": n_reads / dt, "bases_per_s": n_bases / dt,
            "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
            "d2h_note": f"regions: {n_regs * 80} B; the chains path returns u[] + b[] = {n_chains * 8 + 0} B + 16 B per chain anchor",
            "stage_ms_per_step": stage, "gpu_launches": int(launches),
            "stage_note": ("the call runs as a pipeline of chunks (MM2B200_MAP_STAGES, default 2): the sketching of a chunk overlaps the seeding and "
                           "chaining of the chunk before it, so these event spans overlap and add up to more than ms_per_step; sketch = first "
                           "kernel to last kernel on the sketching stream, i.e. it includes waiting for the upload of the bases"),
            "sketch": {"kernel": "k_sketch_odd / k_sketch (+ scan, pack)", "ms": sk_ms, "bases_per_s": n_bases / (sk_ms * 1e-3), "minimizers_per_s": n_mv / (sk_ms * 1e-3),
                       "roofline": {"bound": "hbm", "achieved": sk_bytes / (sk_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                                    "frac": sk_bytes / (sk_ms * 1e-3) / 1e9 / hbm_peak, "peak_source": peak_src,
                                    "algorithmic_bytes": sk_bytes,
                                    "note": "1 B per base in, 16 B per minimizer out (written to the read's slot, then packed); the kernel is bound by "
                                            "integer issue / shared memory (k-mer assembly, window scans), not by HBM"}},
            "parity_with_oracle": {"ok": bool(all(oks)), "reads": ns, "checker": "reference sketch.c + chain.c + hit.c (compiled) + collect_seed_hits port" if use_ref
                                   else "oracle restatements"},
            "cpu_baseline": {"reads_per_s": ns / cpu_s, "cores": threads, "kind": "reference" if use_ref else "port",
                             "sample": f"first {ns} reads, per read mm_sketch -> collect_seed_hits -> mm_chain_dp -> mm_gen_regs + chain_post through ctypes "
                                       f"on {threads} threads (stage shares of the CPU time: sketch {share[0]:.2f}, seeding {share[1]:.2f}, chain {share[2]:.2f}, post {share[3]:.2f})"},
            "driver_boundary": shim,
            "index_build_s": build_s,
        }
        ix.close(); ctx.close(); pin.free(); W.close()
        return blk

    # ---- seeding on the device: minimizers in, chains out, anchors never on the host
    def run_seeded(self, n_reads, steps):
        """One GPU, rank 0: mm2b200_seed_chain_batch on a G-idx workload (synthetic index streams far larger than L2 + reads
        given as minimizers), with the oracle's collect_seed_hits timed beside it on a bounded sample and parity on a sample."""
        from concurrent.futures import ThreadPoolExecutor
        from minimap2_chaindp_b200 import seed as mseed
        api, workload, torch = self.api, self.workload, self.torch
        spec = workload.SEED_CONFIGS["map-ont-100k-seeded"]
        t0 = time.perf_counter()
        W = workload.SeedWorkload(spec)
        build_s = time.perf_counter() - t0
        ctx = api.ChainContext(device=self.local, profile=True)
        ix = mseed.SeedIndex(ctx).load_streams(W.B, W.H, W.V, W.P)
        # sizes first, then the minimizers straight into pinned memory
        off0 = np.zeros(n_reads + 1, np.int64)
        ql = np.zeros(n_reads, np.int32)
        bd = np.zeros(n_reads, np.uint32)
        m = workload.lib().wls_reads_count(W._h, C.byref(W._rs), 0, n_reads, off0.ctypes.data, ql.ctypes.data, bd.ctypes.data)
        pin = api.PinnedArray((int(m), 2), np.uint64)
        workload.lib().wls_reads_fill(W._h, C.byref(W._rs), 0, n_reads, off0.ctypes.data, pin.array.ctypes.data)
        prm = mseed.seed_params(api.params(spec.preset), spec.flag, spec.max_occ)
        L = mseed._lib()
        res = mseed._SeedResult()

        def call(want_anchors=0):
            rc = L.mm2b200_seed_chain_batch(ix._h, C.byref(prm), n_reads, off0.ctypes.data, pin.array.ctypes.data, bd.ctypes.data,
                                            ql.ctypes.data, want_anchors, C.byref(res))
            if rc:
                raise api.ChainError(rc, api.last_error())
        call()                                   # warm-up: allocations, first touch
        call()
        ctx.stage_reset()
        ms0 = ix.stage_ms()
        l0 = ix.launch_count() + ctx.launch_count()
        t0 = time.perf_counter()
        for _ in range(steps):
            call()
        dt = (time.perf_counter() - t0) / steps
        ms1 = ix.stage_ms()
        n_anchors = int(res.n_anchors)
        n_chains = int(np.frombuffer((C.c_char * 8).from_address(res.chains.u_off + 8 * n_reads), np.int64)[0])
        n_b = int(np.frombuffer((C.c_char * 8).from_address(res.chains.b_off + 8 * n_reads), np.int64)[0])
        stage = {k: (ms1[k] - ms0[k]) / steps for k in ms1}
        launches = (ix.launch_count() + ctx.launch_count() - l0) // steps
        ctx.set_split(False)                     # chaining kernels alone on the stream for their own durations (see run())
        ctx.stage_reset()
        call()
        stage.update({k: ctx.stage_ms(k)[0] for k in ("prep", "fill", "compact", "tail", "gather")})
        ctx.set_split(True)
        h2d = int(m) * 16 + (n_reads + 1) * 8 + n_reads * 8
        d2h = n_reads * 4 + 4 * (n_reads + 1) * 8 + n_reads * 4 + n_chains * 8 + n_b * 16 + int(off0[-1]) * 0
        mini_total = int(np.frombuffer((C.c_char * 8).from_address(res.mini_off + 8 * n_reads), np.int64)[0])
        d2h += mini_total * 8
        # parity on a sample: anchors, rep_len, mini_pos and chains against the oracle
        import oracle
        from oracle import seed as oseed
        oix = oseed.SeedIndex(W.B, W.H, W.V, W.P)
        ns = min(2000, n_reads)
        out = ix.seed_chain_batch(prm, off0[:ns + 1], pin.array, bd[:ns], ql[:ns], want_anchors=True)
        ok = True
        want_a, want_off = [], [0]
        for r in range(ns):
            a, rep, mini = oseed.collect_seed_hits(oix, spec.flag, spec.max_occ, pin.array[off0[r]:off0[r + 1]], int(bd[r]), int(ql[r]))
            ok = ok and np.array_equal(out.a[out.a_off[r]:out.a_off[r + 1]], a) and int(out.rep_len[r]) == rep and \
                np.array_equal(out.mini_pos[out.mini_off[r]:out.mini_off[r + 1]], mini)
            want_a.append(a); want_off.append(want_off[-1] + a.shape[0])
        want = oracle.chain_batch(oracle.params(spec.preset), np.array(want_off, np.int64), np.concatenate(want_a))
        ok = bool(ok and np.array_equal(out.chains.n_u, want.n_u) and np.array_equal(out.chains.u, want.u) and np.array_equal(out.chains.b, want.b))
        # CPU beside it: the oracle's collect_seed_hits (a port of map.c:112-236) on a bounded sample, all host threads
        threads = host_cores()
        nc = min(20000, n_reads)

        def part(k):
            tot = 0
            for r in range(k, nc, threads):
                a, _, _ = oseed.collect_seed_hits(oix, spec.flag, spec.max_occ, pin.array[off0[r]:off0[r + 1]], int(bd[r]), int(ql[r]))
                tot += a.shape[0]
            return tot
        with ThreadPoolExecutor(threads) as ex:
            list(ex.map(part, range(threads)))          # warm-up: page in the index
            t0 = time.perf_counter()
            list(ex.map(part, range(threads)))
            cpu_s = time.perf_counter() - t0
        blk = {
            "workload": f"{spec.name}: G-idx synthetic index ({spec.n_sites} minimizer sites, {W.n_keys} keys, streams B/H/V/P "
                        f"{sum(x.size for x in W.streams) / 2**20:.0f} MiB, max occurrence {W.max_occ_seen}, mid_occ {spec.max_occ}), "
                        f"{n_reads} reads x {spec.read_len} bases given as minimizers",
            "api": "mm2b200_seed_chain_batch: pinned host minimizers in, pinned host chains + rep_len + mini_pos out (one call, one sub-batch)",
            "reads": n_reads, "minimizers": int(m), "anchors": n_anchors, "chains": n_chains,
            "ms_per_step": 1e3 * dt, "reads_per_s": n_reads / dt, "minimizers_per_s": int(m) / dt,
            "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
            "stage_ms_per_step": stage, "gpu_launches": int(launches),
            "parity_with_oracle": {"ok": ok, "reads": ns},
            "cpu_seeding_only": {"reads_per_s": nc / cpu_s, "cores": threads, "kind": "port",
                                 "sample": f"first {nc} reads, oracle collect_seed_hits (lookup + expand + radix_sort_128x) over {threads} threads via ctypes"},
            "index_build_s_host": build_s,
        }
        ix.close(); ctx.close(); pin.free(); W.close()
        torch.cuda.empty_cache()
        return blk

    def run_multi(self, spec, prm, n_reads, steps, solo_s):
        """Rank 0 alone owns N contexts (one per GPU) and N x n_reads reads in pinned memory and calls
        mm2b200_chain_batch_multi -- the call INTEGRATION.md level 2 puts behind map.c's submit side.  The other ranks
        wait on a CPU barrier (their GPUs are being driven by rank 0)."""
        api, workload, torch = self.api, self.workload, self.torch
        out = None
        self.cpu_barrier()
        if self.rank == 0:
            N = self.world
            offs = workload.count(spec, 0, N * n_reads)
            pin = api.PinnedArray((int(offs[-1]), 2), np.uint64)
            workload.fill(spec, offs, pin.array, 0)
            ctxs = [api.ChainContext(device=k) for k in range(N)]
            cuts = np.zeros(N + 1, dtype=np.int64)
            resv = (api._Result * N)()
            handles = (C.c_void_p * N)(*[c._h for c in ctxs])

            def step():
                api._check(api.lib().mm2b200_chain_batch_multi(handles, N, C.byref(prm), N * n_reads, offs.ctypes.data, pin.ptr,
                                                               cuts.ctypes.data, resv))
            step()
            t0 = time.perf_counter()
            for _ in range(steps):
                step()
            dt = (time.perf_counter() - t0) / steps
            with api.ChainContext(device=0, count_visited=True) as cc:
                pairs0 = int(cc.chain_batch_raw(prm, n_reads, offs.ctypes.data, pin.ptr).n_visited)
            # visited pairs of the other shards: every rank's own count was summed in the headline; here the per-shard
            # figure is estimated from anchors (identical generator, read ranges differ) -- reads/s is exact
            est_pairs = pairs0 * float(offs[-1]) / float(offs[n_reads])
            out = {"api": "mm2b200_chain_batch_multi: one host process, one context + host thread per GPU, pinned host buffers",
                   "n_gpus": N, "reads_per_step": int(N * n_reads), "anchors_per_step": int(offs[-1]),
                   "ms_per_step": 1e3 * dt, "reads_per_s": N * n_reads / dt,
                   "value": est_pairs / dt, "unit": UNIT, "value_note": "visited pairs scaled from shard 0's count by anchors",
                   "devices": sorted({c.device for c in ctxs}),
                   "scaling_efficiency": (solo_s / dt) if solo_s else None,
                   "scaling_efficiency_note": "reads/s of this call / (N x reads/s of rank 0 alone through mm2b200_chain_batch)",
                   "shard_reads": np.diff(cuts).tolist()}
            for c in ctxs:
                c.close()
            pin.free()
        self.cpu_barrier()
        return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="map-ont-100k")
    ap.add_argument("--reads", type=int, default=0, help="override reads per GPU (default: the config's)")
    ap.add_argument("--configs", default="all", help="'all', 'none' or a comma list of the other BASELINE.json configs to add as the configs block")
    ap.add_argument("--config-steps", type=int, default=2, help="timed device-resident passes per chunk in the configs block")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--slot-anchors", type=int, default=0)
    ap.add_argument("--slots", type=int, default=0)
    ap.add_argument("--wire", action="store_true", help="e2e: upload through the host batcher's 8-byte wire format (default: raw mm128_t)")
    ap.add_argument("--wire-threads", type=int, default=0, help="e2e: host encoder threads (0 = library default)")
    ap.add_argument("--seeded-reads", type=int, default=100000, help="reads of the seeded leg (minimizers in, chains out; N=1 only); 0 = skip")
    ap.add_argument("--only-seeded", action="store_true", help="run the seeded leg alone and print its block")
    ap.add_argument("--mapped-reads", type=int, default=100000, help="reads of the composed leg (bases in, regions out; N=1 only); 0 = skip")
    ap.add_argument("--only-mapped", action="store_true", help="run the composed leg alone and print its block")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    if args.impl == "reference":
        return run_reference(args)
    if int(os.environ.get("RANK", "0")) == 0:
        import __graft_entry__ as ge
        ge.build()
    if args.only_seeded:
        b = Bench(args)
        print(json.dumps(b.run_seeded(args.seeded_reads, max(1, min(args.steps, 3)))))
        return 0
    if args.only_mapped:
        b = Bench(args)
        print(json.dumps(b.run_mapped(args.mapped_reads, max(1, min(args.steps, 3)))))
        return 0
    return Bench(args).run()


if __name__ == "__main__":
    sys.exit(main())
