#!/usr/bin/env python
"""bench.py — GCUPS of the batched local (SW-like) alignment hot path, BASELINE config 3.

  python bench.py --gpus N --steps K --warmup W            (our CUDA path)
  python bench.py --impl reference --gpus N --steps K ...  (CPU oracle on the host cores)

A "step" = one pass of the hot path (validate + DP fill + traceback) over one batch of
synthetic 150 bp DNA read pairs per GPU.  `value` is device-timed with the batch resident
in HBM; `e2e` goes through the public C-ABI call alnr_align_batch with pinned HOST buffers
(H2D and D2H inside the timed region).  One JSON line on stdout (rank 0).
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

METRIC = "GCUPS (cell updates/s) batched SW/NW at 1/2/4/8 B200 vs reference CPU"
OPS_PER_CELL = 10  # ALU ops per cell update of the full recurrence
NCU_DRAM_BYTES_PER_CELL = (1.474577e9 + 7.807713e9) / (1e6 * 150 * 150)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--pairs", type=int, default=10_000_000, help="pairs per GPU (weak scaling)")
    ap.add_argument("--len", type=int, default=150)
    ap.add_argument("--del", dest="dele", type=float, default=10.0)
    ap.add_argument("--ext", type=float, default=2.0)
    ap.add_argument("--mode", default="local", choices=["local", "global"])
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="profiling runs: skip the host-buffer leg")
    return ap.parse_args()


def workload_name(a):
    return "C3: batched %s alignment, %d synthetic %d bp DNA read pairs per GPU, +5/-4, del=%g ext=%g" % (
        "SW-like local" if a.mode == "local" else "NW global", a.pairs, a.len, a.dele, a.ext)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.lines = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.proc.kill()
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx = max(mx, float(f[2]))
            except ValueError:
                continue
            for nm, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        top = sorted(sm)[len(sm) // 2:] if sm else []  # samples under load dominate the upper half
        return {"sm_mhz": float(np.median(top)) if top else None, "sm_max_mhz": mx or None,
                "reasons": sorted(reasons), "samples": len(sm)}


_ALL_CPUS = None


def bind_to_gpu_node(props):
    """Run this rank on the cores of the GPU's own NUMA node, so that the pinned buffers (first
    touch) and the library's host threads sit next to the PCIe root the copies go through.
    Best effort: sysfs may hide the topology inside a container; then nothing changes."""
    try:
        bdf = "%04x:%02x:%02x.0" % (props.pci_domain_id, props.pci_bus_id, props.pci_device_id)
        base = "/sys/bus/pci/devices/" + bdf
        node = int(open(base + "/numa_node").read())
        cpus = set()
        for part in open(base + "/local_cpulist").read().strip().split(","):
            if part:
                lo, _, hi = part.partition("-")
                cpus.update(range(int(lo), int(hi or lo) + 1))
        global _ALL_CPUS
        _ALL_CPUS = os.sched_getaffinity(0)
        cpus &= _ALL_CPUS
        if node < 0 or len(cpus) < 4 or cpus == os.sched_getaffinity(0):
            return {"node": node, "bound": False, "cpus": len(os.sched_getaffinity(0))}
        os.sched_setaffinity(0, cpus)
        return {"node": node, "bound": True, "cpus": len(cpus)}
    except (OSError, ValueError, AttributeError):
        return {"node": None, "bound": False, "cpus": len(os.sched_getaffinity(0))}


def pinned_array(lib, shape, dtype):
    n = int(np.prod(shape)) * np.dtype(dtype).itemsize
    ptr = lib.alnr_host_alloc(max(n, 1))
    if not ptr:
        raise MemoryError("alnr_host_alloc(%d)" % n)
    buf = (C.c_uint8 * max(n, 1)).from_address(ptr)
    arr = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
    return arr


def cpu_baseline(a, threads, seconds):
    """Oracle (C++ restatement; rustc unavailable) timed on a bounded sample of the workload."""
    import oracle_lib as O
    from aligner_b200 import dna_matrix, workloads as W
    M = dna_matrix(5, -4)
    per = 256 * threads
    done_cells, t_used, first = 0.0, 0.0, 0
    O.align_batch(a.mode, W.dna_read_batch(0, 64, a.len), a.dele, a.ext, M, threads=threads)  # warm
    while t_used < seconds:
        pairs = W.dna_read_batch(first, per, a.len)
        t0 = time.perf_counter()
        O.align_batch(a.mode, pairs, a.dele, a.ext, M, threads=threads)
        t_used += time.perf_counter() - t0
        done_cells += pairs.cells()
        first += per
    return done_cells / t_used / 1e9, first, t_used


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    import oracle_lib as O
    from aligner_b200 import dna_matrix, workloads as W
    M = dna_matrix(5, -4)
    per = 512 * threads
    sample = W.dna_read_batch(0, per, a.len)
    for _ in range(max(a.warmup, 1)):
        O.align_batch(a.mode, W.dna_read_batch(0, 64 * threads, a.len), a.dele, a.ext, M, threads=threads)
    t0 = time.perf_counter()
    for _ in range(a.steps):
        O.align_batch(a.mode, sample, a.dele, a.ext, M, threads=threads)
    dt = time.perf_counter() - t0
    gcups = sample.cells() * a.steps / dt / 1e9
    desc = "%d pairs x %d bp per step (bounded sample of the %d-pair workload)" % (per, a.len, a.pairs)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": gcups, "unit": "GCUPS", "n_gpus": a.gpus,
        "steps": a.steps, "warmup": a.warmup, "ms_per_step": dt / a.steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(a), "note": "C++ restatement of aligner-core (rustc unavailable), "
                   "faithful layout, std::thread over pairs"},
        "cpu_baseline": {"value": gcups, "unit": "GCUPS", "cores": threads, "kind": "port", "sample": desc},
        "e2e": {"value": gcups, "unit": "GCUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def main():
    a = parse()
    if a.impl == "reference":
        return run_reference(a)

    import torch
    import torch.distributed as dist
    import aligner_b200 as ab
    from aligner_b200 import _native as N
    from aligner_b200 import workloads as W

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU fallback")
    numa = bind_to_gpu_node(torch.cuda.get_device_properties(local))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib = N.lib()
    mode = N.LOCAL if a.mode == "local" else N.GLOBAL
    M = ab.dna_matrix(5, -4)

    # ---- inputs: this rank's shard [rank*pairs, (rank+1)*pairs) in pinned host memory -------
    cnt, n = a.pairs, a.len
    q2 = pinned_array(lib, (cnt, n), np.uint8)
    t2 = pinned_array(lib, (cnt, n), np.uint8)
    W.dna_read_pairs_torch(rank * cnt, cnt, n, device="cuda:%d" % local, out_q=q2, out_t=t2)
    off = pinned_array(lib, (cnt + 1,), np.uint64)
    off[:] = np.arange(cnt + 1, dtype=np.uint64) * np.uint64(n)
    pairs = ab.PackedPairs(q2.reshape(-1), off, t2.reshape(-1), off)
    cells = float(cnt) * n * n
    torch.cuda.empty_cache()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident timing (`value`) ------------------------------------------------
    db = ab.DeviceBatch(mode, pairs, M, a.dele, a.ext, device=local)
    l0 = lib.alnr_kernel_launch_count()
    db.run_timed(max(a.warmup, 3), profile_fill=False)
    launches_per_step = (lib.alnr_kernel_launch_count() - l0) // max(a.warmup, 3)
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    ms_total, _ = db.run_timed(a.steps, profile_fill=False)
    barrier()
    clocks = sampler.stop()
    # roofline leg: the same steps again with every fill launch bracketed by CUDA events on its
    # own stream.  Profiling serialises the two phases of a chunk (no overlap of the re-run
    # passes / traceback with the next chunk's main pass), so it is kept out of `value`.
    _, fill_total = db.run_timed(a.steps, profile_fill=True)
    numeric = db.numeric
    t = torch.tensor([ms_total, fill_total], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t[0]) / a.steps
    fill_ms_step = float(t[1]) / a.steps
    value = cells * world / (ms_step * 1e-3) / 1e9

    # parity spot check of what was just timed (oracle as checker, first 512 pairs of the shard)
    if rank == 0:
        import oracle_lib as O
        res = db.download()
        reruns = int(db.rerun_count)
        sub = W.dna_read_batch(0, 512, n)
        ref = O.align_batch(a.mode, sub, a.dele, a.ext, M, threads=4)
        okp = bool((res.f[:512] == ref["f"]).all() and (res.coords[:512] == ref["coords"]).all()
                   and (res.aln_len[:512] == ref["aln_len"]).all())
        for i in range(0, 512, 37):
            qa, ta = res.alignment(i)
            L = int(ref["aln_len"][i])
            okp = okp and bool((qa == ref["q_aln"][i, :L]).all() and (ta == ref["t_aln"][i, :L]).all())
        del res
    db.close()

    # ---- end to end through the public C-ABI call, host buffers in and out -------------------
    out = ab.BatchResult.__new__(ab.BatchResult)
    out.count = cnt
    out.f = pinned_array(lib, (cnt,), np.float64)
    out.corner = pinned_array(lib, (cnt,), np.float64)
    out.coords = pinned_array(lib, (cnt, 4), np.uint64)
    out.aln_off = pinned_array(lib, (cnt,), np.uint64)
    out.aln_len = pinned_array(lib, (cnt,), np.uint32)
    out.status = pinned_array(lib, (cnt,), np.int32)
    out.q_aln = pinned_array(lib, (pairs.aln_capacity(),), np.uint8)
    out.t_aln = pinned_array(lib, (pairs.aln_capacity(),), np.uint8)
    # bytes that actually cross PCIe per step: the symbol codes in (equal-length batches get their
    # offsets written on the device), and out the per-pair arrays plus the alignments back to back
    h2d = int(pairs.q.nbytes + pairs.t.nbytes)
    e2e_s = float("nan")
    d2h = 0
    if not a.no_e2e:
        for _ in range(3):  # warm-up
            ab.align_batch(mode, pairs, M, a.dele, a.ext, device=local, out=out)
        barrier()
        t0 = time.perf_counter()
        for _ in range(a.e2e_steps):
            ab.align_batch(mode, pairs, M, a.dele, a.ext, device=local, out=out)
        torch.cuda.synchronize()
        e2e_s = (time.perf_counter() - t0) / a.e2e_steps
        d2h = int(2 * int(out.aln_len.astype(np.int64).sum()) + out.f.nbytes + out.corner.nbytes + out.coords.nbytes
                  + out.aln_off.nbytes + out.aln_len.nbytes + out.status.nbytes)
    te = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_val = cells * world / float(te[0]) / 1e9

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        int_peak = measure_int_peak(lib, local)
        achieved = cells * OPS_PER_CELL / (fill_ms_step * 1e-3) / 1e12 if fill_ms_step > 0 else None
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        d_bytes = cells * 0.25
        line = {
            "metric": METRIC, "value": value, "unit": "GCUPS", "n_gpus": world, "steps": a.steps,
            "warmup": max(a.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": {1: "i16x2", 2: "i32", 3: "f64"}.get(numeric, "?"),
            "data": "synthetic",
            "config": {"workload": workload_name(a), "pairs_per_gpu": cnt, "cells_per_step_per_gpu": cells,
                       "l2": "inputs+scratch (>3 GB) exceed the 126 MB L2; no explicit flush",
                       "parity_spot_check": okp, "host_numa": numa, "speculation_reruns_per_step": reruns},
            "clocks": clocks, "gpu_launches": int(launches_per_step * a.steps),
            "e2e": {"value": e2e_val, "unit": "GCUPS", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "s_per_step": float(te[0]), "timer": "host wall clock around the synchronous C-ABI call"},
            "roofline": {"bound": "int_alu", "kernel": "fill", "achieved": achieved, "peak": int_peak.get("mixed_tops"),
                         "unit": "TOP/s", "frac": (achieved / int_peak["mixed_tops"]) if achieved and int_peak.get("mixed_tops") else None,
                         "ops_per_cell": OPS_PER_CELL, "fill_ms_per_step": fill_ms_step,
                         "fill_share_of_step": fill_ms_step / ms_step if ms_step else None,
                         "peak_source": "measured in this run (alnr_measure_int_peak: IADD3/VIMNMX + IMAD chains, 148 SMs)",
                         "alu_only_peak_tops": int_peak.get("alu_tops"),
                         # dram__bytes_read.sum + dram__bytes_write.sum of the main fill launch, scaled
                         # from the one --set full capture (profiles/r01_fill_strip_ncu.txt:
                         # 1.47 + 7.81 GB for 10^6 pairs = 0.412 B/cell, algorithmic 0.25 B/cell)
                         "traffic": cells * NCU_DRAM_BYTES_PER_CELL,
                         "hbm": {"achieved": d_bytes / (fill_ms_step * 1e-3) / 1e9 if fill_ms_step else None,
                                 "peak": hbm_peak, "unit": "GB/s", "algorithmic_bytes": d_bytes,
                                 "note": "2-bit direction matrix writes, 0.25 B/cell; peak of measured" if peaks else "of fallback"}},
        }
        if not a.no_cpu_baseline:
            if _ALL_CPUS:
                os.sched_setaffinity(0, _ALL_CPUS)  # the CPU arm gets every core of the box
            threads = os.cpu_count() or 1
            g, npairs, secs = cpu_baseline(a, threads, a.cpu_seconds)
            line["cpu_baseline"] = {"value": g, "unit": "GCUPS", "cores": threads, "kind": "port",
                                    "sample": "%d pairs x %d bp in %.1f s (C++ restatement; rustc unavailable)" % (npairs, n, secs)}
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def measure_int_peak(lib, device):
    alu, mixed = C.c_double(), C.c_double()
    if lib.alnr_measure_int_peak(device, C.byref(alu), C.byref(mixed)) != 0:
        return {}
    return {"alu_tops": alu.value / 1e12, "mixed_tops": mixed.value / 1e12}


if __name__ == "__main__":
    main()
