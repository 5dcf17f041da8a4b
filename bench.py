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
    ap.add_argument("--no-e2e", action="store_true", help="profiling runs: skip the host-buffer legs")
    ap.add_argument("--no-e2e-strings", action="store_true", help="skip the string-format host-buffer leg")
    ap.add_argument("--no-configs", action="store_true", help="skip the side configs (C1, C2, C3 linear, C4, C5)")
    ap.add_argument("--parity-pairs", type=int, default=4096, help="pairs of the timed batch checked against the oracle")
    ap.add_argument("--c4-len", type=int, default=100000)
    ap.add_argument("--c5-pairs", type=int, default=2000)
    ap.add_argument("--protein-pairs", type=int, default=100000, help="side config: size of the protein batch")
    ap.add_argument("--c5-mats", type=int, default=100)
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


def oracle_threads():
    return max(1, min(os.cpu_count() or 1, 32))


def check_against_oracle(W, O, mode_name, d, e, M, n, get_alignment, f, coords, aln_len, first, count):
    """Bit parity of `count` pairs starting at `first` of the C3 workload against the oracle."""
    sub = W.dna_read_batch(first, count, n)
    ref = O.align_batch(mode_name, sub, d, e, M, threads=oracle_threads())
    ok = bool((f[:count] == ref["f"]).all() and (coords[:count] == ref["coords"]).all()
              and (aln_len[:count] == ref["aln_len"]).all())
    for i in range(count):
        qa, ta = get_alignment(i)
        L = int(ref["aln_len"][i])
        ok = ok and len(qa) == L and bool((qa == ref["q_aln"][i, :L]).all() and (ta == ref["t_aln"][i, :L]).all())
    return ok


def side_configs(a, lib, local, rank, world, int_peak, dist, torch):
    """The BASELINE configs that are not the headline, each {gcups, ms, parity (vs the oracle on the
    same inputs or a sample), roofline, cpu_port_gcups}.  C1 / C2 / C4 on rank 0; C5 sharded over
    all ranks with the NCCL fitness all-reduce."""
    import hashlib
    import aligner_b200 as ab
    from aligner_b200 import _native as N
    from aligner_b200 import sharding as S
    from aligner_b200 import workloads as W
    import oracle_lib as O
    out = {}
    alu = int_peak.get("mixed_tops")  # alu + fma pipes, see the main roofline

    def roof(cells, secs, bound="int_alu", peak=None):
        peak = alu if peak is None else peak
        ach = cells * OPS_PER_CELL / secs / 1e12
        return {"bound": bound, "achieved": ach, "peak": peak, "unit": "TOP/s", "frac": ach / peak if peak else None}

    def one_pair(name, mode, q, t, d, e, M, symbols, reps, lean=False):
        cls = ab.SimpleGlobalAligner if mode == "global" else ab.SimpleLocalAligner
        al = cls.from_seqs(q, t, symbols=symbols)
        r = al.perform_alignment(d, e, M, want_matrices=False)
        best = 1e9
        for _ in range(reps):
            t0 = time.perf_counter()
            r = al.perform_alignment(d, e, M, want_matrices=False)
            best = min(best, time.perf_counter() - t0)
        t0 = time.perf_counter()
        ref = O.align_pair(mode, q, t, float(d), float(e), M, want_matrices=False, lean=lean)
        cpu = time.perf_counter() - t0
        ok = bool(r.corner_score == ref["corner"] and r.alignment.f == ref["f"] and r.alignment.coords == ref["coords"]
                  and len(r.alignment.query) == len(ref["q_aln"]) and (r.alignment.query == ref["q_aln"]).all()
                  and (r.alignment.target == ref["t_aln"]).all())
        cells = float(len(q)) * len(t)
        out[name] = {"gcups": cells / best / 1e9, "ms": best * 1e3, "parity": ok, "roofline": roof(cells, best),
                     "cpu_port_gcups": cells / cpu / 1e9, "what": "one C-ABI call (alnr_align_pair), host buffers"}
        return r

    if rank == 0:
        ex = json.load(open(os.path.join(ROOT, "tests", "golden", "examples.json")))
        book = [ab.Protein.str_to_vec(r["seq"]) for r in ex["book_example_1.fasta"]]
        prot = [ab.Protein.str_to_vec(r["seq"]) for r in ex["protein.fasta"]]
        gene = [ab.DNA.from_u8_vec(r["seq"].encode()) for r in ex["human_gene_example.fasta"]]
        Md = ab.dna_matrix(5, -4)
        one_pair("c1", "global", book[0], book[1], 8, 8, ab.get_blosum62(), ab.Protein, 20)
        one_pair("c2", "local", prot[0], prot[1], 11, 2, ab.get_blosum62(), ab.Protein, 10)
        one_pair("c4_file", "global", gene[0], gene[1], 8, 8, Md, ab.DNA, 10)
        # C4 synthetic 10^5 x 10^5: the oracle finishes 2 x 10^4 (lean layout) in seconds; the full size
        # is checked through properties (the path covers both sequences and re-scores to the corner)
        q, t = W.dna_long_pair(20000, seed=4)
        one_pair("c4_2e4", "global", q, t, 8, 8, Md, ab.DNA, 3, lean=True)
        q, t = W.dna_long_pair(a.c4_len, seed=4)
        pp = ab.PackedPairs.from_lists([q], [t])
        db = ab.DeviceBatch(N.GLOBAL, pp, Md, 8, 8, device=local)
        db.run_timed(2, profile_fill=False)
        ms, _ = db.run_timed(3, profile_fill=False)
        _, fill = db.run_timed(3, profile_fill=True)
        res = db.download()
        db.close()
        qa, ta = res.alignment(0)
        both = (qa[:-1] != 98) & (ta[:-1] != 98)
        score = Md[ta[:-1][both].astype(np.intp), qa[:-1][both].astype(np.intp)].sum() - 8.0 * float((~both).sum())
        prop = bool((qa[:-1][qa[:-1] != 98] == q).all() and (ta[:-1][ta[:-1] != 98] == t).all()
                    and score == res.corner[0])
        cells = float(len(q)) * len(t)
        out["c4_1e5"] = {"gcups": cells / (ms / 3e3) / 1e9, "ms": ms / 3, "fill_ms": fill / 3,
                         "parity": bool(prop and out["c4_2e4"]["parity"]),
                         "parity_how": "oracle (lean layout) bit-exact at 2x10^4; full size: path covers both "
                                       "sequences and re-scores to the corner score",
                         "roofline": roof(cells, fill / 3e3), "cpu_port_gcups": out["c4_2e4"]["cpu_port_gcups"],
                         "hbm_d_write_gbs": cells * 0.25 / (fill / 3e3) / 1e9,
                         "what": "device-resident fill + traceback of one %d x %d pair" % (len(q), len(t))}

        # ---- beyond BASELINE's five: a large PROTEIN batch (VERDICT: "no large-batch protein number
        # exists"): 10^5 pairs of 270-330 aa, BLOSUM62, local, del=11 ext=2 — the INT32 one-warp-per-pair
        # kernel (the packed 16-bit strip kernel is DNA-only), device-resident, with traceback
        base = W.protein_pairs(0, min(a.protein_pairs, 20000), 270, 330)  # generated pair by pair on the host:
        reps = max(1, a.protein_pairs // base.count)                       # 20 000 distinct pairs, repeated
        lists = [base.pair(i) for i in range(base.count)] * reps
        pp = ab.PackedPairs.from_lists([x[0] for x in lists], [x[1] for x in lists])
        Mp = ab.get_blosum62()
        db = ab.DeviceBatch(N.LOCAL, pp, Mp, 11, 2, device=local)
        db.run_timed(1, profile_fill=False)
        ms, _ = db.run_timed(3, profile_fill=False)
        res = db.download()
        db.close()
        ns = min(256, pp.count)
        sub = ab.PackedPairs.from_lists([lists[i][0] for i in range(ns)], [lists[i][1] for i in range(ns)])
        t0 = time.perf_counter()
        ref = O.align_batch("local", sub, 11.0, 2.0, Mp, threads=os.cpu_count() or 1)
        cpu = time.perf_counter() - t0
        okp = bool((res.f[:ns] == ref["f"]).all() and
                   (np.asarray(res.coords[:ns]).reshape(ns, -1) == np.asarray(ref["coords"]).reshape(ns, -1)).all())
        for i in range(0, ns, 7):
            L = int(ref["aln_len"][i])
            qa, ta = res.alignment(i)
            okp = okp and len(qa) == L and bool((qa == ref["q_aln"][i, :L]).all() and (ta == ref["t_aln"][i, :L]).all())
        qo, to = np.diff(pp.q_off), np.diff(pp.t_off)
        cells = float((qo.astype(np.float64) * to).sum())
        out["protein_batch"] = {"gcups": cells / (ms / 3e3) / 1e9, "ms": ms / 3, "parity": okp,
                                "parity_how": "oracle on the first %d pairs (f, coords; every 7th string)" % ns,
                                "roofline": roof(cells, ms / 3e3),
                                "cpu_port_gcups": float((qo[:ns].astype(np.float64) * to[:ns]).sum()) / cpu / 1e9,
                                "what": "not a BASELINE config: %d protein pairs of 270-330 aa, BLOSUM62, local, del=11 "
                                        "ext=2, device-resident fill + traceback (INT32 warp kernel)" % pp.count}

    # ---- C5: heuristic matrix search, pairs sharded over the ranks, NCCL all-reduce of the block sums
    P, K = a.c5_pairs, a.c5_mats
    pairs = W.protein_pairs(0, P, 270, 330)
    mats = W.candidate_matrices(K)
    freqs = W.pair_frequencies(pairs, 24)
    dev = torch.device("cuda", local)
    S.matrix_search_sharded(S.slice_pairs(pairs, 0, min(P, 8)), mats[:2], 5, 5, -1.0, 576.0, freqs[:min(P, 8)],
                            device=local, collective_device=dev)
    best = None
    for _ in range(3):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        fit, (lo, hi), bk, bf, fg, sg = S.matrix_search_sharded(pairs, mats, 5, 5, -1.0, 576.0, freqs, device=local,
                                                                collective_device=dev)
        torch.cuda.synchronize()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        best = float(dt[0]) if best is None else min(best, float(dt[0]))
    sub = S.slice_pairs(pairs, lo, hi)
    res = ab.heuristic_batch(sub, mats, 5, 5, -1.0, 576.0, freqs[lo:hi], want_matrices=False, device=local)
    n = np.diff(sub.q_off).astype(np.float64)
    m = np.diff(sub.t_off).astype(np.float64)
    cells_t = torch.tensor([float((res.passes.reshape(hi - lo, K) * (n * m)[:, None]).sum())], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(cells_t)
    if rank == 0:
        cells = float(cells_t[0])
        f64 = measure_fp64_peak(lib, local)
        # oracle on a sample of this rank's items (threads over items)
        rng = np.random.default_rng(5)
        pick = rng.choice((hi - lo) * K, size=min(48, (hi - lo) * K), replace=False)
        t0 = time.perf_counter()
        ok, cpu_cells = True, 0.0
        for it in pick:
            i, k = int(it) // K, int(it) % K
            qv, tv = sub.pair(i)
            r = O.heuristic(qv, tv, 5.0, 5.0, mats[k], -1.0, 576.0, freqs[lo + i])
            ok = ok and O.STATUS_TO_ALNR[r["status"]] == int(res.status[it])
            if r["status"] == 0:
                ok = ok and r["f"] == res.f[it] and r["passes"] == int(res.passes[it])
                cpu_cells += r["passes"] * float(len(qv)) * len(tv)
        cpu = time.perf_counter() - t0
        ach = cells * OPS_PER_CELL / best / 1e12 / world  # per GPU, like the peak
        out["c5_%dx%d" % (P, K)] = {
            "gcups": cells / best / 1e9, "ms": best * 1e3, "parity": bool(ok), "n_gpus": world,
            "parity_how": "%d random (pair, matrix) items against the oracle's loop: f, passes, status" % len(pick),
            "fitness_sha256_16": hashlib.sha256(fit.tobytes()).hexdigest()[:16],
            "collective": "one NCCL all-reduce of the [blocks x K] f64 block sums" if world > 1 else "none (1 GPU)",
            "roofline": {"bound": "fp64_add", "per": "GPU", "achieved": ach, "peak": f64.get("add_tops"), "unit": "TOP/s",
                         "frac": ach / f64["add_tops"] if f64.get("add_tops") else None},
            "cpu_port_gcups": cpu_cells / cpu / 1e9, "dtype": "f64",
            "what": "sharded alnr_heuristic_batch calls (host buffers in, fitness out), max over ranks"}
    return out


def load_traffic(cells):
    """dram__bytes_read.sum + dram__bytes_write.sum of the main fill launch from the committed ncu
    capture of THIS round's build (profiles/r02_fill_strip_traffic.json, written by
    scripts/ncu_summary.py), scaled per cell; None when no capture is committed."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "r02_fill_strip_traffic.json")))
        return {"bytes": cells * float(t["dram_bytes_per_cell"]), "per_cell": float(t["dram_bytes_per_cell"]),
                "source": "profiles/r02_fill_strip_traffic.json (" + t.get("capture", "ncu --set full") + ")"}
    except (OSError, KeyError, ValueError):
        return None


def main():
    a = parse()
    if a.impl == "reference":
        return run_reference(a)

    import torch
    import torch.distributed as dist
    import aligner_b200 as ab
    from aligner_b200 import _native as N
    from aligner_b200 import workloads as W
    import oracle_lib as O

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

    # ---- inputs: this rank's shard [rank*pairs, (rank+1)*pairs) in pinned host memory, in both
    # input formats (byte codes; 2-bit packed words) --------------------------------------------
    cnt, n = a.pairs, a.len
    nw = (n + 15) // 16
    q2 = pinned_array(lib, (cnt, n), np.uint8)
    t2 = pinned_array(lib, (cnt, n), np.uint8)
    qp = pinned_array(lib, (cnt, nw), np.uint32)
    tp = pinned_array(lib, (cnt, nw), np.uint32)
    W.dna_read_pairs_torch(rank * cnt, cnt, n, device="cuda:%d" % local, out_q=q2, out_t=t2, out_q2=qp, out_t2=tp)
    off = pinned_array(lib, (cnt + 1,), np.uint64)
    off[:] = np.arange(cnt + 1, dtype=np.uint64) * np.uint64(n)
    pairs = ab.PackedPairs(q2.reshape(-1), off, t2.reshape(-1), off)
    packed = ab.Packed2Pairs(qp.reshape(-1), off, tp.reshape(-1), off)
    cells = float(cnt) * n * n
    torch.cuda.empty_cache()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(vals):
        t = torch.tensor(vals, dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return [float(v) for v in t]

    def device_timed(src, d, e, compact, steps, sample_clocks=False, count=None):
        """K passes over the device-resident batch between CUDA events on the library's stream."""
        if count is not None and count < src.count:
            src = (ab.Packed2Pairs(src.q2[:count * nw], src.q_off[:count + 1], src.t2[:count * nw], src.t_off[:count + 1])
                   if isinstance(src, ab.Packed2Pairs) else
                   ab.PackedPairs(src.q[:count * n], src.q_off[:count + 1], src.t[:count * n], src.t_off[:count + 1]))
        db = ab.DeviceBatch(mode, src, M, d, e, device=local, compact=compact)
        l0 = lib.alnr_kernel_launch_count()
        db.run_timed(max(a.warmup, 3), profile_fill=False)
        lps = (lib.alnr_kernel_launch_count() - l0) // max(a.warmup, 3)
        sampler = ClockSampler(local) if sample_clocks else None
        barrier()
        if sampler:
            sampler.start()
        ms_total, _ = db.run_timed(steps, profile_fill=False)
        barrier()
        clocks = sampler.stop() if sampler else None
        # roofline leg: the same steps again with every fill launch bracketed by CUDA events on its
        # own stream.  Profiling serialises the two phases of a chunk (no overlap of the re-run
        # passes / traceback with the next chunk's main pass), so it is kept out of `value`.
        _, fill_total = db.run_timed(steps, profile_fill=True)
        ms, fill = allmax([ms_total / steps, fill_total / steps])
        return db, ms, fill, lps, clocks

    # ---- device-resident timing (`value`: the reference's result format, two byte strings) ------
    db, ms_step, fill_ms_step, launches_per_step, clocks = device_timed(pairs, a.dele, a.ext, False, a.steps, True)
    numeric = db.numeric
    value = cells * world / (ms_step * 1e-3) / 1e9
    okp, reruns = True, 0
    if rank == 0:  # parity of what was just timed: oracle on the first pairs of the shard, every string
        res = db.download()
        reruns = int(db.rerun_count)
        okp = check_against_oracle(W, O, a.mode, a.dele, a.ext, M, n, res.alignment, res.f, res.coords, res.aln_len,
                                   0, min(a.parity_pairs, cnt))
        del res
    db.close()
    # the same pass in the compact formats (2-bit inputs, move strings out)
    db, ms_c, fill_c, launches_c, _ = device_timed(packed, a.dele, a.ext, True, a.steps)
    okc = True
    if rank == 0:
        cr = db.download()
        npar = min(a.parity_pairs, cnt)
        ex = cr.expand(0, npar)
        okc = check_against_oracle(W, O, a.mode, a.dele, a.ext, M, n, ex.alignment, cr.f, ex.coords, ex.aln_len, 0, npar)
        del cr, ex
    db.close()
    value_compact = cells * world / (ms_c * 1e-3) / 1e9

    # ---- end to end through the public C-ABI call, host buffers in and out -------------------
    def e2e_leg(src, out, steps, count=None):
        for _ in range(3):  # warm-up
            ab.align_batch(mode, src, M, a.dele, a.ext, device=local, out=out)
        barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            ab.align_batch(mode, src, M, a.dele, a.ext, device=local, out=out)
        torch.cuda.synchronize()
        return allmax([(time.perf_counter() - t0) / steps])[0]

    alloc = lambda shape, dt: pinned_array(lib, shape, dt)
    e2e = {"value": None}
    e2e_strings = None
    strong = None
    if not a.no_e2e:
        # compact formats: 2-bit codes in; f, start cell, length, status, move-string offsets + moves out
        cout = ab.CompactResult(packed, mode, 4, want_corner=False, want_coords=False, alloc=alloc)
        s_c = e2e_leg(packed, cout, a.e2e_steps)
        e2e = {"value": cells * world / s_c / 1e9, "unit": "GCUPS", "h2d_bytes_per_step": packed.bytes_in(),
               "d2h_bytes_per_step": cout.bytes_out(), "s_per_step": s_c,
               "format": "compact: 2-bit DNA codes in (alnr_batch.q2/.t2); f, start cell, length, status, move "
                         "offsets and 2-bit move strings out (alnr_batch_out.moves)",
               "timer": "host wall clock around the synchronous C-ABI call (alnr_align_batch), max over ranks"}
        if rank == 0:
            # what a caller does with it: expand on the host cores (lazily, or in bulk as timed here)
            npar = min(a.parity_pairs, cnt)
            ex = cout.expand(0, npar)
            e2e["parity_spot_check"] = check_against_oracle(W, O, a.mode, a.dele, a.ext, M, n, ex.alignment, cout.f,
                                                            ex.coords, ex.aln_len, 0, npar)
            nexp = min(cnt, 1_000_000)
            t0 = time.perf_counter()
            cout.expand(0, nexp)
            e2e["host_expand_pairs_per_s"] = nexp / (time.perf_counter() - t0)
        # strong scaling (BASELINE config 3: 10M pairs sharded across N): pairs / N per GPU
        if world > 1:
            sc = cnt // world
            sub_p = ab.Packed2Pairs(packed.q2[:sc * nw], off[:sc + 1], packed.t2[:sc * nw], off[:sc + 1])
            s_s = e2e_leg(sub_p, cout, a.e2e_steps)
            dbs, ms_s, _, _, _ = device_timed(packed, a.dele, a.ext, True, a.steps, count=sc)
            dbs.close()
            strong = {"pairs_total": sc * world, "pairs_per_gpu": sc, "e2e": float(sc) * world * n * n / s_s / 1e9,
                      "e2e_s_per_step": s_s, "value_compact": float(sc) * world * n * n / (ms_s * 1e-3) / 1e9,
                      "ms_per_step": ms_s, "unit": "GCUPS"}
        del cout
        if not a.no_e2e_strings:
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
            out.end_yx = None
            s_s = e2e_leg(pairs, out, a.e2e_steps)
            d2h = int(2 * int(out.aln_len.astype(np.int64).sum()) + out.f.nbytes + out.corner.nbytes + out.coords.nbytes
                      + out.aln_off.nbytes + out.aln_len.nbytes + out.status.nbytes)
            e2e_strings = {"value": cells * world / s_s / 1e9, "unit": "GCUPS",
                           "h2d_bytes_per_step": int(pairs.q.nbytes + pairs.t.nbytes), "d2h_bytes_per_step": d2h,
                           "s_per_step": s_s, "format": "byte codes in, the reference's two aligned byte strings out"}
            del out
    if strong is None:
        strong = {"pairs_total": cnt, "pairs_per_gpu": cnt, "e2e": e2e.get("value"), "value_compact": value_compact,
                  "unit": "GCUPS", "note": "N = 1: identical to the weak-scaling numbers"}

    int_peak = measure_int_peak(lib, local) if rank == 0 else {}
    configs = {}
    if not a.no_configs:
        # C3 with the linear penalty (del == ext == 6): the other penalty setting SURVEY.md 8d names
        dbl, ms_l, fill_l, _, _ = device_timed(pairs, 6.0, 6.0, False, max(2, a.steps // 2))
        if rank == 0:
            res = dbl.download()
            npar = min(a.parity_pairs, cnt)
            okl = check_against_oracle(W, O, a.mode, 6.0, 6.0, M, n, res.alignment, res.f, res.coords, res.aln_len, 0, npar)
            del res
            ach = cells * OPS_PER_CELL / (fill_l * 1e-3) / 1e12
            tl = time.perf_counter()
            O.align_batch(a.mode, W.dna_read_batch(0, 2048, n), 6.0, 6.0, M, threads=oracle_threads())
            configs["c3_linear"] = {"gcups": cells * world / (ms_l * 1e-3) / 1e9, "ms": ms_l, "parity": okl,
                                    "roofline": {"bound": "int_alu", "achieved": ach, "peak": int_peak.get("mixed_tops"),
                                                 "unit": "TOP/s", "frac": ach / int_peak["mixed_tops"] if int_peak.get("mixed_tops") else None},
                                    "cpu_port_gcups": 2048.0 * n * n / (time.perf_counter() - tl) / 1e9, "n_gpus": world,
                                    "what": "device-resident, del = ext = 6, string format"}
        dbl.close()
        del q2, t2, qp, tp, pairs, packed
        lib.alnr_release_cached_memory()
        configs.update(side_configs(a, lib, local, rank, world, int_peak, dist, torch))

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        achieved = cells * OPS_PER_CELL / (fill_ms_step * 1e-3) / 1e12 if fill_ms_step > 0 else None
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        d_bytes = cells * 0.25
        traffic = load_traffic(cells)
        # the strip kernel issues on BOTH integer pipes on purpose (alu: max / min / PRMT; fma: the
        # borrow-free subtractions and the plane accumulation as IMAD), so its ceiling is the measured
        # alu + fma mix, not the alu pipe alone (against which it would exceed 1)
        peak = int_peak.get("mixed_tops")
        line = {
            "metric": METRIC, "value": value, "unit": "GCUPS", "n_gpus": world, "steps": a.steps,
            "warmup": max(a.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": {1: "i16x2", 2: "i32", 3: "f64"}.get(numeric, "?"),
            "data": "synthetic",
            "config": {"workload": workload_name(a), "pairs_per_gpu": cnt, "cells_per_step_per_gpu": cells,
                       "l2": "inputs+scratch (>3 GB) exceed the 126 MB L2; no explicit flush",
                       "parity_spot_check": bool(okp and okc), "parity_pairs_vs_oracle": min(a.parity_pairs, cnt),
                       "host_numa": numa, "speculation_reruns_per_step": reruns,
                       "value_format": "device-resident; byte codes in HBM -> the reference's two aligned byte "
                                       "strings in HBM (value_compact: 2-bit codes -> move strings)"},
            "value_compact": value_compact, "ms_per_step_compact": ms_c,
            "clocks": clocks, "gpu_launches": int(launches_per_step * a.steps),
            "e2e": e2e, "e2e_strings": e2e_strings, "strong_scaling": strong,
            "roofline": {"bound": "int_alu", "kernel": "fill_strip_kernel (+ re-run passes)", "achieved": achieved,
                         "peak": peak, "unit": "TOP/s",
                         "frac": (achieved / peak) if achieved and peak else None,
                         "ops_per_cell": OPS_PER_CELL, "fill_ms_per_step": fill_ms_step,
                         "fill_share_of_step": fill_ms_step / ms_step if ms_step else None,
                         "peak_source": "measured in this run (alnr_measure_int_peak: independent IADD3/VIMNMX/LOP3 "
                                        "chains on the alu pipe next to IMAD chains on the fma pipe, 1:1, 148 SMs; "
                                        "SURVEY.md 8d); MEASURED_PEAKS.json holds no integer peak",
                         "alu_only_peak_tops": int_peak.get("alu_tops"),
                         "frac_of_alu_only_peak": (achieved / int_peak["alu_tops"]) if achieved and int_peak.get("alu_tops") else None,
                         "traffic": traffic["bytes"] if traffic else None,
                         "traffic_source": traffic["source"] if traffic else "no ncu capture of this build committed",
                         "hbm": {"achieved": d_bytes / (fill_ms_step * 1e-3) / 1e9 if fill_ms_step else None,
                                 "peak": hbm_peak, "unit": "GB/s", "algorithmic_bytes": d_bytes,
                                 "note": "2-bit direction matrix writes, 0.25 B/cell; peak " +
                                         ("from MEASURED_PEAKS.json" if peaks else "fallback of B200_PROFILING.md")}},
            "configs": configs,
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


def measure_fp64_peak(lib, device):
    add, mix = C.c_double(), C.c_double()
    if lib.alnr_measure_fp64_peak(device, C.byref(add), C.byref(mix)) != 0:
        return {}
    return {"add_tops": add.value / 1e12, "addmax_tops": mix.value / 1e12}


def measure_int_peak(lib, device):
    alu, mixed = C.c_double(), C.c_double()
    if lib.alnr_measure_int_peak(device, C.byref(alu), C.byref(mixed)) != 0:
        return {}
    return {"alu_tops": alu.value / 1e12, "mixed_tops": mixed.value / 1e12}


if __name__ == "__main__":
    main()
