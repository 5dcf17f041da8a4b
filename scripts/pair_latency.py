"""Latency of single-pair calls (the reference's per-call API shape): C1 / C2 / C4-file sized pairs."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, aligner_b200 as ab
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ex = json.load(open(os.path.join(ROOT, "tests", "golden", "examples.json")))
book = [ab.Protein.str_to_vec(r["seq"]) for r in ex["book_example_1.fasta"]]
prot = [ab.Protein.str_to_vec(r["seq"]) for r in ex["protein.fasta"]]
gene = [ab.DNA.from_u8_vec(r["seq"].encode()) for r in ex["human_gene_example.fasta"]]
for name, cls, q, t, d, e, M, sym in [("C1 book global", ab.SimpleGlobalAligner, book[0], book[1], 8, 8, ab.get_blosum62(), ab.Protein),
                                     ("C2 protein local", ab.SimpleLocalAligner, prot[0], prot[1], 11, 2, ab.get_blosum62(), ab.Protein),
                                     ("C4 file global", ab.SimpleGlobalAligner, gene[0], gene[1], 8, 8, ab.dna_matrix(), ab.DNA)]:
    al = cls.from_seqs(q, t, symbols=sym)
    for wm in (False, True):
        al.perform_alignment(d, e, M, want_matrices=wm)
        ts = []
        for _ in range(50):
            t0 = time.perf_counter(); al.perform_alignment(d, e, M, want_matrices=wm); ts.append(time.perf_counter() - t0)
        ts.sort()
        print("%-18s want_matrices=%-5s median %.3f ms  min %.3f ms" % (name, wm, ts[25] * 1e3, ts[0] * 1e3), flush=True)
