#!/usr/bin/env python
"""Generate tests/golden/*.json from the read-only reference tree (run in the build
container only; /root/reference does not exist on the GPU box, the committed JSON does).

  python tests/golden/make_fixtures.py [/root/reference]

Sources (relative to the reference root):
  examples/*.fasta                        -> examples.json   (input sequences only)
  src/tests/test_alignment.rs:14-39       -> legacy_global.json (the only golden H/D)
  aligner-core/src/lib.rs:61-90           -> blosum62.json   (24x24 table text)
"""
import json
import os
import re
import sys

ROOT = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
OUT = os.path.dirname(os.path.abspath(__file__))


def read_fasta(path):
    recs, name, seq = [], None, []
    with open(path) as fh:
        for line in fh:
            line = line.rstrip("\n")
            if line.startswith(">"):
                if name is not None:
                    recs.append((name, "".join(seq)))
                name, seq = line[1:], []
            else:
                seq.append(line)  # keep stray spaces: DNA::from_u8_vec skips them, str_to_vec rejects
    if name is not None:
        recs.append((name, "".join(seq)))
    return recs


def main():
    ex = {}
    for fn in ("book_example_1.fasta", "protein.fasta", "human_gene_example.fasta"):
        ex[fn] = [{"name": n, "seq": s} for n, s in read_fasta(os.path.join(ROOT, "examples", fn))]
    with open(os.path.join(OUT, "examples.json"), "w") as fh:
        json.dump(ex, fh, indent=0)

    # legacy golden global test: first two array![...] blocks of test_global_alignment
    src = open(os.path.join(ROOT, "src/tests/test_alignment.rs")).read()
    body = src[src.index("fn test_global_alignment"):src.index("fn test_local_alignment")]
    blocks = []
    for mt in re.finditer(r"array!\[", body):
        i = mt.end()
        depth, j = 1, i
        while depth:
            depth += {"[": 1, "]": -1}.get(body[j], 0)
            j += 1
        blocks.append(body[i:j - 1])
    hrows = [[int(v) for v in re.findall(r"-?\d+", row)] for row in re.findall(r"\[([^\[\]]*)\]", blocks[0])]
    code = {"Top": 0, "Left": 1, "Diagonal": 2, "Beginning": 3}
    drows = [[code[v] for v in re.findall(r"[A-Za-z]+", row)] for row in re.findall(r"\[([^\[\]]*)\]", blocks[1])]
    assert len(hrows) == 8 and all(len(r) == 11 for r in hrows), hrows
    assert len(drows) == 8 and all(len(r) == 11 for r in drows), drows
    # BLOSUM50 entries for the book pair (standard public BLOSUM50; the reference's
    # load_blosum50 is missing from its tree).  rows = target letter, cols = query letter.
    b50 = {
        "P": {"H": -2, "E": -1, "A": -1, "G": -2, "W": -4},
        "A": {"H": -2, "E": -1, "A": 5, "G": 0, "W": -3},
        "W": {"H": -3, "E": -3, "A": -3, "G": -3, "W": 15},
        "H": {"H": 10, "E": 0, "A": -2, "G": -2, "W": -3},
        "E": {"H": 0, "E": 6, "A": -1, "G": -3, "W": -3},
    }
    with open(os.path.join(OUT, "legacy_global.json"), "w") as fh:
        json.dump({"source": "src/tests/test_alignment.rs:14-39", "query": "HEAGAWGHEE",
                   "target": "PAWHEAE", "gap": 8, "blosum50_sub": b50, "H": hrows, "D": drows}, fh)

    lib = open(os.path.join(ROOT, "aligner-core/src/lib.rs")).read()
    txt = lib[lib.index('convert_csv_to_matrix(') :]
    txt = txt[txt.index('"') + 1:]
    txt = txt[:txt.index('"')]
    rows = [[int(v) for v in line.split()] for line in txt.strip().splitlines()]
    assert len(rows) == 24 and all(len(r) == 24 for r in rows)
    with open(os.path.join(OUT, "blosum62.json"), "w") as fh:
        json.dump({"source": "aligner-core/src/lib.rs:61-90", "order": "ARNDCQEGHILKMFPSTWYVBJZX",
                   "rows": rows}, fh)
    print("wrote examples.json legacy_global.json blosum62.json")


if __name__ == "__main__":
    main()
