import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_golden(name):
    with open(os.path.join(GOLDEN, name)) as fh:
        return json.load(fh)


@pytest.fixture(scope="session")
def examples():
    """Input sequences of the reference's examples/*.fasta (tests/golden/examples.json)."""
    from aligner_b200 import DNA, Protein
    ex = load_golden("examples.json")
    book = [Protein.str_to_vec(r["seq"]) for r in ex["book_example_1.fasta"]]
    prot = [Protein.str_to_vec(r["seq"]) for r in ex["protein.fasta"]]
    # DNA::from_u8_vec skips the stray space in the second record (enums.rs:454-467)
    gene = [DNA.from_u8_vec(r["seq"].encode()) for r in ex["human_gene_example.fasta"]]
    return {"book": book, "protein": prot, "gene": gene}


def sha16(arr):
    import hashlib
    data = arr if isinstance(arr, (bytes, bytearray)) else np.ascontiguousarray(arr, dtype=np.uint8).tobytes()
    return hashlib.sha256(data).hexdigest()[:16]
