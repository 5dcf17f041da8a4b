"""The dispersed-repeat search (aligner-core/src/bin/latent-repeat-search).
CPU tests: the host logic against the reference's own test vector (engine/test.rs:5-64) and literal
restatements.  GPU test: the whole engine loop with the GPU backend against the same loop driven by
the oracle (same seed): identical tasks, coordinates, z values, alignments and matrices."""
import json
import os

import numpy as np
import pytest

import oracle_lib as O

from aligner_b200 import repeat_search as R


def test_filter_reference_vector():
    """engine/test.rs:5-64 (Task equality is by left_coord, task.rs:12-16)."""
    mk = lambda z, l, r: R.Task(None, l, r, z)
    tasks = [mk(12.240966, 300, 630), mk(12.378159, 360, 690), mk(11.762683, 1080, 1410),
             mk(10.471823, 1740, 2070), mk(11.392030, 1860, 2190)]
    got = R.filter_tasks(tasks)
    assert [(t.left_coord, t.right_coord, t.z) for t in got] == [(360, 690, 12.378159), (1080, 1410, 11.762683),
                                                                  (1860, 2190, 11.392030)]
    assert R.filter_tasks([]) == [] and len(R.filter_tasks(tasks[:1])) == 1
    # unsorted input, a chain of three overlapping windows, equal z keeps the later one
    got = R.filter_tasks([mk(5.0, 60, 390), mk(5.0, 0, 330), mk(4.0, 30, 360), mk(9.0, 1000, 1330)])
    assert [(t.left_coord, t.z) for t in got] == [(60, 5.0), (1000, 9.0)]


def test_check_intersection_cases():
    assert R.check_intersection((0, 10), (5, 20)) and R.check_intersection((5, 20), (0, 10))
    assert R.check_intersection((0, 100), (10, 20)) and R.check_intersection((10, 20), (0, 100))
    assert R.check_intersection((0, 10), (10, 20)) and not R.check_intersection((0, 9), (10, 20))


def _import_literal(raw):
    """enums.rs:489-527 written out as the reference's loop."""
    codes, freqs, indices = [], [0.0] * 4, []
    pas, count, local = True, 0, 0
    for i, b in enumerate(raw):
        ch = chr(b)
        if ch in "ATCG":
            v = "ATCG".index(ch)
            freqs[v] += 1.0
            if not pas:
                indices.append(R.Index(i - count, count, local))
                local = 0
                pas = True
            codes.append(v)
        else:
            pas = False
            count += 1
            local += 1
    indices.reverse()
    return codes, [f / len(codes) for f in freqs], indices


def test_dna_import_with_indices_and_coordinates():
    rng = np.random.default_rng(0)
    for _ in range(30):
        raw = bytes(rng.choice(list(b"ATCGNnx\n"), size=int(rng.integers(1, 300)), p=[.2, .2, .2, .2, .1, .04, .03, .03]))
        if not any(c in b"ATCG" for c in raw):
            continue
        codes, freqs, idx = R.dna_from_u8_with_freqs_and_indices(raw)
        lc, lf, li = _import_literal(raw)
        assert codes.tolist() == lc and freqs.tolist() == lf and idx == li
        # index_coord maps a cleaned position at or after an index to its raw position
        kept = [i for i, b in enumerate(raw) if chr(b) in "ATCG"]
        for pos in range(len(codes)):
            if idx and pos >= idx[-1].coord:
                assert R.index_coord(pos, idx) == kept[pos]
        rot = R.rotate_indices(idx, len(codes))
        assert len(rot) == len(idx)
    assert R.index_coord(7, []) == 7 and R.rotate_indices([], 10) == []
    with pytest.raises(R.Error):
        R.dna_from_u8_with_freqs_and_indices(b"AT_G")


def test_files_round_trip(tmp_path):
    fa = tmp_path / "x.fasta"
    fa.write_text(">chr1 test\nATCGNN\nATCG\n>chr2\nGGGG\n")
    recs = R.import_fasta(str(fa))
    assert recs == [("chr1 test", b"ATCGNNATCG"), ("chr2", b"GGGG")]
    (tmp_path / "e.fasta").write_text("")
    with pytest.raises(ValueError):
        R.import_fasta(str(tmp_path / "e.fasta"))
    res = {"chr1": ([R.Task(None, 10, 340, 3.5), R.Task(None, 400, 730, 12.378159)], np.arange(8.0).reshape(4, 2))}
    out, mats = R.write_outputs(res, str(tmp_path / "o.csv"))
    assert open(out).read() == "name,z_value,left_coord,right_coord\nchr1,3.5,10,340\nchr1,12.378159,400,730\n"
    assert json.load(open(mats)) == {"chr1": {"v": 1, "dim": [4, 2], "data": [0.0, 1.0, 2.0, 3.0, 4.0, 5.0, 6.0, 7.0]}}
    assert R.read_csv(out) == {"chr1": [(3.5, 10, 340), (12.378159, 400, 730)]}
    assert R.prepare_sequence(b"ATCGATCG", [(1.0, 2, 5)]) == b"ATNNNTCG"
    o = R.Options()
    assert (o.repeat_length, o.query_offset, o.deletions, o.extension, o.rsquared, o.kd, o.threads, o.repeats) == \
        (300, 30, 30.0, 7.0, 100000.0, 0.0, 1, 10)


class OracleBackend:
    """The engine's aligner calls answered by the CPU oracle (test infrastructure)."""

    def __init__(self, codes):
        self.codes = np.ascontiguousarray(codes, dtype=np.uint8)

    def scores(self, pwm, d, e, starts, lens):
        return O.pwm_windows(self.codes, starts, lens, pwm, d, e, want_alignments=False)["f"]

    def alignments(self, pwm, d, e, starts, lens):
        from aligner_b200.pwm import PWMAlignment
        from aligner_b200.enums import DNA
        ref = O.pwm_windows(self.codes, starts, lens, pwm, d, e)
        out = []
        for i in range(len(starts)):
            n = int(ref["aln_len"][i])
            c = ref["coords"][i]
            out.append(PWMAlignment(DNA, ref["numbered"][i, :n].copy(), ref["q_aln"][i, :n].copy(), pwm.shape[1],
                                    ((int(c[0]), int(c[1])), (int(c[2]), int(c[3]))), float(ref["f"][i])))
        return out

    def close(self):
        pass


def _raw_with_repeats(seed=5, length=50000, L=100, copies=30):
    from aligner_b200 import workloads as W
    seq, repeat, planted = W.dna_chromosome_with_repeats(length, L, copies, seed=seed, sub_rate=0.15)
    raw = bytearray(bytes(np.frombuffer(b"ATCG", dtype=np.uint8)[seq]))
    raw[1000:1000] = b"N" * 137            # runs of skipped symbols -> indices
    raw[30000:30000] = b"\nnnn\n"
    return bytes(raw), repeat, planted


@pytest.mark.gpu
@pytest.mark.parametrize("threads,reverse", [(1, True), (3, False)])
def test_engine_loop_gpu_equals_oracle_driven_loop(threads, reverse):
    from aligner_b200 import workloads as W
    raw, repeat, planted = _raw_with_repeats()
    opts = R.Options(repeat_length=100, query_offset=20, deletions=6.0, extension=2.0, threads=threads,
                     repeats=3, reverse=reverse)
    start = W.pwm_from_repeat(repeat, 1.0, -1.0)
    a = R.perform_calculation_per_sequence(opts, raw, "chr", np.random.default_rng(1), R.GpuBackend, matrix=start)
    b = R.perform_calculation_per_sequence(opts, raw, "chr", np.random.default_rng(1), OracleBackend, matrix=start)
    assert a.keys() == b.keys() and ("inverse" in a) == reverse
    for key in a:
        ta, ma = a[key]
        tb, mb = b[key]
        assert (ma == mb).all()
        assert [(t.left_coord, t.right_coord, t.z) for t in ta] == [(t.left_coord, t.right_coord, t.z) for t in tb]
        for x, y in zip(ta, tb):
            assert x.alignment.f == y.alignment.f and x.alignment.coords == y.alignment.coords
            assert (x.alignment.numbered == y.alignment.numbered).all() and (x.alignment.query == y.alignment.query).all()
    found = a["direct"][0]
    assert len(found) >= 20  # the planted copies (30, some overlapping or too diverged)
    # every strong hit overlaps a planted copy (raw coordinates: shifted by the skipped symbols);
    # z >= 3 alone also lets a few chance windows through
    raw_planted = [R.index_coord(int(p), R.dna_from_u8_with_freqs_and_indices(raw)[2]) for p in planted]
    strong = [t for t in found if t.z >= 8.0]
    assert len(strong) >= 15
    for t in strong:
        assert any(t.left_coord - 150 <= p <= t.right_coord + 150 for p in raw_planted), t


@pytest.mark.gpu
def test_commands_and_outputs(tmp_path):
    raw, repeat, _ = _raw_with_repeats(seed=6, length=20000, copies=12)
    fa = tmp_path / "in.fasta"
    fa.write_bytes(b">one\n" + raw + b"\n>two\n" + raw[:8000] + b"\n")
    opts = R.Options(repeat_length=100, query_offset=20, deletions=6.0, extension=2.0, repeats=2, reverse=True)
    res = R.run_exploring(opts, str(fa), rng=np.random.default_rng(2))
    assert set(res) == {"one", "one-reversed", "two", "two-reversed"}
    out, mats = R.write_outputs(res, str(tmp_path / "o.csv"))
    assert set(json.load(open(mats))) == set(res)
    assert all(m.shape == (4, 100) for _t, m in res.values())
    t = R.run_testing(R.Options(repeat_length=60, query_offset=20, deletions=6.0, extension=2.0, rsquared=240.0,
                                kd=-1.0), np.random.default_rng(3), sequence_length=20000, descendants=5)
    assert set(t) == {"test"} and t["test"][1].shape == (4, 60)
    assert R.main(["-i", str(fa), "-o", str(tmp_path / "cli.csv"), "-r", "100", "-q", "20", "-d", "6", "-e", "2",
                   "--repeats", "1", "--seed", "4"]) == 0
    assert os.path.exists(str(tmp_path / "cli.csv.matrices.json"))
