"""Product host builder (ntcomp_b200/csrc/index_build.cpp, linear merges) against the oracle's
independent construction (sort + binary search), plus the flat file format and from_arrays."""
import numpy as np
import pytest

import helpers
import ntcomp_b200 as nt
from oracle import binding as ob


def _same(o, p):
    assert o.n == p.n_sets and o.k == p.k
    assert (o.C() == p.C()).all()
    for a, b in zip(o.rows(), p.rows()):
        assert (a == b).all()
    assert (o.lcs() == p.lcs()).all()


@pytest.mark.parametrize("k", [12, 13, 16, 31, 32, 33, 64, 91, 96, 97, 128, 200, 255])
def test_builder_matches_oracle(k):
    rng = np.random.default_rng(k)
    seqs = [helpers.random_seq(rng, int(n)) for n in rng.choice([0, 5, k - 1, k, k + 3, 500, 3000], size=4)]
    seqs.append(np.frombuffer(b"ACGTNNACGT" * 30 + b"N" + b"ACGGT" * 60, dtype=np.uint8))
    for rc in (True, False):
        o = ob.OracleIndex.build(seqs, k, rc)
        p = nt.Index.build(seqs, k, rc, n_threads=int(rng.integers(1, 9)))
        _same(o, p)


def test_repetitive_and_low_complexity_inputs():
    seqs = [np.frombuffer(b"A" * 300, dtype=np.uint8), np.frombuffer(b"ACACACACAC" * 40, dtype=np.uint8),
            np.frombuffer(b"ACGT" * 100 + b"TTTTTTTTTTTTTTTTTTTTTTTTT" + b"ACGT" * 50, dtype=np.uint8)]
    for k in (12, 31, 91):
        _same(ob.OracleIndex.build(seqs, k, True), nt.Index.build(seqs, k, True))


def test_save_load_and_from_arrays(tmp_path):
    rng = np.random.default_rng(1)
    p = nt.Index.build([helpers.random_seq(rng, 5000)], 31)
    path = tmp_path / "x.ntcidx"
    p.save(path)
    q = nt.Index.load(path)
    assert q.k == p.k and q.n_sets == p.n_sets and (q.lcs() == p.lcs()).all()
    for a, b in zip(p.rows(), q.rows()):
        assert (a == b).all()
    r = nt.Index.from_arrays(p.k, p.n_sets, p.rows(), p.lcs())
    assert (r.C() == p.C()).all()
    with pytest.raises(nt.NtcError):
        nt.Index.load(tmp_path / "missing")
    rows = p.rows()
    rows[0] = rows[0].copy()
    rows[0][0] ^= np.uint64(2)                       # breaks "n-1 edges"
    with pytest.raises(nt.NtcError):
        nt.Index.from_arrays(p.k, p.n_sets, rows, p.lcs())


def test_k_limits():
    g = np.frombuffer(b"ACGT" * 100, dtype=np.uint8)
    for k in (11, 256):
        with pytest.raises(nt.NtcError) as e:
            nt.Index.build([g], k)
        assert e.value.code == 9


def test_synthetic_generator_matches_numpy_restatement():
    """SURVEY 8d formulae restated with numpy uint64 arithmetic (oracle/synth_np.py, what the reference arm of bench.py
    uses so that it never loads the product library) against the product's C++ generator."""
    from ntcomp_b200 import synth
    from oracle import synth_np

    g = synth.genome(5000)
    assert (g == synth_np.genome(5000)).all()
    for (seed, r0, n, L, err) in [(4242, 7, 20, 150, 50_000), (99, 0, 300, 31, 0), (20250801 + 5000, 1234567, 3, 4000, 300_000)]:
        got, off = synth.reads(g, seed, r0, n, L, err)
        want, woff = synth_np.reads(g, seed, r0, n, L, err)
        assert (got == want).all() and (off == woff).all()
    hit = synth.reads(g, 1, 0, 100, 150, 50_000)[0] != synth.reads(g, 1, 0, 100, 150, 0)[0]
    assert 0.03 < float(hit.mean()) < 0.07                       # substitutions always change the base
