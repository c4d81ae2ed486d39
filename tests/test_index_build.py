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


# ---- row f4: the GPU construction (ntc_index_build_gpu; build_pipeline.h + build_kernels.cuh) ---------------------------
# "emul" runs the very same functors and order of operations with loops on the CPU (tests/emul, CpuExec: a stable counting
# sort per digit in place of the radix kernels); "gpu" is the product call.  Both must equal the host builder bit for bit,
# and (first test) the oracle's independent construction.

def _gpu_builders():
    return [pytest.param("emul", id="emul"), pytest.param("gpu", id="gpu", marks=pytest.mark.gpu)]


def _build_with(backend, seqs, k, rc, stats=None):
    if backend == "emul":
        import emul_binding
        return emul_binding.index_build(seqs, k, rc, stats)
    return nt.Index.build_gpu(seqs, k, rc, stats=stats)


def _same_index(a, b):
    assert a.n_sets == b.n_sets and a.k == b.k
    assert (a.C() == b.C()).all()
    for x, y in zip(a.rows(), b.rows()):
        assert (x == y).all()
    assert (a.lcs() == b.lcs()).all()


def _as_arrays(seqs):
    return [np.frombuffer(bytes(s), dtype=np.uint8) if isinstance(s, (bytes, bytearray)) else s for s in seqs]


@pytest.mark.parametrize("backend", _gpu_builders())
@pytest.mark.parametrize("k", [12, 13, 16, 31, 32, 33, 64, 91, 96, 97, 128, 200, 255])
def test_gpu_construction_matches_host_builder_and_oracle(backend, k):
    rng = np.random.default_rng(k)
    seqs = [helpers.random_seq(rng, int(n)) for n in rng.choice([0, 5, k - 1, k, k + 3, 500, 3000], size=4)]
    seqs.append(np.frombuffer(b"ACGTNNACGT" * 30 + b"N" + b"ACGGT" * 60, dtype=np.uint8))
    for rc in (True, False):
        g = _build_with(backend, seqs, k, rc)
        _same_index(g, nt.Index.build(seqs, k, rc))
        _same(ob.OracleIndex.build(seqs, k, rc), g)


@pytest.mark.parametrize("backend", _gpu_builders())
def test_gpu_construction_degenerate_and_repetitive_inputs(backend):
    rep = [b"A" * 300, b"ACACACACAC" * 40, b"ACGT" * 100 + b"T" * 25 + b"ACGT" * 50]
    for k in (12, 31, 91):
        for rc in (True, False):
            _same_index(_build_with(backend, _as_arrays(rep), k, rc), nt.Index.build(_as_arrays(rep), k, rc))
    # nothing to index (root only), a single k-mer, k + 1 bases, only separators, lower case (not ACGT for the builder)
    for seqs in ([], [b""], [b"ACGT"], [b"ACGTACGTACGT"], [b"ACGTACGTACGTA"], [b"NNNN"], [b"acgtacgtacgtacgt"]):
        for rc in (True, False):
            _same_index(_build_with(backend, _as_arrays(seqs), 12, rc), nt.Index.build(_as_arrays(seqs), 12, rc))


@pytest.mark.parametrize("backend", _gpu_builders())
def test_gpu_construction_predecessor_found_only_by_search(backend):
    """A k-mer seen only at the start of a sequence is a source unless some other k-mer ENDS with its (k-1)-prefix -- e.g.
    the last k-mer of another sequence, which is not its neighbour in any text: the binary-search path of PredBySearch."""
    rng = np.random.default_rng(3)
    for k in (12, 31, 33, 91):
        W, y, V = (bytes(helpers.random_seq(rng, n)) for n in (200, k - 1, 150))
        cases = [[W + y, y + b"C" + V], [W + y, y + b"C" + V, y + b"G" + V[:40]], [W + y + b"N" + y + b"T" + V], [W, W[5:]],
                 [W, W[5:5 + k]], [y + b"A", b"C" + y]]
        for seqs in cases:
            for rc in (True, False):
                _same_index(_build_with(backend, _as_arrays(seqs), k, rc), nt.Index.build(_as_arrays(seqs), k, rc))


@pytest.mark.parametrize("backend", _gpu_builders())
def test_gpu_construction_many_short_sequences(backend):
    rng = np.random.default_rng(7)
    for it in range(24):
        k = int(rng.integers(12, 80))
        alpha = list(b"ACGT" if it % 3 else b"AC") + ([ord("N")] if it % 5 == 0 else [])
        seqs = [bytes(rng.choice(alpha, size=int(rng.integers(0, 3 * k))).astype(np.uint8)) for _ in range(int(rng.integers(1, 40)))]
        if it % 4 == 0:
            seqs += seqs[:3]
        _same_index(_build_with(backend, _as_arrays(seqs), k, bool(it & 1)), nt.Index.build(_as_arrays(seqs), k, bool(it & 1)))


@pytest.mark.gpu
@pytest.mark.parametrize("genome_len,k", [(300_000, 31), (1_000_000, 91), (700_000, 255)])
def test_gpu_construction_larger_genomes(genome_len, k):
    """Many radix blocks per pass (the emulation never runs the radix kernels themselves): one random contig plus a copy of a
    piece of it (duplicate k-mers across sequences) and a few short ones."""
    from ntcomp_b200 import synth
    g = synth.genome(genome_len)
    seqs = [g, g[1000:60_000].copy(), g[:k].copy(), g[5:5 + 2 * k].copy()]
    st = {}
    a = nt.Index.build_gpu(seqs, k, True, stats=st)
    _same_index(a, nt.Index.build(seqs, k, True))
    assert st["launches"] > 0 and st["occurrences"] >= 2 * (genome_len - k + 1)


@pytest.mark.gpu
def test_gpu_construction_many_sequences_at_scale():
    """20,000 sequences of 60..900 bases (a fragmented assembly): tens of thousands of source k-mers, more than a million dummy
    candidates to sort, de-duplicate and merge; a quarter of the sequences are copies or substrings of others (sources that
    have a predecessor elsewhere), some hold non-ACGT bytes."""
    rng = np.random.default_rng(2024)
    g = helpers.random_seq(rng, 20_000_000)
    seqs = []
    for i in range(20_000):
        a = int(rng.integers(0, len(g) - 1000))
        s = g[a:a + int(rng.integers(60, 900))].copy()
        if i % 50 == 7:
            s[len(s) // 2] = ord("N")
        seqs.append(s)
    seqs += [s[10:].copy() for s in seqs[:2500]] + [s.copy() for s in seqs[2500:5000]]
    for k in (31, 91):
        st = {}
        a = nt.Index.build_gpu(seqs, k, True, stats=st)
        _same_index(a, nt.Index.build(seqs, k, True))
        assert st["sources"] > 5_000 and st["dummies"] > 100_000


def test_gpu_construction_needs_a_gpu_or_succeeds():
    """no CPU fallback: without a device the call fails with NTC_ERR_CUDA"""
    import torch
    g = np.frombuffer(b"ACGT" * 100, dtype=np.uint8)
    if torch.cuda.is_available():
        assert nt.Index.build_gpu([g], 12).n_sets == nt.Index.build([g], 12).n_sets
    else:
        with pytest.raises(nt.NtcError) as e:
            nt.Index.build_gpu([g], 12)
        assert e.value.code == 2
    with pytest.raises(nt.NtcError) as e:
        nt.Index.build_gpu([g], 11)
    assert e.value.code == 9


# ---- derived device layouts built by kernels at context creation (layout_kernels.cuh) ---------------------------------------
def _layout_cases():
    rng = np.random.default_rng(5)
    yield [helpers.random_seq(rng, 700), helpers.random_seq(rng, 90)], 12
    yield [helpers.random_seq(rng, 20000), helpers.random_seq(rng, 500), np.frombuffer(b"ACGTNNACGT" * 30, dtype=np.uint8)], 31
    yield [helpers.random_seq(rng, 9000)], 91
    yield [helpers.random_seq(rng, 1500), helpers.random_seq(rng, 900)], 255
    yield [np.frombuffer(b"A" * 2000 + b"ACACACACAC" * 300, dtype=np.uint8)], 16        # LCS runs longer than 255 nodes: saturated distances
    g = helpers.ALPHA[rng.integers(0, 3, size=30000)]
    g[[5000, 12000, 12001, 20000, 29000]] = ord("T")
    yield [g], 16                                                                        # skewed: almost no T


def test_layout_kernel_bodies_match_host_builders():
    import emul_binding
    for seqs, k in _layout_cases():
        for rc in (True, False):
            assert emul_binding.layouts_check(nt.Index.build(seqs, k, rc)) == 0, k


@pytest.mark.gpu
def test_device_built_layouts_match_host_built_ones(monkeypatch):
    """ntc_ctx_create derives rank blocks, LCS words, jump and prefix tables with kernels; NTC_HOST_LAYOUTS=1 takes the host builders
    (the arrays every parity test of rounds 1-2 ran on).  Byte for byte the same, for both rank-block layouts and for
    encode-only / decode-only contexts."""
    from ntcomp_b200 import synth
    cases = list(_layout_cases()) + [([synth.genome(1_500_000)], 91)]
    for seqs, k in cases:
        index = nt.Index.build(seqs, k)
        for nodes in ("32", "64"):
            monkeypatch.setenv("NTC_RANK_NODES", nodes)
            monkeypatch.delenv("NTC_HOST_LAYOUTS", raising=False)
            dev = nt.Context(index, 0)
            monkeypatch.setenv("NTC_HOST_LAYOUTS", "1")
            host = nt.Context(index, 0)
            for which in range(4):
                a, b = dev.layout(which), host.layout(which)
                assert len(a) == len(b) and len(a) > 0 and (a == b).all(), (k, nodes, which)
            del dev, host
    monkeypatch.delenv("NTC_HOST_LAYOUTS", raising=False)
    monkeypatch.delenv("NTC_RANK_NODES", raising=False)
    enc = nt.Context(index, 0, uses=1) if "uses" in nt.Context.__init__.__code__.co_varnames else None
    if enc is not None:
        assert len(enc.layout(0)) > 0 and len(enc.layout(2)) == 0
