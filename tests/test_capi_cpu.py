"""CPU-side checks of the C-ABI library (no compute call needs a GPU here):
every symbol include/crest_cuda.h declares is exported, the ctypes table matches the header, and
compute entry points FAIL LOUDLY without a device (there is no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT, has_gpu


def header_symbols():
    txt = open(os.path.join(ROOT, "include", "crest_cuda.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(crest_cuda_[a-z0-9_]+)\s*\(", txt)))


def test_header_symbols_exported(cuda_lib):
    syms = header_symbols()
    assert len(syms) >= 19
    for s in syms:
        assert hasattr(cuda_lib, s), f"{s} declared in include/crest_cuda.h but not exported"


def test_ctypes_table_covers_header():
    from crest_b200 import _lib
    assert sorted(_lib.SYMBOLS) == header_symbols()


def test_thresholds_struct_layout():
    from crest_b200 import _lib
    assert C.sizeof(_lib.Thresholds) == 5 * 8
    t = _lib.Thresholds()
    # defaults of the reference: src/confparse.f90:162-170
    assert (t.ethr, t.rthr, t.bthr, t.bthrmax, t.bthrshift) == (0.05, 0.125, 0.01, 0.025, 0.5)


@pytest.mark.skipif(has_gpu(), reason="checks the no-device behaviour")
def test_no_cpu_fallback(cuda_lib):
    import crest_b200 as cb
    assert cuda_lib.crest_cuda_device_count() == 0
    at = np.array([6, 1, 1], np.int32)
    xyz = np.zeros((2, 3, 3))
    with pytest.raises(cb.CrestCudaError):
        cb.cregen_dup(cb.Env(), 3, 2, at, xyz, np.zeros(2))
    with pytest.raises(cb.CrestCudaError):
        cb.rotconst(at, xyz)
    with pytest.raises(cb.CrestCudaError):
        cb.CregenContext()
    assert b"CUDA" in cuda_lib.crest_cuda_last_error() or b"cuda" in cuda_lib.crest_cuda_last_error()


def test_host_e_threshold_matches_float32_rule(cuda_lib):
    """crest_cuda_e_threshold (host-only): e < E  <=>  real(sqrt(max(0,e)/n),sp) < thr, the
    reference's comparison of the float32-stored RMSD (cregen.f90:1176,1286,1359-1367)."""
    f = cuda_lib.crest_cuda_e_threshold
    rng = np.random.default_rng(0)
    for thr in (0.125, 0.25, 0.1, 0.2, 0.3333, 1e-3):
        for n in (3, 12, 46, 60, 300):
            E = f(thr, n)
            assert E > 0
            # neighbours of E in double precision sit on either side of the float32 decision
            lo = np.nextafter(E, 0.0)
            assert float(np.float32(np.sqrt(lo / n))) < thr
            assert not (float(np.float32(np.sqrt(E / n))) < thr)
            for e in np.concatenate([E * (1 + rng.normal(0, 1e-7, 200)), rng.uniform(0, 2 * E, 200)]):
                assert (e < E) == (float(np.float32(np.sqrt(max(0.0, e) / n))) < thr)
    assert f(0.0, 10) == -np.inf


def test_host_band_tiles(cuda_lib):
    """crest_cuda_band_tiles (host-only): the in-band tile list covers every pair that passes the
    reference's energy gate (cregen.f90:1245-1255) exactly once, and shards partition it."""
    from crest_b200 import cregen as cg
    rng = np.random.default_rng(1)
    for n, ethr in ((1, 1.0), (2, 1e9), (33, 0.05), (257, 0.3), (700, 1e9), (700, 0.02)):
        er = np.sort(rng.uniform(0, 6.0, n)) / 627.50947428 - 50.0
        ti, tj, rowstart, jfirst = cg.band_tiles(er, ethr)
        TI, TJ = cg.TILE_I, cg.TILE_J
        cover = np.zeros((n, n), np.int32)
        for a, b in zip(ti, tj):
            cover[a * TI:(a + 1) * TI, b * TJ:(b + 1) * TJ] += 1
        i, j = np.tril_indices(n, -1)
        gate = (er[i] - er[j]) * 627.50947428 < ethr
        assert np.all(cover[i[gate], j[gate]] == 1)
        assert cover.max() <= 1
        T = len(ti)
        for world in (1, 2, 3, 8):
            parts = [cg.shard_range(T, r, world) for r in range(world)]
            assert parts[0][0] == 0 and parts[-1][1] == T
            assert all(parts[k][1] == parts[k + 1][0] for k in range(world - 1))
            sizes = [e - b for b, e in parts]
            assert max(sizes) - min(sizes) <= 1


def test_host_tf32_items(cuda_lib):
    """crest_cuda_tf32_items (host-only): the (i-tile, j-group range) work list of the tensor-core
    screen covers every pair that passes the reference's energy gate (cregen.f90:1245-1255,1275)
    exactly once, and pairs in groups at or above g_gate_end need no per-pair gate."""
    from crest_b200 import cregen as cg
    rng = np.random.default_rng(3)
    for n, ethr in ((1, 1.0), (2, 1e9), (17, 1e9), (129, 0.05), (300, 0.3), (700, 1e9), (700, 0.02),
                    (1500, 0.5), (600, 0.0)):
        er = np.sort(rng.uniform(0, 6.0, n)) / 627.50947428 - 50.0
        items = cg.tf32_items(er, ethr)
        cover = np.zeros((n, n), np.int32)
        nogate = np.zeros((n, n), bool)
        for I, gb, ge, gg in items:
            assert 0 <= gb < ge
            i0, i1 = I * 128, min(I * 128 + 128, n)
            j0, j1 = gb * 16, min(ge * 16, n)
            cover[i0:i1, j0:j1] += 1
            jn = max(j0, gg * 16)
            if jn < j1:
                nogate[i0:i1, jn:j1] = True
        i, j = np.tril_indices(n, -1)
        gate = (er[i] - er[j]) * 627.50947428 < ethr
        assert np.all(cover[i[gate], j[gate]] == 1), (n, ethr)
        assert cover.max() <= 1
        assert np.all(gate[nogate[i, j]]), (n, ethr)          # un-gated groups hold in-band pairs only
        if n >= 2 and ethr > 1e8:
            assert len(items) >= 1


def test_host_tf32_items_rank_sized(cuda_lib):
    """The lists `nranks` ranks build (smaller items, >= 16 per CTA and rank) cover the same pairs, and
    every rank owns whole rows (interleaved i-tiles)."""
    from crest_b200 import cregen as cg
    rng = np.random.default_rng(4)
    n = 40000
    er = np.sort(rng.uniform(0, 6.0, n)) / 627.50947428 - 50.0
    base = cg.tf32_items(er, 1e9)

    def area(items):
        i1 = np.minimum(items[:, 0] * 128 + 128, n) - items[:, 0] * 128
        return (i1.astype(np.int64) * (np.minimum(items[:, 2] * 16, n) - items[:, 1] * 16)).sum()

    for nranks in (2, 3, 8):
        items = cg.tf32_items(er, 1e9, nranks)
        assert len(items) > len(base)                      # finer cut
        assert area(items) == area(base)
        # per i-tile the j-group ranges tile [g_begin, g_end) of the single-rank list without overlap
        for I in (0, 7, n // 128 - 1):
            def spans(lst):
                r = lst[lst[:, 0] == I][:, 1:3]
                return r[np.argsort(r[:, 0])]
            a, b = spans(items), spans(base)
            assert a[0, 0] == b[0, 0] and a[-1, 1] == b[-1, 1]
            assert np.all(a[1:, 0] == a[:-1, 1])
        # ranks share the pair space by rows: rank r owns the i-tiles I % nranks == r, all their columns
        off = 0
        for r in range(nranks):
            mine = cg.tf32_items_rank(er, 1e9, nranks, r)
            assert np.all(mine[:, 0] % nranks == r)
            assert np.array_equal(mine, items[off:off + len(mine)])
            off += len(mine)
        assert off == len(items)
        work = [area(cg.tf32_items_rank(er, 1e9, nranks, r)) for r in range(nranks)]
        assert max(work) < 1.05 * min(work)                # interleaved rows balance the triangle


def test_argument_checks(cuda_lib):
    # NULL thresholds / bad sizes are rejected before any device work
    rc = cuda_lib.crest_cuda_cregen_dup(3, 2, None, None, None, None, 0, 1, None, None, None, None,
                                        None)
    assert rc != 0 and len(cuda_lib.crest_cuda_last_error()) > 0


def test_host_esort_matches_oracle(cuda_lib):
    """crest_cuda_cregen_esort (host code of the product) reproduces the reference's unstable
    quicksort permutation incl. the order of EQUAL energies (quicksort.f90:140-165)
    and the ewin cut (cregen.f90:1096-1113)."""
    import oracle_lib as O
    from crest_b200 import cregen as cg
    rng = np.random.default_rng(4)
    for n, ties, ewin in ((0, False, 6.0), (1, False, 6.0), (2, True, 6.0), (17, True, 1e9),
                          (1000, False, 3.0), (1000, True, 2.0), (5000, True, 1e9)):
        en = rng.uniform(0, 6.0, n) / 627.50947428 - 40.0
        if ties and n > 1:
            en = np.round(en * 2000) / 2000           # many exact duplicates
        xyz = rng.normal(size=(n, 3, 3))
        es_ref, pos_ref, nout_ref = O.esort(en, ewin=ewin)
        xs, es, order, nout = cg.cregen_esort(xyz, en, ewin)
        assert np.array_equal(es, es_ref)
        assert np.array_equal(order, np.asarray(pos_ref) + (1 - int(np.min(pos_ref, initial=1))))
        assert nout == nout_ref
        if n:
            assert np.array_equal(xs[order - 1], xyz)          # structure i now sits at order[i]
            assert np.all(np.diff(es) >= 0)


def test_host_esort_threads_keep_the_reference_order(cuda_lib):
    """Large ensembles: sub-ranges of the reference quicksort are sorted by several threads and the
    geometry is permuted by threads with non-temporal copies (capi.cu) - the permutation, incl. the order of
    equal energies, is still the serial reference's; the in-place mirror equals the copying one."""
    import oracle_lib as O
    from crest_b200 import cregen as cg
    rng = np.random.default_rng(11)
    n = 400000
    en = rng.uniform(0, 6.0, n) / 627.50947428 - 40.0
    en = np.round(en * 20000) / 20000                  # ~120 000 distinct values: ties everywhere
    xyz = rng.normal(size=(n, 14, 3))                  # 134 MB: the threaded permutation with kept scratch
    es_ref, pos_ref, nout_ref = O.esort(en, ewin=4.0)
    for rep in range(2):                               # second call re-uses the scratch buffer
        xs, es, order, nout = cg.cregen_esort(xyz, en, 4.0)
        assert np.array_equal(es, es_ref) and nout == nout_ref
        assert np.array_equal(order, np.asarray(pos_ref) + (1 - int(np.min(pos_ref))))
        assert np.array_equal(xs[order - 1], xyz)
    x2, e2 = xyz.copy(), en.copy()
    o2, n2 = cg.cregen_esort_inplace(x2, e2, 4.0)
    assert np.array_equal(x2, xs) and np.array_equal(e2, es) and np.array_equal(o2, order) and n2 == nout
