"""GPU parity tests: the CUDA path, called through the C ABI (libcrest_cuda.so via ctypes), against
the oracle and the committed golden vectors.

Tolerances (BASELINE.json north_star): per-pair RMSD within 1e-6 A, rot. constants within 1e-8
relative, duplicate / kept / group sets IDENTICAL except for explicitly listed pairs whose
reference RMSD lies within 1e-6 A of rthr or 2*rthr (helper `boundary_rows`).
"""
import glob
import os

import numpy as np
import pytest

from conftest import GOLDEN
from crest_b200 import synth

pytestmark = pytest.mark.gpu

RMSD_TOL = 1e-6      # Angstrom
ROT_TOL = 1e-8       # relative
AUTOKCAL = 627.50947428


@pytest.fixture(scope="module")
def cb(cuda_lib):
    import crest_b200 as m
    assert cuda_lib.crest_cuda_device_count() > 0, "GPU tests need a CUDA device (no CPU fallback)"
    return m


def gold_kwargs(g):
    kw = {}
    for k in g.files:
        if k.startswith("kw_"):
            v = g[k]
            kw[k[3:]] = v if v.ndim else v.item()
    return kw


def env_from_kwargs(cb, kw, nat):
    env = cb.Env(ethr=kw.get("ethr", 0.05), rthr=kw.get("rthr", 0.125),
                 heavyrmsd=bool(kw.get("heavy", False)), entropic=not kw.get("enantio", True))
    if "include_rmsd" in kw:
        env.subRMSD = True
        env.includeRMSD = np.asarray(kw["include_rmsd"])
        env.rednat = int((env.includeRMSD > 0).sum())
    return env


def boundary_rows(rmat_rows, rthr):
    """Rows whose reference RMSDs touch a threshold within tolerance: the only rows where the
    duplicate decision may legitimately differ (they are LISTED by the tests that use this)."""
    out = set()
    for i, row in rmat_rows.items():
        r = row.astype(np.float64)
        if np.any(np.abs(r - rthr) <= RMSD_TOL) or np.any(np.abs(r - 2 * rthr) <= RMSD_TOL):
            out.add(i)
    return out


def check_axis_frame(got, ref, rot_ref):
    """Principal-axis coordinates agree to 1e-8 A; for (near-)symmetric tops the in-plane axes are
    mathematically arbitrary (degenerate inertia eigenvalues: any rounding difference rotates
    them), so there the frames must agree up to a proper rotation instead."""
    for k in range(len(got)):
        if np.abs(got[k] - ref[k]).max() < 1e-8:
            continue
        r = np.sort(rot_ref[k])
        degenerate = (r[0] == 0) or min(abs(r[1] / r[0] - 1), abs(r[2] / r[1] - 1)) < 1e-5
        assert degenerate, (k, np.abs(got[k] - ref[k]).max())
        H = got[k].T @ ref[k]
        U, S, Vt = np.linalg.svd(H)
        assert np.linalg.det(U @ Vt) > 0                       # same handedness
        e = (got[k] ** 2).sum() + (ref[k] ** 2).sum() - 2 * S.sum()
        assert e / len(got[k]) < 1e-14, (k, e)


def rows_of_allpairs(rmat, n):
    return {i: rmat[i * (i - 1) // 2: i * (i - 1) // 2 + i] for i in range(1, n)}


GOLD_FILES = sorted(p for p in glob.glob(os.path.join(GOLDEN, "*.npz"))
                    if "pairs_kat" not in p and not os.path.basename(p).startswith(("f", "c3_full")))


@pytest.mark.parametrize("path", GOLD_FILES, ids=lambda p: os.path.basename(p)[:-4])
def test_golden(cb, path):
    g = np.load(path)
    kw = gold_kwargs(g)
    at, xyz, er = g["at"], g["xyz"], g["er"]
    nall, nat, _ = xyz.shape
    env = env_from_kwargs(cb, kw, nat)
    nosort = bool(kw.get("nosort", False))
    # stage A
    a = cb.cregen_dup(env, nat, nall, at, xyz, er)
    assert a["npairs"] == int(g["klong"])
    inv = np.argsort(g["orderref"])
    rot_ref = np.empty_like(g["rot"])
    rot_ref[inv] = g["rot"]                     # back to input order
    ok = rot_ref != 0
    assert np.all((a["rot"] == 0) == ~ok)
    assert np.abs(a["rot"][ok] / rot_ref[ok] - 1).max() < ROT_TOL
    xa_ref = np.empty_like(g["xyz_out"])
    xa_ref[inv] = g["xyz_out"]
    check_axis_frame(a["xyz_axis"], xa_ref, rot_ref)
    listed = set()
    all_pairs = kw.get("ethr", 0.05) > 1e8
    if all_pairs and "rmat" in g.files:
        rm = cb.rmat_allpairs(env, at, xyz)
        d = np.abs(rm.astype(np.float64) - g["rmat"].astype(np.float64))
        assert d.max() <= RMSD_TOL, (d.max(), int(d.argmax()))
        listed = boundary_rows(rows_of_allpairs(g["rmat"], nall), env.rthr)
    bad = set(np.where(a["dup"] != g["dup"])[0].tolist())
    assert bad <= listed, f"duplicate decisions differ off-threshold in rows {sorted(bad - listed)}"
    if bad:
        print("LISTED threshold-boundary rows with differing decisions:", sorted(bad))
        return
    # whole cregen_CRE incl. partition and grouping
    r = cb.cregen_CRE(env, nat, nall, at, xyz, er, nosort=nosort)
    assert r.nallout == int(g["nallout"])
    assert np.array_equal(r.order, g["orderref"])
    assert np.array_equal(r.group, g["group"])
    assert np.array_equal(r.er, g["er_out"])
    check_axis_frame(r.xyz, g["xyz_out"], g["rot"])


def test_pairs_kat(cb):
    g = np.load(os.path.join(GOLDEN, "pairs_kat.npz"))
    r, rm = cb.rmsd_pairs(g["packed"], g["i"], g["j"])
    assert np.abs(r - g["rec"][:, 0]).max() < RMSD_TOL
    assert np.abs(rm - g["rec"][:, 1]).max() < RMSD_TOL


@pytest.mark.parametrize("cfg,nall,ethr", [("C1", 1000, 1e9), ("C1", 1000, 0.05), ("C2", 700, 1e9),
                                           ("C4", 130, 1e9), ("C5", 900, 0.1), ("C3", 777, 1e9)])
def test_vs_oracle(cb, oracle, cfg, nall, ethr):
    """Fresh seeded ensembles of the BASELINE shapes at oracle-seconds sizes (C1 at full size)."""
    ens = synth.make_config(cfg, nall=nall, seed=4242)
    at, xyz, er = ens["at"], ens["xyz"], ens["er"]
    nall, nat, _ = xyz.shape
    env = cb.Env(ethr=ethr, heavyrmsd=ens["heavy"])
    ref = oracle.cregen_cre(at, xyz, er, ethr=ethr, heavy=ens["heavy"], want_rmat=True)
    assert ref["nfail"] == 0
    a = cb.cregen_dup(env, nat, nall, at, xyz, er)
    assert a["npairs"] == ref["klong"]
    inv = np.argsort(ref["orderref"])
    rot_ref = np.empty_like(ref["rot"])
    rot_ref[inv] = ref["rot"]
    assert np.abs(a["rot"] / rot_ref - 1).max() < ROT_TOL
    listed = set()
    if ethr > 1e8:
        rm = cb.rmat_allpairs(env, at, xyz)
        d = np.abs(rm.astype(np.float64) - ref["rmat"].astype(np.float64))
        assert d.max() <= RMSD_TOL, (d.max(), int(d.argmax()))
        listed = boundary_rows(rows_of_allpairs(ref["rmat"], nall), env.rthr)
    bad = set(np.where(a["dup"] != ref["dup"])[0].tolist())
    assert bad <= listed, sorted(bad - listed)
    if bad:
        print("LISTED threshold-boundary rows:", sorted(bad))
        return
    r = cb.cregen_CRE(env, nat, nall, at, xyz, er)
    assert r.nallout == ref["nallout"]
    assert np.array_equal(r.order, ref["orderref"])
    assert np.array_equal(r.group, ref["group"])


def test_rotconst_entry(cb, oracle):
    ens = synth.make_config("C4", nall=64, seed=11)
    got = cb.rotconst(ens["at"], ens["xyz"])
    ref = oracle.rotconst(ens["at"], ens["xyz"])
    assert np.abs(got / ref - 1).max() < ROT_TOL


def test_group_entry(cb, oracle):
    """crest_cuda_cregen_group alone on an already partitioned kept set (K5+K6)."""
    ens = synth.make_config("C1", nall=500, seed=5)
    ref = oracle.cregen_cre(ens["at"], ens["xyz"], ens["er"], ethr=0.05)
    n = ref["nallout"]
    group, enuc = cb.cregen_group(cb.Env(), 50, n, ens["at"], ref["xyz"][:n], ref["er"][:n],
                                  ref["rot"][:n])
    assert np.array_equal(group, ref["group"][:n + 1])
    assert np.abs(enuc / ref["enuc"][:n] - 1).max() < 1e-12


@pytest.mark.parametrize("ethr,frac_rotamer", [(0.05, 0.05), (1e9, 0.05), (1e9, 0.6), (0.2, 0.6)])
def test_group_paths_agree(cb, oracle, ethr, frac_rotamer):
    """K6 three ways: two-phase sweep (match lists on the whole GPU + host resolve, the default), the
    same with a list cap of 1 (rows whose single listed match is not a head fall back to the
    single-CTA sweep) and the single-CTA sweep alone -- identical to the oracle's group()."""
    from crest_b200 import cregen
    ens = synth.make_ensemble(2000, 50, 23, seed=17, frac_rotamer=frac_rotamer)
    ref = oracle.cregen_cre(ens["at"], ens["xyz"], ens["er"], ethr=ethr)
    n = ref["nallout"]
    assert ref["group"][0] < n                  # rotamers present: some groups hold > 1 structure
    env = cb.Env(ethr=ethr)
    for tune in (None, "groupcap=1", "seqgroup"):
        cregen.set_tune(tune)
        try:
            group, _ = cb.cregen_group(env, 50, n, ens["at"], ref["xyz"][:n], ref["er"][:n],
                                       ref["rot"][:n])
        finally:
            cregen.set_tune(None)
        assert np.array_equal(group, ref["group"][:n + 1]), tune


def test_group_nontransitive_matches(cb, oracle):
    """Rot. constants spread over a few thresholds make 'equal' dense and NOT transitive: rows whose
    first matches are non-heads exercise the list resolve, and with a list cap of 1-2 the fall-back
    to the single-CTA sweep.  Reference = the K6 loop (cregen.f90:1437-1469) on the oracle's
    equalrotaniso."""
    from crest_b200 import cregen
    rng = np.random.default_rng(23)
    n, nat = 300, 12
    ens = synth.make_ensemble(1, nat, nat, seed=3)
    xyz = np.repeat(ens["xyz"][:1], n, axis=0)            # identical geometry: identical enuc
    er = np.full(n, -100.0)
    base = np.array([3000.0, 1200.0, 900.0])
    rot = base[None, :] * (1.0 + rng.uniform(0.0, 0.10, size=(n, 1)) + rng.uniform(0, 2e-3, size=(n, 3)))
    env = cb.Env(ethr=0.05)
    head = np.zeros(n, bool)
    ref = np.zeros(n + 1, np.int32)
    nonhead_first = 0
    for i in range(n):
        parent, seen_nonhead = -1, False
        for j in range(i):
            if oracle.equalrotaniso(rot[i], rot[j], env.bthr2, env.bthrmax, env.bthrshift):
                if head[j]:
                    parent = j
                    break
                seen_nonhead = True
        nonhead_first += seen_nonhead
        if parent < 0:
            head[i] = True
            ref[0] += 1
            ref[i + 1] = ref[0]
        else:
            ref[i + 1] = ref[parent + 1]
    assert 3 < ref[0] < n // 4 and nonhead_first > 20      # the interesting regime
    for tune in (None, "groupcap=1", "groupcap=2", "seqgroup"):
        cregen.set_tune(tune)
        try:
            group, enuc = cb.cregen_group(env, nat, n, ens["at"], xyz, er, rot)
        finally:
            cregen.set_tune(None)
        assert np.array_equal(group, ref), tune
    assert np.all(enuc == enuc[0])


def test_staged_pageable_copies(cb):
    """Transfers >= 32 MB from / to pageable caller memory go through the threaded pinned staging
    (capi.cu big_copy): same bytes as plain cudaMemcpy (copythr=0), H2D and D2H, odd sizes."""
    from crest_b200 import cregen
    ens = synth.make_ensemble(30011, 50, 23, seed=31)          # 36 MB of xyz, not a chunk multiple
    env = cb.Env(ethr=0.02)
    nall, nat, _ = ens["xyz"].shape
    out = {}
    for tune in ("copythr=0", None, "copythr=3"):
        cregen.set_tune(tune)
        try:
            d = cb.cregen_dup(env, nat, nall, ens["at"], ens["xyz"], ens["er"], want_xyz=True)
            # in/out geometry at an address that is NOT 32-byte aligned: the staging copies and the permuted
            # gather use non-temporal 32-byte stores with unaligned heads and tails (hostcopy.cu fast_copy)
            buf = np.empty(ens["xyz"].size + 1)
            x = buf[1:].reshape(ens["xyz"].shape)
            assert x.ctypes.data % 32 != 0 or buf.ctypes.data % 32 != 0
            x[...] = ens["xyz"]
            e = ens["er"].copy()
            r = cb.cregen_CRE_inplace(env, nat, nall, ens["at"], x, e)
        finally:
            cregen.set_tune(None)
        out[tune] = (d, r)
    d0, r0 = out["copythr=0"]
    for tune in (None, "copythr=3"):
        d, r = out[tune]
        assert np.array_equal(d["dup"], d0["dup"]) and np.array_equal(d["rot"], d0["rot"])
        assert np.array_equal(d["xyz_axis"], d0["xyz_axis"])
        assert np.array_equal(r.xyz, r0.xyz) and np.array_equal(r.order, r0.order)
        assert np.array_equal(r.group, r0.group) and r.nallout == r0.nallout


@pytest.mark.parametrize("nall", [0, 1, 2, 3, 31, 32, 33, 64, 65])
def test_ragged_sizes(cb, oracle, nall):
    """Empty / tiny / tile-edge ensemble sizes (tile 32 x 64)."""
    ens = synth.make_config("C1", nall=max(nall, 1), seed=9)
    at, xyz, er = ens["at"], ens["xyz"][:nall], ens["er"][:nall]
    env = cb.Env(ethr=1e9)
    r = cb.cregen_CRE(env, 50, nall, at, xyz, er)
    if nall == 0:
        assert r.nallout == 0 and r.group[0] == 0
        return
    ref = oracle.cregen_cre(at, xyz, er, ethr=1e9)
    assert r.nallout == ref["nallout"]
    assert np.array_equal(r.order, ref["orderref"])
    assert np.array_equal(r.group, ref["group"])
    assert r.npairs == nall * (nall - 1) // 2


@pytest.mark.parametrize("nat", [1, 2, 3, 7, 8, 9, 40, 41, 97])
def test_ragged_atoms(cb, oracle, nat):
    """Atom counts around the MMA K-block (8) and slab (40) edges; nat=1,2 are degenerate."""
    rng = np.random.default_rng(nat)
    nall = 50
    at = rng.choice([1, 6, 7, 8], size=nat).astype(np.int32)
    base = rng.normal(size=(5, nat, 3)) * 2.0
    xyz = base[rng.integers(0, 5, nall)] + rng.normal(size=(nall, nat, 3)) * 0.03
    er = np.sort(rng.uniform(0, 1e-3, nall))
    ref = oracle.cregen_cre(at, xyz, er, ethr=1e9, want_rmat=True)
    env = cb.Env(ethr=1e9)
    rm = cb.rmat_allpairs(env, at, xyz)
    assert np.abs(rm.astype(np.float64) - ref["rmat"].astype(np.float64)).max() <= RMSD_TOL
    a = cb.cregen_dup(env, nat, nall, at, xyz, er)
    listed = boundary_rows(rows_of_allpairs(ref["rmat"], nall), 0.125)
    bad = set(np.where(a["dup"] != ref["dup"])[0].tolist())
    if nat >= 3:     # for nat<3 rot. constants are 0/NaN-driven; only the RMSD part is compared
        assert bad <= listed, sorted(bad - listed)


def test_errors(cb):
    at = np.array([6, 1, 1], np.int32)
    xyz = np.random.default_rng(0).normal(size=(4, 3, 3))
    with pytest.raises(cb.CrestCudaError, match="ascend"):
        cb.cregen_dup(cb.Env(), 3, 4, at, xyz, np.array([0.0, -1.0, 0.5, 0.6]))
    env = cb.Env(subRMSD=True, includeRMSD=np.zeros(3, np.int32), rednat=0)
    with pytest.raises(cb.CrestCudaError, match="no atom"):
        cb.cregen_dup(env, 3, 4, at, xyz, np.arange(4.0))


def test_float32_threshold_emulation(cb, oracle):
    """Pairs engineered to sit at RMSD = rthr +- a few float32 ulps: decisions must follow the
    float32-rounded value (cregen.f90:1286,1359-1367), up to the 1e-6 listing rule."""
    rng = np.random.default_rng(21)
    nat = 20
    at = np.full(nat, 6, np.int32)
    a = rng.normal(size=(nat, 3)) * 2
    d = rng.normal(size=(nat, 3))
    d -= d.mean(0)
    xs = [a]
    # scale the displacement so the plain (unaligned) RMSD is near rthr and 2*rthr
    for target in (0.125, 0.25):
        for eps in (-3e-6, -1e-7, 0.0, 1e-7, 3e-6):
            s = (target + eps) / np.sqrt((d * d).sum() / nat)
            xs.append(a + s * d)
    xyz = np.array(xs)
    er = -5.0 + 1e-7 * np.arange(len(xs))
    ref = oracle.cregen_cre(at, xyz, er, ethr=1e9, want_rmat=True)
    got = cb.cregen_dup(cb.Env(ethr=1e9), nat, len(xs), at, xyz, er)
    listed = boundary_rows(rows_of_allpairs(ref["rmat"], len(xs)), 0.125)
    bad = set(np.where(got["dup"] != ref["dup"])[0].tolist())
    assert bad <= listed
    rm = cb.rmat_allpairs(cb.Env(ethr=1e9), at, xyz)
    assert np.abs(rm.astype(np.float64) - ref["rmat"].astype(np.float64)).max() <= RMSD_TOL


def test_context_shards_and_mask(cb, oracle):
    """Tile-space sharding (what N ranks do) accumulates to the single-shard result; the packed
    8x8 bitmask reproduces the row-wise 'largest matching j' (cregen.f90:1353-1373)."""
    ens = synth.make_config("C1", nall=900, seed=77)
    at, xyz, er = ens["at"], ens["xyz"], ens["er"]
    env = cb.Env(ethr=1e9)
    ctx = cb.CregenContext()
    ctx.upload(env, at, xyz, er)
    ctx.pack()
    ctx.pairs(env, 0, 1, want_mask=True)
    one = ctx.download()                       # links before backtrack
    mask = ctx.download_mask()
    n = len(er)
    n8 = mask.shape[0]
    bits = np.unpackbits(mask.view(np.uint8).reshape(n8, n8, 8), axis=-1,
                         bitorder="little").reshape(n8, n8, 8, 8)    # [ti, tj, r, c]
    full = bits.transpose(0, 2, 1, 3).reshape(mask.shape[0] * 8, mask.shape[1] * 8)[:n, :n]
    expect = np.where(full.any(1), n - np.argmax(full[:, ::-1], axis=1), 0)
    assert np.array_equal(expect, one["dup"])
    assert one["npairs"] == n * (n - 1) // 2
    ctx.pack()
    for r in range(5):
        ctx.pairs(env, r, 5)
    many = ctx.download()
    assert np.array_equal(many["dup"], one["dup"]) and many["npairs"] == one["npairs"]
    ctx.sweep()
    fin = ctx.download()["dup"]
    ref = oracle.cregen_cre(at, xyz, er, ethr=1e9)
    assert np.array_equal(fin, ref["dup"])
    ctx.close()


def test_full_size_properties_c3(cb, oracle):
    """BASELINE config C3 at FULL size (100 000 x 60, all pairs, enantiomer check): properties that
    need no O(N^2) oracle run + a bounded oracle sample of complete rows."""
    ens = synth.make_config("C3")
    at, xyz, er = ens["at"], ens["xyz"], ens["er"]
    n, nat, _ = xyz.shape
    env = cb.Env(ethr=1e9)
    ctx = cb.CregenContext()
    ctx.upload(env, at, xyz, er)
    ctx.pack()
    ctx.pairs(env)
    links = ctx.download(want_xyz=True)
    assert links["npairs"] == n * (n - 1) // 2
    dup = links["dup"]
    assert np.all(dup <= np.arange(n))                       # parent index < own index (1-based)
    ctx.sweep()
    root = ctx.download()["dup"]
    assert np.all((root == 0) == (dup == 0))
    assert np.all(root[root[root > 0] - 1] == 0)             # roots are kept structures
    # idempotence: a second sweep changes nothing
    ctx.sweep()
    assert np.array_equal(ctx.download()["dup"], root)
    # rigid-motion + per-structure invariance: rotate every input, same classification
    R = synth._random_rotations(np.random.default_rng(1), n)
    xyz2 = np.einsum("nij,naj->nai", R, xyz) + 1.5
    ctx.upload(env, at, xyz2, er)
    ctx.pack()
    ctx.pairs(env)
    dup2 = ctx.download()["dup"]
    diff = np.where(dup2 != dup)[0]
    # bounded oracle sample: complete rows (all j < i) incl. every row that differed above
    rows = sorted(set(np.random.default_rng(2).integers(n // 2, n, 12).tolist()) | set(diff[:8].tolist()))
    xa = links["xyz_axis"]
    rot = links["rot"]
    listed = set()
    for i in rows:
        _, rr = oracle.pair_rows(xa[: i + 1], er[: i + 1], i, i + 1, ethr=1e9)
        r = rr.astype(np.float64)
        exp = 0
        for j in np.where(r < 0.25)[0]:
            if r[j] < 0.125 or oracle.equalrotaniso(rot[i], rot[j], 0.005):
                exp = j + 1
        if np.any(np.abs(r - 0.125) <= RMSD_TOL) or np.any(np.abs(r - 0.25) <= RMSD_TOL):
            listed.add(i)
        assert dup[i] == exp or i in listed, (i, dup[i], exp)
    assert set(diff.tolist()) <= listed, diff[:20]
    # rot. constants of the sampled rows
    ref_rot = oracle.rotconst(at, xyz[rows])
    assert np.abs(rot[rows] / ref_rot - 1).max() < ROT_TOL
    ctx.close()


def test_full_size_golden_c3(cb):
    """BASELINE config C3 at FULL size against the oracle's complete result (tests/golden/c3_full.npz:
    5.0e9 pairs through oracle/, 6869 s on 7 threads, tests/make_golden_c3_full.py): WHOLE arrays - the
    stage-A links after the root chase, the kept-first order, the rotamer groups, the count - are
    identical; the oracle found no row within 1e-6 A of a threshold, so nothing is exempt."""
    import hashlib
    g = np.load(os.path.join(GOLDEN, "c3_full.npz"))
    ens = synth.make_config("C3")
    at, xyz, er = ens["at"], ens["xyz"], ens["er"]
    assert hashlib.sha256(xyz.tobytes() + er.tobytes() + at.tobytes()).hexdigest() == str(g["input_sha256"])
    n, nat, _ = xyz.shape
    assert len(g["boundary_rows"]) == 0
    env = cb.Env(ethr=1e9)
    a = cb.cregen_dup(env, nat, n, at, xyz, er, want_xyz=False)
    assert a["npairs"] == int(g["klong"]) == n * (n - 1) // 2
    assert np.array_equal(a["dup"], g["dup"])
    full = cb.cregen_CRE(env, nat, n, at, xyz, er)
    assert full.nallout == int(g["nallout"])
    assert np.array_equal(full.order, g["orderref"])
    assert np.array_equal(full.group, g["group"])
    assert np.abs(full.rot / g["rot"] - 1).max() < ROT_TOL
    # and the FP64-only kernels
    from crest_b200 import cregen
    cregen.set_tune("notf32")
    try:
        b = cb.cregen_dup(env, nat, n, at, xyz, er, want_xyz=False)
    finally:
        cregen.set_tune(None)
    assert np.array_equal(b["dup"], g["dup"])


# ------------------------------------------------------------------------------------------------
# Tensor-core screen (pair_tf32.cu): TF32 covariance on tcgen05 + FP32 screen + exact FP64 list.
# Its decisions must be THE SAME as those of the FP64 (DMMA) kernels: a pair leaves the FP32 path
# only with a proof that it is far.
# ------------------------------------------------------------------------------------------------
def _stage_a(cb, ens, env, tune=None, debug=None):
    from crest_b200 import cregen
    cregen.set_tune(tune)
    if debug is not None:
        os.environ["CREST_CUDA_DEBUG"] = str(debug)
    try:
        ctx = cb.CregenContext()
        ctx.upload(env, ens["at"], ens["xyz"], ens["er"])
        ctx.pack()
        ctx.pairs(env)
        links = ctx.download()
        info = ctx.info(env)
        ctx.sweep()
        root = ctx.download()["dup"]
        ctx.close()
    finally:
        os.environ.pop("CREST_CUDA_DEBUG", None)
        cregen.set_tune(None)
    return links, root, info


@pytest.mark.parametrize("cfg,nall,ethr", [("C1", 1000, 1e9), ("C1", 1000, 0.05), ("C1", 1000, 0.0),
                                           ("C2", 3000, 1e9), ("C2", 3000, 0.1), ("C5", 5000, 1e9),
                                           ("C5", 4097, 0.3), ("C3", 20000, 1e9),
                                           ("C4", 1500, 1e9), ("C4", 1100, 0.3)])
def test_tf32_equals_fp64_kernels(cb, cfg, nall, ethr):
    ens = synth.make_config(cfg, nall=nall, seed=99)
    env = cb.Env(ethr=ethr, heavyrmsd=ens["heavy"])
    l_tf, r_tf, i_tf = _stage_a(cb, ens, env)
    l_col, r_col, i_col = _stage_a(cb, ens, env, "notf32")
    l_til, r_til, i_til = _stage_a(cb, ens, env, "notf32,tile")
    # A <= 88: one-slab kernel vs column + row-tile DMMA; A > 88 (C4): K-slab kernel vs row-tile DMMA
    assert (i_tf["variant"], i_col["variant"], i_til["variant"]) == (2, 1 if i_tf["A"] <= 88 else 0, 0)
    assert i_tf["flags"] == 0
    assert l_tf["npairs"] == l_col["npairs"] == l_til["npairs"]
    assert np.array_equal(l_tf["dup"], l_col["dup"]) and np.array_equal(l_tf["dup"], l_til["dup"])
    assert np.array_equal(r_tf, r_col)
    # the screen decides the overwhelming majority of pairs in FP32
    if ethr > 1e8 and l_tf["npairs"] > 100000:
        assert i_tf["candidates"] < 0.01 * l_tf["npairs"]


def test_tf32_screen_selfcheck(cb):
    """CREST_CUDA_DEBUG=3 sends EVERY in-band pair through the exact FP64 pass, tagged with the
    screen's verdict; `violations` counts pairs the screen called far whose exact e is below the
    2*rthr threshold.  Rigour of the error bound means 0 - also for straddler-heavy ensembles,
    the adversarial golden geometries and mirror images."""
    from crest_b200 import cregen
    cases = []
    cases.append((synth.make_config("C1", nall=1000, seed=3), cb.Env(ethr=1e9)))
    ens = synth.make_ensemble(1400, 40, 40, copies=20, seed=8, frac_straddle=0.6, frac_mirror=0.2)
    ens["heavy"] = False
    cases.append((ens, cb.Env(ethr=1e9)))
    cases.append((synth.make_config("C4", nall=500, seed=6), cb.Env(ethr=1e9)))     # K-slab kernel
    big = synth.make_ensemble(260, 520, 520, copies=6, seed=9, frac_straddle=0.3)      # 11 slabs: the
    big["heavy"] = False                                                               # atom-count dependent budget
    cases.append((big, cb.Env(ethr=1e9)))
    for name in ("adv_chiral", "adv_linear", "adv_planar"):
        g = np.load(os.path.join(GOLDEN, name + ".npz"))
        cases.append((dict(at=g["at"], xyz=g["xyz"], er=g["er"], heavy=False),
                      env_from_kwargs(cb, gold_kwargs(g), g["xyz"].shape[1])))
    for ens, env in cases:
        n = len(ens["er"])
        try:
            os.environ["CREST_CUDA_DEBUG"] = "3"
            ctx = cb.CregenContext()
            ctx.upload(env, ens["at"], ens["xyz"], ens["er"])
            ctx.pack()
            ctx.pairs(env)
            links = ctx.download()
            info = ctx.info(env)
            ctx.close()
        finally:
            os.environ.pop("CREST_CUDA_DEBUG", None)
        assert info["variant"] == 2 and info["flags"] == 0
        assert info["candidates"] == links["npairs"]          # every pair was re-checked
        assert info["exact_evals"] == links["npairs"]
        assert info["violations"] == 0
        ref, _, _ = _stage_a(cb, ens, env, "notf32")
        assert np.array_equal(links["dup"], ref["dup"])


@pytest.mark.parametrize("nall,nat", [(700, 60), (400, 300), (200, 1000)])
def test_tf32_accumulation_error_within_budget(cb, nall, nat):
    """The bound's gamma budgets the FP32 accumulation error of tcgen05.mma kind::tf32, which is not
    architecturally specified: measure it (every dot product of the TF32-rounded operands recomputed in
    FP64) and require it inside the budget of this atom count (profiles/r02_tf32_accum.txt: 3.4e9 dot
    products, worst 3.4 % of the budget)."""
    ens = synth.make_ensemble(nall, nat, nat, copies=5, seed=21)
    ctx = cb.CregenContext()
    ctx.upload(cb.Env(ethr=1e9), ens["at"], ens["xyz"], ens["er"])
    ctx.pack()
    r = ctx.tf32_accum()
    ctx.close()
    assert r["ndots"] == 9 * nall * (nall - 1) // 2
    assert 0.0 < r["max_rel_err"] < 0.25 * r["gamma_budget"]
    assert r["kappa"] >= 1.8e-3


@pytest.mark.parametrize("cfg,nall,ethr", [("C1", 1000, 1e9), ("C5", 3000, 0.2), ("C4", 700, 1e9)])
def test_tf32_dense_items_go_to_fp64_tensor_kernel(cb, cfg, nall, ethr):
    """`forcedense` hands EVERY work item of the screen to the FP64 tensor kernel through the
    device-built work list (column items for A <= 88, row tiles above): same links, same pair count."""
    ens = synth.make_config(cfg, nall=nall, seed=5)
    env = cb.Env(ethr=ethr)
    ref, rref, _ = _stage_a(cb, ens, env, "notf32")
    got, rgot, info = _stage_a(cb, ens, env, "forcedense")
    assert info["variant"] == 2 and info["dense_items"] == info["tiles_total"] and info["flags"] == 0
    assert info["exact_evals"] == 0
    assert np.array_equal(got["dup"], ref["dup"]) and np.array_equal(rgot, rref)
    assert got["npairs"] == ref["npairs"]


@pytest.mark.parametrize("nall,nat,copies", [(6000, 40, 300), (4000, 60, 2000), (1500, 300, 500)])
def test_tf32_collapsed_ensembles(cb, nall, nat, copies):
    """Collapsed ensembles (hundreds to thousands of copies per conformer: the iMTD-GC regime): nearly
    every candidate is a match.  The top-down scan decides a row with its first exact match, dense
    items go to the FP64 tensor kernel; links identical to the FP64-only run."""
    ens = synth.make_ensemble(nall, nat, nat, copies=copies, seed=17)
    ens["heavy"] = False
    env = cb.Env(ethr=1e9)
    ref, rref, _ = _stage_a(cb, ens, env, "notf32")
    got, rgot, info = _stage_a(cb, ens, env)
    assert info["variant"] == 2 and info["flags"] == 0
    assert np.array_equal(got["dup"], ref["dup"]) and np.array_equal(rgot, rref)
    assert got["npairs"] == ref["npairs"]
    assert info["candidates"] > 0.001 * got["npairs"]
    # without the dense path every candidate block is decided by the scan alone: still identical, and
    # far fewer exact evaluations than candidates (largest-j-wins pruning)
    got2, rgot2, info2 = _stage_a(cb, ens, env, "nodense")
    assert info2["dense_items"] == 0
    assert np.array_equal(got2["dup"], ref["dup"])
    assert info2["exact_evals"] < 0.5 * info2["candidates"]
    # a zero-ish budget inside dense items: rows that need more than 4 evaluations there hand their unit
    # to the FP64 tensor kernel
    got3, rgot3, info3 = _stage_a(cb, ens, env, "densebudget=1")
    assert np.array_equal(got3["dup"], ref["dup"]) and got3["npairs"] == ref["npairs"]


@pytest.mark.parametrize("cfg,nall", [("C3dense", None), ("C3adv", 30000), ("C4ext", None)])
def test_full_size_unfriendly_equal_fp64(cb, cfg, nall):
    """The unfriendly bench workloads at (or near) full size: collapsed ensemble (1000 copies per
    conformer), near-miss ensemble, extended chain.  Links of the screen path == FP64-only kernels, with
    and without the dense-item policy, in one pass and sharded over 8 virtual ranks."""
    ens = synth.make_config(cfg, nall=nall)
    env = cb.Env(ethr=1e9)
    ref, rref, _ = _stage_a(cb, ens, env, "notf32")
    got, rgot, info = _stage_a(cb, ens, env)
    assert info["variant"] == 2 and info["flags"] == 0
    assert np.array_equal(got["dup"], ref["dup"]) and got["npairs"] == ref["npairs"]
    got2, _, info2 = _stage_a(cb, ens, env, "nodense")
    assert np.array_equal(got2["dup"], ref["dup"])
    from crest_b200 import cregen
    ctx = cb.CregenContext()
    ctx.upload(env, ens["at"], ens["xyz"], ens["er"])
    ctx.pack()
    for r in range(8):
        ctx.pairs(env, r, 8)
    many = ctx.download()
    ctx.close()
    assert np.array_equal(many["dup"], ref["dup"]) and many["npairs"] == ref["npairs"]


def test_tf32_adversarial_near_misses(cb):
    """One conformer, noise of 0.135 A per coordinate: every pair at RMSD ~ 0.33 A, i.e. no match at all
    (nothing for the largest-j-wins pruning to use) but inside the zone the TF32 bound cannot decide.
    Rows spend their evaluation budget inside the dense items and hand their units to the FP64 tensor
    kernel; the result is the FP64-only one and the scan's work stays bounded."""
    ens = synth.make_config("C3adv", nall=3000, seed=77)
    env = cb.Env(ethr=1e9)
    ref, rref, _ = _stage_a(cb, ens, env, "notf32")
    got, rgot, info = _stage_a(cb, ens, env)
    assert info["variant"] == 2 and info["flags"] == 0
    assert np.array_equal(got["dup"], ref["dup"]) and got["npairs"] == ref["npairs"]
    assert (got["dup"] != 0).sum() < 30                       # (almost) nothing matches
    assert info["candidates"] > 0.9 * got["npairs"]           # the screen decides (almost) nothing
    assert info["dense_items"] > 0
    assert info["exact_evals"] <= 3000 * (512 + 64)           # the probe budget per row inside dense items


@pytest.mark.parametrize("cfg,nall,ethr,maskmb", [("C1", 1000, 1e9, 1), ("C5", 9000, 1e9, 1), ("C5", 9000, 0.3, 1),
                                                  ("C4", 1200, 1e9, 1)])
def test_tf32_multi_pass_mask(cb, cfg, nall, ethr, maskmb):
    """A candidate bitmask budget of 1 MB forces several passes (descending column blocks, links of one
    pass pruning the next): same links as one pass and as the FP64 kernels."""
    ens = synth.make_config(cfg, nall=nall, seed=41)
    env = cb.Env(ethr=ethr)
    ref, rref, _ = _stage_a(cb, ens, env, "notf32")
    one, rone, i1 = _stage_a(cb, ens, env)
    got, rgot, info = _stage_a(cb, ens, env, f"maskmb={maskmb},chunk=8")
    assert info["variant"] == 2
    if nall >= 5000:
        assert info["launches"] > i1["launches"]
    assert np.array_equal(got["dup"], ref["dup"]) and np.array_equal(one["dup"], ref["dup"])
    assert np.array_equal(rgot, rref)
    assert got["npairs"] == ref["npairs"] == one["npairs"]


@pytest.mark.parametrize("cfg,nall", [("C1", 700), ("C4", 300)])
def test_tf32_unsafe_magnitudes_fall_back(cb, cfg, nall):
    """The screen's sign tests assume finite FP32 arithmetic (test point < 1e9).  One structure blown
    up beyond that makes the converter raise the overflow flag: the launch is redone by the FP64 kernel
    and the result equals the FP64-only run."""
    ens = synth.make_config(cfg, nall=nall, seed=9)
    ens = dict(ens)
    xyz = ens["xyz"].copy()
    xyz[nall // 2] *= 3.0e4        # |x|^2 ~ 1e9 * (size in A)^2: half norm beyond 5e8
    ens["xyz"] = xyz
    env = cb.Env(ethr=1e9)
    ref, rref, _ = _stage_a(cb, ens, env, "notf32")
    got, rgot, info = _stage_a(cb, ens, env)
    assert info["variant"] == 2 and (info["flags"] & 1) == 1 and info["dense_items"] == info["tiles_total"]
    assert np.array_equal(got["dup"], ref["dup"]) and np.array_equal(rgot, rref)
    assert got["npairs"] == ref["npairs"]
    # and the flag is cleared by the next pack of a tame ensemble on a fresh context
    tame = synth.make_config(cfg, nall=nall, seed=9)
    _, _, info2 = _stage_a(cb, tame, env)
    assert info2["variant"] == 2 and info2["flags"] == 0


@pytest.mark.parametrize("nall", [2, 15, 16, 17, 127, 128, 129, 255, 257, 1025])
def test_tf32_ragged_tiles(cb, oracle, nall):
    """Ensemble sizes around the i-tile (128) and j-group (16) edges, against the oracle."""
    ens = synth.make_config("C5", nall=nall, seed=31)
    at, xyz, er = ens["at"], ens["xyz"], ens["er"]
    env = cb.Env(ethr=1e9)
    links, root, info = _stage_a(cb, ens, env)
    assert info["variant"] == 2
    ref = oracle.cregen_cre(at, xyz, er, ethr=1e9, want_rmat=True)
    assert links["npairs"] == nall * (nall - 1) // 2
    listed = boundary_rows(rows_of_allpairs(ref["rmat"], nall), env.rthr)
    bad = set(np.where(root != ref["dup"])[0].tolist())
    assert bad <= listed, sorted(bad - listed)


def test_tf32_sharded_equals_single(cb):
    """Contiguous shards of the tensor-core work list (what N ranks run) accumulate to the
    single-launch result (atomicMax on dup, atomicAdd on the pair counter)."""
    ens = synth.make_config("C2", nall=2500, seed=12)
    env = cb.Env(ethr=0.5, heavyrmsd=True)
    one, _, info = _stage_a(cb, ens, env)
    assert info["variant"] == 2
    ctx = cb.CregenContext()
    ctx.upload(env, ens["at"], ens["xyz"], ens["er"])
    ctx.pack()
    for r in range(3):
        ctx.pairs(env, r, 3)
    many = ctx.download()
    ctx.close()
    assert np.array_equal(many["dup"], one["dup"]) and many["npairs"] == one["npairs"]


@pytest.mark.parametrize("rthr,bthr", [(0.1, 0.01), (0.2, 0.02), (0.05, 0.005)])
def test_tf32_other_thresholds(cb, oracle, rthr, bthr):
    """Non-default RMSD / rot. thresholds (rthr not exactly representable in float32 shifts the
    float32-emulated boundary, cregen.f90:1286,1359-1367) through the tensor-core path vs the oracle."""
    ens = synth.make_ensemble(900, 40, 40, copies=10, seed=17, frac_straddle=0.3)
    ens["heavy"] = False
    at, xyz, er = ens["at"], ens["xyz"], ens["er"]
    env = cb.Env(ethr=1e9, rthr=rthr, bthr2=bthr)
    links, root, info = _stage_a(cb, ens, env)
    assert info["variant"] == 2
    ref = oracle.cregen_cre(at, xyz, er, ethr=1e9, rthr=rthr, bthr=bthr, want_rmat=True)
    listed = boundary_rows(rows_of_allpairs(ref["rmat"], len(er)), rthr)
    bad = set(np.where(root != ref["dup"])[0].tolist())
    assert bad <= listed, sorted(bad - listed)


def test_tf32_far_from_origin(cb):
    """Centring happens in FP64 before the TF32 rounding: a 5 km offset of every structure must not
    change a single decision."""
    ens = synth.make_config("C1", nall=600, seed=23)
    env = cb.Env(ethr=1e9)
    a, ra, _ = _stage_a(cb, ens, env)
    ens2 = dict(ens)
    ens2["xyz"] = ens["xyz"] + np.array([5.0e4, -3.0e4, 1.0e4])
    b, rb, info = _stage_a(cb, ens2, env)
    assert info["variant"] == 2
    # the FP64 reference of the shifted ensemble (cancellation in the centroid costs ~1e-11 A)
    c, rc, _ = _stage_a(cb, ens2, env, "notf32")
    assert np.array_equal(rb, rc)
    assert np.mean(ra != rb) < 0.01


def test_full_size_properties_c5(cb, oracle):
    """BASELINE config C5 at FULL size (1 000 000 x 40, 5.0e11 pairs) on one GPU: pair count,
    structure of the link array, and complete oracle rows for sampled conformers."""
    ens = synth.make_config("C5")
    at, xyz, er = ens["at"], ens["xyz"], ens["er"]
    n = len(er)
    env = cb.Env(ethr=1e9)
    ctx = cb.CregenContext()
    ctx.upload(env, at, xyz, er)
    ctx.pack()
    ctx.pairs(env)
    links = ctx.download(want_xyz=True)
    info = ctx.info(env)
    assert info["variant"] == 2 and info["flags"] == 0
    assert links["npairs"] == n * (n - 1) // 2
    dup = links["dup"]
    assert np.all(dup <= np.arange(n))
    ctx.sweep()
    root = ctx.download()["dup"]
    ctx.close()
    assert np.all((root == 0) == (dup == 0))
    assert np.all(root[root[root > 0] - 1] == 0)
    xa, rot = links["xyz_axis"], links["rot"]
    rows = sorted(np.random.default_rng(5).integers(n // 2, n, 6).tolist())
    for i in rows:
        _, rr = oracle.pair_rows(xa[: i + 1], er[: i + 1], i, i + 1, ethr=1e9)
        r = rr.astype(np.float64)
        exp = 0
        for j in np.where(r < 0.25)[0]:
            if r[j] < 0.125 or oracle.equalrotaniso(rot[i], rot[j], 0.005):
                exp = j + 1
        listed = np.any(np.abs(r - 0.125) <= RMSD_TOL) or np.any(np.abs(r - 0.25) <= RMSD_TOL)
        assert dup[i] == exp or listed, (i, dup[i], exp)


def test_rmsdalign_entry(cb, oracle):
    """crest_cuda_rmsdalign (cregen.f90:2229-2283) against the oracle: aligned
    coordinates within 1e-8 A; the first structure is untouched; hydrogens do not steer."""
    for cfg, nall in (("C1", 300), ("C4", 40), ("C2", 500)):
        ens = synth.make_config(cfg, nall=nall, seed=41)
        at, xyz = ens["at"], ens["xyz"]
        ref, nfail = oracle.rmsdalign(at, xyz)
        assert nfail == 0
        got = cb.rmsdalign(at, xyz)
        assert np.array_equal(got[0], xyz[0])
        assert np.abs(got - ref).max() < 1e-8
    # only hydrogens / helium: nothing to align on, input returned unchanged
    at = np.array([1, 1, 2, 1], np.int32)
    xyz = np.random.default_rng(0).normal(size=(5, 4, 3))
    assert np.array_equal(cb.rmsdalign(at, xyz), xyz)
    # nall = 0 / 1
    assert cb.rmsdalign(np.array([6, 8], np.int32), np.zeros((1, 2, 3))).shape == (1, 2, 3)


# ------------------------------------------------------------------------------------------------
# Engine of the one-call entry points (driver.cu): row-block upload pipeline, single-process multi-GPU
# ------------------------------------------------------------------------------------------------
def _one_call(cb, ens, env, want_xyz=True):
    at, xyz, er = ens["at"], ens["xyz"], ens["er"]
    nall, nat, _ = xyz.shape
    return cb.cregen_dup(env, nat, nall, at, xyz, er, want_xyz=want_xyz)


@pytest.mark.parametrize("cfg,nall,ethr", [("C3", 30000, 1e9), ("C5", 40000, 0.3), ("C4", 6000, 1e9)])
def test_engine_upload_pipeline_blocks(cb, cfg, nall, ethr):
    """The one-call ABI streams xyz in row blocks and starts the pair kernels of block k while block
    k+1 is on the bus: any number of blocks gives the links of the resident single-stage run."""
    from crest_b200 import cregen
    ens = synth.make_config(cfg, nall=nall, seed=12)
    env = cb.Env(ethr=ethr)
    ref, rref, _ = _stage_a(cb, ens, env)                   # context API: everything resident, one stage
    try:
        outs = []
        for tune in ("blocks=1", "blocks=3", "blocks=7", None):
            cregen.set_tune(tune)
            outs.append(_one_call(cb, ens, env))
    finally:
        cregen.set_tune(None)
    for o in outs:
        assert np.array_equal(o["dup"], rref)
        assert o["npairs"] == ref["npairs"]
        assert np.array_equal(o["rot"], outs[0]["rot"])
        assert np.array_equal(o["xyz_axis"], outs[0]["xyz_axis"])


def test_engine_finalize_and_reinit(cb):
    ens = synth.make_config("C1", nall=500, seed=2)
    env = cb.Env(ethr=1e9)
    a = _one_call(cb, ens, env)
    cb.finalize()
    assert cb.ngpu() == 0
    cb.finalize()                                            # idempotent
    b = _one_call(cb, ens, env)                              # lazily re-initialised
    assert cb.ngpu() == 1
    assert np.array_equal(a["dup"], b["dup"]) and np.array_equal(a["rot"], b["rot"])


@pytest.mark.parametrize("ngpu", [2, 4, 8])
def test_engine_single_process_multi_gpu(cb, cuda_lib, ngpu):
    """crest_cuda_init(ngpu) + the ONE-CALL ABI (what the Fortran shim binds) on 2/4/8 GPUs driven by
    one host thread: xyz split over the PCIe links, NVLink peer exchange, sharded pair work, links
    combined by atomicMax_system into GPU 0's dup[].  Identical to one GPU, for stage A alone and for
    the whole cregen_CRE."""
    if cuda_lib.crest_cuda_device_count() < ngpu:
        pytest.skip(f"needs {ngpu} GPUs")
    cases = [(synth.make_config("C3", nall=40000, seed=4), cb.Env(ethr=1e9)),
             (synth.make_config("C5", nall=30000, seed=4), cb.Env(ethr=0.2)),
             (synth.make_config("C4", nall=5000, seed=4), cb.Env(ethr=1e9)),
             (synth.make_ensemble(20000, 60, 60, copies=2000, seed=4), cb.Env(ethr=1e9)),
             (synth.make_config("C1", nall=300, seed=4), cb.Env(ethr=1e9))]
    try:
        for ens, env in cases:
            ens.setdefault("heavy", False)
            at, xyz, er = ens["at"], ens["xyz"], ens["er"]
            nall, nat, _ = xyz.shape
            assert cb.init(1) == 1
            one = _one_call(cb, ens, env)
            full1 = cb.cregen_CRE(env, nat, nall, at, xyz, er)
            assert cb.init(ngpu) == ngpu
            many = _one_call(cb, ens, env)
            fulln = cb.cregen_CRE(env, nat, nall, at, xyz, er)
            assert np.array_equal(one["dup"], many["dup"])
            assert one["npairs"] == many["npairs"]
            assert np.array_equal(one["rot"], many["rot"]) and np.array_equal(one["xyz_axis"], many["xyz_axis"])
            assert full1.nallout == fulln.nallout and np.array_equal(full1.order, fulln.order)
            assert np.array_equal(full1.group, fulln.group) and np.array_equal(full1.xyz, fulln.xyz)
    finally:
        cb.finalize()
