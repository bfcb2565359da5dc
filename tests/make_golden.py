#!/usr/bin/env python
"""Generate tests/golden/*.npz — golden input/output vectors of the CREGEN sorting path.

The reference (Fortran 90) cannot be compiled in this image and its own test-suite holds no vector
for this path, so the vectors are minted by the C++ restatement under oracle/ —
PARITY UNPINNED with respect to a reference binary.  What pins the oracle instead is
tests/test_oracle.py (independent numpy/LAPACK cross-checks of every stage).  The one input
fixture the reference ships for this path, examples/expl-3/xtb.trj (50 structures x 12 atoms,
`crest struc.xyz -cregen xtb.trj -ewin 100.0`), is parsed here when /root/reference is present and
stored together with the oracle's outputs for it.

Run in the build container:   python tests/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
import oracle_lib as O  # noqa: E402
from crest_b200 import synth  # noqa: E402

GOLD = os.path.join(HERE, "golden")
SYM = {"h": 1, "c": 6, "n": 7, "o": 8}


def read_trj(path):
    """xyz trajectory: natoms / comment / atoms.  Energy = value after 'energy:' (the keyword rule
    of reference src/strucreader.f90:2225-2260)."""
    at, xyz, en = None, [], []
    with open(path) as f:
        lines = f.read().splitlines()
    p = 0
    while p < len(lines) and lines[p].strip():
        nat = int(lines[p].split()[0])
        toks = lines[p + 1].replace("energy:", "energy: ").split()
        en.append(float(toks[toks.index("energy:") + 1]))
        rec = [ln.split() for ln in lines[p + 2:p + 2 + nat]]
        at = np.array([SYM[r[0].lower()] for r in rec], np.int32)
        xyz.append([[float(v) for v in r[1:4]] for r in rec])
        p += 2 + nat
    return at, np.array(xyz), np.array(en)


def run_case(at, xyz, er, **kw):
    """Oracle outputs for an energy-sorted ensemble; everything a parity test compares."""
    want_rmat = xyz.shape[0] <= 400
    r = O.cregen_cre(at, xyz, er, want_rmat=want_rmat, **kw)
    out = dict(at=at, xyz=xyz, er=er, rot=r["rot"], xyz_out=r["xyz"], er_out=r["er"], dup=r["dup"],
               orderref=r["orderref"], group=r["group"], enuc=r["enuc"],
               nallout=np.int64(r["nallout"]), klong=np.int64(r["klong"]),
               nfail=np.int64(r["nfail"]))
    if want_rmat:
        out["rmat"] = r["rmat"]
    for k, v in kw.items():
        out["kw_" + k] = np.asarray(v)
    return out


def adversarial():
    """Small hand-made ensembles hitting the degenerate branches."""
    rng = np.random.default_rng(7)
    cases = {}
    # planar hexagon + H (benzene-like): lambda_max degenerate for in-plane flips
    ang = np.arange(6) * np.pi / 3
    ring = np.stack([1.39 * np.cos(ang), 1.39 * np.sin(ang), np.zeros(6)], 1)
    hs = ring * (2.48 / 1.39)
    base = np.concatenate([ring, hs])
    at = np.array([6] * 6 + [1] * 6, np.int32)
    X = [base, base.copy(), base[:, [1, 0, 2]], base + rng.normal(0, 0.01, base.shape),
         base + rng.normal(0, 0.05, base.shape), base * [1, 1, -1],
         base * [1.25, 1.0, 1.0], base * [1.25, 1.0, 1.0] + rng.normal(0, 0.08, base.shape)]
    cases["planar"] = (at, np.array(X))
    # linear triatomic (CO2-like): one moment ~0 => rot(1)=0 => Inf/NaN in equalrotaniso
    lin = np.array([[0, 0, -1.16], [0, 0, 0.0], [0, 0, 1.16]])
    X = [lin, lin + rng.normal(0, 0.01, lin.shape) * [0, 0, 1], lin * 1.02, lin[::-1].copy(),
         lin + [1.0, 2.0, 3.0], lin * 1.4,
         lin + np.array([[0, 0, 0], [0.5, 0, 0], [0, 0, 0]])]
    cases["linear"] = (np.array([8, 6, 8], np.int32), np.array(X))
    # identical copies, rotated/translated copies, mirror copies of a chiral 5-atom centre
    c = np.array([[0, 0, 0], [1.09, 0, 0], [-0.5, 1.3, 0], [-0.5, -0.6, 1.2], [-0.4, -0.7, -1.4]])
    R = synth._random_rotations(rng, 4)
    X = [c, c.copy(), c @ R[0].T + 1.0, (c * [-1, 1, 1]) @ R[1].T, c + rng.normal(0, 0.03, c.shape),
         c + rng.normal(0, 0.09, c.shape), (c + rng.normal(0, 0.09, c.shape)) @ R[2].T,
         c + np.array([[0, 0, 0], [0, 0, 0], [0, 0, 0], [0, 0, 0], [1.0, 0.5, 0]]), c * 1.2]
    cases["chiral"] = (np.array([6, 1, 9, 17, 35], np.int32), np.array(X))
    return cases


def main():
    os.makedirs(GOLD, exist_ok=True)
    O.build()
    # 1. the reference's own example ensemble
    trj = "/root/reference/examples/expl-3/xtb.trj"
    if os.path.exists(trj):
        at, xyz, en = read_trj(trj)
        es, pos, nout = O.esort(en, ewin=100.0)          # `-ewin 100.0` of examples/expl-3/run.sh
        xs = np.empty_like(xyz)
        xs[pos - 1] = xyz
        xs, es = xs[:nout], es[:nout]
        out = run_case(at, np.ascontiguousarray(xs), np.ascontiguousarray(es))
        out.update(raw_xyz=xyz, raw_energy=en, esort_pos=pos, esort_nout=np.int64(nout))
        np.savez_compressed(os.path.join(GOLD, "expl3.npz"), **out)
        print("expl3: nall", nout, "kept", int(out["nallout"]), "groups", int(out["group"][0]))
    # 2. synthetic shapes of BASELINE.json configs, cut to oracle-seconds sizes
    specs = {
        "c1_allpairs": ("C1", 240, dict(ethr=1e9)),
        "c1_band": ("C1", 240, dict(ethr=0.05)),
        "c1_noenantio": ("C1", 160, dict(ethr=1e9, enantio=False)),
        "c1_nosort": ("C1", 160, dict(ethr=0.05, nosort=True)),
        "c2_heavy": ("C2", 160, dict(ethr=1e9, heavy=True)),
        "c3_shape": ("C3", 200, dict(ethr=1e9)),
        "c5_shape": ("C5", 200, dict(ethr=0.2, rthr=0.1)),
    }
    for name, (cfg, nall, kw) in specs.items():
        ens = synth.make_config(cfg, nall=nall)
        out = run_case(ens["at"], ens["xyz"], ens["er"], **kw)
        np.savez_compressed(os.path.join(GOLD, name + ".npz"), **out)
        print(name, "klong", int(out["klong"]), "kept", int(out["nallout"]), "groups",
              int(out["group"][0]), "nfail", int(out["nfail"]))
    # substructure RMSD (includeRMSD mask; rot from the masked atoms, cregen.f90:1225-1230)
    ens = synth.make_config("C1", nall=120)
    inc = (np.arange(50) % 3 != 1).astype(np.int32)
    out = run_case(ens["at"], ens["xyz"], ens["er"], ethr=1e9, include_rmsd=inc)
    np.savez_compressed(os.path.join(GOLD, "c1_subrmsd.npz"), **out)
    # 3. adversarial geometries
    for name, (at, xyz) in adversarial().items():
        er = -10.0 + 1e-6 * np.arange(xyz.shape[0])
        out = run_case(at, np.ascontiguousarray(xyz, np.float64), er, ethr=1e9)
        np.savez_compressed(os.path.join(GOLD, "adv_" + name + ".npz"), **out)
        print("adv", name, "dup", out["dup"].tolist(), "group", out["group"].tolist(),
              "nfail", int(out["nfail"]))
    # 4. pair-level known answers: rmsd / lambda / covariance of explicit pairs
    ens = synth.make_config("C3", nall=64)
    xa, _ = O.axis(ens["at"], ens["xyz"])
    ii, jj = np.tril_indices(64, -1)
    sel = np.random.default_rng(3).choice(len(ii), 300, replace=False)
    rec = []
    for i, j in zip(ii[sel], jj[sel]):
        a = O.rmsd(xa[i], xa[j])
        ym = xa[j].copy()
        ym[:, 0] *= -1
        b = O.rmsd(xa[i], ym)
        rec.append([a["rmsd"], b["rmsd"], a["lam"], b["lam"], a["xn"], a["yn"]])
    np.savez_compressed(os.path.join(GOLD, "pairs_kat.npz"), packed=xa, i=ii[sel].astype(np.int32),
                        j=jj[sel].astype(np.int32), rec=np.array(rec))
    print("golden vectors written to", GOLD)


if __name__ == "__main__":
    main()
