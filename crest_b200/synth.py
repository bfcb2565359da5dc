"""Seeded synthetic conformer ensembles for tests and bench.py.

A chain-like organic molecule is built from internal coordinates (heavy-atom backbone, bond 1.5 A,
angle ~110 deg, Z in {C,N,O}; hydrogens at 1.09 A).  K "true conformers" differ by torsion changes;
each is emitted several times as
  - near-duplicates   (Gaussian noise sigma = 0.02 A  => RMSD << rthr = 0.125 A),
  - mirror images     (10 %; exercises the enantiomer check, reference src/cregen.f90:1279-1284),
  - threshold straddlers (5 %; noise tuned so the RMSD falls in [0.10, 0.30] A around rthr / 2*rthr),
  - rotamers          (5 %; two hydrogens on one heavy atom swapped: same rot. constants and enuc,
                       large RMSD => grouped by K6, reference src/cregen.f90:1437-1456),
every structure with its own random rigid rotation + translation.  Energies: conformer base energy
uniform in a 6 kcal/mol window plus N(0, 0.01 kcal/mol) per copy, sorted ascending (the
cregen_esort post-condition).  Pure numpy, deterministic for a given seed.
"""
import numpy as np

AUTOKCAL = 627.50947428  # reference src/crest_pars.f90:23
SEED = 20261104


def _place(a, b, c, bond, angle, dihedral):
    """NeRF: position d with |cd|=bond, angle(b,c,d)=angle, dihedral(a,b,c,d). Vectorised (K,3)."""
    bc = c - b
    bc /= np.linalg.norm(bc, axis=-1, keepdims=True)
    n = np.cross(b - a, bc)
    nn = np.linalg.norm(n, axis=-1, keepdims=True)
    n = np.where(nn > 1e-9, n / np.maximum(nn, 1e-30), np.array([0.0, 0.0, 1.0]))
    m = np.cross(n, bc)
    d2 = np.stack([-bond * np.cos(angle) * np.ones_like(dihedral),
                   bond * np.sin(angle) * np.cos(dihedral),
                   bond * np.sin(angle) * np.sin(dihedral)], axis=-1)
    return c + d2[..., 0:1] * bc + d2[..., 1:2] * m + d2[..., 2:3] * n


def _random_rotations(rng, n):
    q = rng.normal(size=(n, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    w, x, y, z = q.T
    R = np.empty((n, 3, 3))
    R[:, 0, 0] = 1 - 2 * (y * y + z * z); R[:, 0, 1] = 2 * (x * y - z * w); R[:, 0, 2] = 2 * (x * z + y * w)
    R[:, 1, 0] = 2 * (x * y + z * w); R[:, 1, 1] = 1 - 2 * (x * x + z * z); R[:, 1, 2] = 2 * (y * z - x * w)
    R[:, 2, 0] = 2 * (x * z - y * w); R[:, 2, 1] = 2 * (y * z + x * w); R[:, 2, 2] = 1 - 2 * (x * x + y * y)
    return R


def make_molecule(rng, nat, nheavy, extended=False):
    """Topology: returns (at[nat], parent[nat] for hydrogens, base dihedrals).  extended: an
    all-trans backbone (dihedrals pi +- 0.2) instead of a random coil - a large radius of gyration,
    the unfriendly case for the TF32 screen (its undecided zone grows with Rg)."""
    nheavy = max(3, min(nheavy, nat))
    at = np.ones(nat, dtype=np.int32)
    at[:nheavy] = rng.choice([6, 7, 8], size=nheavy, p=[0.75, 0.10, 0.15])
    nh = nat - nheavy
    # hydrogens distributed round-robin over heavy atoms (up to 3 per heavy atom, then wrap)
    parent = np.arange(nh) % nheavy
    slot = np.arange(nh) // nheavy
    base_dih = rng.uniform(-np.pi, np.pi, size=nheavy)
    if extended:
        base_dih = np.pi + 0.2 * base_dih / np.pi
    return at, parent.astype(np.int64), slot.astype(np.int64), base_dih, nheavy


def build_conformers(rng, K, nat, nheavy, n_torsion_changes=4, extended=False):
    """K conformers (K,nat,3) of one molecule differing in a few backbone torsions."""
    at, parent, slot, base_dih, nheavy = make_molecule(rng, nat, nheavy, extended)
    dih = np.tile(base_dih, (K, 1))
    for _ in range(n_torsion_changes):
        which = rng.integers(3, max(4, nheavy), size=K)
        dih[np.arange(K), np.minimum(which, nheavy - 1)] += rng.choice(
            [2 * np.pi / 3, -2 * np.pi / 3, np.pi], size=K) + rng.normal(0, 0.15, size=K)
    xyz = np.zeros((K, nat, 3))
    ang = np.deg2rad(110.0)
    xyz[:, 0] = 0.0
    xyz[:, 1] = [1.5, 0.0, 0.0]
    xyz[:, 2] = xyz[:, 1] + [1.5 * np.cos(np.pi - ang), 1.5 * np.sin(np.pi - ang), 0.0]
    for i in range(3, nheavy):
        xyz[:, i] = _place(xyz[:, i - 3], xyz[:, i - 2], xyz[:, i - 1], 1.5, ang, dih[:, i])
    hang = np.deg2rad(109.5)
    for h in range(nat - nheavy):
        p = int(parent[h])
        if p >= 2:
            a, b = xyz[:, p - 2], xyz[:, p - 1]
        else:
            a, b = xyz[:, p + 2], xyz[:, p + 1]
        d = np.full(K, (slot[h] % 3) * 2 * np.pi / 3 + 0.3 + 0.5 * (slot[h] // 3))
        xyz[:, nheavy + h] = _place(a, b, xyz[:, p], 1.09, hang, d)
    return at, xyz, parent, slot, nheavy


def make_ensemble(nall, nat, nheavy=None, copies=8, seed=SEED, ewin_kcal=6.0,
                  frac_mirror=0.10, frac_straddle=0.05, frac_rotamer=0.05, sigma=0.02,
                  extended=False, n_torsion_changes=4):
    """Returns dict(at (nat,), xyz (nall,nat,3) C-order == Fortran xyz(3,nat,nall), er (nall,) Eh
    ascending, conf_id (nall,), kind (nall,) 0 dup /1 mirror /2 straddler /3 rotamer)."""
    rng = np.random.default_rng(seed)
    if nheavy is None:
        nheavy = nat
    K = max(1, (nall + copies - 1) // copies)
    at, conf, parent, slot, nheavy = build_conformers(rng, K, nat, nheavy, n_torsion_changes, extended)
    conf_id = rng.integers(0, K, size=nall)
    conf_id[:min(K, nall)] = np.arange(min(K, nall))      # every conformer at least once
    kind = np.zeros(nall, dtype=np.int32)
    u = rng.uniform(size=nall)
    kind[u < frac_mirror] = 1
    kind[(u >= frac_mirror) & (u < frac_mirror + frac_straddle)] = 2
    nh = nat - nheavy
    if nh >= 2:
        kind[(u >= frac_mirror + frac_straddle) & (u < frac_mirror + frac_straddle + frac_rotamer)] = 3
    kind[:min(K, nall)] = 0
    xyz = conf[conf_id].copy()
    sig = np.full(nall, sigma)
    ns = int((kind == 2).sum())
    sig[kind == 2] = rng.uniform(0.10, 0.30, size=ns) / np.sqrt(3.0)
    xyz += rng.normal(size=xyz.shape) * sig[:, None, None]
    xyz[kind == 1, :, 0] *= -1.0
    if nh >= 2:
        # swap the first two hydrogens that share a parent (slot 0 and slot 1 of parent 0)
        h0 = nheavy + 0
        h1 = nheavy + nheavy if nh > nheavy else None
        if h1 is not None and h1 < nat:
            r = np.where(kind == 3)[0]
            tmp = xyz[r, h0].copy()
            xyz[r, h0] = xyz[r, h1]
            xyz[r, h1] = tmp
        else:
            kind[kind == 3] = 0
    R = _random_rotations(rng, nall)
    xyz = np.einsum("nij,naj->nai", R, xyz) + rng.normal(size=(nall, 1, 3)) * 3.0
    e_conf = rng.uniform(0.0, ewin_kcal, size=K)
    er = (e_conf[conf_id] + rng.normal(0, 0.01, size=nall)) / AUTOKCAL - 100.0
    order = np.argsort(er, kind="stable")
    return dict(at=at, xyz=np.ascontiguousarray(xyz[order]), er=np.ascontiguousarray(er[order]),
                conf_id=conf_id[order], kind=kind[order], nheavy=nheavy)


# BASELINE.json configs: name -> kwargs of make_ensemble + mode flags
CONFIGS = {
    "C1": dict(nall=1000, nat=50, nheavy=23, heavy=False),
    "C2": dict(nall=10000, nat=100, nheavy=46, heavy=True),
    "C3": dict(nall=100000, nat=60, nheavy=60, heavy=False),
    "C4": dict(nall=20000, nat=300, nheavy=138, heavy=False),
    "C5": dict(nall=1000000, nat=40, nheavy=40, heavy=False),
    # unfriendly ensembles for the tensor-core screen: a COLLAPSED ensemble (1000 copies
    # per conformer: 100 000 structures -> 100 conformers, the regime after iMTD-GC,
    # reference src/algos/search_conformers.f90:179-180) and an EXTENDED chain (large Rg, conformers
    # one small torsion change apart: many near pairs the TF32 bound cannot decide)
    "C3dense": dict(nall=100000, nat=60, nheavy=60, heavy=False, copies=1000),
    # adversarial for the screen: ONE conformer with noise sigma = 0.135 A per coordinate: every pair sits
    # at RMSD ~ 0.33 A - above 2*rthr (no match, nothing prunes) but inside the zone the TF32 bound cannot
    # decide: all pairs become candidates and the dense items end on the FP64 tensor kernel
    "C3adv": dict(nall=100000, nat=60, nheavy=60, heavy=False, copies=100000, sigma=0.135,
                  frac_mirror=0.0, frac_straddle=0.0, frac_rotamer=0.0),
    "C4ext": dict(nall=20000, nat=300, nheavy=138, heavy=False, extended=True, copies=40,
                  n_torsion_changes=1),
}


def make_config(name, nall=None, seed=SEED):
    cfg = dict(CONFIGS[name])
    heavy = cfg.pop("heavy")
    if nall is not None:
        cfg["nall"] = nall
    ens = make_ensemble(seed=seed, **cfg)
    ens["heavy"] = heavy
    return ens
