"""An INDEPENDENT restatement of cregen_CRE's decisions in numpy (second pin of the oracle).

Nothing here shares code or formulas with oracle/: RMSDs come from a batched Kabsch SVD (proper and
improper superposition), rot. constants from numpy.linalg.eigvalsh of the inertia tensor, erf from the
math module; only the constants of the reference (masses atmasses.f90:27-56, autokcal,
icm2MHz*Aamu2icm) and its decision rules (cregen.f90:1353-1376 K4 on the float32-rounded RMSD,
1397-1419 partition, 1422-1435 enuc, 1437-1474 K6 + group ids; rotcompare.f90:109-169) are restated.
It is O(N^2) python/numpy and meant for the golden ensembles (N <= a few hundred)."""
import math
import numpy as np

AUTOKCAL = 627.50947428
_AMS = [
        1.00794075, 4.00260193, 6.94003660, 9.01218307, 10.81102805, 12.01073590, 14.00670321, 15.99940492,
        18.99840316, 20.18004638, 22.98976928, 24.30505162, 26.98153853, 28.08549871, 30.97376200, 32.06478741,
        35.45293758, 39.94779856, 39.09830091, 40.07802251, 44.95590828, 47.86674496, 50.94146504, 51.99613176,
        54.93804391, 55.84514443, 58.93319429, 58.69334711, 63.54603995, 65.37778253, 69.72306607, 72.62755016,
        74.92159457, 78.95938856, 79.90352778, 83.79800000, 85.46766360, 87.61664447, 88.90584030, 91.22364160,
        92.90637300, 95.95978854, 97.90721240, 101.06494014, 102.90549800, 106.41532751, 107.86814963, 112.41155782,
        114.81808663, 118.71011259, 121.75978367, 127.60312648, 126.90447190, 131.29276145, 132.90545196, 137.32689163,
        138.90546887, 140.11573074, 140.90765760, 144.24159603, 144.91275590, 150.36635571, 151.96437813, 157.25213065,
        158.92535470, 162.49947282, 164.93032880, 167.25908265, 168.93421790, 173.05415017, 174.96681496, 178.48497872,
        180.94787564, 183.84177755, 186.20670455, 190.22485963, 192.21605165, 195.08445686, 196.96656879, 200.59916703,
        204.38341284, 207.21690806, 208.98039910, 208.98243080, 209.98714790, 222.01757820, 223.01973600, 226.02541030,
        227.02775230, 232.03805580, 231.03588420, 238.02891046, 237.04817360, 244.06420530, 243.06138130, 247.07035410,
        247.07030730, 251.07958860, 252.08298000, 257.09510610, 258.09843150, 259.10103000, 262.10961000, 267.12179000,
        269.12791000, 271.13393000, 270.13336000, 276.14846000, 276.15159000, 280.16131000, 282.16912000, 284.17416000,
        284.17873000, 289.19042000, 288.19274000, 293.20449000, 292.20746000, 294.21392000,
]
MASS = {z + 1: m for z, m in enumerate(_AMS)}
ROTCONV = 2.9979245e4 * 16.8576522


def rot_constants(at, x):
    m = np.array([MASS[int(z)] for z in at])
    c = x - (m[:, None] * x).sum(0) / (m.sum() + 1e-20)
    r2 = (c * c).sum(1)
    T = (m * r2).sum() * np.eye(3) - (m[:, None, None] * c[:, :, None] * c[:, None, :]).sum(0)
    w = np.linalg.eigvalsh(T)                        # ascending moments -> rot A >= B >= C
    return np.where(w < 3e-4, 0.0, ROTCONV / np.where(w < 3e-4, 1.0, w))


def batched_min_rmsd(P, i_idx, j_idx, enantio):
    """min(proper, improper) Kabsch RMSD of structure pairs; P (n,A,3) centred."""
    A = P.shape[1]
    H = np.einsum("pak,pal->pkl", P[i_idx], P[j_idx])
    s = np.linalg.svd(H, compute_uv=False)
    d = np.sign(np.linalg.det(H))
    d[d == 0] = 1.0
    nrm = (P[i_idx] ** 2).sum((1, 2)) + (P[j_idx] ** 2).sum((1, 2))
    proper = nrm - 2.0 * (s[:, 0] + s[:, 1] + d * s[:, 2])
    improper = nrm - 2.0 * (s[:, 0] + s[:, 1] - d * s[:, 2])
    e = np.minimum(proper, improper) if enantio else proper
    return np.sqrt(np.maximum(e, 0.0) / A)


def aniso(r):
    av = r.mean()
    return math.sqrt(((r - av) ** 2).sum()) / av / (3.0 * math.sqrt(2.0 / 3.0))


def equal_rot(ri, rj, bthr, bthrmax=0.025, bthrshift=0.5):
    with np.errstate(all="ignore"):
        av = 0.5 * (aniso(ri) + aniso(rj))
        c = (bthrmax * 100.0 - bthr * 100.0) / 2.0
        thr = (math.erf(av * 5.0 - bthrshift / 0.15) * c + (math.erf(2.5) * c + bthr * 100.0)) / 100.0 \
            if av == av else float("nan")
        q = np.abs(ri / rj - 1.0)
        return bool(np.all(q <= thr))


def cregen_cre(at, xyz, er, ethr=0.05, rthr=0.125, bthr=0.01, enantio=True, heavy=False,
               include_rmsd=None, nosort=False, tol=1e-6):
    """Returns dict(dup (roots, 1-based, stage A numbering), orderref, group, nallout, boundary (rows whose
    decision sits within `tol` of a threshold: RMSD vs rthr / 2 rthr, rot ratio vs its threshold))."""
    at = np.asarray(at)
    n, nat, _ = xyz.shape
    substruc = include_rmsd is not None
    sel = np.ones(nat, bool)
    if substruc:
        sel = np.asarray(include_rmsd) > 0
    elif heavy:
        sel = at != 1
    rot = np.array([rot_constants(at[sel] if substruc else at, x[sel] if substruc else x) for x in xyz])
    P = xyz[:, sel, :] - xyz[:, sel, :].mean(1, keepdims=True)
    ii, jj = np.tril_indices(n, -1)
    gate = (er[ii] - er[jj]) * AUTOKCAL < ethr
    ii, jj = ii[gate], jj[gate]
    r64 = np.zeros(len(ii))
    for b in range(0, len(ii), 100000):               # bounded memory for the gathered operands
        r64[b:b + 100000] = batched_min_rmsd(P, ii[b:b + 100000], jj[b:b + 100000], enantio)
    dr = r64.astype(np.float32).astype(np.float64)
    dup = np.zeros(n + 1, np.int64)
    boundary = set()
    near = (np.abs(dr - rthr) <= tol) | (np.abs(dr - 2 * rthr) <= tol)
    boundary.update((ii[near]).tolist())
    order = np.lexsort((jj, ii))                      # row-major: i ascending, j ascending
    for p in order:
        i, j = ii[p] + 1, jj[p] + 1
        if dr[p] < rthr:
            dup[i] = j
        elif dr[p] < 2 * rthr and equal_rot(rot[i - 1], rot[j - 1], 0.5 * bthr):
            dup[i] = j
    for i in range(1, n + 1):                         # backtrack in row order (cregen_old.f90:1396-1410)
        k = i
        while dup[k] > 0:
            k = dup[k]
        if dup[i] > 0:
            dup[i] = k
    kept = [i for i in range(1, n + 1) if dup[i] == 0]
    orderref = np.arange(1, n + 1)
    nallout = n
    if not nosort:
        j, l = 0, n + 1
        for i in range(1, n + 1):
            if dup[i] == 0:
                j += 1; orderref[i - 1] = j
            else:
                l -= 1; orderref[i - 1] = l
        nallout = len(kept)
    inv = np.argsort(orderref)                        # position p holds original inv[p]
    xs, es, rs = xyz[inv], er[inv], rot[inv]
    enuc = np.zeros(nallout)
    for k in range(nallout):
        d2 = ((xs[k][:, None, :] - xs[k][None, :, :]) ** 2).sum(2) + 1e-12
        zz = at[:, None] * at[None, :]
        iu = np.triu_indices(nat, 1)
        enuc[k] = (zz[iu] / d2[iu]).sum()
    head = np.zeros(n + 1, np.int64)
    for i in range(1, nallout + 1):
        for j in range(1, i):
            if head[j] != 0:
                continue
            if abs(es[i - 1] - es[j - 1]) * AUTOKCAL < ethr and equal_rot(rs[i - 1], rs[j - 1], bthr) and \
                    2.0 * abs(enuc[i - 1] - enuc[j - 1]) / (enuc[i - 1] + enuc[j - 1]) < 1e-3:
                k = j
                head[i] = k
                break
    group = np.zeros(n + 1, np.int64)
    ng = 0
    for i in range(1, nallout + 1):
        if head[i] == 0:
            ng += 1; group[i] = ng
        else:
            group[i] = group[head[i]]
    group[0] = ng
    if nosort:
        nallout = ng
    return dict(dup=dup[1:], orderref=orderref, group=group, nallout=nallout, rot=rot, boundary=sorted(boundary),
                rmsd=r64, pairs=(ii, jj))
