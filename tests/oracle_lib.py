"""ctypes access to the CPU oracle (oracle/libcregen_oracle.so).  TEST INFRASTRUCTURE ONLY:
imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs."""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_DIR = os.path.join(os.path.dirname(_HERE), "oracle")
_LIB = None

c_dp = C.POINTER(C.c_double)
c_ip = C.POINTER(C.c_int)
c_fp = C.POINTER(C.c_float)
c_lp = C.POINTER(C.c_longlong)


def build():
    subprocess.run(["make", "-s", "-C", ORACLE_DIR], check=True)


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(ORACLE_DIR, "libcregen_oracle.so")
        if not os.path.exists(so):
            build()
        L = C.CDLL(so)
        L.oracle_rmsd.restype = C.c_double
        L.oracle_rmsd.argtypes = [C.c_int, c_dp, c_dp, c_dp]
        L.oracle_dstmev.restype = C.c_int
        L.oracle_dstmev.argtypes = [c_dp, c_dp, c_dp]
        L.oracle_axis.argtypes = [C.c_int, C.c_int, c_ip, c_dp, c_dp]
        L.oracle_rotconst.argtypes = [C.c_int, C.c_int, c_ip, c_dp, c_dp]
        L.oracle_equalrotaniso.restype = C.c_int
        L.oracle_equalrotaniso.argtypes = [c_dp, c_dp, C.c_double, C.c_double, C.c_double]
        L.oracle_bthrerf.restype = C.c_double
        L.oracle_bthrerf.argtypes = [C.c_double] * 4
        L.oracle_rotaniso.restype = C.c_double
        L.oracle_rotaniso.argtypes = [c_dp]
        L.oracle_rmsdalign.restype = C.c_int
        L.oracle_rmsdalign.argtypes = [C.c_int, C.c_int, c_ip, c_dp]
        L.oracle_esort.restype = C.c_int
        L.oracle_esort.argtypes = [C.c_int, c_dp, c_ip, C.c_double]
        L.oracle_backtrack.argtypes = [C.c_int, c_ip]
        L.oracle_cregen_cre.restype = C.c_int
        L.oracle_cregen_cre.argtypes = [C.c_int, C.c_int, c_ip, c_dp, c_dp, c_ip, C.c_int, C.c_int,
                                        C.c_int, C.c_int] + [C.c_double] * 5 + \
            [c_dp, c_dp, c_ip, c_ip, c_ip, c_dp, c_fp, C.c_longlong, c_lp]
        L.oracle_set_boundary_tol.restype = None
        L.oracle_set_boundary_tol.argtypes = [C.c_double]
        L.oracle_boundary_rows.restype = C.c_int
        L.oracle_boundary_rows.argtypes = [c_ip, C.c_int]
        L.oracle_pair_rows.restype = C.c_longlong
        L.oracle_pair_rows.argtypes = [C.c_int, C.c_int, c_dp, c_dp, C.c_double, C.c_int, C.c_int,
                                       C.c_int, c_fp]
        L.oracle_crossing_dedup.restype = None
        L.oracle_crossing_dedup.argtypes = [C.c_int, C.c_int, c_dp, c_dp, C.c_double, C.c_int, C.c_int, c_ip, c_dp,
                                            c_ip, c_lp, c_dp, c_dp]
        L.oracle_num_threads.restype = C.c_int
        L.oracle_set_num_threads.restype = None
        L.oracle_set_num_threads.argtypes = [C.c_int]
        L.oracle_compare_blocks.restype = None
        L.oracle_compare_blocks.argtypes = [C.c_int, c_dp, C.c_int, c_ip, c_ip, c_dp, C.c_int, c_ip,
                                            c_ip, c_dp]
        L.oracle_compare_correlate.restype = C.c_int
        L.oracle_compare_correlate.argtypes = [C.c_int, C.c_int, c_dp, c_dp, c_dp, C.c_double, c_ip, c_ip]
        L.oracle_intraconf.restype = None
        L.oracle_intraconf.argtypes = [C.c_int, C.c_int, c_dp, c_ip, C.c_int, c_ip, c_ip, C.c_char_p,
                                       c_dp, c_dp, C.c_int, c_dp]
        _LIB = L
    return _LIB


def _dp(a):
    return a.ctypes.data_as(c_dp)


def _ip(a):
    return a.ctypes.data_as(c_ip)


def rmsd(c1, c2):
    """c1,c2: (n,3) float64. Returns dict(rmsd, lam, xn, yn, R(3,3), q(4))."""
    c1 = np.ascontiguousarray(c1, np.float64)
    c2 = np.ascontiguousarray(c2, np.float64)
    out = np.zeros(17)
    lib().oracle_rmsd(c1.shape[0], _dp(c1), _dp(c2), _dp(out))
    return dict(rmsd=out[0], lam=out[1], xn=out[2], yn=out[3], R=out[4:13].reshape(3, 3).copy(),
                q=out[13:17].copy())


def dstmev(S):
    S = np.ascontiguousarray(S, np.float64)
    lam = C.c_double(0)
    ev = np.zeros(4)
    io = lib().oracle_dstmev(_dp(S), C.byref(lam), _dp(ev))
    return io, lam.value, ev


def axis(at, xyz):
    """xyz (n,nat,3) -> (xyz in principal frame, rot (n,3))"""
    xyz = np.array(xyz, np.float64, order="C", copy=True)
    at = np.ascontiguousarray(at, np.int32)
    n, nat, _ = xyz.shape
    rot = np.zeros((n, 3))
    lib().oracle_axis(nat, n, _ip(at), _dp(xyz), _dp(rot))
    return xyz, rot


def rotconst(at, xyz):
    xyz = np.ascontiguousarray(xyz, np.float64)
    at = np.ascontiguousarray(at, np.int32)
    n, nat, _ = xyz.shape
    rot = np.zeros((n, 3))
    lib().oracle_rotconst(nat, n, _ip(at), _dp(xyz), _dp(rot))
    return rot


def equalrotaniso(ri, rj, bthr, bthrmax=0.025, bthrshift=0.5):
    ri = np.ascontiguousarray(ri, np.float64)
    rj = np.ascontiguousarray(rj, np.float64)
    return bool(lib().oracle_equalrotaniso(_dp(ri), _dp(rj), bthr, bthrmax, bthrshift))


def rmsdalign(at, xyz):
    """Oracle cregen_rmsdalign (cregen.f90:2229-2283); returns (aligned copy, nfail)."""
    at = np.ascontiguousarray(at, np.int32)
    out = np.array(xyz, np.float64, order="C", copy=True)
    nall, nat, _ = out.shape
    nfail = lib().oracle_rmsdalign(nat, nall, _ip(at), _dp(out))
    return out, nfail


def esort(energies, ewin=1e9):
    e = np.array(energies, np.float64, copy=True)
    pos = np.zeros(len(e), np.int32)
    nout = lib().oracle_esort(len(e), _dp(e), _ip(pos), ewin)
    return e, pos, nout


def cregen_cre(at, xyz, er, *, ethr=0.05, rthr=0.125, bthr=0.01, bthrmax=0.025, bthrshift=0.5,
               enantio=True, heavy=False, include_rmsd=None, nosort=False, want_rmat=False,
               boundary_tol=None):
    """Run the oracle cregen_CRE.  xyz (nall,nat,3); er ascending (Eh). Returns a dict; xyz is
    returned transformed + permuted like the reference leaves it."""
    xyz = np.array(xyz, np.float64, order="C", copy=True)
    at = np.ascontiguousarray(at, np.int32)
    er = np.ascontiguousarray(er, np.float64)
    nall, nat, _ = xyz.shape
    substruc = include_rmsd is not None
    inc = np.ascontiguousarray(include_rmsd if substruc else np.ones(nat), np.int32)
    rot = np.zeros((nall, 3))
    er_out = np.zeros(nall)
    dup = np.zeros(nall, np.int32)
    orderref = np.zeros(nall, np.int32)
    group = np.zeros(nall + 1, np.int32)
    enuc = np.zeros(nall)
    info = np.zeros(4, np.int64)
    cap = nall * (nall - 1) // 2 if want_rmat else 0
    rmat = np.zeros(max(cap, 1), np.float32) if want_rmat else None
    lib().oracle_set_boundary_tol(-1.0 if boundary_tol is None else float(boundary_tol))
    rc = lib().oracle_cregen_cre(nat, nall, _ip(at), _dp(xyz), _dp(er), _ip(inc), int(substruc),
                                 int(heavy), int(enantio), int(nosort), ethr, rthr, bthr, bthrmax,
                                 bthrshift, _dp(rot), _dp(er_out), _ip(dup), _ip(orderref),
                                 _ip(group), _dp(enuc),
                                 rmat.ctypes.data_as(c_fp) if want_rmat else None, cap,
                                 info.ctypes.data_as(c_lp))
    assert rc == 0
    res = dict(xyz=xyz, rot=rot, er=er_out, dup=dup, orderref=orderref, group=group,
               enuc=enuc[:int(info[0]) if not nosort else nall], nallout=int(info[0]),
               nkept=int(info[1]), klong=int(info[2]), nfail=int(info[3]))
    if want_rmat:
        res["rmat"] = rmat[:int(info[2])]
    if boundary_tol is not None:
        nb = lib().oracle_boundary_rows(None, 0)
        rows = np.zeros(max(nb, 1), np.int32)
        lib().oracle_boundary_rows(_ip(rows), nb)
        res["boundary_rows"] = rows[:nb] - 1      # 0-based rows (stage-A numbering)
        lib().oracle_set_boundary_tol(-1.0)
    return res


def pair_rows(xyz_packed, er, i0, i1, *, ethr=1e9, enantio=True, want=True):
    """Reference-shaped K1 loop on rows [i0,i1) of a packed (nall,n,3) ensemble."""
    xyz = np.ascontiguousarray(xyz_packed, np.float64)
    er = np.ascontiguousarray(er, np.float64)
    nall, n, _ = xyz.shape
    i1 = min(i1, nall)
    tot = sum(range(i0, i1))
    out = np.zeros(max(tot, 1), np.float32) if want else None
    npairs = lib().oracle_pair_rows(n, nall, _dp(xyz), _dp(er), ethr, int(enantio), i0, i1,
                                    out.ctypes.data_as(c_fp) if want else None)
    return npairs, (out[:tot] if want else None)


def num_threads():
    return lib().oracle_num_threads()


def set_num_threads(n):
    """OpenMP threads of the oracle's loops (torchrun exports OMP_NUM_THREADS=1: bench.py overrides it)."""
    lib().oracle_set_num_threads(int(n))


def compare_blocks(xyz1, b1, e1, xyz2, b2, e2):
    """ensemblecomp.f90:213-248; returns rmat (nconf1, nconf2)."""
    xyz1 = np.ascontiguousarray(xyz1, np.float64)
    xyz2 = np.ascontiguousarray(xyz2, np.float64)
    b1, e1, b2, e2 = (np.ascontiguousarray(a, np.int32) for a in (b1, e1, b2, e2))
    out = np.zeros((len(b2), len(b1)))
    lib().oracle_compare_blocks(xyz1.shape[1], _dp(xyz1), len(b1), _ip(b1), _ip(e1), _dp(xyz2), len(b2),
                                _ip(b2), _ip(e2), _dp(out))
    return out.T


def compare_correlate(en1, en2, rmat, rthr):
    en1 = np.ascontiguousarray(en1, np.float64)
    en2 = np.ascontiguousarray(en2, np.float64)
    rm = np.ascontiguousarray(np.asarray(rmat, np.float64).T)
    left = np.zeros(len(en1) + len(en2) + 1, np.int32)
    right = np.zeros_like(left)
    nl = lib().oracle_compare_correlate(len(en1), len(en2), _dp(en1), _dp(en2), _dp(rm), float(rthr),
                                        _ip(left), _ip(right))
    return left[:nl].copy(), right[:nl].copy()


def intraconf(xyz, inc, low, nrot, symsym, enantiofac=None, rmat_conf=-1):
    """entropic.f90:938-1097; returns (fac, enantiofac[, rmat of conformer rmat_conf])."""
    xyz = np.ascontiguousarray(xyz, np.float64)
    inc = np.ascontiguousarray(inc, np.int32)
    low = np.ascontiguousarray(low, np.int32)
    nrot = np.ascontiguousarray(nrot, np.int32)
    nconf = len(low)
    sym = "".join(f"{s:<3.3}" for s in symsym).encode()
    fac = np.ones(max(nconf, 1))
    ef = np.ones(max(nconf, 1)) if enantiofac is None else np.array(enantiofac, np.float64, copy=True)
    rm = None
    if rmat_conf >= 0:
        rm = np.zeros((nrot[rmat_conf], nrot[rmat_conf]))
    lib().oracle_intraconf(xyz.shape[1], xyz.shape[0], _dp(xyz), _ip(inc), nconf, _ip(low), _ip(nrot),
                           sym, _dp(fac), _dp(ef), rmat_conf, _dp(rm) if rm is not None else None)
    if rm is not None:
        return fac[:nconf], ef[:nconf], rm
    return fac[:nconf], ef[:nconf]


def crossing_dedup(cand, xyzref, rthr, maxgen2, huge_number=False):
    """crossing.f90:227-338 with one thread on generated structures cand (ncand,nat,3); returns a dict."""
    cand = np.ascontiguousarray(cand, np.float64)
    ref = np.ascontiguousarray(xyzref, np.float64)
    ncand, nat, _ = cand.shape
    taken = np.zeros(maxgen2, np.int32)
    rms = np.zeros(maxgen2)
    nt = C.c_int(0)
    cnt = np.zeros(4, np.int64)
    rsum = C.c_double(0)
    rval = np.zeros(max(ncand, 1))
    lib().oracle_crossing_dedup(nat, ncand, _dp(cand), _dp(ref), float(rthr), int(maxgen2), int(huge_number),
                                _ip(taken), _dp(rms), C.byref(nt), cnt.ctypes.data_as(c_lp), C.byref(rsum), _dp(rval))
    return dict(taken=taken, rms=rms, ntaken=nt.value, ierr=int(cnt[0]), ident=int(cnt[1]), rcount=int(cnt[2]),
                processed=int(cnt[3]), rmsdsum=rsum.value, rval=rval[:ncand])
