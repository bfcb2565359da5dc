"""Host-side mirror of the two rmsd() callers beside cregen_CRE, over the C ABI.

compare_blocks / compare_correlate: compare_ensembles, /root/reference/src/ensemblecomp.f90:213-300
intraconf:                          intraconfRMSD, /root/reference/src/entropy/entropic.f90:938-1097
Arrays are (nall, nat, 3) C-order = xyz(3,nat,nall) column-major as Fortran passes them."""
import ctypes as C

import numpy as np

from . import _lib
from .cregen import _dp, _ip


def compare_blocks(xyz1, b1, e1, xyz2, b2, e2):
    """rmat(nconf1, nconf2): minimum all-atom RMSD between the rotamer blocks [b1(i), e1(i)] of
    ensemble 1 and [b2(j), e2(j)] of ensemble 2 (1-based inclusive)."""
    L = _lib.lib()
    xyz1 = np.ascontiguousarray(xyz1, np.float64)
    xyz2 = np.ascontiguousarray(xyz2, np.float64)
    if xyz1.shape[1:] != xyz2.shape[1:]:
        raise ValueError("Nat1 /= Nat2")          # ensemblecomp.f90:178-182
    b1, e1, b2, e2 = (np.ascontiguousarray(a, np.int32) for a in (b1, e1, b2, e2))
    out = np.zeros((len(b2), len(b1)))            # column-major (nconf1, nconf2)
    _lib.check(L.crest_cuda_compare_blocks(xyz1.shape[1], xyz1.shape[0], _dp(xyz1), len(b1),
                                           _ip(b1), _ip(e1), xyz2.shape[0], _dp(xyz2), len(b2),
                                           _ip(b2), _ip(e2), _dp(out)))
    return out.T


def compare_correlate(en1, en2, rmat, rthr):
    """The 'Correlation between Conformers' lines as (left, right) 1-based numbers, 0 = blank."""
    L = _lib.lib()
    en1 = np.ascontiguousarray(en1, np.float64)
    en2 = np.ascontiguousarray(en2, np.float64)
    rm = np.ascontiguousarray(np.asarray(rmat, np.float64).T)
    left = np.zeros(len(en1) + len(en2) + 1, np.int32)
    right = np.zeros_like(left)
    nl = L.crest_cuda_compare_correlate(len(en1), len(en2), _dp(en1), _dp(en2), _dp(rm), float(rthr),
                                        _ip(left), _ip(right))
    if nl < 0:
        _lib.check(1)
    return left[:nl].copy(), right[:nl].copy()


def intraconf(xyz, inc, low, nrot, symsym, enantiofac=None):
    """Returns (fac, enantiofac) per conformer; low 1-based first rotamer, symsym list of str."""
    L = _lib.lib()
    xyz = np.ascontiguousarray(xyz, np.float64)
    inc = np.ascontiguousarray(inc, np.int32)
    low = np.ascontiguousarray(low, np.int32)
    nrot = np.ascontiguousarray(nrot, np.int32)
    nconf = len(low)
    sym = "".join(f"{s:<3.3}" for s in symsym).encode()
    fac = np.ones(max(nconf, 1))
    ef = np.ones(max(nconf, 1)) if enantiofac is None else np.array(enantiofac, np.float64, copy=True)
    _lib.check(L.crest_cuda_intraconf(xyz.shape[1], xyz.shape[0], _dp(xyz), _ip(inc), nconf, _ip(low),
                                      _ip(nrot), sym, _dp(fac), _dp(ef)))
    return fac[:nconf], ef[:nconf]


def crossing_dedup(cand, xyzref, rthr, maxgen2, huge_number=False):
    """De-duplication of the genetic crossing (crossing.f90:227-338, single-thread order) on generated
    structures cand (ncand,nat,3) against the reference structure xyzref (nat,3); rthr in the unit of the
    coordinates.  Returns dict(taken (1-based candidate per accepted slot), rms, ntaken, ierr, ident,
    rcount, processed, rmsdsum)."""
    L = _lib.lib()
    cand = np.ascontiguousarray(cand, np.float64)
    ref = np.ascontiguousarray(xyzref, np.float64)
    ncand, nat = cand.shape[0], ref.shape[0]
    taken = np.zeros(maxgen2, np.int32)
    rms = np.zeros(maxgen2)
    nt = C.c_int(0)
    cnt = np.zeros(4, np.int64)
    rsum = C.c_double(0)
    _lib.check(L.crest_cuda_crossing_dedup(nat, ncand, cand.ctypes.data_as(_lib.c_dp), ref.ctypes.data_as(_lib.c_dp),
                                           float(rthr), int(maxgen2), 1 if huge_number else 0,
                                           taken.ctypes.data_as(_lib.c_ip), rms.ctypes.data_as(_lib.c_dp),
                                           C.byref(nt), cnt.ctypes.data_as(_lib.c_llp), C.byref(rsum)))
    return dict(taken=taken, rms=rms, ntaken=nt.value, ierr=int(cnt[0]), ident=int(cnt[1]), rcount=int(cnt[2]),
                processed=int(cnt[3]), rmsdsum=rsum.value)
