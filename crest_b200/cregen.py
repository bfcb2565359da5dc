"""Host-side mirror of the reference's CREGEN sorting interface, on top of the C ABI.

    cregen_CRE(env, nat, nall, at, xyz, energies, nosort)  <->  src/cregen.f90:1125-1506
    CregenContext                                           <->  device-resident path used by
                                                                 bench.py / multi-GPU sharding

`env` carries the same scalars the reference reads from type(systemdata)
(src/classes.f90:332-345, 426, 557, 607, 547): ethr, rthr, bthr2, bthrmax, bthrshift, heavyrmsd,
subRMSD + includeRMSD, entropic.  Indices returned are 1-based like the reference's arrays.
Nothing here computes on the CPU: every number comes from libcrest_cuda.so.
"""
import ctypes as C
from dataclasses import dataclass, field
from typing import Optional

import numpy as np

from . import _lib
from ._lib import Thresholds, CrestCudaError, c_dp, c_ip, c_fp, c_llp  # noqa: F401


@dataclass
class Env:
    """The slice of type(systemdata) that crosses the boundary; defaults = confparse.f90:162-174."""
    ethr: float = 0.05
    rthr: float = 0.125
    bthr2: float = 0.01
    bthrmax: float = 0.025
    bthrshift: float = 0.5
    heavyrmsd: bool = False
    subRMSD: bool = False
    includeRMSD: Optional[np.ndarray] = None   # (nat,) >0 = atom used; with subRMSD
    rednat: Optional[int] = None
    entropic: bool = False

    def thresholds(self):
        return Thresholds(self.ethr, self.rthr, self.bthr2, self.bthrmax, self.bthrshift)


def _dp(a):
    return a.ctypes.data_as(c_dp) if a is not None else None


def _ip(a):
    return a.ctypes.data_as(c_ip) if a is not None else None


def rmsd_mask(env: Env, nat: int, at: np.ndarray):
    """(mask or None, rot_from_mask) as cregen_CRE derives them (cregen.f90:1199-1202,1293-1299)."""
    rednat = env.rednat if env.rednat is not None else nat
    substruc = (nat != rednat) and env.subRMSD
    if substruc:
        m = (np.asarray(env.includeRMSD) > 0).astype(np.int32)
        return np.ascontiguousarray(m), 1
    if env.heavyrmsd:
        return np.ascontiguousarray((np.asarray(at) != 1).astype(np.int32)), 0
    return None, 0


def _tile_shape():
    ti, tj = C.c_int(0), C.c_int(0)
    _lib.lib().crest_cuda_tile_shape(C.byref(ti), C.byref(tj))
    return ti.value, tj.value


def __getattr__(name):            # module-level lazy constants
    if name == "TILE_I":
        return _tile_shape()[0]
    if name == "TILE_J":
        return _tile_shape()[1]
    raise AttributeError(name)


def init(ngpu=0):
    """crest_cuda_init: the engine of 1-8 GPUs behind the one-call entry points (cregen_CRE, cregen_dup,
    ...), driven by this one host thread like CREST's serial region.  ngpu <= 0: every visible GPU,
    1: the current device.  Returns the number of GPUs in use."""
    used = C.c_int(0)
    _lib.check(_lib.lib().crest_cuda_init(int(ngpu), C.byref(used)))
    return used.value


def finalize():
    """crest_cuda_finalize: release the engine (device buffers, streams, pinned staging)."""
    _lib.lib().crest_cuda_finalize()


def ngpu():
    return _lib.lib().crest_cuda_ngpu()


def set_tune(spec=None):
    """Developer knobs of the pair kernels (crest_cuda_set_tune); None restores the defaults."""
    _lib.lib().crest_cuda_set_tune(spec.encode() if spec else None)


def band_tiles(er, ethr):
    """Host-only: the in-band tile list (i-tile, j-tile arrays, rowstart, jfirst) for ascending
    energies er (Eh) and gate ethr (kcal/mol) — reference band cregen.f90:1240-1256 in tiles."""
    L = _lib.lib()
    er = np.ascontiguousarray(er, np.float64)
    n = len(er)
    ti_, _ = _tile_shape()
    nI = (n + ti_ - 1) // ti_
    rowstart = np.zeros(nI + 1, np.int64)
    jfirst = np.zeros(max(nI, 1), np.int32)
    nt = C.c_longlong(0)
    _lib.check(L.crest_cuda_band_tiles(n, _dp(er), float(ethr), rowstart.ctypes.data_as(c_llp),
                                       _ip(jfirst), C.byref(nt)))
    ti, tj = [], []
    for I in range(nI):
        cnt = int(rowstart[I + 1] - rowstart[I])
        ti += [I] * cnt
        tj += list(range(int(jfirst[I]), int(jfirst[I]) + cnt))
    assert len(ti) == nt.value
    return np.array(ti, np.int64), np.array(tj, np.int64), rowstart, jfirst[:nI]


def tf32_items(er, ethr, nranks=1):
    """Host-only: work list of the tensor-core screen kernel, (n,4) int32 rows
    {I, g_begin, g_end, g_gate_end} (see include/crest_cuda.h), as `nranks` ranks build it."""
    L = _lib.lib()
    er = np.ascontiguousarray(er, np.float64)
    n = C.c_longlong(0)
    _lib.check(L.crest_cuda_tf32_items_n(len(er), _dp(er), float(ethr), nranks, None, 0, C.byref(n)))
    items = np.zeros((max(n.value, 1), 4), np.int32)
    _lib.check(L.crest_cuda_tf32_items_n(len(er), _dp(er), float(ethr), nranks, _ip(items), n.value,
                                         C.byref(n)))
    return items[: n.value]


def tf32_items_rank(er, ethr, nranks, rank):
    """Host-only: the screen's work list of ONE rank (its rows: i-tiles I with I % nranks == rank)."""
    L = _lib.lib()
    er = np.ascontiguousarray(er, np.float64)
    n = C.c_longlong(0)
    _lib.check(L.crest_cuda_tf32_items_rank(len(er), _dp(er), float(ethr), nranks, rank, None, 0, C.byref(n)))
    items = np.zeros((max(n.value, 1), 4), np.int32)
    _lib.check(L.crest_cuda_tf32_items_rank(len(er), _dp(er), float(ethr), nranks, rank, _ip(items), n.value,
                                            C.byref(n)))
    return items[: n.value]


def shard_range(ntiles, rank, nranks):
    """Host-only: [begin, end) of the tile list owned by `rank` of `nranks`."""
    b, e = C.c_longlong(0), C.c_longlong(0)
    _lib.check(_lib.lib().crest_cuda_shard_range(ntiles, rank, nranks, C.byref(b), C.byref(e)))
    return b.value, e.value


@dataclass
class CreResult:
    nallout: int
    group: np.ndarray          # (nall+1,), group[0] = number of groups
    order: np.ndarray          # (nall,) new 1-based position of input structure i
    rot: np.ndarray            # (nall,3) MHz, permuted
    xyz: np.ndarray            # (nall,nat,3) principal-axis frame, permuted
    er: np.ndarray             # permuted energies
    npairs: int


def cregen_CRE(env: Env, nat, nall, at, xyz, er, nosort=False) -> CreResult:
    """Whole cregen_CRE (duplicates + rotamer groups) through crest_cuda_cregen_cre; the inputs are
    copied first (the reference works in place, see cregen_CRE_inplace)."""
    xyz = np.array(xyz, np.float64, order="C", copy=True).reshape(nall, nat, 3)
    er = np.array(er, np.float64, copy=True)
    return cregen_CRE_inplace(env, nat, nall, at, xyz, er, nosort)


def cregen_CRE_inplace(env: Env, nat, nall, at, xyz, er, nosort=False) -> CreResult:
    """As the reference's cregen_CRE (cregen.f90:1125): xyz (nall,nat,3) float64 C-order and er are
    overwritten with the principal-axis-frame geometries / energies in the new order."""
    L = _lib.lib()
    at = np.ascontiguousarray(at, np.int32)
    if not (isinstance(xyz, np.ndarray) and xyz.dtype == np.float64 and xyz.flags.c_contiguous and
            isinstance(er, np.ndarray) and er.dtype == np.float64 and er.flags.c_contiguous):
        raise TypeError("cregen_CRE_inplace needs C-contiguous float64 xyz and er")
    mask, rfm = rmsd_mask(env, nat, at)
    rot = np.zeros((max(nall, 1), 3))
    order = np.zeros(max(nall, 1), np.int32)
    group = np.zeros(nall + 1, np.int32)
    nout = C.c_int(0)
    npairs = C.c_longlong(0)
    thr = env.thresholds()
    _lib.check(L.crest_cuda_cregen_cre(nat, nall, _ip(at), _dp(xyz), _dp(er), _ip(mask), rfm,
                                       0 if env.entropic else 1, 1 if nosort else 0, C.byref(thr),
                                       _dp(rot), _ip(order), _ip(group), C.byref(nout),
                                       C.byref(npairs)))
    return CreResult(nout.value, group, order[:nall], rot[:nall], xyz.reshape(nall, nat, 3), er,
                     npairs.value)


def cregen_dup(env: Env, nat, nall, at, xyz, er, want_xyz=True):
    """Stage A only (K0-K4): returns dict(rot, xyz_axis, dup, npairs)."""
    L = _lib.lib()
    at = np.ascontiguousarray(at, np.int32)
    xyz = np.ascontiguousarray(xyz, np.float64).reshape(nall, nat, 3)
    er = np.ascontiguousarray(er, np.float64)
    mask, rfm = rmsd_mask(env, nat, at)
    rot = np.zeros((max(nall, 1), 3))
    xa = np.zeros_like(xyz) if want_xyz else None
    dup = np.zeros(max(nall, 1), np.int32)
    npairs = C.c_longlong(0)
    thr = env.thresholds()
    _lib.check(L.crest_cuda_cregen_dup(nat, nall, _ip(at), _dp(xyz), _dp(er), _ip(mask), rfm,
                                       0 if env.entropic else 1, C.byref(thr), _dp(rot), _dp(xa),
                                       _ip(dup), C.byref(npairs)))
    return dict(rot=rot[:nall], xyz_axis=xa, dup=dup[:nall], npairs=npairs.value)


def cregen_group(env: Env, nat, nkept, at, xyz, er, rot):
    L = _lib.lib()
    at = np.ascontiguousarray(at, np.int32)
    xyz = np.ascontiguousarray(xyz, np.float64)
    er = np.ascontiguousarray(er, np.float64)
    rot = np.ascontiguousarray(rot, np.float64)
    group = np.zeros(nkept + 1, np.int32)
    enuc = np.zeros(max(nkept, 1))
    thr = env.thresholds()
    _lib.check(L.crest_cuda_cregen_group(nat, nkept, _ip(at), _dp(xyz), _dp(er), _dp(rot),
                                         C.byref(thr), _ip(group), _dp(enuc)))
    return group, enuc[:nkept]


def rmsd_pairs(packed, i_idx, j_idx):
    """packed (n,A,3); returns (rmsd, rmsd_mirror) for the listed 0-based pairs."""
    L = _lib.lib()
    packed = np.ascontiguousarray(packed, np.float64)
    n, A, _ = packed.shape
    i_idx = np.ascontiguousarray(i_idx, np.int32)
    j_idx = np.ascontiguousarray(j_idx, np.int32)
    r = np.zeros(len(i_idx))
    rm = np.zeros(len(i_idx))
    _lib.check(L.crest_cuda_rmsd_pairs(A, n, _dp(packed), len(i_idx), _ip(i_idx), _ip(j_idx),
                                       _dp(r), _dp(rm)))
    return r, rm


def rotconst(at, xyz):
    L = _lib.lib()
    at = np.ascontiguousarray(at, np.int32)
    xyz = np.ascontiguousarray(xyz, np.float64)
    n, nat, _ = xyz.shape
    rot = np.zeros((n, 3))
    _lib.check(L.crest_cuda_rotconst(nat, n, _ip(at), _dp(xyz), _dp(rot)))
    return rot


def cregen_esort(xyz, energies, ewin=1e9):
    """cregen_esort (cregen.f90:1038-1121): returns (xyz_sorted, er_sorted, order, nallout);
    order[i] = new 1-based position of input structure i.  xyz may be None."""
    L = _lib.lib()
    er = np.array(energies, np.float64, copy=True)
    nall = len(er)
    xs, nat = None, 1
    if xyz is not None:
        xs = np.array(xyz, np.float64, order="C", copy=True)
        nat = xs.shape[1]
    order = np.zeros(max(nall, 1), np.int32)
    nout = C.c_int(0)
    _lib.check(L.crest_cuda_cregen_esort(nat, nall, _dp(xs), _dp(er), float(ewin), _ip(order),
                                         C.byref(nout)))
    return xs, er, order[:nall], nout.value


def cregen_esort_inplace(xyz, energies, ewin=1e9):
    """As the reference (cregen.f90:1038-1121 sorts its arrays in place): xyz (nall,nat,3) float64 C-order and
    energies float64 are permuted in place; returns (order, nallout)."""
    L = _lib.lib()
    if not (isinstance(energies, np.ndarray) and energies.dtype == np.float64 and energies.flags.c_contiguous):
        raise TypeError("cregen_esort_inplace needs a C-contiguous float64 energy array")
    nall = len(energies)
    nat = 1
    if xyz is not None:
        if not (isinstance(xyz, np.ndarray) and xyz.dtype == np.float64 and xyz.flags.c_contiguous):
            raise TypeError("cregen_esort_inplace needs C-contiguous float64 xyz")
        nat = xyz.shape[1]
    order = np.zeros(max(nall, 1), np.int32)
    nout = C.c_int(0)
    _lib.check(L.crest_cuda_cregen_esort(nat, nall, _dp(xyz) if xyz is not None else None, _dp(energies),
                                         float(ewin), _ip(order), C.byref(nout)))
    return order[:nall], nout.value


def rmsdalign(at, xyz):
    """cregen_rmsdalign (cregen.f90:2229-2283) on the GPU; returns the aligned copy of xyz."""
    L = _lib.lib()
    at = np.ascontiguousarray(at, np.int32)
    out = np.array(xyz, np.float64, order="C", copy=True)
    nall, nat, _ = out.shape
    _lib.check(L.crest_cuda_rmsdalign(nat, nall, _ip(at), _dp(out)))
    return out


def rmat_allpairs(env: Env, at, xyz):
    """float32 band as the reference's rmat with ethr=inf: row i at offset i*(i-1)/2."""
    L = _lib.lib()
    at = np.ascontiguousarray(at, np.int32)
    xyz = np.ascontiguousarray(xyz, np.float64)
    n, nat, _ = xyz.shape
    mask, _ = rmsd_mask(env, nat, at)
    rmat = np.zeros(max(n * (n - 1) // 2, 1), np.float32)
    _lib.check(L.crest_cuda_rmat_allpairs(nat, n, _ip(at), _dp(xyz), _ip(mask),
                                          0 if env.entropic else 1, rmat.ctypes.data_as(c_fp)))
    return rmat[: n * (n - 1) // 2]


class CregenContext:
    """Device-resident ensemble; stages are launched on a caller-supplied CUDA stream so that
    bench.py can time them with events and shard the tile space over ranks."""

    def __init__(self, device=-1):
        self._L = _lib.lib()
        self._h = C.c_void_p()
        _lib.check(self._L.crest_cuda_ctx_create(C.byref(self._h), device))
        self.nall = 0
        self.nat = 0

    def close(self):
        if self._h:
            self._L.crest_cuda_ctx_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def upload(self, env: Env, at, xyz, er):
        at = np.ascontiguousarray(at, np.int32)
        xyz = np.ascontiguousarray(xyz, np.float64)
        er = np.ascontiguousarray(er, np.float64)
        nall, nat, _ = xyz.shape
        mask, rfm = rmsd_mask(env, nat, at)
        _lib.check(self._L.crest_cuda_ctx_upload(self._h, nat, nall, _ip(at), _dp(xyz), _dp(er),
                                                 _ip(mask), rfm))
        self.nall, self.nat = nall, nat

    def upload_rows(self, env: Env, at, xyz, er, rank, world):
        """Multi-GPU upload: reserve the ensemble, copy only this rank's slice of xyz host->device.
        Returns (chunk, first, count); the caller all-gathers `chunk` rows per rank into
        xyz_devptr() (crest_b200/multi.py)."""
        at = np.ascontiguousarray(at, np.int32)
        xyz = np.ascontiguousarray(xyz, np.float64)
        er = np.ascontiguousarray(er, np.float64)
        nall, nat, _ = xyz.shape
        mask, rfm = rmsd_mask(env, nat, at)
        _lib.check(self._L.crest_cuda_ctx_upload(self._h, nat, nall, _ip(at), None, _dp(er),
                                                 _ip(mask), rfm))
        self.nall, self.nat = nall, nat
        if not 1 <= world <= 64:
            # the device xyz buffer has 64 rows of slack (crest_cuda.h): equal slices of ceil(nall/world) rows
            # overshoot nall by < world rows
            raise ValueError(f"upload_rows: world = {world}, the all-gather layout supports 1..64 ranks")
        chunk = (nall + world - 1) // world
        first = min(rank * chunk, nall)
        count = min(chunk, nall - first)
        rows = xyz[first:first + count]
        _lib.check(self._L.crest_cuda_ctx_upload_xyz_rows(self._h, first, count, _dp(rows)))
        return chunk, first, count

    def xyz_devptr(self):
        return self._L.crest_cuda_ctx_xyz_devptr(self._h)

    def pack(self, stream=0):
        _lib.check(self._L.crest_cuda_ctx_pack(self._h, C.c_void_p(stream)))

    def pairs(self, env: Env, rank=0, nranks=1, want_mask=False, stream=0):
        thr = env.thresholds()
        _lib.check(self._L.crest_cuda_ctx_pairs(self._h, 0 if env.entropic else 1, C.byref(thr),
                                                rank, nranks, 1 if want_mask else 0,
                                                C.c_void_p(stream)))

    def sweep(self, stream=0):
        _lib.check(self._L.crest_cuda_ctx_sweep(self._h, C.c_void_p(stream)))

    def dup_devptr(self):
        return self._L.crest_cuda_ctx_dup_devptr(self._h)

    def npairs_devptr(self):
        return self._L.crest_cuda_ctx_npairs_devptr(self._h)

    def info(self, env: Env, rank=0, nranks=1):
        thr = env.thresholds()
        buf = (C.c_longlong * 16)()
        _lib.check(self._L.crest_cuda_ctx_info(self._h, C.byref(thr), rank, nranks, buf))
        keys = ["tiles_shard", "tiles_total", "A", "A_padded", "grid", "smem", "launches",
                "variant", "candidates", "dense_items", "violations", "exact_evals", "flags",
                "dense_rows_tried", "dense_rows_matched"]
        return dict(zip(keys, list(buf)[:15]))

    def download(self, want_xyz=False, stream=0):
        rot = np.zeros((max(self.nall, 1), 3))
        dup = np.zeros(max(self.nall, 1), np.int32)
        xa = np.zeros((self.nall, self.nat, 3)) if want_xyz else None
        npairs = C.c_longlong(0)
        _lib.check(self._L.crest_cuda_ctx_download(self._h, _dp(rot), _dp(xa), _ip(dup),
                                                   C.byref(npairs), C.c_void_p(stream)))
        return dict(rot=rot[: self.nall], dup=dup[: self.nall], xyz_axis=xa, npairs=npairs.value)

    def tf32_accum(self):
        """Probe of the tensor core's FP32 accumulation on the packed ensemble (crest_cuda_ctx_tf32_accum):
        dict(max_rel_err, ndots, gamma_budget, kappa)."""
        e, g, k = C.c_double(0), C.c_double(0), C.c_double(0)
        n = C.c_longlong(0)
        _lib.check(self._L.crest_cuda_ctx_tf32_accum(self._h, C.byref(e), C.byref(n), C.byref(g), C.byref(k)))
        return dict(max_rel_err=e.value, ndots=n.value, gamma_budget=g.value, kappa=k.value)

    def download_mask(self, stream=0):
        n8 = (self.nall + 7) // 8
        m = np.zeros((n8, n8), np.uint64)
        _lib.check(self._L.crest_cuda_ctx_download_mask(
            self._h, m.ctypes.data_as(_lib.c_ullp), C.c_void_p(stream)))
        return m
