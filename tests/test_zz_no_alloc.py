"""crest_cuda_cregen_group_noalloc: K6 on device arrays of the caller with the kernel's scratch in a static
device array (no cudaMalloc anywhere in the call).  It has no counterpart in the reference; it exists for
callers that cannot allocate (CUDA graph capture, serving loops).  Checked through the C ABI against the
existing entry point crest_cuda_cregen_group (group ids are integers: identical), on two calls in a row
that share the scratch array - the second with a different, smaller kept set, so stale contents of the
first call would show.  Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

from crest_b200 import _lib, synth

pytestmark = pytest.mark.gpu


def _group_existing(L, thr, nat, at, xyz, er, rot):
    n = len(er)
    at = np.ascontiguousarray(at, np.int32)
    xyz = np.ascontiguousarray(xyz, np.float64)
    er = np.ascontiguousarray(er, np.float64)
    rot = np.ascontiguousarray(rot, np.float64)
    group = np.zeros(n + 1, np.int32)
    enuc = np.zeros(n, np.float64)
    _lib.check(L.crest_cuda_cregen_group(nat, n, at.ctypes.data_as(_lib.c_ip), xyz.ctypes.data_as(_lib.c_dp),
                                         er.ctypes.data_as(_lib.c_dp), rot.ctypes.data_as(_lib.c_dp),
                                         C.byref(thr), group.ctypes.data_as(_lib.c_ip),
                                         enuc.ctypes.data_as(_lib.c_dp)))
    return group, enuc


def _group_noalloc(L, thr, er, rot, enuc, stream=None):
    import torch
    dev = torch.device("cuda:0")
    n = len(er)
    d_er = torch.from_numpy(np.ascontiguousarray(er, np.float64)).to(dev)
    d_rot = torch.from_numpy(np.ascontiguousarray(rot, np.float64)).to(dev)
    d_enuc = torch.from_numpy(np.ascontiguousarray(enuc, np.float64)).to(dev)
    d_group = torch.full((n + 1,), -7, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    before = torch.cuda.memory_stats(dev)["num_device_alloc"] if hasattr(torch.cuda, "memory_stats") else None
    s = torch.cuda.current_stream(dev) if stream is None else stream
    _lib.check(L.crest_cuda_cregen_group_noalloc(n, d_er.data_ptr(), d_rot.data_ptr(), d_enuc.data_ptr(),
                                                 C.byref(thr), d_group.data_ptr(), s.cuda_stream))
    s.synchronize()
    if before is not None:      # the call went through no torch allocation either
        assert torch.cuda.memory_stats(dev)["num_device_alloc"] == before
    return d_group.cpu().numpy()


def _kept_set(oracle, nall, seed, frac_rotamer, ethr):
    ens = synth.make_ensemble(nall, 50, 23, seed=seed, frac_rotamer=frac_rotamer)
    ref = oracle.cregen_cre(ens["at"], ens["xyz"], ens["er"], ethr=ethr)
    n = ref["nallout"]
    return ens["at"], ref["xyz"][:n], ref["er"][:n], ref["rot"][:n], ref["group"][:n + 1]


def test_group_noalloc_matches_existing_entry_point(cuda_lib, oracle):
    import torch
    assert torch.cuda.is_available()
    torch.cuda.set_device(0)
    L = cuda_lib
    assert L.crest_cuda_noalloc_max_kept() == 32768
    # two calls in a row on the shared scratch: a larger kept set with many rotamers, then a smaller one
    for nall, seed, frac, ethr in ((2000, 17, 0.6, 1e9), (700, 5, 0.05, 0.05)):
        at, xyz, er, rot, gref = _kept_set(oracle, nall, seed, frac, ethr)
        thr = _lib.Thresholds(ethr=ethr)
        g_old, enuc = _group_existing(L, thr, 50, at, xyz, er, rot)
        g_new = _group_noalloc(L, thr, er, rot, enuc)
        assert np.array_equal(g_new, g_old)            # integers: identical, no tolerance
        assert np.array_equal(g_new, gref)             # and both equal the oracle's K6
        assert g_new[0] < len(er)                      # some group holds more than one structure
    # same inputs again, on a side stream: the array keeps the last call's contents and that must not matter
    g_again = _group_noalloc(L, thr, er, rot, enuc, stream=torch.cuda.Stream())
    assert np.array_equal(g_again, g_old)


def test_group_noalloc_limits(cuda_lib):
    import torch
    torch.cuda.set_device(0)
    L = cuda_lib
    thr = _lib.Thresholds()
    d = torch.zeros(8, dtype=torch.float64, device="cuda:0")
    g = torch.zeros(8, dtype=torch.int32, device="cuda:0")
    nmax = L.crest_cuda_noalloc_max_kept()
    rc = L.crest_cuda_cregen_group_noalloc(nmax + 1, d.data_ptr(), d.data_ptr(), d.data_ptr(), C.byref(thr),
                                           g.data_ptr(), None)
    assert rc != 0 and b"exceeds the static scratch" in L.crest_cuda_last_error()
    rc = L.crest_cuda_cregen_group_noalloc(4, None, d.data_ptr(), d.data_ptr(), C.byref(thr), g.data_ptr(), None)
    assert rc != 0
    # nkept = 0 writes group[0] = 0 and nothing else
    g.fill_(5)
    _lib.check(L.crest_cuda_cregen_group_noalloc(0, None, None, None, C.byref(thr), g.data_ptr(), None))
    torch.cuda.synchronize()
    assert g.cpu().tolist() == [0, 5, 5, 5, 5, 5, 5, 5]
