"""Host-side mirror of the ensemble reader / writer of /root/reference/src/strucreader.f90
(rdensembleparam 268-309, rdensemble 433-481, wrxyz 1679-1698, grepenergy 2225-2260) over the C ABI.
Host code in the library as well: text parsing does not go to the GPU."""
import ctypes as C
import os

import numpy as np

from . import _lib
from .cregen import _dp, _ip

COMMENT_LEN = 128          # the reference carries comments as character(len=128) (cregen.f90:2120)


def rdensembleparam(fname):
    """Returns (nat, nall); (0, 0) for a missing file like the reference."""
    nat, nall = C.c_int(0), C.c_int(0)
    _lib.check(_lib.lib().crest_cuda_rdensemble_param(os.fsencode(fname), C.byref(nat), C.byref(nall)))
    return nat.value, nall.value


def rdensemble(fname, want_comments=True):
    """Returns dict(at (nat,), xyz (nall,nat,3), er (nall,), comments [str])."""
    nat, nall = rdensembleparam(fname)
    if nat == 0 or nall == 0:
        raise _lib.CrestCudaError(f"{fname}: no readable structure")
    at = np.zeros(nat, np.int32)
    xyz = np.zeros((nall, nat, 3))
    er = np.zeros(nall)
    buf = C.create_string_buffer(COMMENT_LEN * nall) if want_comments else None
    nread = C.c_int(0)
    _lib.check(_lib.lib().crest_cuda_rdensemble(os.fsencode(fname), nat, nall, _ip(at), _dp(xyz), _dp(er),
                                                buf, COMMENT_LEN, C.byref(nread)))
    n = nread.value
    comments = None
    if want_comments:
        raw = buf.raw
        comments = [raw[COMMENT_LEN * i:COMMENT_LEN * (i + 1)].decode(errors="replace").rstrip()
                    for i in range(n)]
    return dict(at=at, xyz=xyz[:n], er=er[:n], comments=comments)


def wrensemble(fname, at, xyz, er=None, comments=None):
    at = np.ascontiguousarray(at, np.int32)
    xyz = np.ascontiguousarray(xyz, np.float64)
    nall, nat, _ = xyz.shape
    buf = None
    if comments is not None:
        buf = b"".join(c.encode()[:COMMENT_LEN].ljust(COMMENT_LEN) for c in comments)
    erp = None if er is None else _dp(np.ascontiguousarray(er, np.float64))
    _lib.check(_lib.lib().crest_cuda_wrensemble(os.fsencode(fname), nat, nall, _ip(at), _dp(xyz), erp,
                                                buf, COMMENT_LEN))


def grepenergy(line):
    return _lib.lib().crest_cuda_grepenergy(line.encode())


def e2i(label):
    return _lib.lib().crest_cuda_e2i(label.encode())
