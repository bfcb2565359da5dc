"""ctypes binding of libcrest_cuda.so (include/crest_cuda.h).  There is no CPU fallback: if the
library has not been built, or no CUDA device is present, every compute call raises."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("CREST_CUDA_LIB", os.path.join(_HERE, "libcrest_cuda.so"))
_LIB = None

c_dp = C.POINTER(C.c_double)
c_ip = C.POINTER(C.c_int)
c_fp = C.POINTER(C.c_float)
c_llp = C.POINTER(C.c_longlong)
c_ullp = C.POINTER(C.c_ulonglong)


class Thresholds(C.Structure):
    """crest_cuda_thresholds; defaults = reference src/confparse.f90:162-170."""
    _fields_ = [("ethr", C.c_double), ("rthr", C.c_double), ("bthr", C.c_double),
                ("bthrmax", C.c_double), ("bthrshift", C.c_double)]

    def __init__(self, ethr=0.05, rthr=0.125, bthr=0.01, bthrmax=0.025, bthrshift=0.5):
        super().__init__(ethr, rthr, bthr, bthrmax, bthrshift)


class CrestCudaError(RuntimeError):
    pass


# every symbol include/crest_cuda.h declares: (restype, argtypes)
_THR = C.POINTER(Thresholds)
SYMBOLS = {
    "crest_cuda_init": (C.c_int, [C.c_int, c_ip]),
    "crest_cuda_finalize": (None, []),
    "crest_cuda_ngpu": (C.c_int, []),
    "crest_cuda_cregen_dup": (C.c_int, [C.c_int, C.c_int, c_ip, c_dp, c_dp, c_ip, C.c_int, C.c_int,
                                        _THR, c_dp, c_dp, c_ip, c_llp]),
    "crest_cuda_cregen_group": (C.c_int, [C.c_int, C.c_int, c_ip, c_dp, c_dp, c_dp, _THR, c_ip, c_dp]),
    "crest_cuda_cregen_group_noalloc": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, _THR, C.c_void_p,
                                                  C.c_void_p]),
    "crest_cuda_noalloc_max_kept": (C.c_int, []),
    "crest_cuda_cregen_cre": (C.c_int, [C.c_int, C.c_int, c_ip, c_dp, c_dp, c_ip, C.c_int, C.c_int,
                                        C.c_int, _THR, c_dp, c_ip, c_ip, c_ip, c_llp]),
    "crest_cuda_rmsd_pairs": (C.c_int, [C.c_int, C.c_int, c_dp, C.c_longlong, c_ip, c_ip, c_dp, c_dp]),
    "crest_cuda_rotconst": (C.c_int, [C.c_int, C.c_int, c_ip, c_dp, c_dp]),
    "crest_cuda_cregen_esort": (C.c_int, [C.c_int, C.c_int, c_dp, c_dp, C.c_double, c_ip, c_ip]),
    "crest_cuda_rmsdalign": (C.c_int, [C.c_int, C.c_int, c_ip, c_dp]),
    "crest_cuda_rmat_allpairs": (C.c_int, [C.c_int, C.c_int, c_ip, c_dp, c_ip, C.c_int, c_fp]),
    "crest_cuda_rdensemble_param": (C.c_int, [C.c_char_p, c_ip, c_ip]),
    "crest_cuda_rdensemble": (C.c_int, [C.c_char_p, C.c_int, C.c_int, c_ip, c_dp, c_dp, C.c_char_p,
                                        C.c_int, c_ip]),
    "crest_cuda_wrensemble": (C.c_int, [C.c_char_p, C.c_int, C.c_int, c_ip, c_dp, c_dp, C.c_char_p,
                                        C.c_int]),
    "crest_cuda_grepenergy": (C.c_double, [C.c_char_p]),
    "crest_cuda_e2i": (C.c_int, [C.c_char_p]),
    "crest_cuda_compare_blocks": (C.c_int, [C.c_int, C.c_int, c_dp, C.c_int, c_ip, c_ip, C.c_int, c_dp,
                                            C.c_int, c_ip, c_ip, c_dp]),
    "crest_cuda_compare_correlate": (C.c_int, [C.c_int, C.c_int, c_dp, c_dp, c_dp, C.c_double, c_ip, c_ip]),
    "crest_cuda_intraconf": (C.c_int, [C.c_int, C.c_int, c_dp, c_ip, C.c_int, c_ip, c_ip, C.c_char_p,
                                       c_dp, c_dp]),
    "crest_cuda_crossing_dedup": (C.c_int, [C.c_int, C.c_int, c_dp, c_dp, C.c_double, C.c_int, C.c_int, c_ip, c_dp,
                                            c_ip, c_llp, c_dp]),
    "crest_cuda_ctx_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_int]),
    "crest_cuda_ctx_destroy": (None, [C.c_void_p]),
    "crest_cuda_ctx_upload": (C.c_int, [C.c_void_p, C.c_int, C.c_int, c_ip, c_dp, c_dp, c_ip, C.c_int]),
    "crest_cuda_ctx_upload_xyz_rows": (C.c_int, [C.c_void_p, C.c_int, C.c_int, c_dp]),
    "crest_cuda_ctx_xyz_devptr": (C.c_void_p, [C.c_void_p]),
    "crest_cuda_ctx_pack": (C.c_int, [C.c_void_p, C.c_void_p]),
    "crest_cuda_ctx_pairs": (C.c_int, [C.c_void_p, C.c_int, _THR, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "crest_cuda_ctx_dup_devptr": (C.c_void_p, [C.c_void_p]),
    "crest_cuda_ctx_npairs_devptr": (C.c_void_p, [C.c_void_p]),
    "crest_cuda_ctx_sweep": (C.c_int, [C.c_void_p, C.c_void_p]),
    "crest_cuda_ctx_download": (C.c_int, [C.c_void_p, c_dp, c_dp, c_ip, c_llp, C.c_void_p]),
    "crest_cuda_ctx_info": (C.c_int, [C.c_void_p, _THR, C.c_int, C.c_int, c_llp]),
    "crest_cuda_ctx_tf32_accum": (C.c_int, [C.c_void_p, c_dp, c_llp, c_dp, c_dp]),
    "crest_cuda_ctx_download_mask": (C.c_int, [C.c_void_p, c_ullp, C.c_void_p]),
    "crest_cuda_e_threshold": (C.c_double, [C.c_double, C.c_int]),
    "crest_cuda_tile_shape": (None, [c_ip, c_ip]),
    "crest_cuda_band_tiles": (C.c_int, [C.c_int, c_dp, C.c_double, c_llp, c_ip, c_llp]),
    "crest_cuda_tf32_items": (C.c_int, [C.c_int, c_dp, C.c_double, c_ip, C.c_longlong, c_llp]),
    "crest_cuda_tf32_items_rank": (C.c_int, [C.c_int, c_dp, C.c_double, C.c_int, C.c_int, c_ip, C.c_longlong, c_llp]),
    "crest_cuda_tf32_items_n": (C.c_int, [C.c_int, c_dp, C.c_double, C.c_int, c_ip, C.c_longlong, c_llp]),
    "crest_cuda_shard_range": (C.c_int, [C.c_longlong, C.c_int, C.c_int, c_llp, c_llp]),
    "crest_cuda_set_tune": (None, [C.c_char_p]),
    "crest_cuda_last_error": (C.c_char_p, []),
    "crest_cuda_device_count": (C.c_int, []),
}


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise CrestCudaError(
                f"{LIB_PATH} not built (run `python -c 'import __graft_entry__ as g; g.build()'` "
                "or `make -C crest_b200/csrc`); there is no CPU fallback")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            f = getattr(L, name)
            f.restype = res
            f.argtypes = args
        _LIB = L
    return _LIB


def check(rc):
    if rc != 0:
        raise CrestCudaError(lib().crest_cuda_last_error().decode(errors="replace"))
