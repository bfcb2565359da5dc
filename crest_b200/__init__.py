"""crest_b200 — B200-native (sm_100a) CREGEN duplicate / rotamer sorting hot path of crest-lab/crest.

Product code: crest_b200/csrc (CUDA kernels + C ABI, include/crest_cuda.h) and the thin host-side
mirror of the reference interface in crest_b200.cregen.  No CPU fallback exists."""
from .cregen import (Env, CregenContext, cregen_CRE, cregen_dup, cregen_group, rmsd_pairs,  # noqa: F401
                     rotconst, rmat_allpairs, cregen_esort, cregen_esort_inplace, rmsdalign, cregen_CRE_inplace,
                     CrestCudaError, init, finalize, ngpu)
from ._lib import Thresholds  # noqa: F401
from .compare import compare_blocks, compare_correlate, intraconf, crossing_dedup  # noqa: F401
from .ensemble_io import rdensembleparam, rdensemble, wrensemble, grepenergy, e2i  # noqa: F401
