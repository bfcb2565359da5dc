"""One-off: the FULL BASELINE config C3 (100 000 x 60, all pairs, enantiomer check) through the CPU
oracle (oracle/, the C++ restatement of cregen.f90:1125-1506) -> tests/golden/c3_full.npz.
About 20 min on 16 cores / 40 min on 8 (5.0e9 pairs, two rmsd() per pair) and 20 GB for the float32
band.  The GPU test compares whole arrays (tests/test_gpu_parity.py::test_full_size_golden_c3).
Usage: python tests/make_golden_c3_full.py [nall]   (nall < 100000 only for a dry run)"""
import hashlib
import os
import sys
import time
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)
import oracle_lib as O  # noqa: E402
from crest_b200 import synth  # noqa: E402

nall = int(sys.argv[1]) if len(sys.argv) > 1 else None
ens = synth.make_config("C3", nall=nall)
at, xyz, er = ens["at"], ens["xyz"], ens["er"]
digest = hashlib.sha256(xyz.tobytes() + er.tobytes() + at.tobytes()).hexdigest()
t0 = time.time()
ref = O.cregen_cre(at, xyz, er, ethr=1e9, enantio=True, boundary_tol=1e-6)
dt = time.time() - t0
assert ref["nfail"] == 0, ref["nfail"]
out = os.path.join(HERE, "golden", "c3_full.npz" if nall is None else f"c3_full_{nall}.npz")
np.savez_compressed(out, input_sha256=np.array(digest), nall=xyz.shape[0], nat=xyz.shape[1],
                    dup=ref["dup"].astype(np.int32), orderref=ref["orderref"].astype(np.int32),
                    group=ref["group"].astype(np.int32), nallout=ref["nallout"], klong=ref["klong"],
                    rot=ref["rot"], boundary_rows=ref["boundary_rows"].astype(np.int32),
                    oracle_seconds=dt, oracle_threads=O.num_threads())
print(f"{out}: nall {xyz.shape[0]} klong {ref['klong']} kept {ref['nkept']} groups {int(ref['group'][0])} "
      f"boundary rows {len(ref['boundary_rows'])} in {dt:.0f} s on {O.num_threads()} threads")
