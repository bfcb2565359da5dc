"""Golden vectors for the SURVEY.md §8 f2 / f3 rows (tests/golden/f2_compare.npz, f3_intraconf.npz).

Generated from the oracle (oracle/oracle_capi.cpp: oracle_compare_blocks, oracle_compare_correlate,
oracle_intraconf = restatements of /root/reference/src/ensemblecomp.f90:213-300 and
src/entropy/entropic.f90:938-1116); the reference has no fixture for these routines and cannot be
built here (Fortran), so the oracle is cross-checked against numpy in tests/test_compare.py and
frozen here.  Run:  python tests/make_golden_f.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))
import oracle_lib as O  # noqa: E402
from test_compare import rotamer_ensembles  # noqa: E402

GOLD = os.path.join(HERE, "golden")


def main():
    O.build()
    at, x1, b1, e1, x2, b2, e2 = rotamer_ensembles(90, 24, 11, seed=41)
    rmat = O.compare_blocks(x1, b1, e1, x2, b2, e2)
    en1 = -10.0 + 1e-3 * np.arange(len(b1))
    en2 = -10.0 + 1e-3 * np.arange(len(b2)) + 4e-4
    left, right = O.compare_correlate(en1, en2, rmat, 0.125)
    np.savez_compressed(os.path.join(GOLD, "f2_compare.npz"), at=at, xyz1=x1, b1=b1, e1=e1, xyz2=x2,
                        b2=b2, e2=e2, rmat=rmat, en1=en1, en2=en2, left=left, right=right)
    print("f2: rmat", rmat.shape, "matches", int(((left > 0) & (right > 0)).sum()))

    at, x, b, e, *_ = rotamer_ensembles(160, 30, 14, seed=43, copies=40, frac_mirror=0.3,
                                        frac_straddle=0.2)
    nrot = (e - b + 1).astype(np.int32)
    inc = (at != 1).astype(np.int32)
    sym = ["c1" if c % 3 else "cs" for c in range(len(b))]
    fac, ef = O.intraconf(x, inc, b, nrot, sym)
    np.savez_compressed(os.path.join(GOLD, "f3_intraconf.npz"), at=at, xyz=x, inc=inc, low=b,
                        nrot=nrot, sym=np.array(sym), fac=fac, enantiofac=ef)
    print("f3: nrot", nrot.tolist(), "fac", fac.tolist(), "enantiofac", ef.tolist())


if __name__ == "__main__":
    main()
