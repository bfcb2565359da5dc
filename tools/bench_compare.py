"""Throughput of the f2 / f3 rows (blockpair.cu) through the C ABI, host buffers in and out, with the
oracle's OpenMP loop on a bounded sample beside it.  Developer measurement, not the bench line."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import crest_b200 as cb
from crest_b200 import synth
import oracle_lib as O

for n, nat, rot in ((8192, 60, 32), (4096, 100, 16), (2048, 300, 8)):
    ens = synth.make_ensemble(2 * n, nat, nat, seed=5)
    x1, x2 = np.ascontiguousarray(ens["xyz"][0::2]), np.ascontiguousarray(ens["xyz"][1::2])
    b = np.arange(1, n + 1, rot, dtype=np.int32); e = np.minimum(b + rot - 1, n).astype(np.int32)
    cb.compare_blocks(x1, b, e, x2, b, e)
    ts = []
    for _ in range(3):
        t0 = time.perf_counter(); rm = cb.compare_blocks(x1, b, e, x2, b, e); ts.append(time.perf_counter() - t0)
    t = min(ts)
    m = 4                                             # oracle on m x m blocks
    t0 = time.perf_counter(); ref = O.compare_blocks(x1, b[:m], e[:m], x2, b[:m], e[:m]); tc = time.perf_counter() - t0
    pc = (m * rot) ** 2
    assert np.abs(rm[:m, :m] - ref).max() < 1e-6
    print(f"compare {n}x{n} rotamers x {nat} atoms, blocks of {rot}: {t*1e3:.2f} ms = {n*n/t:.3g} pairs/s "
          f"({9*nat*2*n*n/t/1e12:.2f} TFLOP/s FP64 FMA); oracle {pc/tc:.3g} pairs/s on {O.lib().oracle_num_threads()} threads")
    low = b; nrot = (e - b + 1).astype(np.int32)
    inc = np.ones(nat, np.int32); sym = ["c1"] * len(low)
    cb.intraconf(x1, inc, low, nrot, sym)
    t0 = time.perf_counter(); cb.intraconf(x1, inc, low, nrot, sym); t = time.perf_counter() - t0
    npair = int((nrot.astype(np.int64) * (nrot - 1) // 2).sum())
    print(f"intraconf {len(low)} conformers x {rot} rotamers x {nat} atoms: {t*1e3:.2f} ms = {npair/t:.3g} pairs/s")
