"""f1 at scale: wall time of the xyz-ensemble reader, the energy sort and the whole
file -> group[] path at BASELINE sizes.  usage: python tools/bench_io.py [workload] [nall]"""
import os, sys, time, tempfile
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import crest_b200 as cb
from crest_b200 import synth

wl = sys.argv[1] if len(sys.argv) > 1 else "C3"
nall = int(sys.argv[2]) if len(sys.argv) > 2 else None
ens = synth.make_config(wl, nall=nall)
at, xyz, er = ens["at"], ens["xyz"], ens["er"]
n, nat, _ = xyz.shape
rng = np.random.default_rng(0)
perm = rng.permutation(n)                      # the MD/optimisation output is not energy sorted
d = tempfile.mkdtemp(dir=os.environ.get("TMPDIR", "/tmp"))
fn = os.path.join(d, "ens.xyz")
t0 = time.perf_counter(); cb.wrensemble(fn, at, xyz[perm], er[perm]); t_wr = time.perf_counter() - t0
size = os.path.getsize(fn)
for rep in range(2):
    t0 = time.perf_counter(); nat2, nall2 = cb.rdensembleparam(fn); t_par = time.perf_counter() - t0
    t0 = time.perf_counter(); r = cb.rdensemble(fn, want_comments=False); t_rd = time.perf_counter() - t0
at2, xyz2, er2 = r["at"], r["xyz"], r["er"]
assert nall2 == n and np.array_equal(at2, at)
assert np.abs(xyz2 - xyz[perm]).max() < 1e-9 and np.array_equal(er2, er[perm])
t0 = time.perf_counter(); xs, es, order, nout = cb.cregen_esort(xyz2, er2); t_sort = time.perf_counter() - t0
assert np.array_equal(es, er)
print(f"{wl}: {n} x {nat}, file {size/1e6:.0f} MB: write {t_wr*1e3:.0f} ms, rdensembleparam {t_par*1e3:.0f} ms, "
      f"rdensemble {t_rd*1e3:.0f} ms ({size/t_rd/1e6:.0f} MB/s), cregen_esort {t_sort*1e3:.1f} ms", flush=True)
if cb._lib.lib().crest_cuda_device_count() > 0:
    env = cb.Env(ethr=1e9)
    for rep in range(3):
        t0 = time.perf_counter()
        r = cb.rdensemble(fn, want_comments=False)
        at2, xyz2, er2 = r["at"], r["xyz"], r["er"]
        nat2 = len(at2)
        xs, es, order, nout = cb.cregen_esort(xyz2, er2)
        full = cb.cregen_CRE_inplace(env, nat2, nout, at2, xs[:nout], es[:nout])
        t_all = time.perf_counter() - t0
    print(f"   file -> group[]: {t_all*1e3:.0f} ms (kept {full.nallout}, groups {int(full.group[0])})", flush=True)
os.remove(fn); os.rmdir(d)
