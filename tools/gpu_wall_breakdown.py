"""Where the wall time of the whole cregen_CRE call goes (developer tool)."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import crest_b200 as cb
from crest_b200 import synth
ens = synth.make_config("C3")
at, xyz, er = ens["at"], ens["xyz"], ens["er"]
nall, nat, _ = xyz.shape
for ethr in (1e9, 0.05):
    env = cb.Env(ethr=ethr)
    cb.cregen_dup(env, nat, nall, at, xyz, er)
    t0 = time.perf_counter(); a = cb.cregen_dup(env, nat, nall, at, xyz, er, want_xyz=False); t1 = time.perf_counter()
    b = cb.cregen_dup(env, nat, nall, at, xyz, er, want_xyz=True); t2 = time.perf_counter()
    r = cb.cregen_CRE(env, nat, nall, at, xyz, er); t3 = time.perf_counter()
    n = r.nallout
    g, en = cb.cregen_group(env, nat, n, at, r.xyz[:n], r.er[:n], r.rot[:n]); t4 = time.perf_counter()
    print(f"ethr={ethr}: dup(no xyz) {1e3*(t1-t0):.1f} ms, dup(+xyz_axis D2H) {1e3*(t2-t1):.1f} ms, "
          f"cregen_CRE {1e3*(t3-t2):.1f} ms, group alone ({n} kept) {1e3*(t4-t3):.1f} ms, pairs {a['npairs']}")
