"""Developer tool: host/GPU timeline of the pipelined one-call stage A (CREST_CUDA_TIMING=2)."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import crest_b200 as cb
from crest_b200 import synth, cregen
wl = sys.argv[1] if len(sys.argv) > 1 else "C3"
ens = synth.make_config(wl)
at, xyz, er = ens["at"], ens["xyz"], ens["er"]
nall, nat, _ = xyz.shape
env = cb.Env(ethr=1e9)
if len(sys.argv) > 2:
    cb.init(int(sys.argv[2]))
for tune in (None, "blocks=1", "blocks=4", "blocks=6"):
    cregen.set_tune(tune)
    for rep in range(3):
        os.environ["CREST_CUDA_TIMING"] = "2" if rep == 2 else "0"
        t0 = time.perf_counter()
        a = cb.cregen_dup(env, nat, nall, at, xyz, er, want_xyz=False)
        dt = time.perf_counter() - t0
    print(f"tune={tune}: call {dt*1e3:.2f} ms", flush=True)
