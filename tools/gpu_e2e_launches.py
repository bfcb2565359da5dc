"""Developer tool: ONE call of the one-call ABI (crest_cuda_cregen_dup) for `ncu --metrics gpu__time_duration.sum`:
the per-stage kernels of the upload pipeline.  usage: gpu_e2e_launches.py <workload> [tune]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import crest_b200 as cb
from crest_b200 import synth, cregen
ens = synth.make_config(sys.argv[1] if len(sys.argv) > 1 else "C3")
if len(sys.argv) > 2:
    cregen.set_tune(sys.argv[2])
at, xyz, er = ens["at"], ens["xyz"], ens["er"]
nall, nat, _ = xyz.shape
a = cb.cregen_dup(cb.Env(ethr=1e9), nat, nall, at, xyz, er, want_xyz=False)
print("pairs", a["npairs"], "dups", int((a["dup"] != 0).sum()))
