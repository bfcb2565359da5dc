"""Ad-hoc first GPU parity check against the oracle (tools only; tests/ hold the real ones)."""
import sys, os, time
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
import numpy as np
import oracle_lib as O
import crest_b200 as cb
from crest_b200 import synth

def run(name, nall, ethr):
    ens = synth.make_config(name, nall=nall)
    at, xyz, er = ens["at"], ens["xyz"], ens["er"]
    nall, nat, _ = xyz.shape
    env = cb.Env(ethr=ethr, heavyrmsd=ens["heavy"])
    t0 = time.time()
    g = cb.cregen_dup(env, nat, nall, at, xyz, er)
    t1 = time.time()
    ref = O.cregen_cre(at, xyz, er, ethr=ethr, heavy=ens["heavy"], want_rmat=True)
    t2 = time.time()
    print(name, nall, "gpu s", t1 - t0, "oracle s", t2 - t1, "npairs", g["npairs"], ref["klong"])
    rot_ref = ref["rot"][ref["orderref"] - 1]
    print(" rot max rel diff", np.abs(g["rot"] / rot_ref - 1).max())
    xa_ref = ref["xyz"][ref["orderref"] - 1]
    print(" xyz_axis max abs diff", np.abs(g["xyz_axis"] - xa_ref).max())
    print(" dup equal:", np.array_equal(g["dup"], ref["dup"]), "ndup", (ref["dup"] != 0).sum(),
          "mismatch", (g["dup"] != ref["dup"]).sum())
    if ethr > 1e8:
        rm = cb.rmat_allpairs(env, at, xyz)
        d = np.abs(rm.astype(np.float64) - ref["rmat"].astype(np.float64))
        print(" rmat max abs diff", d.max(), "at", d.argmax(), "ref", ref["rmat"][d.argmax()])
    return g, ref

if __name__ == "__main__":
    run("C1", 300, 1e9)
    run("C1", 1000, 1e9)
    run("C1", 1000, 0.05)
    run("C2", 2000, 1e9)
