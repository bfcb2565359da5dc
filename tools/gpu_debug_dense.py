import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import crest_b200 as cb
from crest_b200 import synth, cregen
name = sys.argv[1] if len(sys.argv) > 1 else "C3dense"
nall = int(sys.argv[2]) if len(sys.argv) > 2 else None
ens = synth.make_config(name, nall=nall)
at, xyz, er = ens["at"], ens["xyz"], ens["er"]
n, nat, _ = xyz.shape
env = cb.Env(ethr=1e9)
def ctx_run(tune=None, ranks=1):
    cregen.set_tune(tune)
    c = cb.CregenContext(); c.upload(env, at, xyz, er); c.pack()
    for r in range(ranks):
        c.pairs(env, r, ranks)
    d = c.download(); info = c.info(env); c.close(); cregen.set_tune(None)
    return d, info
ref, _ = ctx_run("notf32")
def cmp(tag, d, info=None):
    bad = np.where(d["dup"] != ref["dup"])[0]
    print(f"{tag:28s} npairs ok {d['npairs'] == ref['npairs']}  differing rows {len(bad)}", (info or ""), flush=True)
    for i in bad[:5]:
        print("     row", i, "got", d["dup"][i], "ref", ref["dup"][i])
d, i = ctx_run(); cmp("ctx 1 rank", d, {k: i[k] for k in ("candidates", "exact_evals", "dense_items", "flags")})
d, i = ctx_run(None, 8); cmp("ctx 8 virtual ranks", d, {k: i[k] for k in ("candidates", "exact_evals", "dense_items", "flags")})
d, i = ctx_run("nodense", 8); cmp("ctx 8 virtual ranks nodense", d)
for ng in (1, 2):
    if ng > cb._lib.lib().crest_cuda_device_count():
        continue
    cb.init(ng)
    for tune in ("blocks=1", None, "blocks=4,nodense"):
        cregen.set_tune(tune)
        # links before the root chase are not returned by the one-call ABI: compare roots
        c = cb.CregenContext(); c.upload(env, at, xyz, er); c.pack(); c.pairs(env); c.sweep(); root = c.download()["dup"]; c.close()
        o = cb.cregen_dup(env, nat, n, at, xyz, er, want_xyz=False)
        bad = np.where(o["dup"] != root)[0]
        print(f"engine ngpu={ng} tune={tune}: npairs ok {o['npairs'] == ref['npairs']} differing roots {len(bad)}", flush=True)
        for i in bad[:5]:
            print("     row", i, "got", o["dup"][i], "ref", root[i])
    cregen.set_tune(None)
cb.finalize()
