"""GPU check of the tensor-core screen path against the FP64 column kernel (developer tool)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import crest_b200 as cb
from crest_b200 import synth, cregen

def run(ens, env, tune=None, debug=None):
    cregen.set_tune(tune)
    if debug is None:
        os.environ.pop("CREST_CUDA_DEBUG", None)
    else:
        os.environ["CREST_CUDA_DEBUG"] = str(debug)
    ctx = cb.CregenContext(0)
    ctx.upload(env, ens["at"], ens["xyz"], ens["er"])
    ctx.pack()
    t0 = time.perf_counter()
    ctx.pairs(env)
    ctx.sweep()
    out = ctx.download()
    dt = time.perf_counter() - t0
    info = ctx.info(env)
    ctx.close()
    os.environ.pop("CREST_CUDA_DEBUG", None)
    cregen.set_tune(None)
    return out, info, dt

def case(name, nall, ethr, heavy=None):
    ens = synth.make_config(name, nall=nall)
    env = cb.Env(ethr=ethr, heavyrmsd=ens["heavy"])
    ref, iref, tref = run(ens, env, "notf32")
    got, ig, tg = run(ens, env)
    same = np.array_equal(ref["dup"], got["dup"]) and ref["npairs"] == got["npairs"]
    print(f"{name} nall={nall} ethr={ethr}: variant {iref['variant']}->{ig['variant']} same={same} "
          f"npairs {ref['npairs']} vs {got['npairs']} dups {int((ref['dup']!=0).sum())} vs {int((got['dup']!=0).sum())} "
          f"cand={ig['candidates']} ovf={ig['overflow']} t_ref={tref*1e3:.1f}ms t_tf32={tg*1e3:.1f}ms", flush=True)
    if not same:
        bad = np.nonzero(ref["dup"] != got["dup"])[0]
        print("  first mismatches:", [(int(i), int(ref["dup"][i]), int(got["dup"][i])) for i in bad[:10]], "count", len(bad))
    return ens, env, ref, same

ok = True
ens, env, ref, s = case("C1", 1000, 1e9); ok &= s
# self-check: every pair through the exact kernel, count pairs the screen wrongly called far
got, ig, _ = run(ens, env, None, debug=3)
print("  all_cand self-check: violations", ig["violations"], "cand", ig["candidates"], "same", np.array_equal(ref["dup"], got["dup"]), flush=True)
ok &= ig["violations"] == 0 and np.array_equal(ref["dup"], got["dup"])
got, ig, _ = run(ens, env, "cap=64")
print("  overflow fallback: ovf", ig["overflow"], "same", np.array_equal(ref["dup"], got["dup"]), flush=True)
ok &= ig["overflow"] == 1 and np.array_equal(ref["dup"], got["dup"])
_, _, _, s = case("C1", 1000, 0.05); ok &= s
_, _, _, s = case("C2", 3000, 1e9); ok &= s
_, _, _, s = case("C2", 3000, 0.1); ok &= s
_, _, _, s = case("C5", 5000, 1e9); ok &= s
ens, env, ref, s = case("C3", 20000, 1e9); ok &= s
got, ig, _ = run(ens, env, None, debug=3)
print("  C3/20000 all_cand self-check: violations", ig["violations"], "ovf", ig["overflow"], flush=True)
_, _, _, s = case("C3", 100000, 1e9); ok &= s
print("ALL OK" if ok else "FAILED")
