"""Evidence for the TF32 screen's error budget: measures, on the B200, the largest
|R~_ab - R_ab(exact on the TF32-rounded operands)| / (|x_a||y_b|) that tcgen05.mma kind::tf32 with FP32
accumulation produces, over >= 1e9 dot products, including cancellation-heavy and wide-dynamic-range
inputs, and compares it with the gamma the bound budgets (pair_tf32.cu: tf32_kappa).  Also runs the
CREST_CUDA_DEBUG=3 self-check (every pair through the exact pass, tagged with the screen's verdict) at
2e8 pairs of C3 and on the C5 / C4 shapes.  Output: profiles/r02_tf32_accum.txt"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import crest_b200 as cb
from crest_b200 import synth


def probe(name, at, xyz, er):
    ctx = cb.CregenContext()
    ctx.upload(cb.Env(ethr=1e9), at, xyz, er)
    ctx.pack()
    r = ctx.tf32_accum()
    ctx.close()
    print(f"{name:34s} n={xyz.shape[0]:6d} A={xyz.shape[1]:4d}  dots={r['ndots']:.3e}  max|dR|/(|x||y|)={r['max_rel_err']:.3e}  "
          f"gamma budget={r['gamma_budget']:.2e}  ratio={r['max_rel_err'] / r['gamma_budget']:.4f}  kappa={r['kappa']:.4e}",
          flush=True)
    return r


def cancel_ensemble(n, nat, seed, spread):
    """Atoms in (+p, -p) pairs in half of the structures and (+q, +q) pairs in the others, so that the
    dot products of a mixed pair cancel term by term; coordinate magnitudes spread over `spread` decades."""
    rng = np.random.default_rng(seed)
    mag = 10.0 ** rng.uniform(-spread / 2, spread / 2, size=(n, nat // 2, 1))
    p = rng.normal(size=(n, nat // 2, 3)) * mag
    sign = np.where(rng.uniform(size=(n, 1, 1)) < 0.5, -1.0, 1.0)
    xyz = np.concatenate([p, sign * p + 1e-3 * rng.normal(size=p.shape)], axis=1)
    xyz = xyz[:, rng.permutation(xyz.shape[1])]
    at = np.full(xyz.shape[1], 6, np.int32)
    er = np.sort(rng.uniform(-1, 0, n))
    return at, np.ascontiguousarray(xyz), er


total = 0
worst = 0.0
e = synth.make_config("C3", nall=16000)
r = probe("C3 synthetic (A=60)", e["at"], e["xyz"], e["er"]); total += r["ndots"]; worst = max(worst, r["max_rel_err"] / r["gamma_budget"])
e = synth.make_config("C5", nall=12000)
r = probe("C5 synthetic (A=40)", e["at"], e["xyz"], e["er"]); total += r["ndots"]; worst = max(worst, r["max_rel_err"] / r["gamma_budget"])
at, xyz, er = cancel_ensemble(12000, 64, 1, 0.0)
r = probe("cancelling +-pairs (A=64)", at, xyz, er); total += r["ndots"]; worst = max(worst, r["max_rel_err"] / r["gamma_budget"])
at, xyz, er = cancel_ensemble(12000, 88, 2, 4.0)
r = probe("cancelling, 4 decades (A=88)", at, xyz, er); total += r["ndots"]; worst = max(worst, r["max_rel_err"] / r["gamma_budget"])
e = synth.make_config("C4", nall=6000)
r = probe("C4 synthetic (A=300, K slabs)", e["at"], e["xyz"], e["er"]); total += r["ndots"]; worst = max(worst, r["max_rel_err"] / r["gamma_budget"])
at, xyz, er = cancel_ensemble(5000, 600, 3, 3.0)
r = probe("cancelling, 3 decades (A=600)", at, xyz, er); total += r["ndots"]; worst = max(worst, r["max_rel_err"] / r["gamma_budget"])
at, xyz, er = cancel_ensemble(3000, 1000, 4, 2.0)
r = probe("cancelling, 2 decades (A=1000)", at, xyz, er); total += r["ndots"]; worst = max(worst, r["max_rel_err"] / r["gamma_budget"])
print(f"TOTAL dot products checked: {total:.3e}; worst measured/budget = {worst:.4f}  ({'WITHIN' if worst < 1 else 'EXCEEDS'} the budget)")

# self-check of the screen's verdicts at scale: every pair through the exact pass
os.environ["CREST_CUDA_DEBUG"] = "3"
for name, nall in (("C3", 20000), ("C5", 20000), ("C4", 6000), ("C4ext", 5000)):
    e = synth.make_config(name, nall=nall)
    env = cb.Env(ethr=1e9)
    ctx = cb.CregenContext()
    ctx.upload(env, e["at"], e["xyz"], e["er"])
    ctx.pack(); ctx.pairs(env)
    info = ctx.info(env); links = ctx.download(); ctx.close()
    print(f"self-check {name} nall={nall}: pairs {links['npairs']:.3e}, all re-evaluated exactly: {info['exact_evals'] == links['npairs']}, "
          f"screen violations: {info['violations']}", flush=True)
os.environ.pop("CREST_CUDA_DEBUG")
