"""N-GPU check (torchrun): DistributedCregen over NCCL gives the single-GPU result (developer tool).
usage: python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/multi_gpu_check.py"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
import crest_b200 as cb
from crest_b200 import synth, multi

rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
ok = True
for cfg, nall, ethr in (("C1", 1000, 1e9), ("C2", 3000, 0.1), ("C3", 20000, 1e9), ("C4", 300, 1e9)):
    ens = synth.make_config(cfg, nall=nall, seed=7)
    env = cb.Env(ethr=ethr, heavyrmsd=ens["heavy"])
    eng = multi.CudaEngine(local)
    out = multi.DistributedCregen(eng).dup(env, ens["at"], ens["xyz"], ens["er"])
    eng.close()
    ctx = cb.CregenContext(local)
    ctx.upload(env, ens["at"], ens["xyz"], ens["er"]); ctx.pack(); ctx.pairs(env); ctx.sweep()
    one = ctx.download(); ctx.close()
    same = np.array_equal(out["dup"], one["dup"]) and out["npairs"] == one["npairs"]
    flag = torch.tensor([1 if same else 0], device="cuda"); dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(f"{cfg} nall={nall} ethr={ethr}: world={dist.get_world_size()} identical on all ranks: {bool(flag.item())} "
              f"(npairs {out['npairs']}, dups {int((out['dup'] != 0).sum())})", flush=True)
    ok &= bool(flag.item())
if rank == 0:
    print("MULTI OK" if ok else "MULTI FAILED", flush=True)
dist.destroy_process_group()
