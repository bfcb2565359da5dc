"""Multi-GPU driver of the duplicate-sorting path: one process per GPU (torch.distributed), the
in-band tile list of the pair triangle split into contiguous shards, every rank holding the whole
(small) packed ensemble.  The only exchange step is the per-row "largest matching j"
(reference src/cregen.f90:1353-1373: `double(i)=j` without early exit), i.e. ONE all-reduce(MAX)
over int32[nall] plus an all-reduce(SUM) of the pair counter; the root chase (backtrack,
cregen_old.f90:1396-1410) then runs redundantly on every rank.

The compute engine is the device-resident CregenContext (crest_cuda_ctx_*).  `engine` may be
replaced by an object with the same methods — the CPU tests do that to exercise this host logic
over gloo without a GPU; nothing in this package provides a CPU engine.
"""
import numpy as np

from .cregen import CregenContext, Env


class _DevArray:
    """Raw device pointer -> torch tensor through __cuda_array_interface__ (no copy)."""

    def __init__(self, ptr, n, typestr):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False),
                                         "version": 2}


class CudaEngine:
    """CregenContext plus torch views of its dup / npairs device buffers for the collectives."""

    def __init__(self, device):
        import torch
        self.torch = torch
        self.device = torch.device("cuda", device)
        torch.cuda.set_device(device)
        self.ctx = CregenContext(device)
        self._dup_t = None
        self._np_t = None

    def upload(self, env, at, xyz, er):
        self.ctx.upload(env, at, xyz, er)
        t = self.torch
        self._dup_t = t.as_tensor(_DevArray(self.ctx.dup_devptr(), max(self.ctx.nall, 1), "<i4"),
                                  device=self.device)
        self._np_t = t.as_tensor(_DevArray(self.ctx.npairs_devptr(), 1, "<i8"), device=self.device)

    def upload_rows(self, env, at, xyz, er, rank, world):
        """This rank's slice host->device; returns (full view, my slice view) of the device xyz
        buffer for the in-place all-gather (equal slices of `chunk` rows)."""
        chunk, first, count = self.ctx.upload_rows(env, at, xyz, er, rank, world)
        t = self.torch
        rec = 3 * self.ctx.nat
        full = t.as_tensor(_DevArray(self.ctx.xyz_devptr(), chunk * world * rec, "<f8"),
                           device=self.device)
        self._dup_t = t.as_tensor(_DevArray(self.ctx.dup_devptr(), max(self.ctx.nall, 1), "<i4"),
                                  device=self.device)
        self._np_t = t.as_tensor(_DevArray(self.ctx.npairs_devptr(), 1, "<i8"), device=self.device)
        return full, full[rank * chunk * rec:(rank + 1) * chunk * rec]

    def stream(self):
        return self.torch.cuda.current_stream().cuda_stream

    def pack(self):
        self.ctx.pack(self.stream())

    def pairs(self, env, rank, world):
        self.ctx.pairs(env, rank, world, False, self.stream())

    def dup_tensor(self):
        return self._dup_t

    def npairs_tensor(self):
        return self._np_t

    def sweep(self):
        self.ctx.sweep(self.stream())

    def download(self):
        return self.ctx.download(stream=self.stream())

    def close(self):
        self.ctx.close()


class DistributedCregen:
    """dup-stage (K0-K4) over the ranks of a torch.distributed process group."""

    def __init__(self, engine, group=None):
        import torch.distributed as dist
        self.dist = dist
        self.group = group
        self.engine = engine
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1

    def step(self, env: Env):
        """pack -> this rank's tile shard -> all-reduce(MAX) -> sweep, on the resident ensemble."""
        e = self.engine
        e.pack()
        e.pairs(env, self.rank, self.world)
        if self.world > 1:
            self.dist.all_reduce(e.dup_tensor(), op=self.dist.ReduceOp.MAX, group=self.group)
        e.sweep()

    def npairs_total(self):
        t = self.engine.npairs_tensor().clone()
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM, group=self.group)
        return int(t.item())

    def dup(self, env: Env, at, xyz, er):
        """Upload (every rank holds the full ensemble), run one step, return rank-identical results."""
        if self.world > 1 and hasattr(self.engine, "upload_rows"):
            # each rank pulls 1/world of the geometry over PCIe; NVLink all-gather replicates it
            full, mine = self.engine.upload_rows(env, np.asarray(at), np.asarray(xyz),
                                                 np.asarray(er), self.rank, self.world)
            self.dist.all_gather_into_tensor(full, mine, group=self.group)
        else:
            self.engine.upload(env, np.asarray(at), np.asarray(xyz), np.asarray(er))
        self.step(env)
        out = self.engine.download()
        out["npairs"] = self.npairs_total()
        return out
