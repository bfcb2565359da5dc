"""world_size-2 (and 3) gloo tests of the multi-rank host logic in crest_b200/multi.py:
tile-list sharding (crest_cuda_band_tiles / crest_cuda_shard_range, host-only C-ABI),
all-reduce(MAX) of the per-row links, all-reduce(SUM) of the pair counter, redundant sweep.

No GPU here, and the product has no CPU engine: the test injects an engine that answers each
tile from a precomputed ORACLE rmat (checker only), so what is under test is the sharding and
reduction logic, not the arithmetic."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from crest_b200 import synth

AUTOKCAL = 627.50947428


class OracleTileEngine:
    def __init__(self, oracle_result, er, rthr, bthr, O):
        self.ref, self.er, self.rthr, self.bthr, self.O = oracle_result, er, rthr, bthr, O
        self.n = len(er)

    def upload(self, env, at, xyz, er):
        pass

    def pack(self):
        self._dup = torch.zeros(self.n, dtype=torch.int32)
        self._np = torch.zeros(1, dtype=torch.int64)

    def pairs(self, env, rank, world):
        from crest_b200 import cregen as cg
        ti, tj, _, _ = cg.band_tiles(self.er, env.ethr)
        b, e = cg.shard_range(len(ti), rank, world)
        TI, TJ = cg.TILE_I, cg.TILE_J
        inv = np.argsort(self.ref["orderref"])
        rot = np.empty_like(self.ref["rot"])
        rot[inv] = self.ref["rot"]
        rmat = self.ref["rmat"]
        for I, J in zip(ti[b:e], tj[b:e]):
            for i in range(I * TI, min((I + 1) * TI, self.n)):
                for j in range(J * TJ, min((J + 1) * TJ, i)):
                    if not ((self.er[i] - self.er[j]) * AUTOKCAL < env.ethr):
                        continue
                    self._np += 1
                    dr = float(rmat[i * (i - 1) // 2 + j])
                    if dr < self.rthr or (dr < 2 * self.rthr and
                                          self.O.equalrotaniso(rot[i], rot[j], 0.5 * self.bthr)):
                        self._dup[i] = max(int(self._dup[i]), j + 1)

    def dup_tensor(self):
        return self._dup

    def npairs_tensor(self):
        return self._np

    def sweep(self):
        d = self._dup.numpy()
        out = d.copy()
        for i in range(self.n):
            p = d[i]
            while p > 0 and d[p - 1] > 0:
                p = d[p - 1]
            out[i] = p
        self._dup.copy_(torch.from_numpy(out))

    def download(self):
        return dict(dup=self._dup.numpy().copy())


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import oracle_lib as O
    import crest_b200 as cb
    from crest_b200.multi import DistributedCregen
    ens = synth.make_config("C1", nall=150, seed=31)
    ref = O.cregen_cre(ens["at"], ens["xyz"], ens["er"], ethr=1e9, want_rmat=True)
    ok = True
    for ethr in (1e9, 0.05):
        env = cb.Env(ethr=ethr)
        eng = OracleTileEngine(ref, ens["er"], env.rthr, env.bthr2, O)
        drv = DistributedCregen(eng)
        assert drv.world == world and drv.rank == rank
        got = drv.dup(env, ens["at"], ens["xyz"], ens["er"])
        want = O.cregen_cre(ens["at"], ens["xyz"], ens["er"], ethr=ethr)
        ok &= bool(np.array_equal(got["dup"], want["dup"]))
        ok &= got["npairs"] == want["klong"]
    q.put((rank, ok))
    dist.barrier()
    dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_dup_over_gloo(world, cuda_lib, oracle):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(r for r, _ in res) == list(range(world))
    assert all(ok for _, ok in res)
