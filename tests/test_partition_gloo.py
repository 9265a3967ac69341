"""Multi-rank host logic on CPU (gloo, world_size 2): the owner-computes row partition tiles the dof range, balances the
(row, element) pairs, every rank lists exactly the elements touching its rows, and the scalar all-reduce of per-rank
element-block integrals reproduces the single-rank sum."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from ngsxfem_b200 import meshes, partition
    from oracle import orc
    m = meshes.MakeStructured3DMesh(6)
    ls = meshes.sphere_levelset(m.vert, (0.5, 0.5, 0.5), 0.35)
    e2d, ndof = m.h1_dofs(2)
    haspos = (ls[m.el2v.long()] > 0).any(dim=1)
    hasneg = (ls[m.el2v.long()] < 0).any(dim=1)
    active = torch.nonzero(hasneg).reshape(-1)
    blocks = partition.row_blocks(e2d[active], ndof, world)
    lo, hi = blocks[rank]
    mine = partition.elements_touching(e2d, active, lo, hi)
    # pairs owned by this rank
    d = e2d[active].reshape(-1)
    owned = int(((d >= lo) & (d < hi)).sum())
    t = torch.tensor([owned], dtype=torch.int64)
    allp = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(allp, t)
    # per-rank element block integral of the NEG volume with the CPU oracle, then the path's scalar all-reduce
    elo, ehi = partition.element_block(m.ne, rank, world)
    part, _ = orc.integrate_mesh(3, m.el2v[elo:ehi].numpy(), m.vert.numpy(), ls.numpy(), 0, 4)
    s = torch.tensor([part], dtype=torch.float64)
    dist.all_reduce(s)
    # cost-weighted element blocks of the sweeps (a cut element = 40 sign tests): they tile [0, ne), every rank derives the same
    # boundaries, the weights balance, and the all-reduced interface integral over them is the single-rank one
    wts = 1.0 + 40.0 * (haspos & hasneg).to(torch.float64)
    wb = partition.element_blocks_weighted(wts, world)
    assert wb[0][0] == 0 and wb[-1][1] == m.ne and all(wb[r][1] == wb[r + 1][0] for r in range(world - 1))
    share = float(wts[wb[rank][0]:wb[rank][1]].sum() / wts.sum())
    assert abs(share - 1.0 / world) < 0.05
    wpart, _ = orc.integrate_mesh(3, m.el2v[wb[rank][0]:wb[rank][1]].numpy(), m.vert.numpy(), ls.numpy(), 2, 4)
    ws = torch.tensor([wpart], dtype=torch.float64)
    dist.all_reduce(ws)
    wfull, _ = orc.integrate_mesh(3, m.el2v.numpy(), m.vert.numpy(), ls.numpy(), 2, 4)
    assert abs(float(ws) - wfull) <= 1e-13 * abs(wfull)
    if rank == 0:
        full, _ = orc.integrate_mesh(3, m.el2v.numpy(), m.vert.numpy(), ls.numpy(), 0, 4)
        out.put(dict(blocks=blocks, pairs=[int(x) for x in allp], total=int(d.numel()), sum=float(s), full=float(full), ndof=ndof))
    # element-chunk ownership (the partition bench.py uses): every dof of an active element has exactly one owner, the rank's
    # list holds its own chunk, and the lists of all ranks cover the active elements
    owner = partition.row_owners(e2d[active], ndof, world)
    touched = torch.zeros(ndof, dtype=torch.bool)
    touched[e2d[active].reshape(-1).long()] = True
    assert bool(((owner >= 0) == touched).all()) and int(owner.max()) == world - 1
    mine_o = partition.elements_of_owner(e2d, active, owner, rank)
    n = active.numel()
    idx = torch.arange(n)
    own_chunk = active[torch.div(idx * world, n, rounding_mode="floor") == rank]
    assert bool(torch.isin(own_chunk, mine_o).all())
    cnt = torch.tensor([int((owner == rank).sum()), mine_o.numel()], dtype=torch.int64)
    allc = [torch.zeros_like(cnt) for _ in range(world)]
    dist.all_gather(allc, cnt)
    assert sum(int(c[0]) for c in allc) == int(touched.sum())
    assert all(int(c[1]) <= 0.85 * n for c in allc)                                # a chunk + a halo layer, not everything
    # every element touching an owned row is listed, and nothing else
    touch = torch.zeros(m.ne, dtype=torch.bool)
    dd = e2d[active]
    touch[active[((dd >= lo) & (dd < hi)).any(dim=1)]] = True
    got = torch.zeros(m.ne, dtype=torch.bool)
    got[mine] = True
    assert torch.equal(touch, got)
    dist.barrier()
    dist.destroy_process_group()


def test_owner_computes_partition_world2():
    world = 2
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, world, port, out)) for r in range(world)]
    for p in procs:
        p.start()
    res = out.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    blocks = res["blocks"]
    assert blocks[0][0] == 0 and blocks[-1][1] == res["ndof"] and all(blocks[r][1] == blocks[r + 1][0] for r in range(world - 1))
    assert sum(res["pairs"]) == res["total"]
    assert max(res["pairs"]) <= 0.55 * res["total"]                     # balanced by pairs
    assert abs(res["sum"] - res["full"]) <= 1e-13 * abs(res["full"])
