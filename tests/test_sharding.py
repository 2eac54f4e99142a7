"""Multi-process path on CPU: world_size-2 gloo run of the sharded hot path (host build of the kernels),
checking that shard -> integrate -> gather reproduces the single-process result bit for bit."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from cocofest_b200 import sharding


def test_shard_bounds_cover_the_batch_exactly():
    for batch in (1, 7, 64, 1000003):
        for world in (1, 2, 3, 8):
            spans = [sharding.shard_bounds(batch, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == batch
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from cocofest_b200 import _lib, fes, protocols
        from tests.conftest import _load_emul

        _lib._LIB = _load_emul()  # test-only injection of the host build (no GPU in this container)
        B = 37  # not divisible by 2: unequal shards
        proto = protocols.fatigue_protocol(final_time=2)
        params, x0 = protocols.fatigue_population(B)
        win = torch.tensor(proto["win"])
        run = lambda p, x: fes.ivp(fes.FES_DING2003, True, p, x, win, proto["dt"], n_steps=10, n_shooting=proto["ns"],
                                   dirs=fes.FES_DIR_X0)
        mine = run(sharding.shard_last(params, rank, world), sharding.shard_last(x0, rank, world))
        lo, hi = sharding.shard_bounds(B, rank, world)
        assert mine.x_final.shape == (5, hi - lo)
        full_x = sharding.gather_last(mine.x_final, batch=B)
        full_s = sharding.gather_last(mine.sens_final, batch=B)
        assert full_x.shape == (5, B) and full_s.shape == (5, 5, B)
        # without `batch` the shard sizes are exchanged instead of derived; equal shards take the single-copy path
        assert torch.equal(sharding.gather_last(mine.x_final), full_x)
        even = sharding.gather_last(mine.sens_final[..., :18].contiguous())
        assert even.shape == (5, 5, 36) and torch.equal(even[..., :18], full_s[..., :18]) and torch.equal(even[..., 18:], full_s[..., 19:])
        # the other two entry points shard the same way (the batch is the last axis everywhere): scenarios of a
        # multiple-shooting evaluation with the force factor as an input, candidates of an identification
        ns = 6
        winN = win[:ns + 1]
        X = (x0[None] * torch.linspace(1.0, 0.8, ns + 1, dtype=torch.float64)[:, None, None]).contiguous()
        fc = (0.6 + 0.5 * torch.rand((ns, B), generator=torch.Generator().manual_seed(1), dtype=torch.float64))
        shoot = lambda p, x, f: fes.shoot(fes.FES_DING2003, True, p, x, winN, proto["dt"], n_steps=10, force_factor=f)
        sh = shoot(sharding.shard_last(params, rank, world), sharding.shard_last(X, rank, world), sharding.shard_last(fc, rank, world))
        G, Jx, Jc = (sharding.gather_last(t.contiguous(), batch=B) for t in (sh.G, sh.Jx, sh.Jc))
        # ... and the second-order blocks of the same scenarios (fes_shoot_hess_f64: the batch is the last axis there too)
        lam = torch.rand((ns, 5, B), generator=torch.Generator().manual_seed(2), dtype=torch.float64)
        hess = lambda p, x, f, l: fes.shoot_hessian(fes.FES_DING2003, True, p, x, winN, proto["dt"], n_steps=10, force_factor=f, multipliers=l)
        hs = hess(*(sharding.shard_last(t, rank, world) for t in (params, X, fc, lam)))
        D2, H = sharding.gather_last(hs.D2, batch=B), sharding.gather_last(hs.H, batch=B)
        proto_h, cand, x0_h, ctrl_h = protocols.hmed2018_candidates(B)
        win_h = torch.tensor(proto_h["win"])
        truth = torch.tensor(fes.default_params(fes.FES_HMED2018))
        target = fes.ivp(fes.FES_HMED2018, False, truth, x0_h[:, :1].contiguous(), win_h, proto_h["dt"], ctrl=ctrl_h,
                         n_shooting=proto_h["ns"], trace="nodes").trace[:, 1, 0].contiguous()
        ident = lambda p: fes.identify(fes.FES_HMED2018, p, win_h, proto_h["dt"], target, ctrl=ctrl_h, max_iter=2)
        idm = ident(sharding.shard_last(cand, rank, world))
        P, C = sharding.gather_last(idm.params, batch=B), sharding.gather_last(idm.cost, batch=B)
        if rank == 0:
            ref = run(params, x0)
            assert torch.equal(full_x, ref.x_final) and torch.equal(full_s, ref.sens_final)
            rs = shoot(params, X, fc)
            assert torch.equal(G, rs.G) and torch.equal(Jx, rs.Jx) and torch.equal(Jc, rs.Jc)
            rh = hess(params, X, fc, lam)
            assert torch.equal(D2, rh.D2) and torch.equal(H, rh.H) and D2.shape == (ns, 5, 21, B)
            ri = ident(cand)
            assert torch.equal(P, ri.params) and torch.equal(C, ri.cost)
            open(os.path.join(out_dir, "ok"), "w").write("ok")
    finally:
        dist.destroy_process_group()


def test_world_size_2_gloo_shard_integrate_gather(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    assert (tmp_path / "ok").exists()
