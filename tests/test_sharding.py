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
        if rank == 0:
            ref = run(params, x0)
            assert torch.equal(full_x, ref.x_final) and torch.equal(full_s, ref.sens_final)
            open(os.path.join(out_dir, "ok"), "w").write("ok")
    finally:
        dist.destroy_process_group()


def test_world_size_2_gloo_shard_integrate_gather(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    assert (tmp_path / "ok").exists()
