"""world_size-2 gloo tests (CPU) of the host-side logic of the point-sharded path: shard ranges tile the
points, the all-gather packs the rank partials rank-major, and the rank-ordered log-sum-exp merge of
partials (restated with the oracle's arithmetic) reproduces the single-rank slice state."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from haet_b200.distributed import gather_partials, shard_range


def test_shard_ranges_tile_the_points():
    for n in (1, 7, 32186, 1_000_000, 8_000_001):
        for w in (1, 2, 3, 4, 8):
            r = [shard_range(n, k, w) for k in range(w)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[k][1] == r[k + 1][0] for k in range(w - 1))
            sizes = [b - a for a, b in r]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)


def _partial(xp, own, lo, hi):
    """One rank's packed (m2, l, S) over points [lo, hi) -- the definition in include/haet_b200.h."""
    import math
    import torch.nn.functional as F
    x = xp[:, :, lo:hi]
    tau = 0.5 + F.linear(x, own["ada_temp_linear.weight"], own["ada_temp_linear.bias"]).clamp(-0.4, 0.4)
    c = float(torch.log(-torch.log(torch.tensor(1e-6))))   # fp32 constant, as the reference computes it
    z2 = (F.linear(x, own["in_project_slice.weight"], own["in_project_slice.bias"]) - c) / tau * math.log2(math.e)
    m2 = z2.amax(2)                                   # [B,H,G]
    e = torch.exp2(z2 - m2[:, :, None])
    l = e.sum(2)
    S = torch.einsum("bhnc,bhng->bhgc", x, e)
    return torch.cat([m2[..., None], l[..., None], S], dim=-1).reshape(-1)


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import haet_oracle as O
        from tests.helpers import load_layer_golden, oracle_kwargs
        cfg, sd, x, out = load_layer_golden("layer_irregular_small")
        kw = oracle_kwargs(cfg)
        own, _ = O.split_state_dict(sd, torch.float64)
        xp = O.in_project(own, x.double(), kw["heads"], kw["dim_head"])
        N = x.shape[1]
        lo, hi = shard_range(N, rank, world)
        mine = _partial(xp, own, lo, hi)
        parts, R = gather_partials(mine)
        assert R == world and parts.numel() == world * mine.numel()
        # rank-major packing
        assert torch.equal(parts.view(world, -1)[rank], mine)
        # rank-ordered LSE merge == single rank
        p = parts.view(world, -1, kw["dim_head"] + 2)
        m = p[:, :, 0].amax(0)
        sc = torch.exp2(p[:, :, 0] - m)
        l = (p[:, :, 1] * sc).sum(0)
        S = (p[:, :, 2:] * sc[..., None]).sum(0) / l[:, None] / (1 + 1e-5)
        _, S_ref, _ = O.slice_weights_and_states(own, xp)
        err = float((S - S_ref.reshape(S.shape)).abs().max() / S_ref.abs().max())
        ret[rank] = err
    finally:
        dist.destroy_process_group()


def test_gloo_world2_gather_and_merge():
    world = 2
    port = 29500 + (os.getpid() % 2000)
    with mp.Manager() as mgr:
        ret = mgr.dict()
        mp.spawn(_worker, args=(world, port, ret), nprocs=world, join=True)
        assert len(ret) == world
        assert all(e < 1e-9 for e in ret.values()), dict(ret)


# ------------------------------------------------------------------------------------------------
# slab sharding of the structured meshes: halo exchange of the conv front-end
# ------------------------------------------------------------------------------------------------
def test_plane_shards_tile_the_grid():
    from haet_b200.distributed import shard_grid, shard_planes
    for H in (2, 7, 64, 85):
        for w in (1, 2, 3, 4):
            if H < w:
                continue
            r = [shard_planes(H, k, w) for k in range(w)]
            assert r[0][0] == 0 and r[-1][1] == H and all(r[k][1] == r[k + 1][0] for k in range(w - 1))
    x = torch.arange(2 * 6 * 5 * 4 * 3, dtype=torch.float32).view(2, 6 * 5 * 4, 3)
    parts = [shard_grid(x, (6, 5, 4), k, 3) for k in range(3)]
    assert torch.equal(torch.cat(parts, 1), x) and all(p.shape[1] == 2 * 20 for p in parts)


def _halo_worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from haet_b200.distributed import exchange_halo, shard_grid, shard_planes
        from oracle import haet_oracle as O
        from tests.helpers import load_layer_golden, oracle_kwargs
        errs = []
        for name in ("layer_3d_small", "layer_2d_small"):
            cfg, sd, x, out = load_layer_golden(name)
            kw = oracle_kwargs(cfg)
            grid = kw["grid"]
            plane = 1
            for g in grid[1:]:
                plane *= g
            own, _ = O.split_state_dict(sd, torch.float64)
            xp_full = O.in_project(own, x.double(), kw["heads"], kw["dim_head"], grid=grid)     # [B, H, N, C]: the global conv
            mine = shard_grid(x, grid, rank, world)
            ext, lo, hi = exchange_halo(mine, plane, 1, None)
            p_lo, p_hi = shard_planes(grid[0], rank, world)
            assert lo == (1 if rank > 0 else 0) and hi == (1 if rank < world - 1 else 0)
            # the extended slab is exactly the rows of the global mesh around this rank's planes
            assert torch.equal(ext, x[:, (p_lo - lo) * plane:(p_hi + hi) * plane])
            # conv over the extended slab (zero padding beyond it), halo rows dropped == the global conv on this rank's rows
            g_ext = (p_hi - p_lo + lo + hi,) + tuple(grid[1:])
            xp_ext = O.in_project(own, ext.double(), kw["heads"], kw["dim_head"], grid=g_ext)
            got = xp_ext[:, :, lo * plane:lo * plane + mine.shape[1]]
            want = xp_full[:, :, p_lo * plane:p_hi * plane]
            errs.append(float((got - want).abs().max()))
        ret[rank] = max(errs)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_gloo_halo_exchange_reproduces_the_global_conv(world):
    port = 31500 + (os.getpid() % 2000) + world
    with mp.Manager() as mgr:
        ret = mgr.dict()
        mp.spawn(_halo_worker, args=(world, port, ret), nprocs=world, join=True)
        assert len(ret) == world
        assert all(e < 1e-12 for e in ret.values()), dict(ret)
