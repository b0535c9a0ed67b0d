"""CPU tests of the block decomposition + ghost exchange (gloo, world_size 2) and of the box
partitions the overlapped RHS uses.  No CUDA needed."""
import os
import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
from htr_solver_b200 import halo


def _global_field(N, comp):
    z, y, x = np.meshgrid(np.arange(N[2]), np.arange(N[1]), np.arange(N[0]), indexing="ij")
    base = (x + 100.0 * y + 10000.0 * z)[..., None] + 0.25 * np.arange(comp)
    return base


def _worker(rank, world, tiles, n, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        h = halo.HALO
        N = tuple(n[d] * tiles[d] for d in range(3))
        tc = halo.tile_coords(rank, tiles)
        G = _global_field(N, 3)
        Gs = (G[..., 0] % 2).astype(np.uint8)
        a = torch.full((n[2] + 2 * h, n[1] + 2 * h, n[0] + 2 * h, 3), -1.0, dtype=torch.float64)
        b = torch.full((n[2] + 2 * h, n[1] + 2 * h, n[0] + 2 * h), 7, dtype=torch.uint8)
        o = [tc[d] * n[d] for d in range(3)]
        a[h:-h, h:-h, h:-h] = torch.from_numpy(G[o[2]:o[2] + n[2], o[1]:o[1] + n[1], o[0]:o[0] + n[0]])
        b[h:-h, h:-h, h:-h] = torch.from_numpy(Gs[o[2]:o[2] + n[2], o[1]:o[1] + n[1], o[0]:o[0] + n[0]])
        halo.exchange([a, b], n, rank, tiles, dist)
        # expected: periodic global field sampled on the block + ghosts
        ix = [(np.arange(-h, n[d] + h) + o[d]) % N[d] for d in range(3)]
        exp = G[np.ix_(ix[2], ix[1], ix[0])]
        exps = Gs[np.ix_(ix[2], ix[1], ix[0])]
        ok = bool(np.array_equal(a.numpy(), exp) and np.array_equal(b.numpy(), exps))
        q.put((rank, ok))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("tiles", [(1, 1, 2), (2, 1, 1), (1, 2, 1)])
def test_exchange_world2_gloo(tiles):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() + sum(i * t for i, t in enumerate(tiles))) % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, tiles, (6, 5, 4), port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert all(ok for _, ok in res), res


def test_exchange_single_rank_wraps_locally():
    n = (5, 4, 6); h = halo.HALO
    G = _global_field(n, 2)
    a = torch.zeros((n[2] + 2 * h, n[1] + 2 * h, n[0] + 2 * h, 2), dtype=torch.float64)
    a[h:-h, h:-h, h:-h] = torch.from_numpy(G)
    halo.exchange([a], n, 0, (1, 1, 1), None)
    ix = [np.arange(-h, n[d] + h) % n[d] for d in range(3)]
    assert np.array_equal(a.numpy(), G[np.ix_(ix[2], ix[1], ix[0])])


def test_neighbours_and_tile_coords_roundtrip():
    tiles = (2, 2, 2)
    for r in range(8):
        assert halo.rank_of(halo.tile_coords(r, tiles), tiles) == r
    nb = halo.neighbours(0, tiles)
    assert nb == [(1, 1), (2, 2), (4, 4)]
    assert halo.neighbours(5, (1, 2, 4)) == [(5, 5), (4, 4), (3, 7)]


def _count(boxes, shape, origin):
    c = np.zeros(shape, dtype=np.int32)
    for lo, hi in boxes:
        if any(l > u for l, u in zip(lo, hi)):
            continue
        sl = tuple(slice(lo[d] - origin[d], hi[d] - origin[d] + 1) for d in (2, 1, 0))
        c[sl] += 1
    return c


@pytest.mark.parametrize("n", [(16, 12, 10), (7, 6, 8), (6, 6, 6)])
def test_boundary_boxes_tile_the_block_exactly_once(n):
    origin = (32, 0, 16)
    inner, shells = halo.boundary_rects(origin, n)
    c = _count([inner] + shells, (n[2], n[1], n[0]), origin)
    assert c.min() == 1 and c.max() == 1


def test_ghost_boxes_cover_the_ghost_region_exactly_once():
    n = (8, 6, 5); origin = (8, 12, 0); h = halo.HALO
    boxes = halo.shell_rects(origin, n)
    full_origin = tuple(o - h for o in origin)
    c = _count(boxes, (n[2] + 2 * h, n[1] + 2 * h, n[0] + 2 * h), full_origin)
    interior = np.zeros_like(c, dtype=bool)
    interior[h:-h, h:-h, h:-h] = True
    assert np.all(c[interior] == 0) and np.all(c[~interior] == 1)


@pytest.mark.parametrize("n", [(96, 40, 33), (64, 64, 64), (20, 18, 16)])
def test_tile_thick_shells_tile_the_block_exactly_once(n):
    """The shell widths the exchanger uses (one staged tile thick: SHELL_W, clipped for small blocks)."""
    origin = (0, 64, 8)
    w = tuple(min(halo.SHELL_W[d], max(halo.HALO, n[d] // 4)) for d in range(3))
    assert min(w) >= halo.HALO
    inner, shells = halo.boundary_rects(origin, n, w)
    c = _count([inner] + shells, (n[2], n[1], n[0]), origin)
    assert c.min() == 1 and c.max() == 1
    # the inner box keeps the stencil reach to every block face
    for d in range(3):
        assert inner[0][d] - origin[d] >= halo.HALO and origin[d] + n[d] - 1 - inner[1][d] >= halo.HALO


def test_rk_coefficients_match_the_reference():
    """RK_C of src/prometeo_const.rg:54-58 (SSP-RK3): each stage's state weights sum to 1."""
    from htr_solver_b200.core import RK_C
    assert RK_C == ((1.0, 0.0, 1.0), (0.75, 0.25, 0.25), (1.0/3.0, 2.0/3.0, 2.0/3.0))
    for a, b, _ in RK_C:
        assert abs(a + b - 1.0) < 1e-15


def test_exchange_policy_by_remote_directions():
    """HaloExchanger overlaps the exchange with the inner box from two remote directions on; with one (or none) the exchange
    runs first and the block is one htr_rhs call (htr-solver_b200/halo.py: overlap_default).  HTR_OVERLAP forces either."""
    assert halo.overlap_default((1, 1, 1), env="") is False
    assert halo.overlap_default((1, 1, 2), env="") is False
    assert halo.overlap_default((1, 2, 2), env="") is True
    assert halo.overlap_default((2, 2, 2), env="") is True
    assert halo.overlap_default((1, 1, 2), env="1") is True
    assert halo.overlap_default((2, 2, 2), env="0") is False
