"""World-size-2 / 3 / 4 gloo tests of the multi-GPU host logic on CPU: the direct-send exchange of
owlexabrick_b200/multigpu.py (every rank receives all ranks' partials of ITS pixel span) followed by the
ordered composite must reproduce the sequential front-to-back composite of all ranks' partial images,
including the early-termination marker; pixel spans, frame strips and image-tile ownership must partition
the frame.  (The CUDA composite kernel is checked against multigpu.composite_reference by the -m gpu tests.)"""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from owlexabrick_b200 import multigpu  # noqa: E402


def _partials(world, npix, seed=3):
    rng = np.random.RandomState(seed)
    out = []
    for _ in range(world):
        a = rng.uniform(0, 0.9, (npix, 1)).astype(np.float32)
        a[rng.uniform(size=(npix, 1)) < 0.3] = 0.0                  # empty pixels
        rgb = rng.uniform(0, 1, (npix, 3)).astype(np.float32) * a   # premultiplied
        term = (rng.uniform(size=(npix, 1)) < 0.1) & (a > 0.5)      # rays that met the termination threshold in this rank
        out.append(np.concatenate([rgb, a + 2.0 * term], axis=1).astype(np.float32))
    return out


def _sequential(parts, order):
    """the single-GPU ray: accumulate rank after rank, stop (and scale by alpha, exabrick.cu:1694-1697) at a marked one"""
    n = parts[0].shape[0]
    out = np.zeros((n, 4), np.float32)
    for i in range(n):
        pc = np.zeros(4, np.float32)
        for k in order:
            a = parts[k][i].copy()
            term = a[3] >= 2.0
            if term:
                a[3] -= 2.0
            pc = pc + (np.float32(1.0) - pc[3]) * a
            if term:
                pc = np.array([pc[0] * pc[3], pc[1] * pc[3], pc[2] * pc[3], 1.0], np.float32)
                break
        out[i] = pc
    return out


def _worker(rank, world, port, npix, order, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = torch.from_numpy(_partials(world, npix)[rank].copy())
    got = multigpu.exchange_spans(dist, torch, mine, rank, world)
    comp = multigpu.composite_reference(np, [g.numpy() for g in got], order)
    ret[rank] = (comp, multigpu.span_of(rank, world, npix))
    dist.destroy_process_group()


@pytest.mark.parametrize("world,order", [(2, [0, 1]), (2, [1, 0]), (3, [2, 0, 1]), (4, [1, 0, 3, 2])])
def test_direct_send_matches_sequential_composite(world, order):
    npix = 1001                                                      # not a multiple of the rank count
    mgr = mp.Manager()
    ret = mgr.dict()
    port = 29500 + (os.getpid() % 2000) + 7 * world + order[0]
    mp.spawn(_worker, args=(world, port, npix, order, ret), nprocs=world, join=True)
    ref = _sequential(_partials(world, npix), order)
    covered = np.zeros(npix, bool)
    for rank in range(world):
        comp, (lo, hi) = ret[rank]
        assert comp.shape[0] == hi - lo
        assert np.allclose(comp, ref[lo:hi], atol=1e-6)
        assert not covered[lo:hi].any()
        covered[lo:hi] = True
    assert covered.all()


def test_composite_reference_applies_threshold_between_parts():
    a = np.array([[0.5, 0.0, 0.0, 0.99]], np.float32)               # first part alone crosses 0.98 without a marker
    b = np.array([[0.0, 0.5, 0.0, 0.5]], np.float32)
    out = multigpu.composite_reference(np, [a, b], [0, 1], ert=0.98)
    assert np.allclose(out, [[0.5 * 0.99, 0, 0, 1.0]])
    out = multigpu.composite_reference(np, [a, b], [0, 1], ert=2.0)  # threshold off: plain over
    assert np.allclose(out, a + (1 - 0.99) * b)


def test_spans_strips_and_tiles_partition_the_frame():
    for world in (1, 2, 3, 4, 8):
        npix = 3840 * 2160
        spans = [multigpu.span_of(r, world, npix) for r in range(world)]
        assert spans[0][0] == 0 and spans[-1][1] == npix and all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
        for height in (2160, 1024, 116, 5):
            strips = [multigpu.strip_rows(r, world, height) for r in range(world)]
            assert strips[0][0] == 0 and strips[-1][1] == height and all(strips[i][1] == strips[i + 1][0] for i in range(world - 1))
            assert all(y0 % 8 == 0 for y0, _ in strips)
            rows = multigpu.weighted_strip_rows([1.0 + (r % 2) for r in range(world)], height)
            assert rows[0] == 0 and rows[-1] == height and all(a <= c for a, c in zip(rows, rows[1:])) and all(y % 8 == 0 or y == height for y in rows)
            assert multigpu.weighted_strip_rows([3.0] * world, height) == [s[0] for s in strips] + [height]
        tiles = [set(multigpu.owned_tiles(r, world, 1024, 1024)) for r in range(world)]
        assert sum(len(t) for t in tiles) == 64 * 128 and len(set.union(*tiles)) == 64 * 128


def test_cost_balanced_tile_deal_and_strip_rows_in_the_library():
    """exa_balance_tiles / exa_strip_rows are host-only entry points of the C ABI (no GPU needed): the LPT deal is a
    partition of the tiles, deterministic, balanced to within the heaviest tile, capped at ~1.25x a rank's share of the
    tile COUNT (zero-cost tails), and every rank's list is sorted by descending cost; the strips equal the Python mirror"""
    import ctypes as C
    from owlexabrick_b200 import build, capi
    build.build_lib()
    lib = capi.lib
    rng = np.random.RandomState(7)
    nt = 240 * 270                                                   # 4K frame
    costs = (rng.pareto(1.5, nt) * 2000).astype(np.uint32)           # heavy tail, like the frame's tile costs
    costs[rng.uniform(size=nt) < 0.4] = 0                            # tiles outside the scene
    for world in (1, 2, 3, 8, 16):
        lists = []
        for k in range(world):
            mine, n = np.zeros(nt, np.uint32), C.c_int()
            capi.check(lib.exa_balance_tiles(costs.ctypes.data_as(C.c_void_p), nt, world, k, mine.ctypes.data_as(C.c_void_p), C.byref(n)))
            again, n2 = np.zeros(nt, np.uint32), C.c_int()
            capi.check(lib.exa_balance_tiles(costs.ctypes.data_as(C.c_void_p), nt, world, k, again.ctypes.data_as(C.c_void_p), C.byref(n2)))
            assert n.value == n2.value and np.array_equal(mine[:n.value], again[:n.value])
            lists.append(mine[:n.value].copy())
        allt = np.concatenate(lists)
        assert len(allt) == nt and len(np.unique(allt)) == nt       # a partition of the frame's tiles
        loads = np.array([costs[l].astype(np.int64).sum() for l in lists], np.float64)
        assert loads.max() - loads.min() <= costs.max() + len(lists[0])
        share = (nt + world - 1) // world
        assert max(len(l) for l in lists) <= share + nt // (4 * world) + 1
        for l in lists:
            c = costs[l].astype(np.int64)
            assert (np.diff(c) <= 0).all()                           # most expensive first: the launch order of the rank
    assert lib.exa_balance_tiles(costs.ctypes.data_as(C.c_void_p), nt, 4, 4, costs.ctypes.data_as(C.c_void_p), C.byref(C.c_int())) == capi.EXA_ERR_INVALID
    for world in (1, 2, 3, 5, 8, 16):
        for height in (2160, 1080, 116, 9, 5):
            for k in range(world):
                y0, y1 = C.c_int(), C.c_int()
                capi.check(lib.exa_strip_rows(k, world, height, C.byref(y0), C.byref(y1)))
                assert (y0.value, y1.value) == multigpu.strip_rows(k, world, height)
    assert lib.exa_strip_rows(3, 3, 100, C.byref(C.c_int()), C.byref(C.c_int())) == capi.EXA_ERR_INVALID
