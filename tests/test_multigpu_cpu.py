"""World-size-2 (and 4) gloo tests of the multi-GPU host logic on CPU: the binary-swap schedule of
owlexabrick_b200/multigpu.py with numpy "over"/finish callbacks must reproduce the sequential
front-to-back composite of all ranks' partial images; image-tile ownership must partition the
frame.  (The CUDA composite/finish kernels are covered by the -m gpu tests.)"""
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
        out.append(np.concatenate([rgb, a], axis=1))
    return out


def _over(front, back):
    return front + (1.0 - front[:, 3:4]) * back


def _front_order(world, front_bits):
    """depth order of the ranks (front first) when at tree level l the child with bit b == front_bits[l] is in front"""
    levels = world.bit_length() - 1

    def key(rank):
        # most significant rank bit = first kd level; round r pairs ranks differing in bit r
        return [0 if ((rank >> (levels - 1 - l)) & 1) == front_bits[l] else 1 for l in range(levels)]
    return sorted(range(world), key=key)


def _worker(rank, world, port, npix, front_bits, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    levels = world.bit_length() - 1
    mine = torch.from_numpy(_partials(world, npix)[rank].copy())

    def composite(a, b, front):
        res = _over(a.numpy(), b.numpy()) if front else _over(b.numpy(), a.numpy())
        a.copy_(torch.from_numpy(res))

    def i_am_front(rnd):
        lvl = levels - 1 - rnd                                       # partner differs in bit `rnd` = tree level lvl
        return ((rank >> rnd) & 1) == front_bits[lvl]

    def finish(pixels, first):
        return (pixels[:, 3] * 1000).to(torch.int32) + first * 0     # any per-pixel function; keep alpha*1000

    strips, span, calls = multigpu.binary_swap(dist, torch, mine, rank, world, composite, i_am_front, finish)
    lo, hi = span
    ret[rank] = (mine[lo:hi].numpy().copy(), span, torch.cat(strips).numpy().copy(), calls)
    dist.destroy_process_group()


@pytest.mark.parametrize("world,front_bits", [(2, [0]), (2, [1]), (4, [1, 0])])
def test_binary_swap_matches_sequential_composite(world, front_bits):
    npix = 1000
    mgr = mp.Manager()
    ret = mgr.dict()
    port = 29500 + (os.getpid() % 2000) + world
    mp.spawn(_worker, args=(world, port, npix, front_bits, ret), nprocs=world, join=True)
    parts = _partials(world, npix)
    ref = np.zeros((npix, 4), np.float32)
    for rk in _front_order(world, front_bits):                       # front to back
        ref = _over(ref, parts[rk])
    covered = np.zeros(npix, bool)
    for rank in range(world):
        strip, (lo, hi), gathered, calls = ret[rank]
        assert np.allclose(strip, ref[lo:hi], atol=1e-6)
        assert calls == world.bit_length()                           # log2(N) composites + 1 finish
        covered[lo:hi] = True
        assert len(gathered) == npix
    assert covered.all()
    # the gathered image is ordered by each rank's final span
    spans = sorted((ret[r][1], r) for r in range(world))
    assert [s for s, _ in spans][0][0] == 0 and [s for s, _ in spans][-1][1] == npix


def test_swap_schedule_and_tiles():
    for world in (1, 2, 4, 8):
        spans = [multigpu.final_span(r, world, 3840 * 2160) for r in range(world)]
        assert sorted(spans)[0][0] == 0 and sum(h - l for l, h in spans) == 3840 * 2160
        tiles = [set(multigpu.owned_tiles(r, world, 1024, 1024)) for r in range(world)]
        assert sum(len(t) for t in tiles) == 64 * 128 and len(set.union(*tiles)) == 64 * 128
    with pytest.raises(ValueError):
        multigpu.swap_schedule(0, 3, 100)
