"""Multi-GPU plumbing of the rendering path: one process per GPU, torch.distributed for the one
real exchange step.  The reference only replicates the launch on every device
(programs/LaunchParams.h:31-32 deviceIndex/deviceCount are never read), so this is new.

binary_swap   sort-last compositing of premultiplied float4 partial images: log2(N) rounds, in
              round r rank k exchanges half of its current pixel span with rank k ^ (1 << r) and
              composites ("over") the half it keeps; afterwards rank k owns a fully composited
              1/N strip, which `finish` turns into RGBA8 before the strips are all-gathered.
              The per-pixel operators are callbacks: the product passes the CUDA kernels behind
              exa_composite_over / exa_renderer_finish_pixels, the CPU tests pass numpy.
tile_owner    image-tile mode: tile t (16x8 pixels, scan order) belongs to rank t mod N.
"""


def swap_schedule(rank, world, npix):
    """[(partner, keep_span, send_span)] for each binary-swap round; spans are [lo, hi) pixel ranges"""
    if world & (world - 1):
        raise ValueError("binary swap needs a power-of-two rank count")
    lo, hi = 0, npix
    out = []
    for rnd in range(world.bit_length() - 1):
        mid = (lo + hi) // 2
        keep_low = ((rank >> rnd) & 1) == 0
        keep = (lo, mid) if keep_low else (mid, hi)
        send = (mid, hi) if keep_low else (lo, mid)
        out.append((rank ^ (1 << rnd), keep, send))
        lo, hi = keep
    return out


def final_span(rank, world, npix):
    sched = swap_schedule(rank, world, npix)
    return sched[-1][1] if sched else (0, npix)


def binary_swap(dist, torch, partial, rank, world, composite, i_am_front, finish):
    """partial: (npix, 4) float32 tensor of this rank (modified in place).
    composite(mine, other, mine_is_front): mine = over(front, back) in place.
    i_am_front(round) -> bool: is this rank's sub-volume in front of its partner's.
    finish(pixels, first) -> int32 tensor of packed RGBA8 for pixels [first, first+len).
    Returns (list of all ranks' strips, (lo, hi) of this rank, number of composite/finish calls)."""
    npix = partial.shape[0]
    calls = 0
    for rnd, (partner, keep, send) in enumerate(swap_schedule(rank, world, npix)):
        recv_buf = torch.empty((keep[1] - keep[0], 4), dtype=torch.float32, device=partial.device)
        ops = [dist.P2POp(dist.isend, partial[send[0]:send[1]], partner), dist.P2POp(dist.irecv, recv_buf, partner)]
        for req in dist.batch_isend_irecv(ops):
            req.wait()
        composite(partial[keep[0]:keep[1]], recv_buf, i_am_front(rnd))
        calls += 1
    lo, hi = final_span(rank, world, npix)
    strip = finish(partial[lo:hi], lo)
    calls += 1
    strips = [torch.empty_like(strip) for _ in range(world)]
    dist.all_gather(strips, strip)
    return strips, (lo, hi), calls


def tile_owner(tile, world):
    return tile % world


def owned_tiles(rank, world, width, height, tile_w=16, tile_h=8):
    """tiles of rank `rank` in image-tile mode (render.cu:ownedTile): blocks of `world` consecutive
    tiles, the deal rotating by one per block"""
    n = ((width + tile_w - 1) // tile_w) * ((height + tile_h - 1) // tile_h)
    out = []
    for j in range((n + world - 1) // world):
        t = j * world + (rank + j) % world
        if t < n:
            out.append(t)
    return out
