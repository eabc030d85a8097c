"""Multi-GPU plumbing of the rendering path: one process per GPU, torch.distributed for rendezvous and
the exchange steps.  The reference only replicates the launch on every device
(programs/LaunchParams.h:31-32 deviceIndex/deviceCount are never read), so all of this is new.

Sort-last ("direct send"): every rank renders its kd sub-volume to a premultiplied float4 partial of the
whole image; rank k then composites pixel span k of ALL partials in visibility order and finishes those
pixels (exa_renderer_composite_partials).  Two exchange paths:
  * peer memory (default on one NVSwitch box): the partial buffers are CUDA-IPC mapped into every rank
    and the composite kernel reads them directly over NVLink - the exchange is fused into the kernel;
  * exchange_spans(): an all-to-all of spans over the process group (NCCL / gloo), used when peer
    mapping is unavailable and by the CPU tests.
Image tiles: the scene is replicated, rank r renders owned_tiles(r); the frame lives as N row strips,
strip k on GPU k, written by every rank's frame kernel through peer pointers and read back by rank k
over its own PCIe link.
"""
import ctypes as C


def span_of(rank, world, npix):
    """pixel span [lo, hi) that `rank` composites / owns"""
    return (npix * rank) // world, (npix * (rank + 1)) // world


def strip_rows(rank, world, height, tile_h=8):
    """row strip [y0, y1) of the frame that lives on GPU `rank` (image-tile mode): whole tile rows"""
    rows = (height + tile_h - 1) // tile_h
    return min(height, tile_h * ((rows * rank) // world)), min(height, tile_h * ((rows * (rank + 1)) // world))


def weighted_strip_rows(weights, height, tile_h=8):
    """first pixel row of every strip (len(weights) + 1 entries) when strip k gets a share of the tile rows proportional
    to weights[k] (a GPU's measured device-to-host bandwidth): largest-remainder rounding, deterministic, every rank
    computes the same table.  Equal weights give strip_rows()."""
    n = len(weights)
    rows = (height + tile_h - 1) // tile_h
    w = [max(float(x), 0.0) for x in weights]
    tot = sum(w)
    if tot <= 0 or max(w) - min(w) <= 1e-6 * tot:                 # equal: the kernel's default split, exactly
        return [strip_rows(k, n, height, tile_h)[0] for k in range(n)] + [height]
    exact = [rows * x / tot for x in w]
    base = [int(e) for e in exact]
    rest = rows - sum(base)
    for k in sorted(range(n), key=lambda k: (-(exact[k] - base[k]), k))[:rest]:
        base[k] += 1
    out, acc = [0], 0
    for k in range(n):
        acc += base[k]
        out.append(min(height, tile_h * acc))
    out[-1] = height
    return out


def exchange_spans(dist, torch, partial, rank, world):
    """partial: (npix, C) tensor of this rank.  Returns a list of `world` tensors: every rank's partial
    restricted to THIS rank's span (index = source rank)."""
    npix = partial.shape[0]
    lo, hi = span_of(rank, world, npix)
    out = [None] * world
    ops = []
    for k in range(world):
        if k == rank:
            out[k] = partial[lo:hi]
            continue
        klo, khi = span_of(k, world, npix)
        out[k] = torch.empty((hi - lo,) + tuple(partial.shape[1:]), dtype=partial.dtype, device=partial.device)
        ops.append(dist.P2POp(dist.isend, partial[klo:khi].contiguous(), k))
        ops.append(dist.P2POp(dist.irecv, out[k], k))
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()
    return out


def composite_reference(np, parts, order, ert=2.0):
    """numpy restatement of render.cu:k_compositeFinish's compositing loop (tests): parts[k] = (n, 4) premultiplied
    partial of rank k, alpha + 2 marks a ray that met the termination threshold inside that rank"""
    n = parts[0].shape[0]
    pc = np.zeros((n, 4), np.float32)
    done = np.zeros(n, bool)
    for k in order:
        a = parts[k].astype(np.float32).copy()
        term = a[:, 3] >= 2.0
        a[term, 3] -= 2.0
        kk = (1.0 - pc[:, 3:4]).astype(np.float32)
        new = pc + kk * a
        pc = np.where(done[:, None], pc, new)
        stop = ~done & (term | (pc[:, 3] >= ert))
        pc[stop] = np.concatenate([pc[stop, :3] * pc[stop, 3:4], np.ones((int(stop.sum()), 1), np.float32)], axis=1)
        done |= stop
    return pc


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


class PeerBuffers:
    """`count` device buffers of `nbytes` per rank, each its own allocation (exa_device_malloc), CUDA-IPC
    mapped into every other rank of the node.  ptr[k][i] = device pointer, valid in THIS process, of buffer i of
    rank k.  Raises if a mapping fails (caller falls back to the collective path)."""

    def __init__(self, dist, torch, capi, rank, world, device, nbytes, count):
        self.capi, self.rank, self.world = capi, rank, world
        self.ptr = [[None] * count for _ in range(world)]
        mine = torch.zeros((count, 64), dtype=torch.uint8)
        for i in range(count):
            p = C.c_void_p()
            capi.check(capi.lib.exa_device_malloc(nbytes, device, C.byref(p)))
            self.ptr[rank][i] = p.value
            capi.check(capi.lib.exa_ipc_export(C.c_void_p(p.value), C.c_void_p(mine[i].data_ptr())))
        allh = [torch.zeros((count, 64), dtype=torch.uint8, device="cuda") for _ in range(world)]
        dist.all_gather(allh, mine.cuda())
        ok = 1
        for k in range(world):
            if k == rank:
                continue
            h = allh[k].cpu()
            for i in range(count):
                p = C.c_void_p()
                if capi.lib.exa_ipc_import(C.c_void_p(h[i].data_ptr()), device, C.byref(p)) != 0:
                    ok = 0
                    break
                self.ptr[k][i] = p.value
        flag = torch.tensor([ok], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        self.ok = bool(flag.item())

    def close(self, dist):
        dist.barrier()
        for k in range(self.world):
            if k != self.rank:
                for p in self.ptr[k]:
                    if p:
                        self.capi.lib.exa_ipc_close(C.c_void_p(p))
        dist.barrier()
        for p in self.ptr[self.rank]:
            if p:
                self.capi.lib.exa_device_free(C.c_void_p(p))
        self.ptr = None


class PeerBarrier:
    """stream-ordered barrier over peer memory (exa_peer_barrier); nccl=True: a 4-byte all-reduce instead"""

    def __init__(self, dist, torch, capi, rank, world, device, nccl=False):
        self.dist, self.torch, self.capi, self.rank, self.world, self.nccl = dist, torch, capi, rank, world, nccl
        self.seq = 0
        self.flag = torch.zeros(1, dtype=torch.int32, device="cuda")
        self.bufs = None
        if not nccl:
            self.bufs = PeerBuffers(dist, torch, capi, rank, world, device, 128, 1)
            if not self.bufs.ok:
                self.nccl = True
        if not self.nccl:
            self.ptrs = (C.c_void_p * world)(*[self.bufs.ptr[k][0] for k in range(world)])

    def wait(self, torch_stream):
        """enqueue the barrier on `torch_stream`"""
        if self.nccl:
            with self.torch.cuda.stream(torch_stream):
                self.dist.all_reduce(self.flag)
            return
        self.seq += 1
        self.capi.check(self.capi.lib.exa_peer_barrier(self.ptrs, self.rank, self.world, self.seq, C.c_void_p(torch_stream.cuda_stream)))

    def timed_out(self):
        if self.nccl or self.bufs is None:
            return False
        w = (C.c_uint32 * 1)()
        self.capi.check(self.capi.lib.exa_copy_async(w, C.c_void_p(self.bufs.ptr[self.rank][0] + 64), 4, None))
        self.torch.cuda.synchronize()
        return bool(w[0])

    def close(self):
        if self.bufs is not None:
            self.bufs.close(self.dist)
            self.bufs = None


class HostFrames:
    """`count` RGBA8 host frames shared by the ranks of the node: one POSIX shared-memory file mapped and page-locked
    (exa_host_register) by every rank, so that each rank copies ITS strip / span of a frame over its own PCIe link and
    the consumer (rank 0) finds the whole frame in ordinary scan-line layout."""

    def __init__(self, dist, torch, capi, np, rank, nbytes, count, tag):
        import mmap
        import os
        self.capi, self.rank = capi, rank
        self.path = "/dev/shm/exa_hostframes_%s" % tag
        total = nbytes * count
        if rank == 0:
            with open(self.path, "wb") as f:
                f.truncate(total)
        dist.barrier()
        self.f = open(self.path, "r+b")
        self.mm = mmap.mmap(self.f.fileno(), total)
        self.arr = np.frombuffer(self.mm, dtype=np.uint8).reshape(count, nbytes)
        self.base = self.arr.ctypes.data
        self.nbytes = nbytes
        capi.check(capi.lib.exa_host_register(C.c_void_p(self.base), total))
        dist.barrier()
        if rank == 0:
            os.unlink(self.path)                                   # the mappings keep it alive

    def ptr(self, i, offset=0):
        return self.base + i * self.nbytes + offset

    def close(self):
        self.capi.lib.exa_host_unregister(C.c_void_p(self.base))
        self.arr = None
        try:
            self.mm.close()
        except BufferError:
            pass
        self.f.close()


class _Pipeline:
    """Common part of the two multi-GPU frame pipelines: triple-buffered peer buffers, one barrier per frame on
    a side stream.  Frame j uses buffer j % 3 and may start once the barrier of frame j-2 has completed on this
    rank: by then every rank has finished reading (and copying out) what frame j-3 left in that buffer, because
    those reads precede the rank's barrier j-2 in stream order."""

    def __init__(self, dist, torch, capi, np, r, rank, world, device, W, H, stream, nccl_barrier, tag):
        self.dist, self.torch, self.capi, self.np, self.r = dist, torch, capi, np, r
        self.rank, self.world, self.device, self.W, self.H, self.stream = rank, world, device, W, H, stream
        self.bar = PeerBarrier(dist, torch, capi, rank, world, device, nccl=nccl_barrier)
        self.host = HostFrames(dist, torch, capi, np, rank, W * H * 4, 3, tag)
        self.side = torch.cuda.Stream()
        self.ev_render = [torch.cuda.Event() for _ in range(3)]
        self.ev_bar = [torch.cuda.Event() for _ in range(3)]
        self.j = 0

    def _begin(self, s=None):
        j = self.j
        self.j += 1
        if j >= 2:
            (s or self.stream).wait_event(self.ev_bar[(j - 2) % 3])
        return j % 3

    def _exchange(self, b, s=None):
        self.ev_render[b].record(s or self.stream)
        self.side.wait_event(self.ev_render[b])
        self.bar.wait(self.side)
        self.ev_bar[b].record(self.side)

    def drain(self):
        """the main stream waits for every barrier still in flight (end of a device-timed region)"""
        for e in self.ev_bar:
            self.stream.wait_event(e)

    def finish(self):
        """host waits for the side stream: the last frames' copies have landed in the shared host frames"""
        self.side.synchronize()

    def host_frame(self, b):
        return self.host.arr[b].view(self.np.uint32).reshape(self.H, self.W)

    def close(self):
        self.torch.cuda.synchronize()
        self.dist.barrier()
        self._close_buffers()
        self.bar.close()
        self.host.close()


class TilePipeline(_Pipeline):
    """Image tiles over a replicated scene: the RGBA8 frame is N row strips, strip k in GPU k's memory and peer-mapped
    everywhere; every rank's frame kernel stores its tiles into the strips (the gather is fused into the kernel, over
    NVLink), one peer-memory barrier says "frame complete", then every rank copies ITS strip to the shared pinned host
    frame over its own PCIe link."""

    def __init__(self, dist, torch, capi, np, r, rank, world, device, W, H, stream, nccl_barrier=False, tag="t", weighted=True):
        super().__init__(dist, torch, capi, np, r, rank, world, device, W, H, stream, nccl_barrier, tag)
        r.set_partition(capi.PARTITION_IMAGE_TILES, rank, world)
        # strips sized by what each GPU's link to host memory delivers while ALL ranks copy (GPUs behind a shared PCIe
        # uplink get less): the N copies of a frame then finish together
        self.copy_gbs = self._measure_copy_bandwidth() if (weighted and world > 1) else [1.0] * world
        self.rows = weighted_strip_rows(self.copy_gbs, H)
        self.y0, self.y1 = self.rows[rank], self.rows[rank + 1]
        rows = max(self.rows[k + 1] - self.rows[k] for k in range(world))
        self.strips = PeerBuffers(dist, torch, capi, rank, world, device, max(rows, 1) * W * 4, 3)
        if not self.strips.ok:
            raise RuntimeError("CUDA IPC mapping of the frame strips failed")
        self.tables = [(C.c_void_p * world)(*[self.strips.ptr[k][b] for k in range(world)]) for b in range(3)]
        self.row_table = (C.c_int * (world + 1))(*self.rows)
        self.d2h_bytes = (self.y1 - self.y0) * W * 4
        self.balanced, self.alt = False, None

    def _measure_copy_bandwidth(self, reps=8):
        """GB/s of this rank's device-to-host copy into the shared host frame while every rank copies its uniform strip"""
        torch, dist = self.torch, self.dist
        y0, y1 = strip_rows(self.rank, self.world, self.H)
        nbytes = max((y1 - y0) * self.W * 4, 4096)
        src = torch.zeros(nbytes, dtype=torch.uint8, device="cuda")
        dst = self.host.ptr(0, min(y0 * self.W * 4, max(self.W * self.H * 4 - nbytes, 0)))
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        def burst(n):
            for _ in range(n):
                self.capi.check(self.capi.lib.exa_copy_async(C.c_void_p(dst), C.c_void_p(src.data_ptr()), nbytes, C.c_void_p(self.side.cuda_stream)))
        burst(2); torch.cuda.synchronize(); dist.barrier()
        c0.record(self.side); burst(reps); c1.record(self.side); torch.cuda.synchronize()
        mine = torch.tensor([nbytes * reps / (c0.elapsed_time(c1) * 1e-3) / 1e9], dtype=torch.float64, device="cuda")
        allbw = [torch.zeros(1, dtype=torch.float64, device="cuda") for _ in range(self.world)]
        dist.all_gather(allbw, mine)
        bw = [float(x.item()) for x in allbw]
        # two classes of links, not eight noisy numbers: snap to 5 % of the fastest
        top = max(bw)
        return [round(x / top * 20) / 20 * top for x in bw]

    def step(self, readback=False, overlap=False):
        """overlap: consecutive frames go to two alternating streams, so that the head of frame j+1 (its most expensive
        tiles) fills the SMs that the tail of frame j (a few tiles of long rays) leaves idle; the frames write different
        strip buffers.  Only with an installed tile list (the per-frame cost-order feedback of the single-GPU schedule is a
        frame-to-frame dependency) and only while NO device state of the renderer changes between the frames (camera
        changes are kernel parameters and fine; transfer-function / iso / trace updates are not): call drain() first."""
        s = None
        if overlap and self.balanced:
            if self.alt is None:
                self.alt = self.torch.cuda.Stream()
            s = self.alt if (self.j & 1) else self.stream
            self.r.set_stream(s.cuda_stream)
        b = self._begin(s)
        self.capi.check(self.capi.lib.exa_renderer_set_color_strips_rows(self.r.h, self.tables[b], self.row_table, self.world))
        self.r.render(sync=False)                                  # own tiles -> the strips of all GPUs
        self._exchange(b, s)
        if s is not None:
            self.r.set_stream(self.stream.cuda_stream)
        if readback and self.d2h_bytes:
            self.capi.check(self.capi.lib.exa_copy_async(C.c_void_p(self.host.ptr(b, self.y0 * self.W * 4)), C.c_void_p(self.strips.ptr[self.rank][b]),
                                                         self.d2h_bytes, C.c_void_p(self.side.cuda_stream)))
        return b

    def rebalance(self, rounds=2):
        """Deal the tiles by measured cost: render a frame with per-tile cycle counters on, combine the ranks' counters
        (each tile is rendered by exactly one rank), run the deterministic LPT deal on every rank and install this rank's
        list.  The second round measures under the balanced deal.  A viewer calls this every few dozen frames (the cost
        is one frame plus a 130 KB all-reduce); bench.py calls it during warm-up.  Returns max / mean rank load."""
        np, torch, capi = self.np, self.torch, self.capi
        nt = ((self.W + 15) // 16) * ((self.H + 7) // 8)
        ratio = None
        for _ in range(rounds):
            capi.check(capi.lib.exa_renderer_enable_tile_costs(self.r.h, 1))
            self.step(); self.drain(); torch.cuda.synchronize()
            costs = np.zeros(nt, np.uint32)
            capi.check(capi.lib.exa_renderer_get_tile_costs(self.r.h, costs.ctypes.data_as(C.c_void_p), nt))
            capi.check(capi.lib.exa_renderer_enable_tile_costs(self.r.h, 0))
            t = torch.from_numpy(np.minimum(costs, 0x7fffffff).astype(np.int32)).cuda()
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
            costs = t.cpu().numpy().astype(np.uint32)
            mine, n = np.zeros(nt, np.uint32), C.c_int()
            capi.check(capi.lib.exa_balance_tiles(costs.ctypes.data_as(C.c_void_p), nt, self.world, self.rank, mine.ctypes.data_as(C.c_void_p), C.byref(n)))
            capi.check(capi.lib.exa_renderer_set_tile_list(self.r.h, mine.ctypes.data_as(C.c_void_p), n.value))
            self.balanced = True
            load = torch.tensor([float(costs[mine[:n.value]].astype(np.float64).sum())], dtype=torch.float64, device="cuda")
            loads = [torch.zeros(1, dtype=torch.float64, device="cuda") for _ in range(self.world)]
            self.dist.all_gather(loads, load)
            v = [float(x.item()) for x in loads]
            ratio = max(v) / max(sum(v) / len(v), 1e-9)
        return ratio

    def _close_buffers(self):
        self.capi.lib.exa_renderer_set_tile_list(self.r.h, None, 0)
        self.capi.lib.exa_renderer_set_color_strips(self.r.h, None, 0)
        self.strips.close(self.dist)


class SortLastPipeline(_Pipeline):
    """Sort-last over a data-partitioned scene (renderer created with part=(rank, world)): every rank renders its kd
    sub-volume to a float4 partial of the whole image in a peer-mapped buffer, one barrier, then rank k composites pixel
    span k reading ALL ranks' partials through the peer pointers (the exchange happens inside the composite kernel) and
    copies its span of RGBA8 to the shared host frame."""

    def __init__(self, dist, torch, capi, np, r, rank, world, device, W, H, stream, nccl_barrier=False, surf=False, tag="s"):
        super().__init__(dist, torch, capi, np, r, rank, world, device, W, H, stream, nccl_barrier, tag)
        npix = W * H
        self.lo, self.hi = span_of(rank, world, npix)
        self.partials = PeerBuffers(dist, torch, capi, rank, world, device, npix * 16, 3)
        self.surfs = PeerBuffers(dist, torch, capi, rank, world, device, npix * 16, 3) if surf else None
        if not self.partials.ok or (surf and not self.surfs.ok):
            raise RuntimeError("CUDA IPC mapping of the partial images failed")
        self.out = torch.zeros(npix, dtype=torch.int32, device="cuda")
        self.d2h_bytes = (self.hi - self.lo) * 4
        self.peer_read_bytes = (world - 1) * (self.hi - self.lo) * 16 * (2 if surf else 1)

    def step(self, readback=False):
        b = self._begin()
        r, rank = self.r, self.rank
        r.render_partial(self.partials.ptr[rank][b], self.surfs.ptr[rank][b] if self.surfs else None)
        self._exchange(b)
        order, _ = r.partition_order()                             # visibility order for the current camera (host, microseconds)
        r.set_stream(self.side.cuda_stream)
        r.composite_partials([self.partials.ptr[k][b] for k in order], [self.surfs.ptr[k][b] for k in order] if self.surfs else None,
                             self.lo, self.hi - self.lo, self.out.data_ptr())
        r.set_stream(self.stream.cuda_stream)
        if readback and self.d2h_bytes:
            self.capi.check(self.capi.lib.exa_copy_async(C.c_void_p(self.host.ptr(b, self.lo * 4)), C.c_void_p(self.out.data_ptr() + self.lo * 4),
                                                         self.d2h_bytes, C.c_void_p(self.side.cuda_stream)))
        return b

    def _close_buffers(self):
        self.partials.close(self.dist)
        if self.surfs:
            self.surfs.close(self.dist)
