"""Synthetic structured-AMR inputs in the reference's "exajet cells" format.

cells   : int32[N,4] = {x, y, z, level}, level 0 = finest, cell width 1<<level
          (reference: builder/builder.cpp:120-134, README.md:76-82)
scalars : float32[N], same order (tools/artificial/artificial.cpp:177-178)

`grids_to_cells` restates tools/artificial/artificial.cpp:58-88,131-165 (exaArtificial) so
the reference's ex0..ex4.grids fixtures can be expanded anywhere; `nested_boxes` makes the
nested-refinement shapes SURVEY.md §8(d) names for configs C1/C2.
"""
import numpy as np

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(x):
    """vectorised splitmix64 finaliser over uint64 counters"""
    with np.errstate(over="ignore"):
        z = (x + np.uint64(0x9E3779B97F4A7C15)) & _M64
        z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M64
        z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M64
        return z ^ (z >> np.uint64(31))


def uniform01(n, seed):
    """n floats in [0,1) from splitmix64(seed + i)"""
    with np.errstate(over="ignore"):
        ctr = np.arange(n, dtype=np.uint64) + np.uint64(seed)
    return (splitmix64(ctr) >> np.uint64(40)).astype(np.float32) * np.float32(1.0 / (1 << 24))


def random_points(n, lo, hi, seed=1):
    u = uniform01(3 * n, seed).reshape(n, 3)
    lo = np.asarray(lo, np.float32)
    hi = np.asarray(hi, np.float32)
    return (lo + u * (hi - lo)).astype(np.float32)


def grids_to_cells(text):
    """exaArtificial: one box per line 'minx miny minz nx ny nz level v0 [v1..v7]'"""
    cells, scalars = [], []
    for line in text.splitlines():
        tok = line.split()
        if len(tok) not in (8, 15):
            continue
        mc = [int(t, 0) for t in tok[:3]]
        nx, ny, nz, level = (int(t, 0) for t in tok[3:7])
        v = [np.float32(t) for t in tok[7:]]
        if len(v) == 1:
            v = v * 8
        cs = 1 << level
        maxc = [mc[0] + (nx - 1) * cs, mc[1] + (ny - 1) * cs, mc[2] + (nz - 1) * cs]
        f32 = np.float32
        lerp = lambda a, b, x: (f32(1) - x) * a + x * b
        for cz in range(mc[2], maxc[2] + 1, cs):
            for cy in range(mc[1], maxc[1] + 1, cs):
                for cx in range(mc[0], maxc[0] + 1, cs):
                    x = f32(cx - mc[0]) / (f32(maxc[0]) - f32(mc[0]) + f32(1))
                    y = f32(cy - mc[1]) / (f32(maxc[1]) - f32(mc[1]) + f32(1))
                    z = f32(cz - mc[2]) / (f32(maxc[2]) - f32(mc[2]) + f32(1))
                    cells.append((cx, cy, cz, level))
                    scalars.append(lerp(lerp(lerp(v[0], v[1], x), lerp(v[2], v[3], x), y),
                                        lerp(lerp(v[4], v[5], x), lerp(v[6], v[7], x), y), z))
    return np.array(cells, np.int32).reshape(-1, 4), np.array(scalars, np.float32)


def _box_cells(lo, hi, level, hole=None):
    """all level-`level` cells tiling [lo,hi)^3 (3-vectors) except those inside `hole`"""
    cw = 1 << level
    ax = [np.arange(lo[d], hi[d], cw, dtype=np.int32) for d in range(3)]
    z, y, x = np.meshgrid(ax[2], ax[1], ax[0], indexing="ij")
    x, y, z = x.ravel(), y.ravel(), z.ravel()
    if hole is not None:
        hl, hh = hole
        inside = ((x >= hl[0]) & (x < hh[0]) & (y >= hl[1]) & (y < hh[1]) & (z >= hl[2]) & (z < hh[2]))
        x, y, z = x[~inside], y[~inside], z[~inside]
    out = np.empty((len(x), 4), np.int32)
    out[:, 0], out[:, 1], out[:, 2], out[:, 3] = x, y, z, level
    return out


def nested_boxes(domain, boxes):
    """Nested refinement: boxes[k] = (lo, hi) (ints or 3-vectors) is the region refined to
    level <= k; level len(boxes) fills the rest of [0,domain)^3.  E.g. SURVEY C1:
    nested_boxes(256, [(72,184), (32,224)]) -> L0 in [72,184)^3, L1 shell to [32,224)^3,
    L2 elsewhere (2 265 600 cells)."""
    def v3(a):
        return np.array([a, a, a] if np.isscalar(a) else a, np.int64)
    dom_lo, dom_hi = v3(0), v3(domain)
    parts, hole = [], None
    for k, (lo, hi) in enumerate(boxes):
        lo, hi = v3(lo), v3(hi)
        parts.append(_box_cells(lo, hi, k, hole))
        hole = (lo, hi)
    parts.append(_box_cells(dom_lo, dom_hi, len(boxes), hole))
    return np.concatenate(parts)


def analytic_field(cells, domain, channel=0, noise=0.05):
    """s = 0.5 + 0.5 sin(2pi x/Lx) cos(2pi y/Ly) sin(2pi z/Lz + phi_c) + noise*U(-1,1)
    evaluated at cell centres; noise from splitmix64 seeded 0xE8AB51C4 + channel."""
    cells = np.asarray(cells)
    L = np.array([domain] * 3 if np.isscalar(domain) else domain, np.float64)
    cw = (1 << cells[:, 3]).astype(np.float64)
    c = cells[:, :3].astype(np.float64) + 0.5 * cw[:, None]
    phi = 0.7 * channel
    s = 0.5 + 0.5 * np.sin(2 * np.pi * c[:, 0] / L[0]) * np.cos(2 * np.pi * c[:, 1] / L[1]) * np.sin(2 * np.pi * c[:, 2] / L[2] + phi)
    u = uniform01(len(cells), 0xE8AB51C4 + channel).astype(np.float64)
    return (s + noise * (2.0 * u - 1.0)).astype(np.float32)


def analytic_field_chunked(cells, domain, channel=0, noise=0.05, chunk=1 << 24, first=0):
    """analytic_field for arrays too large for its float64 temporaries (same values: the noise
    counter is the global cell index; `first` = global index of cells[0] when the caller
    passes a slice)"""
    out = np.empty(len(cells), np.float32)
    L = np.array([domain] * 3 if np.isscalar(domain) else domain, np.float64)
    phi = 0.7 * channel
    for b in range(0, len(cells), chunk):
        c4 = cells[b:b + chunk]
        cw = (1 << c4[:, 3]).astype(np.float64)
        c = c4[:, :3].astype(np.float64) + 0.5 * cw[:, None]
        s = 0.5 + 0.5 * np.sin(2 * np.pi * c[:, 0] / L[0]) * np.cos(2 * np.pi * c[:, 1] / L[1]) * np.sin(2 * np.pi * c[:, 2] / L[2] + phi)
        with np.errstate(over="ignore"):
            ctr = np.arange(first + b, first + b + len(c4), dtype=np.uint64) + np.uint64(0xE8AB51C4 + channel)
        u = ((splitmix64(ctr) >> np.uint64(40)).astype(np.float32) * np.float32(1.0 / (1 << 24))).astype(np.float64)
        out[b:b + chunk] = (s + noise * (2.0 * u - 1.0)).astype(np.float32)
    return out


# exajet-shaped, 4 levels, elongated 4:1:2 domain, one long L0/L1 "fuselage"
# box plus off-axis refinement boxes ("wings", "tail", "wake").  Boxes are (lo, hi, level) in
# finest units of the scale-1 domain 4096 x 1024 x 2048, aligned to the coarsest cell (8);
# a cell's level is the finest level of any box that contains it, else 3.
EXAJET_DOMAIN = (4096, 1024, 2048)
EXAJET_BOXES = [
    # level 0: fuselage, wing, tail plane, fin
    ((768, 352, 800), (3328, 672, 1248), 0),
    ((1664, 448, 192), (2368, 576, 1856), 0),
    ((2944, 480, 576), (3264, 544, 1472), 0),
    ((2880, 640, 992), (3328, 896, 1056), 0),
    # level 1: shells around those + the wake
    ((640, 320, 704), (3456, 704, 1344), 1),
    ((1536, 384, 128), (2496, 640, 1920), 1),
    ((2816, 416, 448), (3392, 960, 1600), 1),
    ((3328, 320, 704), (4032, 704, 1344), 1),
    # level 2: far shells
    ((384, 192, 448), (4096, 896, 1600), 2),
    ((1280, 256, 0), (2752, 768, 2048), 2),
]


def exajet_shaped(scale=1):
    """-> (cells int32[N,4], domain 3-vector).  scale=1: 4096x1024x2048 finest units, ~650 M cells
    (SURVEY C3); scale=k divides every box coordinate by k (all box coordinates are multiples of 32,
    so k = 1, 2, 4 keep them on the coarse grid), cells ~ 650 M / k^3."""
    dom = np.array(EXAJET_DOMAIN, np.int64) // scale
    cg = dom // 8                                                    # coarse (level-3) grid
    lvl = np.full((cg[2], cg[1], cg[0]), 3, np.int8)
    for lo, hi, l in EXAJET_BOXES:
        lo = np.array(lo, np.int64) // scale // 8
        hi = np.array(hi, np.int64) // scale // 8
        sub = lvl[lo[2]:hi[2], lo[1]:hi[1], lo[0]:hi[0]]
        np.minimum(sub, l, out=sub)
    counts = [int((lvl == l).sum()) for l in range(4)]
    total = sum(counts[l] * (8 >> l) ** 3 for l in range(4))
    out = np.empty((total, 4), np.int32)
    pos = 0
    for l in range(4):
        z, y, x = np.nonzero(lvl == l)
        m = 8 >> l                                                   # cells per coarse cell per axis
        cw = 1 << l
        o = np.arange(m, dtype=np.int32) * cw
        oz, oy, ox = (a.ravel() for a in np.meshgrid(o, o, o, indexing="ij"))
        step = max(1, (1 << 22) // (m ** 3))
        for b in range(0, len(x), step):
            n = len(x[b:b + step]) * m ** 3
            blk = out[pos:pos + n].reshape(-1, m ** 3, 4)
            blk[:, :, 0] = (x[b:b + step, None] * 8).astype(np.int32) + ox[None, :]
            blk[:, :, 1] = (y[b:b + step, None] * 8).astype(np.int32) + oy[None, :]
            blk[:, :, 2] = (z[b:b + step, None] * 8).astype(np.int32) + oz[None, :]
            blk[:, :, 3] = l
            pos += n
    assert pos == total
    return out, dom


def linear_field(cells, coef=(0.01, 0.02, 0.03), offset=0.1):
    cells = np.asarray(cells)
    cw = (1 << cells[:, 3]).astype(np.float64)
    c = cells[:, :3].astype(np.float64) + 0.5 * cw[:, None]
    return (offset + c @ np.asarray(coef, np.float64)).astype(np.float32)


# molecular-cloud-shaped — clumpy octree-like refinement from a seeded
# multi-octave noise threshold, 4 levels, 3 channels.  Block-structured like the FLASH data the
# reference shows (tools/fromFlash): an octree whose nodes each hold CLOUD_BLOCK^3 cells.  The
# coarsest nodes (level 3, 8*8 = 64 finest units wide) tile the domain; a node of level l > 0 whose
# centre density exceeds CLOUD_THRESHOLD[l] is replaced by its 8 children (level l-1), which are
# tested in turn.  Density = 4 octaves (3, 6, 12, 24 cycles per domain) of 3 plane waves each,
# directions/phases from splitmix64(CLOUD_SEED).
CLOUD_DOMAIN = 2048
CLOUD_BLOCK = 8
CLOUD_SEED = 0xC10D5EED
CLOUD_THRESHOLD = {3: 0.50, 2: 0.55, 1: 0.60}


def _cloud_waves():
    u = uniform01(4 * 3 * 4, CLOUD_SEED).astype(np.float64).reshape(4, 3, 4)
    waves = []
    for o in range(4):
        f = 3.0 * (1 << o)
        for k in range(3):
            z = 2.0 * u[o, k, 0] - 1.0
            phi = 2.0 * np.pi * u[o, k, 1]
            r = np.sqrt(max(0.0, 1.0 - z * z))
            d = np.array([r * np.cos(phi), r * np.sin(phi), z])
            waves.append((2.0 * np.pi * f * d, 2.0 * np.pi * u[o, k, 2], 0.5 ** o / 3.0))
    return waves


def cloud_density(centres, domain):
    """density in [0,1] at points given in finest units of a cubic domain of that width"""
    p = np.asarray(centres, np.float64) / float(domain)
    s = np.zeros(len(p))
    for kvec, phase, amp in _cloud_waves():
        s += amp * np.sin(p @ kvec + phase)
    return 0.5 + 0.5 * s / sum(0.5 ** o for o in range(4))


def cloud_nodes(scale=1):
    """-> per level l an int32[n,3] array of the lower corners of the octree nodes that stay at
    that level (each is CLOUD_BLOCK^3 cells of width 1<<l)"""
    dom = CLOUD_DOMAIN // scale
    top = CLOUD_BLOCK * 8                                           # width of a level-3 node
    g = dom // top
    assert g >= 1 and g * top == dom, "scale must divide %d" % (CLOUD_DOMAIN // top)
    child = np.array([[i & 1, (i >> 1) & 1, (i >> 2) & 1] for i in range(8)], np.int32)
    zz, yy, xx = np.meshgrid(np.arange(g, dtype=np.int32), np.arange(g, dtype=np.int32), np.arange(g, dtype=np.int32), indexing="ij")
    cur = np.stack([xx.ravel(), yy.ravel(), zz.ravel()], axis=1) * top
    nodes = {}
    for l in (3, 2, 1):
        w = CLOUD_BLOCK << l
        refine = cloud_density(cur + 0.5 * w, dom) > CLOUD_THRESHOLD[l]
        nodes[l] = cur[~refine]
        cur = (cur[refine][:, None, :] + child[None, :, :] * (w // 2)).reshape(-1, 3)
    nodes[0] = cur
    return nodes


def cloud_shaped_counts(scale=1):
    """cells per level without materialising them"""
    n = cloud_nodes(scale)
    return [len(n[l]) * CLOUD_BLOCK ** 3 for l in range(4)]


def cloud_shaped(scale=8):
    """-> (cells int32[N,4], domain).  scale=1: 2048^3 finest units (~2 B cells: 32 GB of
    cells — generate only where that fits); scale=k (a power of two <= 32): domain 2048/k, ~2 B / k^3 cells."""
    nodes = cloud_nodes(scale)
    B = CLOUD_BLOCK
    total = sum(len(nodes[l]) for l in range(4)) * B ** 3
    out = np.empty((total, 4), np.int32)
    o = np.arange(B, dtype=np.int32)
    oz, oy, ox = (a.ravel() for a in np.meshgrid(o, o, o, indexing="ij"))
    pos = 0
    for l in range(4):
        lo = nodes[l]
        for b in range(0, len(lo), 1 << 13):
            blk_lo = lo[b:b + (1 << 13)]
            n = len(blk_lo) * B ** 3
            blk = out[pos:pos + n].reshape(-1, B ** 3, 4)
            blk[:, :, 0] = blk_lo[:, None, 0] + (ox << l)[None, :]
            blk[:, :, 1] = blk_lo[:, None, 1] + (oy << l)[None, :]
            blk[:, :, 2] = blk_lo[:, None, 2] + (oz << l)[None, :]
            blk[:, :, 3] = l
            pos += n
    assert pos == total
    return out, CLOUD_DOMAIN // scale


def cloud_velocity(cells, domain, chunk=1 << 22):
    """3 smooth channels (vx, vy, vz) for the multichannel DVR + streamline config: a swirl around
    the z axis through the domain centre plus an upward drift, modulated by the §8(d) analytic
    field (no noise); values in about [-1.3, 1.3].  Evaluated in chunks (float64 temporaries)."""
    cells = np.asarray(cells)
    out = [np.empty(len(cells), np.float32) for _ in range(3)]
    for b in range(0, len(cells), chunk):
        c4 = cells[b:b + chunk]
        cw = (1 << c4[:, 3]).astype(np.float64)
        c = (c4[:, :3].astype(np.float64) + 0.5 * cw[:, None]) / float(domain)
        a = 0.5 + 0.5 * np.sin(2 * np.pi * c[:, 0]) * np.cos(2 * np.pi * c[:, 1]) * np.sin(2 * np.pi * c[:, 2])
        m = 0.75 + 0.5 * (a.astype(np.float32).astype(np.float64) - 0.5)
        c -= 0.5
        out[0][b:b + chunk] = ((-2.0 * c[:, 1] + 0.1) * m).astype(np.float32)
        out[1][b:b + chunk] = ((2.0 * c[:, 0]) * m).astype(np.float32)
        out[2][b:b + chunk] = ((0.3 + 0.4 * c[:, 2]) * m).astype(np.float32)
    return out
