"""CPU tests of the oracle (oracle/): pinned against the reference's own outputs
(tests/golden/*.npz, made by tests/golden/make_golden.py from oracle/_ref) and against
analytic known answers for the parts the reference gives no vectors for."""
import hashlib
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN, load_golden
from owlexabrick_b200 import synth
import scenes

CASES = ["ex0", "ex1", "ex2", "ex3", "ex4", "nb64", "nb64_mlw8", "nb64_dups"]


@pytest.mark.parametrize("name", CASES)
def test_builder_matches_reference(orc, name):
    g = load_golden(name)
    mine = orc.build_bricks(g["cells"], int(g["max_leaf_width"]))
    assert mine == g["bricks"].tobytes()


@pytest.mark.parametrize("name", CASES)
def test_abr_matches_reference(orc, name):
    g = load_golden(name)
    bricks, ids = orc.parse_bricks(g["bricks"].tobytes())
    sc = orc.flatten_scalars(ids, list(g["fields"]))
    regions, leaf = orc.build_abr(bricks, sc)
    rr, rl = orc.parse_abr(g["abr"].tobytes())
    assert regions.tobytes() == rr.tobytes()
    assert leaf.tobytes() == rl.tobytes()


def test_ex4_known_structure(orc):
    # SURVEY §4 probe: ex4 -> 200 cells -> 3 bricks (4,8,4)@L0,(2,2,2)@L1,(4,4,4)@L0 -> 18 regions
    g = load_golden("ex4")
    bricks, _ = orc.parse_bricks(g["bricks"].tobytes())
    assert [tuple(b["size"]) + (int(b["level"]),) for b in bricks] == [(4, 8, 4, 0), (2, 2, 2, 1), (4, 4, 4, 0)]
    rr, rl = orc.parse_abr(g["abr"].tobytes())
    assert len(rr) == 18 and rr["leafListSize"].max() == 3


def test_c1_hash_pinned(orc):
    """config C1 (2.27 M cells): the reference's output is pinned by SHA-256"""
    hashes = json.load(open(os.path.join(GOLDEN, "golden_hashes.json")))
    cells = synth.nested_boxes(256, [(72, 184), (32, 224)])
    assert len(cells) == hashes["C1"]["cells"] == 2265600
    raw = orc.build_bricks(cells)
    assert hashlib.sha256(raw).hexdigest() == hashes["C1"]["bricks_sha256"]


def _scene(orc, cells, fields):
    raw = orc.build_bricks(cells)
    bricks, ids = orc.parse_bricks(raw)
    sc = orc.flatten_scalars(ids, fields)
    regions, leaf = orc.build_abr(bricks, sc)
    return orc.Scene(bricks, sc, regions, leaf), bricks


def test_single_brick_is_trilinear(orc):
    """KAT: inside one brick the tent basis equals trilinear interpolation of cell-centred
    data; a linear field is reproduced exactly (up to rounding)."""
    n = 8
    cells = synth.nested_boxes(n, [])            # one level-0 brick of 8^3
    f = synth.linear_field(cells, (0.25, -0.5, 0.125), 1.0)
    sc, bricks = _scene(orc, cells, [f])
    assert len(bricks) == 1
    pts = synth.random_points(2000, [0.5] * 3, [n - 0.5] * 3, seed=3)
    val, ok, st = sc.sample(pts)
    assert ok.all()
    expect = 1.0 + pts.astype(np.float64) @ np.array([0.25, -0.5, 0.125])
    assert np.abs(val - expect).max() < 1e-5
    assert st["samples"] == 2000 and st["brickVisits"] == 2000 and st["cornerLoads"] == 16000


def test_constant_field_and_fall_off(orc):
    """KAT: constant field -> constant wherever sum(w) > 0, also in the half-cell apron
    outside the brick; outside the basis domain -> invalid."""
    cells, _, dom = scenes.small_scene("nb32")
    sc, _ = _scene(orc, cells, [np.full(len(cells), 0.375, np.float32)])
    pts = synth.random_points(5000, [-1.9] * 3, [dom + 1.9] * 3, seed=5)
    val, ok, _ = sc.sample(pts)
    assert ok.sum() > 4000
    assert np.abs(val[ok == 1] - 0.375).max() < 1e-6
    out = np.array([[-2.5, 3, 3], [dom + 2.5, 3, 3], [3, 3, -100]], np.float32)
    _, ok2, _ = sc.sample(out)
    assert not ok2.any()


def test_gradient_direction_linear_field(orc):
    """KAT: the (unnormalised) analytic gradient of a linear field is parallel to its slope"""
    cells = synth.nested_boxes(8, [])
    coef = np.array([0.3, -0.2, 0.6])
    sc, _ = _scene(orc, cells, [synth.linear_field(cells, coef, 0.5)])
    pts = synth.random_points(500, [1.0] * 3, [7.0] * 3, seed=9)
    _, ok, g, _ = sc.sample(pts, grad=True)
    assert ok.all()
    gn = g / np.linalg.norm(g, axis=1, keepdims=True)
    assert np.abs(gn - coef / np.linalg.norm(coef)).max() < 1e-4


def test_region_boundary_continuity(orc):
    """KAT: reconstruction is continuous across region (ABR) boundaries"""
    cells, fields, dom = scenes.small_scene("nb32")
    sc, _ = _scene(orc, cells, fields)
    r = sc.regions
    faces = np.unique(r["hi"][:, 0])
    faces = faces[(faces > 2) & (faces < dom - 2)][:6]
    rng = np.random.RandomState(1)
    yz = rng.uniform(2, dom - 2, size=(200, 2)).astype(np.float32)
    for x in faces:
        a = np.concatenate([np.full((200, 1), x - 1e-3, np.float32), yz], axis=1)
        b = np.concatenate([np.full((200, 1), x + 1e-3, np.float32), yz], axis=1)
        va, oka, _ = sc.sample(a)
        vb, okb, _ = sc.sample(b)
        m = (oka == 1) & (okb == 1)
        assert m.sum() > 150
        assert np.abs(va[m] - vb[m]).max() < 5e-3


def test_render_transparent_xf_is_background(orc):
    cells, fields, dom = scenes.small_scene("nb32")
    sc, _ = _scene(orc, cells, fields)
    from owlexabrick_b200.renderer import default_camera
    fd = scenes.FrameDesc(16, 12, default_camera([0] * 3, [dom] * 3, 16, 12))
    fd.xf[0] = (np.zeros(128, np.float32), fd.xf[0][1])
    accum, rgba, st = sc.render(fd.to_orc(orc))
    assert (scenes.unpack_rgba8(rgba)[..., :3] == 0).all()
    assert st["samples"] == 0            # space skipping: every region inactive


def test_render_opaque_constant_colour(orc):
    """KAT: a constant field under an opaque single-colour XF gives, for every ray that hits
    the volume, exactly the reference's terminated colour (rgb*alpha, alpha>=0.98)."""
    cells, _, dom = scenes.small_scene("nb32")
    sc, _ = _scene(orc, cells, [np.full(len(cells), 0.5, np.float32)])
    from owlexabrick_b200.renderer import default_camera
    fd = scenes.FrameDesc(24, 24, default_camera([0] * 3, [dom] * 3, 24, 24))
    fd.grad_dvr = False
    fd.xf[0] = (np.ones(128, np.float32), np.tile(np.array([[0.2, 0.4, 0.8]], np.float32), (128, 1)))
    accum, rgba, st = sc.render(fd.to_orc(orc))
    centre = accum[12, 12]
    # alpha reaches 1 after the first sample: C=(rgb,1) -> terminated -> rgb*1 -> colour = 1*rgb
    assert np.allclose(centre[:3], [0.2, 0.4, 0.8], atol=1e-6)
    assert st["samples"] > 0


def test_streamlines_uniform_flow(orc):
    """KAT: RK4 through a constant vector field advances by steplen*v per step and dies
    when it leaves the world bounds."""
    cells = synth.nested_boxes(16, [])
    n = len(cells)
    fields = [np.full(n, 1.0, np.float32), np.full(n, 0.5, np.float32), np.full(n, 0.0, np.float32)]
    sc, _ = _scene(orc, cells, fields)
    seeds = np.array([[2, 2, 8], [10, 3, 4]], np.float32)
    tr = sc.trace(seeds, 12, (0, 1, 2), 1.0, [0, 0, 0], [16, 16, 16])
    assert np.allclose(tr[0, 5], [7, 4.5, 8], atol=1e-4)
    assert np.allclose(tr[1, 3], [13, 4.5, 4], atol=1e-4)
    assert (tr[1, 8:] >= 2e10).all()     # left the brick's basis domain -> dead


def test_bvh_region_query_equals_brute_force(orc, monkeypatch):
    """the oracle's private BVH for the closest-region ray query (stand-in for the OptiX BVH of
    exabrick.cu:184-238) must return exactly what the brute-force loop over all regions returns"""
    import scenes
    from owlexabrick_b200 import renderer as R
    cells, fields, dom = scenes.small_scene("nb64_off", 2)
    raw = orc.build_bricks(cells)
    bricks, ids = orc.parse_bricks(raw)
    sc = orc.flatten_scalars(ids, fields)
    regions, leaf = orc.build_abr(bricks, sc)
    fd = scenes.FrameDesc(80, 60, R.default_camera([0] * 3, [dom] * 3, 80, 60), 2)
    fd.opacity_scale = 0.1
    fd.iso = [(1, 0.5, 0), (0, 0.0, 0)]
    out = []
    for brute in ("0", "1"):
        monkeypatch.setenv("ORC_BRUTE_FORCE", brute)
        scene = orc.Scene(bricks, sc, regions, leaf)
        acc, rgba, st = scene.render(fd.to_orc(orc))
        out.append((acc.copy(), rgba.copy(), dict(st)))
    assert out[0][0].tobytes() == out[1][0].tobytes() and out[0][1].tobytes() == out[1][1].tobytes()
    assert out[0][2] == out[1][2]


def test_c5_shaped_builder_and_abr_hash_vs_reference(orc):
    """config C5's shape at 1/16 scale (block-structured octree, 4 levels, 3 channels): the oracle's bricks and
    ABRs (domains, value ranges over all channels, leaf lists) hash to what the reference's own exaBuilder and
    ExaBrickRegions::buildFrom produced (golden_hashes.json: C5_scale16)"""
    h = json.load(open(os.path.join(GOLDEN, "golden_hashes.json")))["C5_scale16"]
    cells, dom = synth.cloud_shaped(16)
    assert len(cells) == h["cells"]
    fields = synth.cloud_velocity(cells, dom)
    raw = orc.build_bricks(cells)
    assert hashlib.sha256(raw).hexdigest() == h["bricks_sha256"]
    bricks, ids = orc.parse_bricks(raw)
    regions, leaf = orc.build_abr(bricks, orc.flatten_scalars(ids, fields))
    blob = np.array([len(regions), len(leaf)], "<i4").tobytes() + regions.tobytes() + leaf.tobytes()
    assert hashlib.sha256(blob).hexdigest() == h["abr_sha256"]
