"""Builder and ABR build on random octree AMR (ragged, negative coordinates, shuffled, duplicated cells, leaf
widths 3 / 8 / 17 / 127, 1-2 channels) against the reference's own exaBuilder and ExaBrickRegions::buildFrom
(builder/builder.cpp, exa/Regions.cpp compiled in place): the product's host builder, the oracle's builder and
the oracle's ABR build must be bit-identical.  SHA-256 of the reference outputs for the seeds below are in
tests/golden/golden_hashes.json["fuzz"] (tests/golden/make_golden.py:fuzz_golden); with oracle/_ref present the
reference binaries are also run live on further seeds."""
import hashlib
import json
import os
import tempfile

import pytest

import scenes
from conftest import GOLDEN

SEEDS = list(range(12))
LIVE_SEEDS = list(range(100, 108))


@pytest.fixture(scope="module")
def R(built_lib):
    from owlexabrick_b200 import renderer
    return renderer


def _ours(R, orc, cells, mlw, fields):
    ob = orc.build_bricks(cells, mlw)
    pb = R.ExaBricks.build(cells, max_leaf_width=mlw).to_bytes()
    bricks, ids = orc.parse_bricks(ob)
    regions, leaf = orc.build_abr(bricks, orc.flatten_scalars(ids, fields))
    return ob, pb, regions, leaf


@pytest.mark.parametrize("seed", SEEDS)
def test_random_amr_equals_reference_golden(R, orc, seed):
    h = json.load(open(os.path.join(GOLDEN, "golden_hashes.json")))["fuzz"][str(seed)]
    cells, mlw, fields = scenes.random_amr(seed)
    assert len(cells) == h["cells"]
    ob, pb, regions, leaf = _ours(R, orc, cells, mlw, fields)
    assert hashlib.sha256(ob).hexdigest() == h["bricks_sha256"], "oracle builder"
    assert hashlib.sha256(pb).hexdigest() == h["bricks_sha256"], "product builder"
    assert hashlib.sha256(regions.tobytes() + leaf.tobytes()).hexdigest() == h["abr_sha256"], "oracle ABR"
    assert len(regions) == h["regions"]


@pytest.mark.parametrize("seed", LIVE_SEEDS)
def test_random_amr_equals_reference_live(R, orc, seed):
    if not orc.have_ref():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    cells, mlw, fields = scenes.random_amr(seed)
    with tempfile.TemporaryDirectory() as d:
        rb = orc.ref_build_bricks(cells, d, None if mlw == 127 else mlw)
        rr, rl = orc.parse_abr(orc.ref_build_abr(rb, fields, d))
    ob, pb, regions, leaf = _ours(R, orc, cells, mlw, fields)
    assert ob == rb and pb == rb
    assert regions.tobytes() == rr.tobytes() and leaf.tobytes() == rl.tobytes()
