"""GPU exaBuilder (exa_bricks_build_gpu, csrc/gpu_builder.cu; run with -m gpu): bit-exact with the reference's own
exaBuilder — the same golden .bricks bytes, SHA-256 hashes and fuzz inputs that pin the host builder
(tests/test_host.py, tests/test_fuzz_vs_reference.py) — including the kd file, the spatial-median builder, duplicates,
negative coordinates, ragged domains and the error behaviour."""
import hashlib
import json
import os

import numpy as np
import pytest

import scenes
from conftest import GOLDEN, load_golden
from owlexabrick_b200 import synth

pytestmark = pytest.mark.gpu

CASES = ["ex0", "ex1", "ex2", "ex3", "ex4", "nb64", "nb64_mlw8", "nb64_dups"]


@pytest.fixture(scope="module")
def R(built_lib):
    assert built_lib.lib.exa_cuda_available() == 1, built_lib.last_error()
    from owlexabrick_b200 import renderer
    return renderer


def _same(a, b):
    assert a.to_bytes() == b.to_bytes()
    assert a.kd_nodes.tobytes() == b.kd_nodes.tobytes()
    assert a.headers.tobytes() == b.headers.tobytes()


@pytest.mark.parametrize("name", CASES)
def test_gpu_builder_bit_exact_with_reference_golden(R, name):
    g = load_golden(name)
    b = R.ExaBricks.build_gpu(g["cells"], max_leaf_width=int(g["max_leaf_width"]))
    assert b.to_bytes() == g["bricks"].tobytes()
    _same(b, R.ExaBricks.build(g["cells"], max_leaf_width=int(g["max_leaf_width"])))


@pytest.mark.parametrize("seed", list(range(12)))
def test_gpu_builder_random_amr_equals_reference_hash(R, seed):
    """the 12 fuzz inputs: ragged octrees, negative coordinates, shuffled, duplicated cells, leaf widths 3 / 8 / 17 / 127"""
    h = json.load(open(os.path.join(GOLDEN, "golden_hashes.json")))["fuzz"][str(seed)]
    cells, mlw, _ = scenes.random_amr(seed)
    b = R.ExaBricks.build_gpu(cells, max_leaf_width=mlw)
    assert hashlib.sha256(b.to_bytes()).hexdigest() == h["bricks_sha256"]
    _same(b, R.ExaBricks.build(cells, max_leaf_width=mlw))


@pytest.mark.parametrize("seed", list(range(100, 108)))
def test_gpu_builder_random_amr_equals_host_builder(R, built_lib, seed):
    cells, mlw, _ = scenes.random_amr(seed)
    _same(R.ExaBricks.build_gpu(cells, max_leaf_width=mlw), R.ExaBricks.build(cells, max_leaf_width=mlw))
    _same(R.ExaBricks.build_gpu(cells, builder=built_lib.BUILDER_SPATIAL_MEDIAN, max_leaf_width=mlw),
          R.ExaBricks.build(cells, builder=built_lib.BUILDER_SPATIAL_MEDIAN, max_leaf_width=mlw))


def test_gpu_builder_c1_and_c5_quarter_hashes(R):
    """config C1 (2.27 M cells) and the C5 shape at 1/4 scale (26.6 M cells, 4 021 bricks): SHA-256 of the reference's .bricks"""
    hashes = json.load(open(os.path.join(GOLDEN, "golden_hashes.json")))
    cells = synth.nested_boxes(256, [(72, 184), (32, 224)])
    b = R.ExaBricks.build_gpu(cells)
    assert hashlib.sha256(b.to_bytes()).hexdigest() == hashes["C1"]["bricks_sha256"]
    assert b.total_num_cells == len(cells) and b.timings["levels"] > 1
    cells, dom = synth.cloud_shaped(4)
    b = R.ExaBricks.build_gpu(cells)
    h = hashes["C5_scale4"]
    assert b.num_bricks == h["bricks"] and hashlib.sha256(b.to_bytes()).hexdigest() == h["bricks_sha256"]


def test_gpu_builder_large_slice_counts_and_shuffled_input(R):
    """a node with more coarse slices than the shared-memory histogram holds (global-atomics path), cells in random order"""
    n = 2400                                                       # coarse width 2: 1201 + 1 + 1 slices > 1024
    x = np.arange(n, dtype=np.int32)
    cells = np.stack([np.repeat(x, 4), np.tile(np.array([0, 0, 1, 1], np.int32), n), np.tile(np.array([0, 1, 0, 1], np.int32), n), np.zeros(4 * n, np.int32)], axis=1)
    cells = cells[np.random.RandomState(1).permutation(len(cells))]
    cells = np.concatenate([cells, np.array([[n, 0, 0, 1]], np.int32)])   # one coarser cell: the root is not a leaf
    _same(R.ExaBricks.build_gpu(cells), R.ExaBricks.build(cells))


def test_gpu_builder_errors(R):
    with pytest.raises(Exception, match="level out of range"):
        R.ExaBricks.build_gpu(np.array([[0, 0, 0, 31], [1, 0, 0, 0]], np.int32))
    # overlapping cells of different levels: the reference throws "cell straddles split plane" or "invalid split"; both builders must fail
    bad = np.array([[0, 0, 0, 1], [1, 0, 0, 0], [2, 0, 0, 0], [0, 2, 0, 0]], np.int32)
    with pytest.raises(Exception):
        R.ExaBricks.build(bad)
    with pytest.raises(Exception):
        R.ExaBricks.build_gpu(bad)
    b = R.ExaBricks.build_gpu(np.zeros((0, 4), np.int32))
    assert b.num_bricks == 0
