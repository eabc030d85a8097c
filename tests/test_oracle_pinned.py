"""The CPU oracle PINNED against the reference's own device source.

programs/exabrick.cu — raygen, intersection and bounds programs, tent-basis reconstruction with
analytic gradients, transfer function, integration, iso-surfaces, contour planes, streamline
cones, AO — is compiled as plain C++ in place from /root/reference (oracle/ref_shader_main.cpp +
oracle/ref_shim/shader_shim.h supply the OptiX/OWL device API on the host) and run on the frames of
tests/ref_shader_cases.py.  Its outputs are committed as tests/golden/ref_shader_frames.npz (made
by tests/golden/make_golden.py); when oracle/_ref/ref_shader is present it is also run live.

The oracle must reproduce the accumulation buffer and the RGBA8 frame of every case.  The only
knob: C++ leaves the order of the two rnd() calls in `vec2f(rnd(), rnd())` (exabrick.cu:1594, 1624)
unspecified; g++ (the reference compiled here) evaluates right to left, nvcc left to right, so the
oracle is switched to g++'s order for this comparison (ORC_RND_ARGS_RTL)."""
import os
import tempfile

import numpy as np
import pytest

from conftest import GOLDEN
import ref_shader_cases as RC

CASES = list(RC.cases(extra=True))


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(GOLDEN, "ref_shader_frames.npz"))


def _oracle_frame(orc, case, monkeypatch):
    name, kind, channels, mlw, configure = case
    raw, fields, dom, sc = RC.build(orc, kind, channels, mlw)
    fd = RC.frame(dom, channels, configure)
    monkeypatch.setenv("ORC_RND_ARGS_RTL", "1")
    acc0 = RC.start_accum(name) if fd.frame_id > 0 else None
    acc, rgba, _ = sc.render(fd.to_orc(orc), accum=None if acc0 is None else acc0.copy())
    return raw, fields, fd, acc0, acc, rgba


def _compare(name, acc, rgba, racc, rrgba):
    err = np.abs(acc[..., :3] - racc[..., :3]).max(axis=-1)
    # a surface / region hit decided by two exactly equal ray parameters may resolve differently
    # (closest-hit ties): allow a handful of pixels, everything else to rounding
    bad = int((err > 1e-6).sum())
    assert bad <= 2, "%s: %d pixels differ from the reference shader (max %.3g)" % (name, bad, err.max())
    assert (rgba != rrgba).sum() <= 2, name
    assert racc[..., :3].max() > 0.05, name + ": empty test frame"


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_oracle_equals_reference_shader_golden(orc, golden, case, monkeypatch):
    _, _, _, _, acc, rgba = _oracle_frame(orc, case, monkeypatch)
    _compare(case[0], acc, rgba, golden[case[0] + "_accum"], golden[case[0] + "_rgba"])


@pytest.mark.parametrize("case", CASES[::3], ids=[c[0] for c in CASES[::3]])
def test_oracle_equals_reference_shader_live(orc, case, monkeypatch):
    if not orc.have_ref_shader():
        pytest.skip("oracle/_ref/ref_shader not built (needs /root/reference)")
    raw, fields, fd, acc0, acc, rgba = _oracle_frame(orc, case, monkeypatch)
    with tempfile.TemporaryDirectory() as d:
        racc, rrgba = orc.ref_render(raw, fields, fd.to_orc(orc), d, accum=acc0)
    _compare(case[0], acc, rgba, racc, rrgba)


CD_CASES = list(RC.cd_cases())


@pytest.mark.parametrize("case", CD_CASES, ids=[c[0] for c in CD_CASES])
def test_oracle_central_differences_equal_reference_build(orc, case, monkeypatch):
    """the reference's off-default gradient option (-DANALYTIC_GRADIENTS=0: gradientCD, exabrick.cu:833-880): the oracle
    with OrcFrame.centralDifferences against frames of oracle/_ref/ref_shader_cd (tests/golden/ref_shader_frames_cd.npz;
    live too where the binary is present), and the frames do differ from the analytic-gradient ones"""
    g = np.load(os.path.join(GOLDEN, "ref_shader_frames_cd.npz"))
    raw, fields, fd, _, acc, rgba = _oracle_frame(orc, case, monkeypatch)
    _compare(case[0], acc, rgba, g[case[0] + "_accum"], g[case[0] + "_rgba"])
    fd.central_differences = False
    acc_a, _, _ = RC.build(orc, case[1], case[2], case[3])[3].render(fd.to_orc(orc))
    assert np.abs(acc_a[..., :3] - acc[..., :3]).max() > 0.02
    if orc.have_ref_shader_cd() and case[0] in ("cd_dvr", "cd_iso", "cd_x_iso_ch1_ao"):
        fd.central_differences = True
        with tempfile.TemporaryDirectory() as d:
            racc, rrgba = orc.ref_render(raw, fields, fd.to_orc(orc), d, binary="ref_shader_cd")
        _compare(case[0], acc, rgba, racc, rrgba)


def test_rnd_order_is_the_only_difference(orc, golden, monkeypatch):
    """with nvcc's left-to-right order (the product's and the oracle's default) the same frame differs
    from the g++-compiled reference only through the swapped pixel jitter"""
    case = CASES[0]
    name, kind, channels, mlw, configure = case
    raw, fields, dom, sc = RC.build(orc, kind, channels, mlw)
    fd = RC.frame(dom, channels, configure)
    monkeypatch.delenv("ORC_RND_ARGS_RTL", raising=False)
    acc, _, _ = sc.render(fd.to_orc(orc))
    assert np.abs(acc[..., :3] - golden[name + "_accum"][..., :3]).max() > 1e-3


def test_tracer_equals_reference_compute_traces(orc, golden):
    """RK4 advection: the oracle against the reference's computeTraces (exabrick.cu:1531-1574, with its
    sampleDirection -> traceVolumeRay point query) run for every seed and time step — bit-identical,
    including which vertices die (leave the world bounds / invalid sample)"""
    cells, fields, dom, seeds, T, steplen = RC.tracer_case()
    raw = orc.build_bricks(cells)
    bricks, ids = orc.parse_bricks(raw)
    sc = orc.flatten_scalars(ids, fields)
    regions, leaf = orc.build_abr(bricks, sc)
    o = orc.Scene(bricks, sc, regions, leaf).trace(seeds, T, (0, 1, 2), steplen, [0, 0, 0], [dom] * 3)
    assert np.array_equal(o, golden["tracer"])
    assert 0 < (o[..., 0] >= 2e10).sum() < 0.7 * o.shape[0] * T      # some traces leave the volume, most vertices live
    if orc.have_ref_shader():
        with tempfile.TemporaryDirectory() as d:
            r = orc.ref_trace(raw, fields, seeds, T, (0, 1, 2), steplen, [0, 0, 0], [dom] * 3, d)
        assert np.array_equal(r, golden["tracer"])
    ch, sl, lo, hi = RC.tracer_case_b()                  # permuted channels, backward, bounds inside the volume
    o = orc.Scene(bricks, sc, regions, leaf).trace(seeds, T, ch, sl, lo, hi)
    assert np.array_equal(o, golden["tracer_b"])
    assert 0.2 * o.shape[0] * T < (o[..., 0] < 2e10).sum() < o.shape[0] * T


def test_reconstruction_1m_points_c1_equals_reference(orc):
    """north_star correctness check 2, on the CPU side: the oracle's reconstruction (value, validity and
    analytic gradient) at 1 M random points of config C1 is bit-identical to the reference's own
    samplePoint / samplePointWithDerivative behind its own point query (SHA-256 made by
    tests/golden/make_golden.py from oracle/_ref/ref_shader).  tests/test_gpu_parity.py::
    test_reconstruction_1m_points_c1 holds the CUDA path to within 1e-5 relative of these same values."""
    import hashlib
    import json
    from owlexabrick_b200 import synth
    h = json.load(open(os.path.join(GOLDEN, "golden_hashes.json")))["C1_points_1M"]
    c = synth.nested_boxes(256, [(72, 184), (32, 224)])
    f = synth.analytic_field(c, 256, 0)
    bricks, ids = orc.parse_bricks(orc.build_bricks(c))
    sc = orc.flatten_scalars(ids, [f])
    regions, leaf = orc.build_abr(bricks, sc)
    pts = synth.random_points(1000000, [0.0] * 3, [256.0] * 3, seed=1)
    v, ok, g, _ = orc.Scene(bricks, sc, regions, leaf).sample(pts, 0, grad=True)
    head = np.load(os.path.join(GOLDEN, "c1_points_head.npz"))
    assert np.array_equal(v[:4096], head["values"]) and np.array_equal(ok[:4096], head["valid"]) and np.array_equal(g[:4096], head["grad"])
    assert int(ok.sum()) == h["valid"]
    assert hashlib.sha256(ok.tobytes()).hexdigest() == h["valid_sha256"]
    assert hashlib.sha256(v.tobytes()).hexdigest() == h["values_sha256"]
    assert hashlib.sha256(g.tobytes()).hexdigest() == h["grad_sha256"]
