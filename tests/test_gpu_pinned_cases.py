"""GPU parity on the frames the oracle is pinned on (run with -m gpu): the 15 cases of
tests/ref_shader_cases.py — on which the CPU oracle is bit-identical to the reference's own device
source compiled as C++ (tests/test_oracle_pinned.py) — rendered by the CUDA path through the C ABI and
compared with the oracle (nvcc's rnd() argument order on both sides): RGBA within 1e-3 max absolute per
channel.  Frames with surfaces may have a few pixels whose hit is decided to within rounding."""
import numpy as np
import pytest

import ref_shader_cases as RC
from test_gpu_parity import R, scalar_layout, RGBA_TOL  # noqa: F401 (fixtures)

pytestmark = pytest.mark.gpu
CASES = list(RC.cases())
SURFACE = {"iso", "iso_flat", "two_iso_2ch", "contour_planes", "cones", "cones_ao_iso"}


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_cuda_path_on_pinned_case(R, orc, case):
    name, kind, channels, mlw, configure = case
    raw, fields, dom, sc = RC.build(orc, kind, channels, mlw)
    r = R.Renderer(R.ExaBricks.from_bytes(raw), fields)
    fd = RC.frame(dom, channels, configure)
    fd.apply(r)
    acc0 = RC.start_accum(name) if fd.frame_id > 0 else None
    if acc0 is not None:
        r.write_accum(acc0)
    if fd.traces is not None:                          # the traces are an input here: load them as seeds of a 1-step-per-vertex tracer
        r.load_traces(fd.traces, fd.current_timestep)
    r.render()
    g_acc = r.read_accum()
    o_acc, _, _ = sc.render(fd.to_orc(orc), accum=None if acc0 is None else acc0.copy())
    err = np.abs(g_acc[..., :3] - o_acc[..., :3]).max(axis=-1)
    # AO rays start ON the surface they shade: whether the first two samples of the occlusion march bracket
    # the iso value there is decided by the last bit of the reconstruction, and both rays of a pixel share
    # that origin — a pixel can go from unshadowed to fully shadowed between any two fp32 implementations
    allowed = 0.03 * err.size if name == "cones_ao_iso" else 0.01 * err.size if name in SURFACE else 0
    assert (err > RGBA_TOL).sum() <= allowed, "%s: %d pixels differ (max %.3g)" % (name, int((err > RGBA_TOL).sum()), err.max())
    assert o_acc[..., :3].max() > 0.05


CD_CASES = list(RC.cd_cases())


@pytest.mark.parametrize("case", CD_CASES, ids=[c[0] for c in CD_CASES])
def test_cuda_path_central_difference_gradients(R, orc, case):
    """exa_renderer_set_gradient_mode(CENTRAL_DIFFERENCES): the reference's off-default -DANALYTIC_GRADIENTS=0 build
    (gradientCD, exabrick.cu:833-880).  The oracle with the same option is bit-identical to the reference's own shader
    compiled that way (tests/test_oracle_pinned.py); here the CUDA path against that oracle, and the switch back"""
    name, kind, channels, mlw, configure = case
    raw, fields, dom, sc = RC.build(orc, kind, channels, mlw)
    r = R.Renderer(R.ExaBricks.from_bytes(raw), fields)
    fd = RC.frame(dom, channels, configure)
    fd.apply(r)
    r.render()
    g_acc = r.read_accum().copy()
    o_acc, _, _ = sc.render(fd.to_orc(orc))
    err = np.abs(g_acc[..., :3] - o_acc[..., :3]).max(axis=-1)
    surface = "iso" in name
    # a shading sample that falls within rounding of a region face may be located on the other side (six more point
    # locations per shaded sample): a few pixels on top of the surface-hit flips
    allowed = 0.03 * err.size if "ao" in name else 0.01 * err.size if surface else 0.002 * err.size
    assert (err > RGBA_TOL).sum() <= allowed, "%s: %d pixels differ (max %.3g)" % (name, int((err > RGBA_TOL).sum()), err.max())
    with pytest.raises(RuntimeError, match="central-difference"):
        r.render_stats()
    # back to the analytic gradients: a different frame, equal to the analytic oracle frame
    fd.central_differences = False
    fd.apply(r)
    r.render()
    a_acc = r.read_accum()
    oa_acc, _, _ = sc.render(fd.to_orc(orc))
    assert np.abs(a_acc[..., :3] - g_acc[..., :3]).max() > 0.02
    aerr = np.abs(a_acc[..., :3] - oa_acc[..., :3]).max(axis=-1)
    assert (aerr > RGBA_TOL).sum() <= (0.03 * err.size if "ao" in name else 0.01 * err.size if surface else 0)
