"""CPU tests of the surface geometry of the frame kernel (streamline cones, contour planes, AO
sample frame; programs/exabrick.cu:441-501, 1267-1394, 78-94).

The product's functions live in owlexabrick_b200/csrc/surfaces.cuh as host+device code; the
harness tests/harness/surfaces_host.cpp compiles them with g++ (test-only, never part of the
library) so that the SAME arithmetic the kernel runs is checked here against the oracle's
independent restatement and against analytic known answers."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
F3 = C.c_float * 3


@pytest.fixture(scope="module")
def surf(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("surf") / "libsurf_host.so")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", "-I/usr/local/cuda/include",
                           "-o", out, os.path.join(ROOT, "tests", "harness", "surfaces_host.cpp")])
    lib = C.CDLL(out)
    lib.surf_contour_plane_t.restype = C.c_float
    return lib


@pytest.fixture(scope="module")
def olib(orc):
    lib = orc.lib()
    lib.orc_contour_plane_t.restype = C.c_float
    return lib


def _cone(lib, name, pa, pb, ra, rb, ro, rd):
    t = C.c_float()
    n = F3()
    hit = getattr(lib, name)(F3(*pa), F3(*pb), C.c_float(ra), C.c_float(rb), F3(*ro), F3(*rd), C.byref(t), n)
    return hit, t.value, np.array(list(n), np.float32)


def _plane(lib, name, ro, rd, normal, offset, wlo, whi):
    pts = (C.c_float * 18)()
    cnt = C.c_int()
    t = getattr(lib, name)(F3(*ro), F3(*rd), F3(*normal), C.c_float(offset), F3(*wlo), F3(*whi), pts, C.byref(cnt))
    return t, np.array(list(pts), np.float32).reshape(6, 3)[:cnt.value]


def test_rounded_cone_matches_oracle(surf, olib):
    rng = np.random.default_rng(5)
    hits = 0
    for _ in range(4000):
        pa = rng.uniform(-20, 20, 3).astype(np.float32)
        pb = (pa + rng.uniform(-6, 6, 3)).astype(np.float32)
        ro = rng.uniform(-60, 60, 3).astype(np.float32)
        tgt = pa + rng.uniform(-4, 4, 3)
        rd = (tgt - ro) / np.linalg.norm(tgt - ro)
        rd = rd.astype(np.float32)
        a = _cone(surf, "surf_rounded_cone", pa, pb, 2.0, 2.0, ro, rd)
        b = _cone(olib, "orc_rounded_cone", pa, pb, 2.0, 2.0, ro, rd)
        assert a[0] == b[0]
        if a[0]:
            hits += 1
            assert a[1] == b[1]                       # same expressions, no contraction: bit-identical
            assert np.array_equal(a[2], b[2])
    assert hits > 1000


def test_rounded_cone_known_answers(surf):
    # cylinder of radius 2 along x from (0,0,0) to (10,0,0); ray along -z through its axis at x=5
    hit, t, n = _cone(surf, "surf_rounded_cone", (0, 0, 0), (10, 0, 0), 2.0, 2.0, (5, 0, 30), (0, 0, -1))
    assert hit and abs(t - 28.0) < 1e-4
    assert abs(n[0]) < 1e-3 * np.linalg.norm(n) and n[2] > 0
    # the cap around pa is a sphere of radius ra: ray along +x hits it at x = -2
    hit, t, n = _cone(surf, "surf_rounded_cone", (0, 0, 0), (10, 0, 0), 2.0, 2.0, (-30, 0, 0), (1, 0, 0))
    assert hit and abs(t - 28.0) < 1e-4 and n[0] < 0
    # a ray that passes 3 units from the axis misses
    hit, _, _ = _cone(surf, "surf_rounded_cone", (0, 0, 0), (10, 0, 0), 2.0, 2.0, (5, 3, 30), (0, 0, -1))
    assert not hit


def test_contour_plane_matches_oracle(surf, olib):
    rng = np.random.default_rng(11)
    hits = 0
    for i in range(3000):
        nrm = rng.normal(size=3)
        if i % 5 == 0:
            nrm = np.eye(3)[i % 3]                    # axis-aligned planes (the viewer's default is x)
        nrm = (nrm / np.linalg.norm(nrm)).astype(np.float32)
        off = float(rng.uniform(-0.2, 1.2)) if i % 5 else float(rng.uniform(0.05, 0.95))
        wlo = rng.uniform(-10, 0, 3).astype(np.float32)
        whi = (wlo + rng.uniform(16, 64, 3)).astype(np.float32)
        ro = (wlo - rng.uniform(10, 40, 3)).astype(np.float32)
        tgt = wlo + rng.uniform(0, 1, 3) * (whi - wlo)
        rd = ((tgt - ro) / np.linalg.norm(tgt - ro)).astype(np.float32)
        ta, pa = _plane(surf, "surf_contour_plane_t", ro, rd, nrm, off, wlo, whi)
        tb, pb = _plane(olib, "orc_contour_plane_t", ro, rd, nrm, off, wlo, whi)
        assert len(pa) == len(pb)
        assert np.array_equal(pa, pb)
        assert ta == tb
        hits += ta >= 0
    assert hits > 300


def test_contour_plane_known_answer(surf):
    # plane x = 0.5 of the unit cube scaled to [0,64]^3 -> x = 32; ray along +x from (-10, 20, 30)
    t, pts = _plane(surf, "surf_contour_plane_t", (-10, 20, 30), (1, 0, 0), (1, 0, 0), 0.5, (0, 0, 0), (64, 64, 64))
    assert len(pts) == 4 and np.allclose(pts[:, 0], 32.0)
    assert abs(t - 42.0) < 1e-4
    # outside the polygon: miss
    t, _ = _plane(surf, "surf_contour_plane_t", (-10, 80, 30), (1, 0, 0), (1, 0, 0), 0.5, (0, 0, 0), (64, 64, 64))
    assert t == -1.0
    # a plane that does not cut the cube: no polygon
    t, pts = _plane(surf, "surf_contour_plane_t", (-10, 20, 30), (1, 0, 0), (1, 0, 0), 1.5, (0, 0, 0), (64, 64, 64))
    assert len(pts) == 0 and t == -1.0


def test_ao_direction_matches_oracle(surf, olib):
    rng = np.random.default_rng(3)
    for _ in range(2000):
        nrm = rng.normal(size=3)
        nrm = (nrm / np.linalg.norm(nrm)).astype(np.float32)
        u1, u2 = (float(np.float32(x)) for x in rng.uniform(0, 1, 2))
        a, b = F3(), F3()
        surf.surf_ao_direction(F3(*nrm), C.c_float(u1), C.c_float(u2), a)
        olib.orc_ao_direction(F3(*nrm), C.c_float(u1), C.c_float(u2), b)
        a, b = np.array(list(a)), np.array(list(b))
        assert np.allclose(a, b, atol=2e-7)
        assert abs(np.linalg.norm(a) - 1) < 1e-5
        assert np.dot(a, nrm) >= -1e-6                # in the hemisphere of the normal
