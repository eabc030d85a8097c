"""CPU tests of the product's host side: the bit-exact builder, loaders/file formats, the
C-ABI surface (every symbol of include/exabrick_b200.h is exported) and loud failure
without a GPU.  No compute call needs a device here."""
import ctypes as C
import hashlib
import json
import os
import re
import sys

import numpy as np
import pytest

from conftest import GOLDEN, ROOT, load_golden
from owlexabrick_b200 import synth

CASES = ["ex0", "ex1", "ex2", "ex3", "ex4", "nb64", "nb64_mlw8", "nb64_dups"]


@pytest.fixture(scope="module")
def R(built_lib):
    from owlexabrick_b200 import renderer
    return renderer


def test_header_symbols_exported(built_lib):
    hdr = open(os.path.join(ROOT, "include", "exabrick_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(exa_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) > 50
    for name in sorted(declared):
        assert hasattr(built_lib.lib, name), "library does not export " + name
    assert declared == set(built_lib.EXPORTS), declared ^ set(built_lib.EXPORTS)


def test_pod_sizes(built_lib, R):
    assert R.BRICK_DT.itemsize == 32 and R.REGION_DT.itemsize == 44 and R.KD_DT.itemsize == 12
    assert C.sizeof(built_lib.ExaFrameState) == 408 and C.sizeof(built_lib.ExaLaunchParams) == 240


@pytest.mark.parametrize("name", CASES)
def test_builder_bit_exact_with_reference(R, name):
    g = load_golden(name)
    b = R.ExaBricks.build(g["cells"], max_leaf_width=int(g["max_leaf_width"]))
    assert b.to_bytes() == g["bricks"].tobytes()


def test_builder_c1_hash(R):
    hashes = json.load(open(os.path.join(GOLDEN, "golden_hashes.json")))
    cells = synth.nested_boxes(256, [(72, 184), (32, 224)])
    b = R.ExaBricks.build(cells)
    assert hashlib.sha256(b.to_bytes()).hexdigest() == hashes["C1"]["bricks_sha256"]
    assert b.total_num_cells == len(cells)


def test_builder_matches_oracle_on_ragged_input(R, orc):
    """shuffled order, duplicates, negative coordinates, 4 levels, odd leaf widths"""
    cells, _, _ = __import__("scenes").small_scene("nb64_off")
    cells = cells.copy()
    cells[:, :3] -= np.array([40, 8, 24], np.int32)          # negative coordinates (multiples of 8)
    rng = np.random.RandomState(3)
    cells = np.concatenate([cells, cells[rng.randint(0, len(cells), 300)]])[rng.permutation(len(cells) + 300)]
    for mlw in (127, 5):
        b = R.ExaBricks.build(cells, max_leaf_width=mlw)
        assert b.to_bytes() == orc.build_bricks(cells, mlw)


def test_builder_spatial_median_and_kd(R, built_lib):
    cells = synth.nested_boxes(32, [(10, 22), (4, 28)])
    b = R.ExaBricks.build(cells, builder=built_lib.BUILDER_SPATIAL_MEDIAN)
    assert b.total_num_cells == len(cells)
    assert sorted(b.cell_ids.tolist()) == list(range(len(cells)))
    kd = b.kd_nodes
    leaves = kd[(kd["axis_bits"] >> 30) == 3]
    assert len(leaves) == b.num_bricks and len(kd) == 2 * b.num_bricks - 1
    assert sorted(leaves["split_or_first_prim"].tolist()) == list(range(b.num_bricks))


def test_builder_errors(R, built_lib):
    empty = R.ExaBricks.build(np.zeros((0, 4), np.int32))
    assert empty.num_bricks == 0
    with pytest.raises(built_lib.ExaError):
        R.ExaBricks.build(np.array([[0, 0, 0, 99]], np.int32))


def test_bricks_file_round_trip(R, tmp_path):
    g = load_golden("nb64")
    p = str(tmp_path / "a.bricks")
    open(p, "wb").write(g["bricks"].tobytes())
    b = R.ExaBricks.load(p)
    assert b.to_bytes() == g["bricks"].tobytes()
    q = str(tmp_path / "b.bricks")
    b.save(q)
    assert open(q, "rb").read() == g["bricks"].tobytes()
    lo, hi = b.get_bounds()
    assert lo.tolist() == [0, 0, 0] and hi.tolist() == [64, 64, 64]
    h = b.headers
    n = h["size"].astype(np.int64).prod(axis=1)
    assert (h["begin"] == np.concatenate([[0], np.cumsum(n)[:-1]])).all()
    with pytest.raises(Exception):
        R.ExaBricks.load(str(tmp_path / "missing.bricks"))


def test_scalar_field_loader_quirk(R, tmp_path):
    v = np.linspace(0.25, 2.0, 200).astype(np.float32)
    p = str(tmp_path / "f.scalars")
    v.tofile(p)
    vals, rng = R.load_scalar_field(p, compat=True)          # ScalarField.cpp:27-34: sized by BYTES
    assert len(vals) == 800 and (vals[:200] == v).all() and (vals[200:] == 0).all()
    assert rng[0] == 0.0 and rng[1] == 2.0
    vals, rng = R.load_scalar_field(p, compat=False)
    assert len(vals) == 200 and rng[0] == np.float32(0.25)


def test_colormapper(R):
    cm = R.ColorMapper(colors=[[0, 0, 0], [1, 0, 0], [1, 1, 1]])
    assert np.allclose(cm(0.25), [0.5, 0, 0]) and np.allclose(cm(0.75), [1, .5, .5])
    assert np.allclose(cm(1.0), [1, 1, 1]) and np.allclose(cm(2.0), [1, 1, 1])
    assert cm.table().shape == (128, 3)
    cm2 = R.ColorMapper(control_points=[[1.0, 1, 1, 1], [0.0, 0, 0, 0], [0.1, 1, 0, 0]])   # unsorted -> sorted
    assert np.allclose(cm2(0.05), [0.5, 0, 0])


def test_config_parser(R, tmp_path, built_lib):
    g = load_golden("ex4")
    open(tmp_path / "ex4.bricks", "wb").write(g["bricks"].tobytes())
    g["fields"][0].astype(np.float32).tofile(tmp_path / "ex4.scalars")
    cfg_text = """# comment line
remap_from 0 0 0 8 8 4
remap_to   -1 -1 -1  1 1 0   # trailing comment
scalar density ex4.scalars
value_range 0.1 0.9
vector vel ex4.scalars ex4.scalars ex4.scalars
bricks ex4.bricks
"""
    p = str(tmp_path / "ex4.exa")
    open(p, "w").write(cfg_text)
    cfg = R.Config(p)
    assert [f["name"] for f in cfg.fields] == ["density", "vel"]
    assert np.allclose(cfg.fields[0]["valueRange"], [0.1, 0.9])
    assert np.allclose(cfg.fields[1]["values"][:200], np.sqrt(3.0) * g["fields"][0], rtol=1e-6)
    # voxel = T(from.lo) S(from.span) (T(to.lo) S(to.span))^-1 world: world (-1,-1,-1) -> voxel (0,0,0); (1,1,0) -> (8,8,4)
    m = cfg.voxelSpaceTransform
    xfm = lambda w: m[0:3] * w[0] + m[3:6] * w[1] + m[6:9] * w[2] + m[9:12]
    assert np.allclose(xfm([-1, -1, -1]), [0, 0, 0], atol=1e-6) and np.allclose(xfm([1, 1, 0]), [8, 8, 4], atol=1e-5)
    lo, hi = cfg.get_bounds()
    assert np.allclose(lo, [-1, -1, -1], atol=1e-6) and np.allclose(hi, [1, 1, 0], atol=1e-6)
    open(p, "w").write("bogus_token 1 2 3\n")
    with pytest.raises(built_lib.ExaError, match="unknown token"):
        R.Config(p)
    with pytest.raises(built_lib.ExaError):
        R.Config(str(tmp_path / "nope.exa"))


def test_no_cpu_fallback(R, built_lib):
    """without a usable CUDA device the renderer must refuse, not compute on the host"""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    assert built_lib.lib.exa_cuda_available() == 0
    g = load_golden("ex4")
    b = R.ExaBricks.from_bytes(g["bricks"].tobytes())
    with pytest.raises(built_lib.ExaError) as e:
        R.Renderer(b, [g["fields"][0]])
    assert e.value.code == built_lib.EXA_ERR_CUDA


def test_product_never_touches_the_oracle():
    """the oracle is test infrastructure: nothing in the package, its CUDA/C++ sources, the public headers or
    tools/ imports, includes, links or executes anything under oracle/; bench.py does so only in its two
    CPU-baseline legs"""
    root = ROOT
    pat = re.compile(r"(from\s+oracle|import\s+oracle|import\s+orc\b|oracle/|oracle\.h|liboracle|\borc\.)")
    files = []
    for d in ("owlexabrick_b200", "include", "tools"):
        for base, _, names in os.walk(os.path.join(root, d)):
            files += [os.path.join(base, n) for n in names if n.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".hpp", ".inl"))]
    assert len(files) > 15
    for f in files:
        hits = [ln for ln in open(f, errors="replace") if pat.search(ln) and "#" not in ln.split("oracle")[0] and "//" not in ln.split("oracle")[0] and "*" not in ln.split("oracle")[0][:4]]
        assert not hits, (f, hits[:3])
    # bench.py: the oracle is imported in exactly two places — the reference arm and, in the product arm, the
    # cpu_baseline block that runs AFTER the timed region and is skipped under torchrun / --no-cpu-baseline
    src = open(os.path.join(root, "bench.py")).read().split("\n")
    imports = [i for i, ln in enumerate(src) if re.search(r"from oracle import|import orc\b", ln)]
    assert len(imports) == 2, imports
    owner = lambda i: next(re.match(r"def (\w+)", src[j]).group(1) for j in range(i, -1, -1) if src[j].startswith("def "))
    assert [owner(i) for i in imports] == ["run_reference", "run_product"]
    assert "not args.no_cpu_baseline" in src[imports[1] - 1] and "world == 1" in src[imports[1] - 1]


def test_camera_matches_reference_math(R):
    pos, d00, du, dv = R.simple_camera([0, 0, 5], [0, 0, 0], [0, 1, 0], 90.0, 200, 100)
    # fovy 90, focal 5 -> screen height 10, width 20; lower-left = (-10,-5,-5)
    assert np.allclose(d00, [-10, -5, -5], atol=1e-5)
    assert np.allclose(du * 200, [20, 0, 0], atol=1e-5) and np.allclose(dv * 100, [0, 10, 0], atol=1e-5)


def test_cloud_shaped_c5_generator(R, orc):
    """config C5's shape at 1/16 scale: octree-like refinement, 4 levels, the cells tile the domain
    exactly, and the host builder agrees bit-exactly with the oracle's exaBuilder restatement; with
    the reference's own exaBuilder too when oracle/_ref is present"""
    cells, dom = synth.cloud_shaped(16)
    assert dom == 128 and set(np.unique(cells[:, 3])) == {0, 1, 2, 3}
    w = (1 << cells[:, 3]).astype(np.int64)
    assert int((w ** 3).sum()) == dom ** 3                       # no holes, no overlap (volumes add up ...
    key = np.unique(cells[:, 0].astype(np.int64) + dom * (cells[:, 1].astype(np.int64) + dom * cells[:, 2].astype(np.int64)))
    assert len(key) == len(cells)                                # ... and no two cells share a corner)
    assert (cells[:, :3] % w[:, None] == 0).all()
    counts = synth.cloud_shaped_counts(16)
    assert [int((cells[:, 3] == l).sum()) for l in range(4)] == counts
    mine = R.ExaBricks.build(cells).to_bytes()
    assert mine == orc.build_bricks(cells)
    if orc.have_ref():
        import tempfile
        with tempfile.TemporaryDirectory() as d:
            assert mine == orc.ref_build_bricks(cells, d)
    v = synth.cloud_velocity(cells, dom)
    assert len(v) == 3 and all(abs(f).max() <= 1.5 for f in v)


def test_triangle_mesh_file_and_config(R, built_lib, tmp_path):
    """TriangleMesh::load (exa/TriangleMesh.cpp:21-68): a file is a sequence of meshes; the loader
    returns them concatenated with rebased indices; `triangles <file>` in a config (Config.cpp:165-171)"""
    import scenes
    v1, i1 = scenes.uv_sphere([1, 2, 3], 2.0, 8, 4)
    v2 = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], np.float32)
    i2 = np.array([[0, 1, 2]], np.int32)
    p = tmp_path / "m.tris"
    with open(p, "wb") as f:
        for v, i in ((v1, i1), (v2, i2)):
            f.write(np.int32(len(v)).tobytes() + v.tobytes() + np.int32(len(i)).tobytes() + i.tobytes())
    v, i = R.load_triangle_meshes(str(p))
    assert np.array_equal(v, np.concatenate([v1, v2])) and np.array_equal(i, np.concatenate([i1, i2 + len(v1)]))
    # broken index -> the reference's error text
    with open(tmp_path / "bad.tris", "wb") as f:
        f.write(np.int32(3).tobytes() + v2.tobytes() + np.int32(1).tobytes() + np.array([[0, 1, 3]], np.int32).tobytes())
    with pytest.raises(Exception, match="broken triangle model"):
        R.load_triangle_meshes(str(tmp_path / "bad.tris"))
    with pytest.raises(Exception, match="cannot open file"):
        R.load_triangle_meshes(str(tmp_path / "missing.tris"))


def test_raw_to_cells_matches_reference(R, orc):
    """exaRawToCells (tools/fromVolume/raw2cells.cpp) as a call: same cells and scalars, in the same
    order, as the reference tool wrote for these volumes (tests/golden/raw2cells.npz, made by
    tests/golden/make_golden.py from oracle/_ref/exaRawToCells) — including its byte-sum wrap-around
    and its level = int(log2f(width)) for refinement 3"""
    import sys
    sys.path.insert(0, GOLDEN)
    import make_golden
    g = np.load(os.path.join(GOLDEN, "raw2cells.npz"))
    levels = set()
    for i, (fmt, vol, thr, ml, ref) in enumerate(make_golden.raw2cells_cases()):
        cells, scal = R.raw_to_cells(vol, thr, ml, ref)
        assert np.array_equal(cells, g["cells%d" % i]) and np.array_equal(scal, g["scalars%d" % i]), (i, fmt)
        if ref == 2:                                             # the cells tile the volume (clipped at the far faces)
            lo, hi = cells[:, :3], np.minimum(cells[:, :3] + (1 << cells[:, 3:4]), np.array(vol.shape[::-1]))
            assert int(np.prod(hi - lo, axis=1).sum()) == vol.size
        levels |= set(np.unique(cells[:, 3]).tolist())
    assert len(levels) >= 3
    with pytest.raises(Exception, match="unknown format"):
        R.raw_to_cells(np.zeros((4, 4, 4), np.int16), 1.0, 2)


def test_config_expression_fields_match_reference(R, tmp_path):
    """`scalar <name> expr "<postfix>"` (Config.cpp:113-123, ScalarField::createFromExpression): every field
    of the config — loaded, vector magnitude and four expressions (**, comparison + select, log/abs/sqrt,
    division) — equals what the reference's own Config::parseConfigFile produced (tests/golden/
    config_expr.npz, made from oracle/_ref/ref_config), values and value ranges"""
    import sys
    sys.path.insert(0, GOLDEN)
    import make_golden
    g = np.load(os.path.join(GOLDEN, "config_expr.npz"))
    cfg = R.Config(make_golden.config_inputs(str(tmp_path)))
    assert len(cfg.fields) == 7
    for i, f in enumerate(cfg.fields):
        assert f["name"] == str(g["name%d" % i])
        assert np.array_equal(f["values"], g["values%d" % i], equal_nan=True), f["name"]
        assert np.array_equal(np.asarray(f["valueRange"], np.float32), g["range%d" % i], equal_nan=True), f["name"]
    # Config::finalize / getBounds (Config.cpp:23-55) with a non-trivial remap_from / remap_to
    assert np.array_equal(np.asarray(cfg.voxelSpaceTransform, np.float32), g["xfm"])
    lo, hi = cfg.get_bounds()
    assert np.array_equal(np.concatenate([lo, hi]).astype(np.float32), g["bounds"])
    v, i = cfg.triangles()                                   # `triangles <file>` (Config.cpp:165-171, TriangleMesh::load)
    assert np.array_equal(v, g["mesh_vertices"]) and np.array_equal(i, g["mesh_indices"]) and len(i) > 40
    open(tmp_path / "bad.exa", "w").write('scalar a a.scalars\nscalar e expr "%0 +"\n')
    with pytest.raises(Exception, match="insufficient operands"):
        R.Config(str(tmp_path / "bad.exa"))


def test_colormapper_matches_reference(R):
    """exa_colormap_eval == the reference's own ColorMapper::operator() (exa/ColorMapper.h:64-80, compiled in
    place: oracle/_ref/ref_colormap -> tests/golden/colormap.npz), bit for bit: below / above the range, between
    unevenly spaced points, at a duplicated position, with a single control point"""
    import sys
    sys.path.insert(0, GOLDEN)
    import make_golden
    g = np.load(os.path.join(GOLDEN, "colormap.npz"))
    for i, (cp, t) in enumerate(make_golden.colormap_cases()):
        cm = R.ColorMapper(control_points=cp)
        mine = np.stack([cm(float(x)) for x in t])
        assert np.array_equal(mine, g["rgb%d" % i]), "case %d: max diff %g" % (i, np.abs(mine - g["rgb%d" % i]).max())


def test_builder_types_and_kd_file_match_reference(R, built_lib, orc, tmp_path):
    """both builder types (SAH-alike default, --spatial-median) and the kd-tree file (`exaBuilder -kd`,
    builder.cpp:904-907; 12-byte KdTreeNode, exa/KdTree.h) are byte-identical to the reference exaBuilder's
    output (SHA-256 in golden_hashes.json; run live too when oracle/_ref is present)"""
    import subprocess
    import sys
    sys.path.insert(0, GOLDEN)
    import make_golden
    h = json.load(open(os.path.join(GOLDEN, "golden_hashes.json")))["builder_variants"]
    for name, cells, flags in make_golden.builder_variant_cases():
        bt = built_lib.BUILDER_SPATIAL_MEDIAN if flags else built_lib.BUILDER_SAH_ALIKE
        b = R.ExaBricks.build(cells, builder=bt)
        kd = b.kd_nodes.tobytes()
        assert hashlib.sha256(b.to_bytes()).hexdigest() == h[name]["bricks_sha256"], name
        assert hashlib.sha256(kd).hexdigest() == h[name]["kd_sha256"] and len(kd) // 12 == h[name]["kd_nodes"], name
        b.save(str(tmp_path / "o.bricks"), str(tmp_path / "o.kd"))                    # the files the tool writes
        assert open(tmp_path / "o.kd", "rb").read() == kd and open(tmp_path / "o.bricks", "rb").read() == b.to_bytes()
        if orc.have_ref():
            np.ascontiguousarray(cells, np.int32).tofile(tmp_path / "in.cells")
            subprocess.check_call([os.path.join(orc.REF_DIR, "exaBuilder"), str(tmp_path / "in.cells"), "-o", str(tmp_path / "r.bricks"),
                                   "-kd", str(tmp_path / "r.kd")] + flags, stdout=subprocess.DEVNULL)
            assert open(tmp_path / "r.kd", "rb").read() == kd and open(tmp_path / "r.bricks", "rb").read() == b.to_bytes()


def test_colormapper_embedded_maps_equal_the_reference_decode(R):
    """exa_colormap_embedded == what the reference's PNG constructor (stb_image) decodes from exa/embedded_colormaps.h,
    control point for control point (tests/golden/colormaps_embedded.npz, written by make_colormaps.py from
    oracle/_ref/ref_colormap --embedded; compared live too when the reference binaries are here)"""
    import subprocess
    sys_maps = np.load(os.path.join(GOLDEN, "colormaps_embedded.npz"))
    names = R.ColorMapper.names()
    assert len(names) == 16 and names[0] == "Paraview Cool Warm" and names[14] == "HSV" and names[15] == "Custom"
    sys.path.insert(0, GOLDEN)
    import make_colormaps
    for i, (name, sym) in enumerate(make_colormaps.MAPS):
        cm = R.ColorMapper.embedded(i)
        assert cm.name == name
        assert cm.values_.tobytes() == sys_maps[sym].tobytes()
    custom = R.ColorMapper.embedded(15)
    assert np.array_equal(custom.values_, np.array([[0, 0, 0, 0], [1, 1, 1, 1]], np.float32))
    if os.path.exists(make_colormaps.REF):
        import tempfile
        with tempfile.TemporaryDirectory() as d:
            subprocess.check_call([make_colormaps.REF, "--embedded", os.path.join(d, "e.bin")])
            for (name, sym), cp in zip(make_colormaps.MAPS, make_colormaps.read_maps(os.path.join(d, "e.bin"))):
                assert cp.tobytes() == sys_maps[sym].tobytes()


def test_colormapper_png_constructor(R):
    """the PNG constructor: our decoder on PNGs with stored, fixed-Huffman and dynamic-Huffman streams, every row filter,
    grey / RGB / RGBA / palette; on the re-encoded embedded maps of include/exa/embedded_colormaps.h; and - where the
    reference tree is present - on the reference's own PNG arrays, against its stb_image decode"""
    import re
    import struct
    import zlib
    sys.path.insert(0, GOLDEN)
    import make_colormaps

    def png(rows, ctype, level, filt=0, palette=None):
        h, w = rows.shape[0], rows.shape[1]
        ch = rows.shape[2] if rows.ndim == 3 else 1
        raw = b""
        prev = np.zeros(w * ch, np.int32)
        for y in range(h):
            cur = rows[y].reshape(-1).astype(np.int32)
            a = np.concatenate([np.zeros(ch, np.int32), cur[:-ch]])
            cc = np.concatenate([np.zeros(ch, np.int32), prev[:-ch]])
            if filt == 0: enc = cur
            elif filt == 1: enc = cur - a
            elif filt == 2: enc = cur - prev
            elif filt == 3: enc = cur - ((a + prev) >> 1)
            else:
                p = a + prev - cc
                pa, pb, pc = np.abs(p - a), np.abs(p - prev), np.abs(p - cc)
                pred = np.where((pa <= pb) & (pa <= pc), a, np.where(pb <= pc, prev, cc))
                enc = cur - pred
            raw += bytes([filt]) + (enc & 0xff).astype(np.uint8).tobytes()
            prev = cur

        def chunk(tag, data):
            return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xffffffff)
        if level == "fixed":
            co = zlib.compressobj(9, zlib.DEFLATED, 15, 9, zlib.Z_FIXED)
            z = co.compress(raw) + co.flush()
        else:
            z = zlib.compress(raw, level)
        out = b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, ctype, 0, 0, 0))
        if palette is not None:
            out += chunk(b"PLTE", palette.astype(np.uint8).tobytes())
        half = len(z) // 2                                        # two IDAT chunks: the stream spans chunk boundaries
        return out + chunk(b"IDAT", z[:half]) + chunk(b"IDAT", z[half:]) + chunk(b"IEND", b"")

    rng = np.random.RandomState(5)
    smooth = (np.linspace(0, 255, 300)[None, :, None] * np.array([1.0, 0.5, 0.25])[None, None, :]).astype(np.uint8)
    smooth = np.repeat(smooth, 3, axis=0)
    noise = rng.randint(0, 256, (2, 257, 3)).astype(np.uint8)
    for rows in (smooth, noise):
        w = rows.shape[1]
        want = np.concatenate([(np.arange(w, dtype=np.float32) / np.float32(w - 1))[:, None], rows[0] / np.float32(255.0)], axis=1).astype(np.float32)
        for level in (0, 6, 9, "fixed"):
            for filt in range(5):
                cm = R.ColorMapper(png=png(rows, 2, level, filt))
                assert cm.values_.tobytes() == want.tobytes(), (level, filt)
        rgba = np.concatenate([rows, np.full(rows.shape[:2] + (1,), 200, np.uint8)], axis=2)
        assert R.ColorMapper(png=png(rgba, 6, 6, 4)).values_.tobytes() == want.tobytes()
    grey = rng.randint(0, 256, (1, 100)).astype(np.uint8)
    g = R.ColorMapper(png=png(grey, 0, 6, 1)).values_
    assert np.array_equal(g[:, 1], grey[0] / np.float32(255.0)) and np.array_equal(g[:, 1], g[:, 3])
    pal = rng.randint(0, 256, (256, 3))
    idx = rng.randint(0, 256, (1, 64)).astype(np.uint8)
    pcm = R.ColorMapper(png=png(idx, 3, 6, 0, palette=pal)).values_
    assert np.array_equal(pcm[:, 1:], (pal[idx[0]] / np.float32(255.0)).astype(np.float32))
    with pytest.raises(Exception):
        R.ColorMapper(png=b"not a png at all")
    with pytest.raises(Exception):
        R.ColorMapper(png=png(smooth, 2, 6)[:-40])                 # truncated
    # hostile input: corrupted / truncated streams must be refused or decoded, never crash or balloon (the decoder stops
    # after the first pixel row); a header that claims a huge image over a tiny stream is refused
    base = bytearray(png(noise, 2, 6, 4))
    frng = np.random.RandomState(11)
    for _ in range(1500):
        m = bytearray(base)
        for _k in range(int(frng.randint(1, 6))):
            m[int(frng.randint(8, len(m)))] = int(frng.randint(0, 256))
        if frng.rand() < 0.2:
            m = m[:int(frng.randint(8, len(m)))]
        try:
            R.ColorMapper(png=bytes(m))
        except Exception:
            pass
    huge = bytearray(base); huge[16:24] = struct.pack(">II", 0x7fffffff, 0x7fffffff)
    with pytest.raises(Exception):
        R.ColorMapper(png=bytes(huge))
    # the generated header (what a C++ caller includes): each array decodes to the embedded map
    hdr = open(os.path.join(ROOT, "include", "exa", "embedded_colormaps.h")).read()

    def arrays(text):
        return {m.group(1): bytes(int(x, 16) for x in re.findall(r"0x[0-9a-fA-F]+", m.group(2)))
                for m in re.finditer(r"uint8_t\s+(\w+)\[\]\s*=\s*\{([^}]*)\}", text)}
    ours = arrays(hdr)
    assert len(ours) == 15
    for i, (name, sym) in enumerate(make_colormaps.MAPS):
        assert R.ColorMapper(png=ours[sym]).values_.tobytes() == R.ColorMapper.embedded(i).values_.tobytes()
    ref_hdr = "/root/reference/exa/embedded_colormaps.h"
    if os.path.exists(ref_hdr):                                     # the reference's own (deflate-compressed) PNG bytes, read where they lie
        theirs = arrays(open(ref_hdr).read())
        for i, (name, sym) in enumerate(make_colormaps.MAPS):
            assert theirs[sym] != ours[sym]                         # not a copy
            assert R.ColorMapper(png=theirs[sym]).values_.tobytes() == R.ColorMapper.embedded(i).values_.tobytes()


def test_colormapper_string_constructor_matches_reference(R, tmp_path):
    """ColorMapper(std::string) (exa/ColorMapper.cpp:92-131) vs the reference's own constructor where it is built"""
    import subprocess
    sys.path.insert(0, GOLDEN)
    import make_colormaps
    cases = ["0.0,(0.0,0.1,0.2)\n1.0,(1.0,0.9,0.8)",
             "1.0,(1.0,0.9,0.8)\n0.25,(0.5,0.5,0.5)\n\n0.0,(0.0,0.1,0.2)\n",
             "  0.5,(0.1, 0.2 ,0.3)  \n\tgarbage line\n0.1,(1,2,3)\n0.7,1,2\n",
             ""]
    want = [np.array([[0, 0, .1, .2], [1, 1, .9, .8]], np.float32),
            np.array([[0, 0, .1, .2], [.25, .5, .5, .5], [1, 1, .9, .8]], np.float32),
            np.array([[.1, 1, 2, 3], [.5, .1, .2, .3]], np.float32), np.zeros((0, 4), np.float32)]
    for sidx, (text, w) in enumerate(zip(cases, want)):
        cm = R.ColorMapper(string=text)
        assert np.array_equal(cm.values_, w), sidx
        if os.path.exists(make_colormaps.REF):
            open(tmp_path / "s.txt", "w").write(text)
            subprocess.check_call([make_colormaps.REF, "--string", str(tmp_path / "s.txt"), str(tmp_path / "o.bin")], stdout=subprocess.DEVNULL)
            ref = make_colormaps.read_maps(str(tmp_path / "o.bin"))[0]
            assert ref.tobytes() == cm.values_.tobytes(), sidx


def test_xf_file_roundtrip(R, tmp_path):
    """.xf files (exa/viewer.cpp:139-145): 128 float32 opacities"""
    a = (np.arange(128) / 127.0).astype(np.float32) ** 2
    R.save_xf(str(tmp_path / "t.xf"), a)
    assert os.path.getsize(tmp_path / "t.xf") == 512 and np.array_equal(R.load_xf(str(tmp_path / "t.xf")), a)
    with pytest.raises(RuntimeError):
        R.save_xf(str(tmp_path / "bad.xf"), a[:100])


def test_cpp_class_wrapper_over_the_c_abi(R, built_lib, tmp_path):
    """include/exa/exa.hpp — the reference's C++ class names (Config, ExaBricks, ScalarField, ColorMapper,
    TriangleMesh, OptixRenderer) over the C ABI — compiles with g++ -Wall -Wextra as a reference-side caller would
    use it (tests/harness/exa_hpp_host.cpp), links against the library alone, and sees exactly what the Python
    mirror sees on the config fixture; without a CUDA device its renderer constructor throws (no CPU fallback)"""
    import subprocess
    import sys
    import torch
    sys.path.insert(0, GOLDEN)
    import make_golden
    libdir = os.path.join(ROOT, "owlexabrick_b200")
    exe = str(tmp_path / "exa_hpp_host")
    cc = subprocess.run(["g++", "-std=c++14", "-O1", "-Wall", "-Wextra", "-Wno-comment", "-Werror", "-I" + os.path.join(ROOT, "include"),
                         os.path.join(ROOT, "tests", "harness", "exa_hpp_host.cpp"), "-o", exe,
                         "-L" + libdir, "-lexabrick_b200", "-Wl,-rpath," + libdir], capture_output=True, text=True)
    assert cc.returncode == 0, cc.stderr[-2000:]
    cfg_path = make_golden.config_inputs(str(tmp_path))
    run = subprocess.run([exe, cfg_path], capture_output=True, text=True, timeout=120)
    assert run.returncode == 0, run.stdout + run.stderr
    out = {ln.split(" ", 1)[0]: ln.split(" ", 1)[1] for ln in run.stdout.splitlines() if " " in ln and not ln.startswith("field ")}
    cfg = R.Config(cfg_path)
    fields = [ln.split() for ln in run.stdout.splitlines() if ln.startswith("field ")]
    assert int(out["fields"]) == len(cfg.fields) == len(fields)
    for f, (_, name, n, lo, hi) in zip(cfg.fields, fields):
        assert f["name"] == name and len(f["values"]) == int(n)
        assert np.float32(lo) == np.float32(f["valueRange"][0]) and np.float32(hi) == np.float32(f["valueRange"][1])
    lo, hi = cfg.get_bounds()
    assert np.array_equal(np.array(out["bounds"].split(), np.float32), np.concatenate([lo, hi]).astype(np.float32))
    v, i = cfg.triangles()
    assert out["surfaces"].split() == [str(len(v)), str(len(i))]
    assert np.array_equal(np.array(out["colormap"].split(), np.float32), np.array([0.5, 0, 0], np.float32))
    emb = R.ColorMapper.embedded(0)
    assert out["colormap_png"].split()[0] == str(len(emb.values_))
    assert np.array_equal(np.array(out["colormap_png"].split()[1:], np.float32), emb(0.3))
    st = R.ColorMapper(string="1.0,(1.0,0.9,0.8)\n0.0,(0.0,0.1,0.2)\nnot a point\n")
    assert out["colormap_string"].split()[0] == "2" and np.array_equal(np.array(out["colormap_string"].split()[1:], np.float32), st(0.5))
    a = cfg.fields[0]["valueRange"]
    assert np.array_equal(np.array(out["expr"].split()[1:], np.float32), np.array([2 * a[0], 2 * a[1]], np.float32))
    if torch.cuda.is_available():
        assert out["renderer"] == "ok"
    else:
        assert out["renderer"].startswith("error: no usable CUDA device")


def test_builder_c5_shape_quarter_scale_hash(R, orc):
    """26.6 M cells of the C5 shape (synth.cloud_shaped(4), 4 levels): the OpenMP host builder's .bricks bytes hash to
    what the reference's serial exaBuilder wrote (41 s there; tests/golden/host_stage_baselines.py --config c5s4), and
    the oracle's ABR build over them with 3 channels (251 744 regions) to the reference's ExaBrickRegions::buildFrom"""
    h = json.load(open(os.path.join(GOLDEN, "golden_hashes.json")))["C5_scale4"]
    cells, dom = synth.cloud_shaped(4)
    assert len(cells) == h["cells"] and dom == 512
    b = R.ExaBricks.build(cells)
    assert b.num_bricks == h["bricks"] and b.total_num_cells == len(cells)
    assert hashlib.sha256(b.to_bytes()).hexdigest() == h["bricks_sha256"]
    fields = [np.empty(len(cells), np.float32) for _ in range(3)]
    for s0 in range(0, len(cells), 1 << 22):                 # chunked: the generator's temporaries stay small
        for c, v in enumerate(synth.cloud_velocity(cells[s0:s0 + (1 << 22)], dom)):
            fields[c][s0:s0 + (1 << 22)] = v
    regions, leaf = orc.build_abr(b.headers.view(orc.BRICK_DT), orc.flatten_scalars(b.cell_ids, fields))
    assert len(regions) == h["regions"]
    assert hashlib.sha256(regions.tobytes() + leaf.tobytes()).hexdigest() == h["abr_sha256"]
