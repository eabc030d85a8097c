"""Generate tests/golden/*.npz from the reference's OWN code (oracle/_ref, built by
`make -C oracle ref` from /root/reference in place).  Run in the build container only:

    python tests/golden/make_golden.py

Each fixture holds the input (cells, scalar fields) and the reference's outputs:
the bytes of the .bricks file written by exaBuilder (builder/builder.cpp:895-902) and the
regions/leafList produced by ExaBrickRegions::buildFrom (exa/Regions.cpp:242-320) dumped
by oracle/ref_abr_main.cpp.  Large cases are pinned by SHA-256 only (golden_hashes.json).
"""
import hashlib, json, os, sys, tempfile
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import orc                     # noqa: E402
from owlexabrick_b200 import synth         # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
REF_GRIDS = "/root/reference/tools/artificial"


def cases():
    for i in range(5):
        c, s = synth.grids_to_cells(open("%s/ex%d.grids" % (REF_GRIDS, i)).read())
        yield "ex%d" % i, c, [s], 127, True
    c = synth.nested_boxes(64, [(18, 46), (8, 56)])
    yield "nb64", c, [synth.analytic_field(c, 64, 0), synth.analytic_field(c, 64, 1)], 127, True
    yield "nb64_mlw8", c, [synth.analytic_field(c, 64, 0)], 8, True
    # duplicates + a shuffled order exercise the dedup path (builder.cpp:301-350)
    rng = np.random.RandomState(7)
    c2 = np.concatenate([c, c[rng.randint(0, len(c), 500)]])
    perm = rng.permutation(len(c2))
    c2 = c2[perm]
    yield "nb64_dups", c2, [synth.analytic_field(c2, 64, 0)], 127, True
    c = synth.nested_boxes(256, [(72, 184), (32, 224)])          # SURVEY config C1
    yield "C1", c, [synth.analytic_field(c, 256, 0)], 127, False
    c, dom = synth.cloud_shaped(16)                               # SURVEY config C5's shape at 1/16 scale, 3 channels
    yield "C5_scale16", c, synth.cloud_velocity(c, dom), 127, False


def main():
    assert orc.have_ref(), "run `make -C oracle ref` first (needs /root/reference)"
    hashes = {}
    for name, cells, fields, mlw, store in cases():
        with tempfile.TemporaryDirectory() as d:
            # exaArtificial is also exercised for the ex* cases: its .cells must equal ours
            if name.startswith("ex"):
                import subprocess
                subprocess.check_call([os.path.join(orc.REF_DIR, "exaArtificial"), "%s/%s.grids" % (REF_GRIDS, name),
                                       "-o", os.path.join(d, "a")], stdout=subprocess.DEVNULL)
                assert np.array_equal(np.fromfile(os.path.join(d, "a.cells"), np.int32).reshape(-1, 4), cells)
                assert np.array_equal(np.fromfile(os.path.join(d, "a.scalars"), np.float32), fields[0])
            bricks = orc.ref_build_bricks(cells, d, None if mlw == 127 else mlw)
            abr = orc.ref_build_abr(bricks, fields, d)
        hashes[name] = {"cells": len(cells), "max_leaf_width": mlw,
                        "bricks_sha256": hashlib.sha256(bricks).hexdigest(),
                        "abr_sha256": hashlib.sha256(abr).hexdigest()}
        if store:
            np.savez_compressed(os.path.join(HERE, name + ".npz"), cells=cells, fields=np.stack(fields),
                                bricks=np.frombuffer(bricks, np.uint8), abr=np.frombuffer(abr, np.uint8),
                                max_leaf_width=mlw)
        print(name, hashes[name])
    json.dump(hashes, open(os.path.join(HERE, "golden_hashes.json"), "w"), indent=1, sort_keys=True)
    raw2cells_golden()
    ref_shader_golden()
    ref_shader_cd_golden()
    points_golden(hashes)
    config_golden()
    colormap_golden()
    builder_variants_golden(hashes)
    fuzz_golden(hashes)
    fuzz_frames_golden(hashes)
    fuzz_points_golden(hashes)
    json.dump(hashes, open(os.path.join(HERE, "golden_hashes.json"), "w"), indent=1, sort_keys=True)
    json.dump(hashes, open(os.path.join(HERE, "golden_hashes.json"), "w"), indent=1, sort_keys=True)


def raw2cells_cases():
    """small structured volumes for exaRawToCells (tools/fromVolume/raw2cells.cpp): (format, volume[z,y,x], threshold, maxLevel, refinement)"""
    def smooth(shape, bump):
        z, y, x = np.meshgrid(*[np.arange(n) for n in shape], indexing="ij")
        v = 100 + 40 * np.sin(x / 9.0) * np.cos(y / 11.0) + 20 * np.sin(z / 6.0)
        if bump:
            v[2:4, 3:6, 1:5] += 90
        return np.clip(v, 0, 255)
    yield "byte", smooth((13, 17, 20), True).astype(np.uint8), 8.0, 3, 2
    yield "byte", smooth((16, 16, 32), False).astype(np.uint8), 200.0, 4, 2      # whole root cells collapse: byte sums wrap
    yield "float", smooth((9, 16, 11), True).astype(np.float32), 5.5, 2, 2
    yield "double", smooth((8, 12, 8), False).astype(np.float64), 3.0, 3, 2
    yield "byte", smooth((10, 9, 27), True).astype(np.uint8), 20.0, 2, 3         # refinement 3: level = int(log2f(3^l))


def raw2cells_golden():
    import subprocess
    out = {}
    for i, (fmt, vol, thr, ml, ref) in enumerate(raw2cells_cases()):
        with tempfile.TemporaryDirectory() as d:
            vol.tofile(os.path.join(d, "v.raw"))
            nz, ny, nx = vol.shape
            cmd = [os.path.join(orc.REF_DIR, "exaRawToCells"), str(nx), str(ny), str(nz), fmt, os.path.join(d, "v.raw"),
                   os.path.join(d, "out"), str(thr), str(ml)] + ([str(ref)] if ref != 2 else [])
            subprocess.check_call(cmd, stdout=subprocess.DEVNULL)
            out["cells%d" % i] = np.fromfile(os.path.join(d, "out.cells"), np.int32).reshape(-1, 4)
            out["scalars%d" % i] = np.fromfile(os.path.join(d, "out.scalars"), np.float32)
        print("raw2cells", i, fmt, vol.shape, len(out["scalars%d" % i]), "cells, levels", np.unique(out["cells%d" % i][:, 3]))
    np.savez_compressed(os.path.join(HERE, "raw2cells.npz"), **out)



def ref_shader_cd_golden():
    """the gradient-shaded frames again through the reference built with -DANALYTIC_GRADIENTS=0 (oracle/_ref/ref_shader_cd)"""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import ref_shader_cases as RC
    assert orc.have_ref_shader_cd(), "run `make -C oracle ref` first"
    out = {}
    for name, kind, channels, mlw, configure in RC.cd_cases():
        raw, fields, dom, _ = RC.build(orc, kind, channels, mlw)
        fd = RC.frame(dom, channels, configure)
        with tempfile.TemporaryDirectory() as d:
            acc, rgba = orc.ref_render(raw, fields, fd.to_orc(orc), d, binary="ref_shader_cd")
        out[name + "_accum"], out[name + "_rgba"] = acc, rgba
        print("ref_shader_cd", name, "max colour %.3f" % acc[..., :3].max())
    np.savez_compressed(os.path.join(HERE, "ref_shader_frames_cd.npz"), **out)


def ref_shader_golden():
    """frames rendered by the reference's own device source compiled as C++ (oracle/_ref/ref_shader)"""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import ref_shader_cases as RC
    assert orc.have_ref_shader(), "run `make -C oracle ref` first"
    out = {}
    for name, kind, channels, mlw, configure in RC.cases(extra=True):
        raw, fields, dom, _ = RC.build(orc, kind, channels, mlw)
        fd = RC.frame(dom, channels, configure)
        acc0 = RC.start_accum(name) if fd.frame_id > 0 else None
        with tempfile.TemporaryDirectory() as d:
            acc, rgba = orc.ref_render(raw, fields, fd.to_orc(orc), d, accum=acc0)
        out[name + "_accum"], out[name + "_rgba"] = acc, rgba
        print("ref_shader", name, "max colour %.3f" % acc[..., :3].max(), "lit pixels", int((acc[..., :3].max(axis=-1) > 0).sum()))
    cells, fields, dom, seeds, T, steplen = RC.tracer_case()
    with tempfile.TemporaryDirectory() as d:
        out["tracer"] = orc.ref_trace(orc.build_bricks(cells), fields, seeds, T, (0, 1, 2), steplen, [0, 0, 0], [dom] * 3, d)
        ch, sl, lo, hi = RC.tracer_case_b()
        out["tracer_b"] = orc.ref_trace(orc.build_bricks(cells), fields, seeds, T, ch, sl, lo, hi, d)
    print("ref_shader tracer: dead vertices", int((out["tracer"][..., 0] >= 2e10).sum()), "of", out["tracer"].shape[0] * T)
    np.savez_compressed(os.path.join(HERE, "ref_shader_frames.npz"), **out)


def points_golden(hashes):
    """north_star check 2: the reference's reconstruction at 1 M random points of config C1, by its own
    samplePoint behind its own point query (programs/exabrick.cu compiled as C++) — SHA-256 of the values
    + validity, and the first 4096 values in the clear"""
    c = synth.nested_boxes(256, [(72, 184), (32, 224)])
    f = synth.analytic_field(c, 256, 0)
    pts = synth.random_points(1000000, [0.0] * 3, [256.0] * 3, seed=1)
    with tempfile.TemporaryDirectory() as d:
        v, ok, g = orc.ref_sample_points(orc.build_bricks(c), [f], pts, 0, d)
    hashes["C1_points_1M"] = {"values_sha256": hashlib.sha256(v.tobytes()).hexdigest(), "valid_sha256": hashlib.sha256(ok.tobytes()).hexdigest(),
                              "grad_sha256": hashlib.sha256(g.tobytes()).hexdigest(), "valid": int(ok.sum())}
    np.savez_compressed(os.path.join(HERE, "c1_points_head.npz"), values=v[:4096], valid=ok[:4096], grad=g[:4096])
    print("C1 1M points", hashes["C1_points_1M"])


CONFIG_TEXT = """# expression fields (Config.cpp:113-123, ScalarField.cpp:96-230)
remap_from 0.5 -1 2 8.5 7 6.25
remap_to   -1.25 -1 -3  1 1.5 0.75
scalar a a.scalars
scalar b b.scalars
vector v a.scalars b.scalars a.scalars
scalar e0 expr "%0 2 ** %1 +"
scalar e1 expr "%0 %1 < %0 %1 select"
scalar e2 expr "%0 1 + log %1 abs sqrt * 0.25 -"
scalar e3 expr "%2 %3 / %4 >= 3 *"
bricks ex4.bricks
triangles m.tris
"""


def config_inputs(workdir):
    g = np.load(os.path.join(HERE, "ex4.npz"))
    open(os.path.join(workdir, "ex4.bricks"), "wb").write(g["bricks"].tobytes())
    n = len(g["fields"][0])
    a = (np.sin(np.arange(n) * 0.37) * 2 + 0.5).astype(np.float32)
    b = (np.cos(np.arange(n) * 0.11) * 3 - 0.25).astype(np.float32)
    a.tofile(os.path.join(workdir, "a.scalars"))
    b.tofile(os.path.join(workdir, "b.scalars"))
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import scenes
    with open(os.path.join(workdir, "m.tris"), "wb") as f:                 # two meshes in one file (TriangleMesh.cpp:21-68)
        for v, i in (scenes.uv_sphere([1, 2, 3], 2.0, 8, 4), (np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], np.float32), np.array([[0, 1, 2]], np.int32))):
            f.write(np.int32(len(v)).tobytes() + v.tobytes() + np.int32(len(i)).tobytes() + i.tobytes())
    p = os.path.join(workdir, "c.exa")
    open(p, "w").write(CONFIG_TEXT)
    return p


def parse_ref_config(raw):
    off, out = 4, []
    n = int(np.frombuffer(raw[:4], "<i4")[0])
    for _ in range(n):
        nl = int(np.frombuffer(raw[off:off + 4], "<i4")[0]); off += 4
        name = raw[off:off + nl].decode(); off += nl
        cnt = int(np.frombuffer(raw[off:off + 8], "<i8")[0]); off += 8
        rng = np.frombuffer(raw[off:off + 8], "<f4").copy(); off += 8
        vals = np.frombuffer(raw[off:off + 4 * cnt], "<f4").copy(); off += 4 * cnt
        out.append((name, rng, vals))
    out.append(("__xfm_bounds__", None, np.frombuffer(raw[off:off + 72], "<f4").copy())); off += 72
    nv = int(np.frombuffer(raw[off:off + 4], "<i4")[0]); off += 4
    verts = np.frombuffer(raw[off:off + 12 * nv], "<f4").reshape(-1, 3).copy(); off += 12 * nv
    nt = int(np.frombuffer(raw[off:off + 4], "<i4")[0]); off += 4
    idx = np.frombuffer(raw[off:off + 12 * nt], "<i4").reshape(-1, 3).copy()
    out.append(("__mesh__", verts, idx))
    return out


def config_golden():
    import subprocess
    with tempfile.TemporaryDirectory() as d:
        p = config_inputs(d)
        subprocess.check_call([os.path.join(orc.REF_DIR, "ref_config"), p, os.path.join(d, "out.bin")], stdout=subprocess.DEVNULL)
        fields = parse_ref_config(open(os.path.join(d, "out.bin"), "rb").read())
    out = {}
    out["xfm"], out["bounds"] = fields[-2][2][:12], fields[-2][2][12:]
    out["mesh_vertices"], out["mesh_indices"] = fields[-1][1], fields[-1][2]
    print("ref_config voxelSpaceTransform", out["xfm"], "bounds", out["bounds"], "mesh", out["mesh_vertices"].shape, out["mesh_indices"].shape)
    for i, (name, rng, vals) in enumerate(fields[:-2]):
        out["name%d" % i], out["range%d" % i], out["values%d" % i] = np.array(name), rng, vals
        print("ref_config", name, rng, len(vals))
    np.savez_compressed(os.path.join(HERE, "config_expr.npz"), **out)


def colormap_cases():
    rng = np.random.RandomState(9)
    t = np.concatenate([np.linspace(-0.2, 1.2, 57), rng.uniform(0, 1, 200)]).astype(np.float32)
    yield np.array([[0, 0, 0, 0], [0.5, 1, 0, 0], [1, 1, 1, 1]], np.float32), t
    cp = np.concatenate([np.sort(rng.uniform(0, 1, 9))[:, None], rng.uniform(0, 1, (9, 3))], axis=1).astype(np.float32)
    yield cp, t
    yield np.array([[0.3, 0.2, 0.4, 0.6], [0.3, 0.9, 0.1, 0.5], [0.8, 0.0, 1.0, 0.25]], np.float32), t     # duplicate position
    yield np.array([[0.5, 0.1, 0.2, 0.3]], np.float32), t                                                   # single point


def colormap_golden():
    """the reference's own ColorMapper::operator() (exa/ColorMapper.h:64-80) on a few control-point sets"""
    import subprocess
    out = {}
    for i, (cp, t) in enumerate(colormap_cases()):
        with tempfile.TemporaryDirectory() as d:
            with open(os.path.join(d, "in.bin"), "wb") as f:
                f.write(np.int32(len(cp)).tobytes() + cp.tobytes() + np.int32(len(t)).tobytes() + t.tobytes())
            subprocess.check_call([os.path.join(orc.REF_DIR, "ref_colormap"), os.path.join(d, "in.bin"), os.path.join(d, "out.bin")])
            out["rgb%d" % i] = np.fromfile(os.path.join(d, "out.bin"), np.float32).reshape(-1, 3)
    np.savez_compressed(os.path.join(HERE, "colormap.npz"), **out)
    print("ref_colormap:", len(out), "cases")


def builder_variant_cases():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import scenes
    for kind in ("nb64_off", "nb32"):
        cells = scenes.small_scene(kind)[0]
        yield kind + "_sah", cells, []
        yield kind + "_median", cells, ["--spatial-median"]


def builder_variants_golden(hashes):
    """exaBuilder with its kd-tree output (-kd, builder.cpp:904-907) for both builder types (builder.cpp:31-33,858-860)"""
    import subprocess
    out = {}
    for name, cells, flags in builder_variant_cases():
        with tempfile.TemporaryDirectory() as d:
            np.ascontiguousarray(cells, np.int32).tofile(os.path.join(d, "in.cells"))
            subprocess.check_call([os.path.join(orc.REF_DIR, "exaBuilder"), os.path.join(d, "in.cells"), "-o", os.path.join(d, "o.bricks"),
                                   "-kd", os.path.join(d, "o.kd")] + flags, stdout=subprocess.DEVNULL)
            b, k = open(os.path.join(d, "o.bricks"), "rb").read(), open(os.path.join(d, "o.kd"), "rb").read()
        out[name] = {"bricks_sha256": hashlib.sha256(b).hexdigest(), "kd_sha256": hashlib.sha256(k).hexdigest(), "kd_nodes": len(k) // 12}
        print("builder variant", name, out[name])
    hashes["builder_variants"] = out


def fuzz_golden(hashes):
    """reference exaBuilder + ExaBrickRegions::buildFrom on the random octree AMR of tests/scenes.py:random_amr"""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import scenes
    hashes["fuzz"] = {}
    for seed in range(12):
        cells, mlw, fields = scenes.random_amr(seed)
        with tempfile.TemporaryDirectory() as d:
            rb = orc.ref_build_bricks(cells, d, None if mlw == 127 else mlw)
            rr, rl = orc.parse_abr(orc.ref_build_abr(rb, fields, d))
        hashes["fuzz"][str(seed)] = {"cells": len(cells), "max_leaf_width": mlw, "channels": len(fields), "regions": len(rr),
                                     "bricks_sha256": hashlib.sha256(rb).hexdigest(),
                                     "abr_sha256": hashlib.sha256(rr.tobytes() + rl.tobytes()).hexdigest()}
        print("fuzz", seed, hashes["fuzz"][str(seed)])


def fuzz_frames_golden(hashes):
    """the reference shader's frames on tests/scenes.py:random_amr_frame"""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import scenes
    hashes["fuzz_frames"] = {}
    for seed in range(8):
        cells, mlw, fields, fd = scenes.random_amr_frame(seed)
        with tempfile.TemporaryDirectory() as d:
            acc, rgba = orc.ref_render(orc.ref_build_bricks(cells, d, None if mlw == 127 else mlw), fields, fd.to_orc(orc), d)
        hashes["fuzz_frames"][str(seed)] = {"lit_pixels": int((acc[..., :3].max(axis=-1) > 0.02).sum()),
                                            "frame_sha256": hashlib.sha256(acc.tobytes() + rgba.tobytes()).hexdigest()}
        print("fuzz frame", seed, hashes["fuzz_frames"][str(seed)])


def fuzz_points_golden(hashes):
    """the reference's reconstruction at the points of tests/scenes.py:random_amr_points"""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import scenes
    hashes["fuzz_points"] = {}
    for seed in range(8):
        cells, mlw, fields, pts, ch = scenes.random_amr_points(seed)
        with tempfile.TemporaryDirectory() as d:
            v, ok, g = orc.ref_sample_points(orc.ref_build_bricks(cells, d, None if mlw == 127 else mlw), fields, pts, ch, d)
        hashes["fuzz_points"][str(seed)] = {"valid": int(ok.sum()), "points_sha256": hashlib.sha256(v.tobytes() + ok.tobytes() + g.tobytes()).hexdigest()}
        print("fuzz points", seed, hashes["fuzz_points"][str(seed)])


if __name__ == "__main__":
    main()
