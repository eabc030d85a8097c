#!/usr/bin/env python
"""BASELINE.md rows B1 / B2 on the host: wall time of the reference's own exaBuilder and
ExaBrickRegions::buildFrom (oracle/_ref, compiled in place from /root/reference — so this runs only
where that exists) on config C1, next to the product's host builder (exa_bricks_build) on the same
cells.  The binaries' own file I/O is inside their wall time (tmpfs); outputs are compared bit for bit.
Prints one JSON line.  Test infrastructure (lives beside make_golden.py, which runs the same binaries): a baseline measurement,
not part of the product.

    python tests/golden/host_stage_baselines.py [--config c1|c5s16|c5s4]
"""
import argparse, hashlib, json, os, subprocess, sys, tempfile, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
REF = os.path.join(ROOT, "oracle", "_ref")               # the reference binaries (not the restatement)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="c1", choices=["c1", "c5s16", "c5s4"])
    a = ap.parse_args()
    from owlexabrick_b200 import renderer as R, synth
    if a.config == "c1":
        cells = synth.nested_boxes(256, [(72, 184), (32, 224)])
        fields = [synth.analytic_field(cells, 256, 0)]
    else:
        cells, dom = synth.cloud_shaped(16 if a.config == "c5s16" else 4)
        fields = synth.cloud_velocity(cells, dom)
    with tempfile.TemporaryDirectory(dir="/dev/shm" if os.path.isdir("/dev/shm") else None) as d:
        cf, bf, of = (os.path.join(d, n) for n in ("a.cells", "a.bricks", "a.abr"))
        np.ascontiguousarray(cells, np.int32).tofile(cf)
        t = time.time()
        subprocess.check_call([os.path.join(REF, "exaBuilder"), cf, "-o", bf], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        t_ref_build = time.time() - t
        ref_bricks = open(bf, "rb").read()
        sfs = []
        for i, f in enumerate(fields):
            p = os.path.join(d, "s%d.scalars" % i)
            np.ascontiguousarray(f, np.float32).tofile(p)
            sfs.append(p)
        t = time.time()
        subprocess.check_call([os.path.join(REF, "ref_abr"), bf, of] + sfs, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        t_ref_abr = time.time() - t
        nr = int(np.frombuffer(open(of, "rb").read(8), "<i4")[0])
    ts = []
    for _ in range(3):
        t = time.time()
        b = R.ExaBricks.build(cells)
        ts.append(time.time() - t)
    same = b.to_bytes() == ref_bricks
    out = {"config": a.config, "cells": int(len(cells)), "bricks": int(b.num_bricks), "regions": nr, "host_cores": os.cpu_count(),
           "reference_exaBuilder": {"wall_s": round(t_ref_build, 3), "cells_per_s": round(len(cells) / t_ref_build), "threads": 1, "note": "serial, as shipped; includes reading .cells and writing .bricks on tmpfs"},
           "reference_buildFrom": {"wall_s": round(t_ref_abr, 3), "regions_per_s": round(nr / t_ref_abr), "note": "ExaBrickRegions::buildFrom incl. ExaBricks::load and the scalar gather; serial recursion, value ranges via OWL parallel_for (serial build: -DOWL_DISABLE_TBB)"},
           "product_exa_bricks_build": {"wall_s": round(min(ts), 3), "cells_per_s": round(len(cells) / min(ts)), "threads": os.cpu_count(), "bit_identical_to_reference": bool(same)},
           "bricks_sha256": hashlib.sha256(ref_bricks).hexdigest()}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
