#!/usr/bin/env python
"""Where does a frame's wall time go on a tiny scene: host enqueue vs. GPU (GPU box only)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import bench
from owlexabrick_b200 import capi, renderer as R, synth

domain, boxes, image = bench.WORKLOADS["small"]
cells = synth.nested_boxes(domain, boxes); field = synth.analytic_field(cells, domain, 0)
r = R.Renderer(R.ExaBricks.build(cells), [field])
stream = torch.cuda.current_stream(); r.set_stream(stream.cuda_stream)
for img in ((256, 256), (64, 64), (1024, 1024)):
    fd = bench.frame_desc(R, domain, img, (float(field.min()), float(field.max()))); fd.apply(r)
    for _ in range(20): r.render(sync=False)
    torch.cuda.synchronize()
    n = 500
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t = time.perf_counter(); e0.record(stream)
    for _ in range(n): r.render(sync=False)
    t_enq = time.perf_counter() - t
    e1.record(stream); torch.cuda.synchronize(); t_all = time.perf_counter() - t
    print("image %s: host enqueue %.1f us/frame, wall %.1f us/frame, events %.1f us/frame" % (img, 1e6 * t_enq / n, 1e6 * t_all / n, 1e3 * e0.elapsed_time(e1) / n), flush=True)
    # one frame at a time (latency)
    t = time.perf_counter()
    for _ in range(100): r.render(sync=True)
    print("   synchronous render: %.1f us/frame" % (1e6 * (time.perf_counter() - t) / 100), flush=True)
