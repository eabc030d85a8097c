#!/usr/bin/env python
"""Per-phase cycle counts of a few CTAs of the frame kernel (needs the EXA_PHASE_TIMING variant build)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch, bench
from owlexabrick_b200 import renderer as R, synth
name = sys.argv[1] if len(sys.argv) > 1 else "small"
domain, boxes, image = bench.WORKLOADS[name]
cells = synth.nested_boxes(domain, boxes); field = synth.analytic_field(cells, domain, 0)
r = R.Renderer(R.ExaBricks.build(cells), [field])
fd = bench.frame_desc(R, domain, image, (float(field.min()), float(field.max()))); fd.apply(r)
for _ in range(5): r.render()
r.updateFrameID(77); r.render(); r.updateFrameID(0)
torch.cuda.synchronize()
