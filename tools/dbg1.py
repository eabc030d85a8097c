import sys, numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
from owlexabrick_b200 import renderer as R, synth
from oracle import orc
import scenes
from test_gpu_parity import make_pair
cells, fields, dom = scenes.small_scene("nb32", 1)
r, sc, _ = make_pair(R, orc, cells, fields)
W,H=96,64
fd = scenes.FrameDesc(W,H,R.default_camera([0]*3,[dom]*3,W,H),1)
for cfg in [dict(ert=2.0, ss=False, dt=0.8, op=0.05), dict(ert=2.0, ss=True, dt=0.5, op=0.05), dict(ert=0.98, ss=False, dt=0.8, op=0.05), dict(ert=2.0, ss=False, dt=0.5, op=0.3)]:
    fd.ert=cfg["ert"]; fd.space_skipping=cfg["ss"]; fd.dt=cfg["dt"]; fd.opacity_scale=cfg["op"]
    fd.apply(r); gs = r.render_stats(); g = r.read_accum()
    o, _, os_ = sc.render(fd.to_orc(orc))
    err = np.abs(g[...,:3]-o[...,:3]).max(axis=-1)
    bad = np.argwhere(err>1e-3)
    print(cfg, "maxerr", err.max(), "nbad", len(bad), "gpu", gs, "orc", os_)
    print(" bad px", bad[:8].tolist(), [ (g[y,x,:3].tolist(), o[y,x,:3].tolist()) for y,x in bad[:3]])
