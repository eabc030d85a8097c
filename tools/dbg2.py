import sys, os, numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
from owlexabrick_b200 import renderer as R, synth
from oracle import orc
import scenes
from test_gpu_parity import make_pair
cells, fields, dom = scenes.small_scene("nb32", 1)
r, sc, _ = make_pair(R, orc, cells, fields)
W,H=96,64
fd = scenes.FrameDesc(W,H,R.default_camera([0]*3,[dom]*3,W,H),1)
fd.ert=2.0; fd.space_skipping=False; fd.dt=0.8; fd.opacity_scale=0.05
fd.apply(r); r.render()
sys.stdout.flush()
o,_,_ = sc.render(fd.to_orc(orc), window=(50,47,51,48))
