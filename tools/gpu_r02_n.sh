#!/bin/bash
# round 2, GPU call N (1 GPU): A/B of the frame-kernel prologue (tile look-up by warp 0), the sRGB fast path and the clockScale branch
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
python - <<'PY'
import sys; sys.path.insert(0, ".")
import bench
from owlexabrick_b200 import renderer as R
bench.load_or_build_scene("c3", R, gpu_builder=True)
PY
run() { tag=$1; shift; env "$@" PQ_CLEAN_FLUSH=1 timeout 300 python tools/perf_quick.py --workload c3 --tag "$tag" > gpurun_out/n_pq_$tag.json 2> gpurun_out/n_pq_$tag.err; python - "$tag" <<'PY'
import json, sys
try:
    d = json.load(open("gpurun_out/n_pq_%s.json" % sys.argv[1])); print(sys.argv[1], {k: d[k] for k in ("opaque", "translucent") if k in d})
except Exception as e:
    print(sys.argv[1], "no result", e)
PY
}
for rep in 1 2; do
run cur_$rep
for v in oldpro oldsrgb noclk r1state; do run ${v}_$rep EXA_LIB=owlexabrick_b200/libexabrick_b200.$v.so; done
done
