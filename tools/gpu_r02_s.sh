#!/bin/bash
# round 2, GPU call S (1 GPU): point query with 2 / 3 / 4 interleaved points per thread (descents advanced together) at several occupancies
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
python - <<'PY'
import sys; sys.path.insert(0, ".")
import bench
from owlexabrick_b200 import renderer as R
bench.load_or_build_scene("c3", R, gpu_builder=True)
PY
pq() { tag=$1; shift; env "$@" PQ_CLEAN_FLUSH=1 PQ_DIAG=1 timeout 300 python tools/perf_quick.py --workload c3 --tag "$tag" > gpurun_out/s_pq_$tag.json 2> gpurun_out/s_pq_$tag.err; python - "$tag" <<'PY'
import json, sys
try:
    d = json.load(open("gpurun_out/s_pq_%s.json" % sys.argv[1])); print(sys.argv[1], {k: d[k] for k in ("points", "points_sorted_us", "points_local128_us", "points_host") if k in d})
except Exception as e:
    print(sys.argv[1], "no result", e)
PY
}
pq base
for v in p2_12 p2_16 p2_10 p3_10 p4_8; do pq $v EXA_LIB=owlexabrick_b200/libexabrick_b200.$v.so; done
