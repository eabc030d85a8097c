#!/bin/bash
# round 2, GPU call A (1 GPU): parity suite, kernel-variant A/B on C3, bench line with the parity object.
# Every step has its own timeout and log under gpurun_out/; a failing step does not stop the others.
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/a_smi.txt 2>&1
echo "== pytest -m gpu"
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/a_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/a_pytest.log
echo "== perf_quick variants on c3"
run() { tag=$1; shift; env "$@" timeout 300 python tools/perf_quick.py --workload c3 --tag "$tag" > gpurun_out/a_pq_$tag.json 2> gpurun_out/a_pq_$tag.err; echo "$tag rc=$?"; tail -c 600 gpurun_out/a_pq_$tag.json; echo; }
run base PQ_DIAG=1 PQ_CLEAN_FLUSH=1
run noprefetch EXA_INDEX_PREFETCH=0 PQ_CLEAN_FLUSH=1
run mb6 EXA_LIB=owlexabrick_b200/libexabrick_b200.mb6.so PQ_CLEAN_FLUSH=1
run mb8 EXA_LIB=owlexabrick_b200/libexabrick_b200.mb8.so PQ_CLEAN_FLUSH=1
run ppt2 EXA_LIB=owlexabrick_b200/libexabrick_b200.ppt2.so PQ_CLEAN_FLUSH=1
run ppt4 EXA_LIB=owlexabrick_b200/libexabrick_b200.ppt4.so PQ_CLEAN_FLUSH=1
echo "== bench c3"
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/a_bench_c3.json 2> gpurun_out/a_bench_c3.err; echo "bench rc=$?"; tail -c 1500 gpurun_out/a_bench_c3.json; tail -5 gpurun_out/a_bench_c3.err
