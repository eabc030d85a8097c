#!/bin/bash
# kernel-variant libraries for A/B timing (tools/perf_quick.py with EXA_LIB=...); rebuild after ANY change of the ABI
cd "$(dirname "$0")/.." || exit 1
python -m owlexabrick_b200.build > /tmp/build_main.log 2>&1 || { tail -20 /tmp/build_main.log; exit 1; }
for v in "$@"; do
  n=${v%%:*}; f=${v#*:}
  (EXA_VARIANT=$n EXA_FLAGS="$f" python -m owlexabrick_b200.build > /tmp/build_$n.log 2>&1; echo "$n rc=$?") &
done
wait
ls -la owlexabrick_b200/*.so
