#!/bin/bash
# round 2, GPU call M (N GPUs): config C5 through the data-partitioned sort-last pipeline (tools/c5_multi.py)
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
N=${1:-2}; SCALE=${2:-4}; IMG=${3:-7680x4320}; LIMIT=${4:-900}
free -g | head -2; df -h /dev/shm | tail -1; nproc
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512"
timeout $LIMIT $TR tools/c5_multi.py --scale $SCALE --image $IMG --modes ${5:-sortlast} > gpurun_out/m_c5_s${SCALE}_n${N}_${5:-sortlast}.json 2> gpurun_out/m_c5_s${SCALE}_n${N}_${5:-sortlast}.err; echo "rc=$?"
grep "^\[c5" gpurun_out/m_c5_s${SCALE}_n${N}_${5:-sortlast}.err | cut -c1-400 | head -30
grep -iE "Traceback|Error|error:" gpurun_out/m_c5_s${SCALE}_n${N}_${5:-sortlast}.err | head -8
grep "^{" gpurun_out/m_c5_s${SCALE}_n${N}_${5:-sortlast}.json | cut -c1-3500
rm -f /dev/shm/exa_c5_* /dev/shm/exa_hostframes_*
