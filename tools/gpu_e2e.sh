#!/bin/bash
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
: > gpurun_out/e2e_probe_${N}gpu.log
run() { echo "== $*" >> gpurun_out/e2e_probe_${N}gpu.log; env "$@" timeout 300 $TR --master-port 29521 tools/e2e_probe.py 2>&1 | grep "^E2E\|rror" >> gpurun_out/e2e_probe_${N}gpu.log; }
run E2E_RAW=0 HYD_HOST_GATHER=sm
run E2E_RAW=0 HYD_HOST_GATHER=sm E2E_SM_RESERVE=12
cat gpurun_out/e2e_probe_${N}gpu.log
