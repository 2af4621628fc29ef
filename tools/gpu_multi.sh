#!/bin/bash
# Multi-GPU call (gpurun --gpus N -- 'bash tools/gpu_multi.sh N'): one process per GPU.
#  (1) tools/dist_check.py: every redistribution pair, the reduce side and every Gemm variant against the oracle, on the
#      push transport and on staged NCCL;  (2) bench.py on both transports;  (3) the redistribution bench.
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_${N}gpu.txt 2>&1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29511 tools/dist_check.py > gpurun_out/dist_check_${N}gpu_push.log 2>&1; echo "dist_check(push) rc=$?" | tee -a gpurun_out/dist_check_${N}gpu_push.log
grep -E "grid|DIST_CHECK|MISMATCH" gpurun_out/dist_check_${N}gpu_push.log | tail -6
HYD_TRANSPORT=nccl timeout 600 $TR --master-port 29512 tools/dist_check.py > gpurun_out/dist_check_${N}gpu_nccl.log 2>&1; echo "dist_check(nccl) rc=$?" | tee -a gpurun_out/dist_check_${N}gpu_nccl.log
grep -E "DIST_CHECK|MISMATCH" gpurun_out/dist_check_${N}gpu_nccl.log | tail -3
timeout 900 $TR --master-port 29513 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err; echo "bench rc=$?"
grep -v "OMP_NUM_THREADS\|\*\*\*\*" gpurun_out/bench_${N}gpu.err | tail -c 2000
timeout 600 $TR --master-port 29514 bench.py --gpus $N --steps 5 --warmup 3 --transport nccl --no-extra > gpurun_out/bench_${N}gpu_nccl.json 2> gpurun_out/bench_${N}gpu_nccl.err; echo "bench(nccl) rc=$?"
grep -v "OMP_NUM_THREADS\|\*\*\*\*" gpurun_out/bench_${N}gpu_nccl.err | tail -c 1000
# the C++ client-program form, one process per GPU, world from the environment (no torch, no MPI)
for r in $(seq 0 $((N-1))); do RANK=$r WORLD_SIZE=$N LOCAL_RANK=$r timeout 300 examples/DistGemm 0 1500 128 > gpurun_out/distgemm_${N}gpu_rank$r.log 2>&1 & done; wait
cat gpurun_out/distgemm_${N}gpu_rank*.log | grep -E "DistGemm|Distributed|PASSED|FAILED|rror" | sort | uniq -c | head -12
timeout 600 $TR --master-port 29515 tools/redist_bench.py > gpurun_out/redist_bench_${N}gpu.log 2>&1; echo "redist_bench rc=$?"
python - $N <<'PY'
import json, sys
N = sys.argv[1]
for f in (f"gpurun_out/bench_{N}gpu.json", f"gpurun_out/bench_{N}gpu_nccl.json"):
    try:
        j = json.loads([l for l in open(f) if l.startswith("{")][-1])
        print(f, "value", j["value"], "frac", j["frac_of_fp64_peak_all_gpus"], "e2e", j["e2e"]["value"], "comm", j["config"]["comm"], "exposed", j["roofline"]["exposed_non_gemm_ms_per_step"], "parity", j["parity"]["gemm_rel"], j["parity"]["redist_bit_exact"])
        for x in j["extra"]:
            print("  ", {k: v for k, v in x.items() if k not in ("clocks", "roofline", "name")}, x["name"][:60], (x.get("roofline") or {}).get("frac"))
    except Exception as e:
        print(f, "no line:", e)
try:
    for row in json.load(open(f"gpurun_out/redist_bench_{N}gpu.json")):
        print("  ", row["op"], row["ms"], row.get("frac_of_roofline"), row.get("rank0_ms_total_pack_unpack"))
except Exception as e:
    print("no redist bench:", e)
PY
