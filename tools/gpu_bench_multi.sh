#!/bin/bash
# gpurun --gpus N -- 'bash tools/gpu_bench_multi.sh N': the bench line and one traced host-resident Gemm
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29513 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err; echo "bench rc=$?"
grep -v "OMP_NUM_THREADS\|\*\*\*\*" gpurun_out/bench_${N}gpu.err | tail -c 1500
E2E_RAW=0 timeout 200 $TR --master-port 29521 tools/e2e_probe.py 2>&1 | grep "^E2E\|rror" > gpurun_out/e2e_probe_${N}gpu.log
python - $N <<'PY'
import json, sys
N = sys.argv[1]
try:
    j = json.loads([l for l in open(f"gpurun_out/bench_{N}gpu.json") if l.startswith("{")][-1])
    print("value", j["value"], "frac", j["frac_of_fp64_peak_all_gpus"], "e2e", j["e2e"]["value"], j["e2e"]["ms_per_step"], "comm", j["config"]["comm"], "exposed", j["roofline"]["exposed_non_gemm_ms_per_step"], "parity", j["parity"]["gemm_rel"], j["parity"]["redist_bit_exact"])
    for x in j["extra"]:
        print("  ", {k: v for k, v in x.items() if k not in ("clocks", "roofline", "name")}, x["name"][:60], (x.get("roofline") or {}).get("frac"))
except Exception as e:
    print("no line:", e)
for l in open(f"gpurun_out/e2e_probe_{N}gpu.log"):
    if l.startswith("E2E "):
        j = json.loads(l[4:]); p = j.get("rank0_profile_of_one_call", {})
        print("probe: dev", j["device_resident_ms"], "host", j["host_resident_ms"], "call", p.get("call_ms"))
        for k in ("gemm_intervals_ms", "upload_stream_intervals_ms", "downloads_intervals_ms"):
            print("    ", k, p.get(k))
PY
