#!/bin/bash
# One-GPU call: the whole -m gpu suite, the product against the recompiled reference GPU layer, then the bench line.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --durations=5 > gpurun_out/pytest_gpu.log 2>&1; echo "suite rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -6 gpurun_out/pytest_gpu.log
timeout 120 tools/ref_gpu_check gpurun_out/ref_gpu_check.json 2> gpurun_out/ref_gpu_check.err > /dev/null; echo "ref_gpu_check rc=$?"
grep "timing" gpurun_out/ref_gpu_check.json | grep -E "Gemm|Transpose"
if [ "$1" != "nobench" ]; then
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_1gpu.json 2> gpurun_out/bench_1gpu.err; echo "bench rc=$?"
tail -c 1500 gpurun_out/bench_1gpu.err
python - <<'PY'
import json
try:
    j = json.loads([l for l in open("gpurun_out/bench_1gpu.json") if l.startswith("{")][-1])
    print("value", j["value"], "frac", j["roofline"]["frac"], "e2e", (j.get("e2e") or {}).get("value"), "ms/step", j["ms_per_step"], "parity", (j.get("parity") or {}).get("gemm_rel"))
    for x in j.get("extra", []):
        print({k: v for k, v in x.items() if k not in ("clocks", "roofline")}, (x.get("roofline") or {}).get("frac"))
except Exception as e:
    print("no bench line:", e)
PY
fi
