"""One compact summary of what tools/gpu_first_call.sh left in gpurun_out/ (so the next step can be decided from a single
read): the last line of every pytest log, the first failing test names, and the headline numbers of the JSON tools."""
import glob
import json
import os
import re
import sys

OUT = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out"


def last_json(path):
    try:
        for line in reversed(open(path).read().strip().splitlines()):
            line = line.strip()
            if line.startswith("{"):
                return json.loads(line)
    except Exception as e:  # noqa: BLE001
        return {"error": repr(e)}
    return None


def main():
    s = {}
    for log in sorted(glob.glob(os.path.join(OUT, "pytest_*.log"))):
        txt = open(log, errors="replace").read()
        tail = [l for l in txt.strip().splitlines() if l.strip()][-2:]
        failed = re.findall(r"^(?:FAILED|ERROR) (\S+)", txt, flags=re.M)[:12]
        s[os.path.basename(log)] = {"tail": tail, "failed": failed}
    for name in ("ref_gpu_check.json", "ref_gpu_check_wn2.json", "ref_gpu_check_level1.json", "host_blocked_check.json",
                 "syrk_bench.json", "bench_1gpu.json"):
        j = last_json(os.path.join(OUT, name))
        if j is None:
            s[name] = None
            continue
        if name.startswith("bench"):
            s[name] = {k: j.get(k) for k in ("value", "ms_per_step", "e2e", "roofline", "clocks", "gpu_launches")}
            if isinstance(s[name].get("roofline"), dict):
                s[name]["roofline"] = {k: s[name]["roofline"].get(k) for k in ("achieved", "peak", "frac")}
            if isinstance(s[name].get("e2e"), dict):
                s[name]["e2e"] = {k: s[name]["e2e"].get(k) for k in ("value", "ms_per_step", "error")}
        elif name.startswith("ref_gpu_check"):
            s[name] = {k: v for k, v in j.items() if k in ("ok", "failures", "parity", "summary") or k.startswith("gemm") or k.endswith("_gbs")}
            if not s[name]:
                s[name] = {"keys": sorted(j)[:40]}
        else:
            s[name] = j
    for err in sorted(glob.glob(os.path.join(OUT, "*.err"))):
        txt = open(err, errors="replace").read().strip()
        if txt:
            s[os.path.basename(err)] = txt.splitlines()[-3:]
    print(json.dumps(s, indent=1)[:12000])


if __name__ == "__main__":
    main()
