#!/bin/bash
# quick GPU check: parity tests (fail fast) + short bench without the CPU baseline.  usage: bash tools/gpu_quick.sh <tag> [pytest-k-expr]
TAG=$1; KEXPR=${2:-}
mkdir -p gpurun_out
if [ -n "$KEXPR" ]; then timeout 600 python -m pytest tests -m gpu -q -x -k "$KEXPR" > gpurun_out/${TAG}_pytest.log 2>&1; else timeout 600 python -m pytest tests -m gpu -q -x > gpurun_out/${TAG}_pytest.log 2>&1; fi
tail -4 gpurun_out/${TAG}_pytest.log
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
python - <<P
import json
try:
    d=json.load(open("gpurun_out/${TAG}_bench.json"))
    print("evals/s %.0f  ms/step %.3f  e2e %.0f  path frac %.3f" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["path"]["frac"]))
    print({k: round(v,3) for k,v in d["roofline"]["stages_ms_per_step"].items()})
except Exception as e:
    print("bench failed", e); print(open("gpurun_out/${TAG}_bench.err").read()[-2000:])
P
