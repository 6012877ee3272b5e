#!/bin/bash
# sweep of the sub-chunk pipeline (PIRL_SUBCHUNK) on the bench workload; usage: bash tools/gpu_subchunk_sweep.sh "<sizes>"
for sc in $1; do
  PIRL_SUBCHUNK=$sc timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-grad > gpurun_out/sc_$sc.json 2> gpurun_out/sc_$sc.err
  python - <<P
import json
try:
    d=json.load(open("gpurun_out/sc_$sc.json"))
    print("subchunk $sc: evals/s %.0f  ms/step %.3f" % (d["value"], d["ms_per_step"]), {k: round(v,3) for k,v in d["roofline"]["stages_ms_per_step"].items() if v})
except Exception as e:
    print("subchunk $sc failed", e); print(open("gpurun_out/sc_$sc.err").read()[-1500:])
P
done
