#!/bin/bash
# bench line summary for another order: bash tools/gpu_bench_order.sh <order> <n_cells> <batch> [extra bench args]
O=$1; NC=$2; B=$3; shift 3
timeout 500 python bench.py --order $O --n-cells $NC --batch $B --steps 3 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/bench_N$O.json 2> gpurun_out/bench_N$O.err
python - <<P
import json
try:
    d=json.load(open("gpurun_out/bench_N$O.json"))
    g=d.get("adjoint_grad") or {}
    print("N=$O n=$NC B=$B: evals/s %.0f ms/step %.2f e2e %.0f path frac %.3f grad/s %.0f" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["path"]["frac"], g.get("value", 0)), {k: round(v,2) for k,v in d["roofline"]["stages_ms_per_step"].items() if v})
except Exception as e:
    print("failed", e); print(open("gpurun_out/bench_N$O.err").read()[-1500:])
P
