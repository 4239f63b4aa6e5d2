#!/bin/bash
# quick GPU check used while iterating on kernels: a parity subset, then the config-2 stage times
# usage (under gpurun): bash tools/gpu_quick.sh <tag> [pytest -k expression]
tag=${1:-quick}
kexpr=${2:-"model_vs_oracle or full_size or golden_reference"}
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "$kexpr" 2>&1 | tail -4
python bench.py --steps 5 --no-cpu-baseline --no-share > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err || tail -5 gpurun_out/${tag}_bench.err
python - <<PY
import json
d = json.load(open("gpurun_out/${tag}_bench.json"))
print("ms/step", round(d["ms_per_step"], 4), "value", round(d["value"]), "e2e", round(d["e2e"]["value"]))
print({k: v for k, v in d["stages_ms"].items()})
PY
