#!/bin/bash
# Usage (under gpurun): bash scripts/gpu_quick.sh [bench args]  -- GPU tests, then a short bench summary.
python -m pytest tests -m gpu -x -q 2>&1 | tail -8
python bench.py --steps 100 --warmup 5 --no-cpu-baseline "$@" > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err
tail -3 gpurun_out/bench_quick.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench_quick.json"))
print("Gtris/s %.3f  ms/step %.4f  pipeline frac %.4f" % (d["value"] / 1e9, d["ms_per_step"], d["pipeline_roofline"]["frac"]))
print({k: round(v["ms"], 4) for k, v in d["pipeline_roofline"]["kernels"].items()})
print("roofline", {k: d["roofline"][k] for k in ("achieved", "frac", "traffic")}, "dominant", {k: d["roofline"]["dominant_kernel"][k] for k in ("kernel", "frac", "share_of_step")})
e = d["e2e"]
print("e2e %.1f Mtris/s %.2f ms" % (e["value"] / 1e6, e["ms_per_step"]), {k: round(e[k]["ms_per_step"], 2) for k in ("malloc_results", "single_call", "single_call_pooled", "single_call_pageable_inputs") if k in e}, d["clocks"])
PY
