#!/bin/bash
# Usage (under gpurun): bash scripts/final_bench.sh <tag>  -- the round's bench lines into gpurun_out/<tag>_*.json
tag=${1:-final}
python bench.py > gpurun_out/${tag}_bench_16m.json 2> gpurun_out/${tag}_bench_16m.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${tag}_bench_reference.json 2> gpurun_out/${tag}_bench_reference.err
for w in grid1m highdup16m_perm grid64m batch10k; do
  python bench.py --workload $w --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/${tag}_bench_${w}.json 2> gpurun_out/${tag}_bench_${w}.err
done
python - <<PY
import json, glob
for f in sorted(glob.glob("gpurun_out/${tag}_bench_*.json")):
    try:
        d = json.load(open(f))
    except Exception as e:
        print(f, "unreadable", e); continue
    e = d.get("e2e") or {}
    print(f.split("/")[-1], "value %.4g" % d.get("value", 0), d.get("unit"), "ms/step %.4f" % d.get("ms_per_step", 0),
          "frac", (d.get("roofline") or {}).get("frac"), "e2e %.4g" % e.get("value", 0), e.get("ms_per_step"))
PY
