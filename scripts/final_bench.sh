#!/bin/bash
# Usage (under gpurun): bash scripts/final_bench.sh <tag>  -- the round's bench lines into gpurun_out/<tag>_*.json
tag=${1:-final}
python bench.py > gpurun_out/${tag}_bench_16m.json 2> gpurun_out/${tag}_bench_16m.err
python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/${tag}_bench_reference.json 2> gpurun_out/${tag}_bench_reference.err
for w in grid1m highdup16m_perm grid64m batch10k; do
  python bench.py --workload $w --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/${tag}_bench_${w}.json 2> gpurun_out/${tag}_bench_${w}.err
done
python bench.py --workload adversarial64m --steps 10 --warmup 3 --no-cpu-baseline --e2e-steps 1 > gpurun_out/${tag}_bench_adversarial64m.json 2> gpurun_out/${tag}_bench_adversarial64m.err
python - <<PY
import json, glob
for f in sorted(glob.glob("gpurun_out/${tag}_bench_*.json")):
    try:
        d = json.load(open(f))
    except Exception as e:
        print(f, "unreadable", e); continue
    e = d.get("e2e") or {}
    sc = {k: round(v["ms"], 3) for k, v in (d.get("schedules") or {}).items()}
    print(f.split("/")[-1], "value %.4g" % d.get("value", 0), "ms/step %.4f" % d.get("ms_per_step", 0),
          "frac", round((d.get("roofline") or {}).get("frac") or 0, 4), "e2e %.4g" % e.get("value", 0),
          "dropin %.4g" % ((e.get("dropin_default") or {}).get("value") or 0), sc)
PY
