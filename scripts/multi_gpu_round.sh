#!/bin/bash
# Usage (under gpurun --gpus N): bash scripts/multi_gpu_round.sh N  -- config 4 and the copy ceiling on N GPUs of one box.
N=${1:-2}
o=gpurun_out
python -m pytest tests/test_gpu_parity.py -q -k "two_gpus" 2>&1 | tail -2 > $o/r02_two_gpu_test_n$N.txt
python bench.py --workload batch10k --gpus $N --steps 10 --warmup 3 --no-cpu-baseline > $o/r02_bench_batch10k_n$N.json 2> $o/mg.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 scripts/dma_ceiling.py > $o/r02_dma_ceiling_n$N.json 2>> $o/mg.err
python scripts/bench_batch_native.py --meshes 10000 --gpus $N --repeats 2 > $o/r02_batch_native_10k_g$N.json 2>> $o/mg.err
python bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline > $o/r02_bench_16m_n$N.json 2>> $o/mg.err
cat $o/r02_two_gpu_test_n$N.txt $o/r02_dma_ceiling_n$N.json $o/r02_batch_native_10k_g$N.json
python - <<PY
import json
for f in ("r02_bench_batch10k_n$N", "r02_bench_16m_n$N"):
    d = json.load(open("$o/%s.json" % f)); print(f, "%.4g" % d["value"], d["ms_per_step"], "e2e %.4g" % d["e2e"]["value"])
PY
