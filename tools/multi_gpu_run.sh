#!/bin/bash
# Multi-GPU evidence on one box with N GPUs (gpurun --gpus N -- 'bash tools/multi_gpu_run.sh N tag'):
# the library's NCCL path (tests/test_gpu_comm.py: two-device tile / frame shares, CLI --gpus), then bench.py under torchrun at
# 2 .. N ranks (weak: frame shares + one reduce of the accumulation buffers; strong: the C3 tile leg of every line).
N=${1:-2}; TAG=${2:-r2c}
O=gpurun_out; mkdir -p $O
nvidia-smi -L | head -8
python -m pytest tests/test_gpu_comm.py -m gpu -q 2>&1 | tail -5
for n in 2 4 8; do
  [ $n -le $N ] || continue
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
      bench.py --gpus $n --steps 20 --warmup 5 > $O/${TAG}_bench_c2_n${n}.json 2> $O/${TAG}_bench_c2_n${n}.err
  echo "== N=$n rc=$?"; cut -c1-250 $O/${TAG}_bench_c2_n${n}.json; tail -3 $O/${TAG}_bench_c2_n${n}.err
  python - <<PY
import json
try:
    d = json.loads(open("$O/${TAG}_bench_c2_n${n}.json").read().strip().splitlines()[-1])
    print("   weak:", round(d["value"], 1), "Mrays/s", round(d["ms_per_step"], 4), "ms/step   e2e:", round(d["e2e"]["value"], 1), " strong:", d.get("strong"))
except Exception as e:
    print("   no line:", e)
PY
done
python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline > $O/${TAG}_bench_c2_n1.json 2> $O/${TAG}_bench_c2_n1.err; cut -c1-250 $O/${TAG}_bench_c2_n1.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 tools/multi_gpu_check.py 2>&1 | tail -3
