#!/bin/bash
# r3d: K-raygen with one queue atomic per block against the previous build, then the suite and the C2 bench line on the new build
O=gpurun_out; mkdir -p $O
V=vulkan-pathtracer_b200/variants
one() { if [ "$1" = default ]; then timeout 300 python tools/gpu_sweep.py one $2 | tail -1; else B200PT_LIB=$PWD/$V/libb200pt_$1.so timeout 300 python tools/gpu_sweep.py one $2 | tail -1; fi; }
{
export SWEEP_FRAMES=16
for cfg in c2 c4 c1; do for v in old default old default; do echo "== $v $cfg"; one $v $cfg; done; done
} > $O/r3d_variants.txt 2>&1
timeout 400 python -m pytest tests -m gpu -x -q > $O/r3d_pytest_gpu.log 2>&1
timeout 200 python bench.py > $O/r3d_bench_c2.json 2> $O/r3d_bench_c2.err
cat $O/r3d_variants.txt; tail -3 $O/r3d_pytest_gpu.log; cut -c1-200 $O/r3d_bench_c2.json
