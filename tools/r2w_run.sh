#!/bin/bash
# r2w / r2x: node header word `valid` instead of the node_bits array (5 instead of 7 LDG.128 per node test) against the previous build
TAG=${1:-r2w}
O=gpurun_out; mkdir -p $O
V=vulkan-pathtracer_b200/variants
one() { if [ "$1" = default ]; then timeout 300 python tools/gpu_sweep.py one $2 | tail -1; else B200PT_LIB=$PWD/$V/libb200pt_$1.so timeout 300 python tools/gpu_sweep.py one $2 | tail -1; fi; }
{
export SWEEP_FRAMES=16
for cfg in c2 c4 c1; do for v in old default old default; do echo "== $v $cfg"; one $v $cfg; done; done
export SWEEP_FRAMES=4
for v in old default; do echo "== $v c5"; one $v c5; done
export SWEEP_FRAMES=2
for v in old default; do echo "== $v c3"; one $v c3; done
} > $O/${TAG}_variants.txt 2>&1
timeout 600 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1
cat $O/${TAG}_variants.txt; tail -5 $O/${TAG}_pytest_gpu.log
