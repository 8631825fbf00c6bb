#!/bin/bash
# A/B of the working tree's library against the previous commit's build on ONE box (run under gpurun from the repo root):
#   git worktree add /tmp/wt HEAD && (cd /tmp/wt && sh tools/build_variant.sh old) && cp /tmp/wt/vulkan-pathtracer_b200/variants/libb200pt_old.so \
#       vulkan-pathtracer_b200/variants/     # the previous build travels with the snapshot
#   gpurun -- 'bash tools/ab_run.sh <tag>'
# Per-kernel CUDA-event times of C2 / C4 / C1 (16 frames, twice each, interleaved), C5 and C3 once, then the GPU suite on the new build.
# This is how every late round-2 change was accepted or dropped (profiles/r2w ... r3a_variants.txt).
TAG=${1:-ab}
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
