#!/bin/bash
# A/B of the K-trace variants (tools/build_variant.sh) + a parity pass for the candidates
O=gpurun_out; mkdir -p $O
T0=$(date +%s); lap() { echo "#### $(( $(date +%s) - T0 )) s: $*"; }
lap "variants"
for c in c2 c4 c3_8th; do echo "== variants $c"; python tools/gpu_sweep.py one $c | tail -1; sh tools/gpu_variants.sh $c; done 2>&1 | tee $O/r2c_variants.txt
lap "parity of the candidates"
for v in ts6all trishare6 earlyrejfb; do
  f=$PWD/vulkan-pathtracer_b200/variants/libb200pt_$v.so
  echo "== parity with $v"; B200PT_LIB=$f timeout 600 python -m pytest tests/test_gpu_parity.py -q -x -k "brute_force or golden or sample_and_bounce or instances or c1_cornell or c2_sponza or tie_break or edge_cases" 2>&1 | tail -3
done 2>&1 | tee $O/r2c_variant_parity.txt
lap "done"
