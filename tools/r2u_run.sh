#!/bin/bash
# r2u: second pass over the K-trace live-state variants (hit in memory + 8 CTAs; queue window in shared memory and dominant
# axis in the slot word; L1 prefetch of the node's hit-mask words), refill threshold / chunk re-sweep at 8 CTAs.
O=gpurun_out; mkdir -p $O
V=vulkan-pathtracer_b200/variants
one() { # <lib or "default"> <cfg>
  if [ "$1" = default ]; then timeout 300 python tools/gpu_sweep.py one $2 | tail -1
  else B200PT_LIB=$PWD/$V/libb200pt_$1.so timeout 300 python tools/gpu_sweep.py one $2 | tail -1; fi
}
{
export SWEEP_FRAMES=16
for cfg in c2 c4; do
  for v in default hm8 lean8 lean8pf hm7pf lean9 lean9pf hm8r2 hm8r6 hm8r8 hm8 default; do echo "== $v $cfg"; one $v $cfg; done
done
echo "== chunk sweep, hm8, c2"; B200PT_LIB=$PWD/$V/libb200pt_hm8.so timeout 300 python tools/gpu_sweep.py chunk c2
echo "== chunk sweep, lean8pf, c2"; B200PT_LIB=$PWD/$V/libb200pt_lean8pf.so timeout 300 python tools/gpu_sweep.py chunk c2
export SWEEP_FRAMES=4
for cfg in c5 c1; do
  for v in default hm8 lean8 lean8pf; do echo "== $v $cfg"; one $v $cfg; done
done
export SWEEP_FRAMES=2
for v in default hm8 lean8pf; do echo "== $v c3"; one $v c3; done
} > $O/r2u_variants.txt 2>&1
{
for v in lean8pf lean8; do
  echo "== parity with $v"
  B200PT_LIB=$PWD/$V/libb200pt_$v.so timeout 400 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
done
} > $O/r2u_variant_parity.txt 2>&1
cat $O/r2u_variants.txt; cat $O/r2u_variant_parity.txt
