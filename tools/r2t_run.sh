#!/bin/bash
# r2t: K-trace live-state variants (closest hit in memory, shear rows in shared memory -> 8 / 9 CTAs per SM without spills),
# evict-first ray loads, SAH exchange rate / leaf size re-sweep.  Run under gpurun from the repo root.
O=gpurun_out; mkdir -p $O
V=vulkan-pathtracer_b200/variants
one() { # <lib or "default"> <cfg>
  if [ "$1" = default ]; then timeout 300 python tools/gpu_sweep.py one $2 | tail -1
  else B200PT_LIB=$PWD/$V/libb200pt_$1.so timeout 300 python tools/gpu_sweep.py one $2 | tail -1; fi
}
{
export SWEEP_FRAMES=16
for cfg in c2 c4; do
  for v in default hm7 hm8 rs8 both8 both9 both8s stream ctri05 ctri09 leaf3 default; do echo "== $v $cfg"; one $v $cfg; done
done
export SWEEP_FRAMES=4
for cfg in c5 c1; do
  for v in default hm8 both8 both9; do echo "== $v $cfg"; one $v $cfg; done
done
export SWEEP_FRAMES=2
for v in default both8 both9; do echo "== $v c3"; one $v c3; done
} > $O/r2t_variants.txt 2>&1
{
for v in both8 both9 hm8; do
  echo "== parity with $v"
  B200PT_LIB=$PWD/$V/libb200pt_$v.so timeout 400 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
done
} > $O/r2t_variant_parity.txt 2>&1
cat $O/r2t_variants.txt; cat $O/r2t_variant_parity.txt
