#!/bin/bash
# First gpurun of round 2 (from the repo root): the opt-in builder stages of csrc/bvh_sah.cuh have only run as a serial CPU
# model so far.  (1) the default suite must still be green, (2) the gated hardware tests of the new stages, (3) A/B of the
# builders on C2 / C4 / C5 with per-ray counts, kernel times and image identity, (4) bench lines with and without them.
#   gpurun --timeout 1500 -- 'bash tools/r2_first_gpurun.sh > gpurun_out/r2a_first.log 2>&1'
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/r2a_pytest_gpu.log 2>&1; tail -3 $O/r2a_pytest_gpu.log
python tools/gpu_peaks.py > $O/r2a_peaks.json 2> $O/r2a_peaks.err; cat $O/r2a_peaks.json
B200PT_TEST_UNVERIFIED=1 timeout 900 python -m pytest tests/test_zz_gpu_bvh_sah.py -q > $O/r2a_pytest_sah.log 2>&1; tail -15 $O/r2a_pytest_sah.log
for c in c2 c4 c5 c1; do echo "== $c"; timeout 600 python tools/gpu_sweep.py builders $c; done 2>&1 | tee $O/r2a_builders.txt
python bench.py --no-cpu-baseline > $O/r2a_bench_c2_default.json 2> /dev/null
python bench.py --no-cpu-baseline --flags 192 > $O/r2a_bench_c2_sah_opt.json 2> $O/r2a_bench_c2_sah_opt.err
for f in $O/r2a_bench_c2_*.json; do echo $f; cut -c1-300 $f; echo; done
# kernel variants built beforehand by tools/r2_build_variants.sh (refill threshold of K-trace's dynamic fetch)
if ls vulkan-pathtracer_b200/variants/*.so > /dev/null 2>&1; then
  for c in c2 c4; do echo "== variants $c"; python tools/gpu_sweep.py one $c | tail -1; sh tools/gpu_variants.sh $c; done 2>&1 | tee $O/r2a_variants.txt
fi
# tile shares on one GPU (pt_create_sibling_part): does overlapping the wavefront tails of 2 / 4 shares pay?  C3 is where the
# tails cost most (449 iterations per frame); C2 for the headline.  Compare ms_per_step.
for sh in 2 4; do
  python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu-baseline --partition tiles --tile-shares $sh > $O/r2a_bench_c3_shares$sh.json 2> $O/r2a_bench_c3_shares$sh.err
  python bench.py --no-cpu-baseline --partition tiles --tile-shares $sh > $O/r2a_bench_c2_shares$sh.json 2> $O/r2a_bench_c2_shares$sh.err
done
python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu-baseline > $O/r2a_bench_c3_default.json 2> /dev/null
for f in $O/r2a_bench_c3_*.json $O/r2a_bench_c2_shares*.json; do echo $f; cut -c1-200 $f; echo; done
# parity of the variants whose results are not bit-identical by construction (occlusion-query last segments: proven in the CPU
# restatement, kernels never run; sincos: last-place differences possible)
for v in skipdead occl occlskip sincos; do
  f=$PWD/vulkan-pathtracer_b200/variants/libb200pt_$v.so
  [ -f $f ] && { echo "== parity with $v"; B200PT_LIB=$f timeout 900 python -m pytest tests/test_gpu_parity.py -q 2>&1 | tail -5; } | tee -a $O/r2a_variant_parity.txt
done
