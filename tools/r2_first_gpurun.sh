#!/bin/bash
# First gpurun of round 2 (from the repo root): everything that was staged without hardware in round 1.
#   gpurun --timeout 2400 -- 'bash tools/r2_first_gpurun.sh > gpurun_out/r2a_first.log 2>&1'
O=gpurun_out; mkdir -p $O
T0=$(date +%s); lap() { echo "#### $(( $(date +%s) - T0 )) s: $*"; }
lap "default suite (with the config-size C3/C4/C5 tile-sample tests)"
python -m pytest tests -m gpu -x -q > $O/r2a_pytest_gpu.log 2>&1; tail -5 $O/r2a_pytest_gpu.log
lap "peaks"
python tools/gpu_peaks.py > $O/r2a_peaks.json 2> $O/r2a_peaks.err; cat $O/r2a_peaks.json
lap "gated tests of the staged builder stages + tile shares"
B200PT_TEST_UNVERIFIED=1 timeout 900 python -m pytest tests/test_zz_gpu_bvh_sah.py -q > $O/r2a_pytest_sah.log 2>&1; tail -15 $O/r2a_pytest_sah.log
lap "builders A/B"
for c in c2 c4 c5; do echo "== $c"; timeout 600 python tools/gpu_sweep.py builders $c; done 2>&1 | tee $O/r2a_builders.txt
lap "bench default / SAH+optimal"
python bench.py --no-cpu-baseline > $O/r2a_bench_c2_default.json 2> /dev/null
python bench.py --no-cpu-baseline --flags 192 > $O/r2a_bench_c2_sah_opt.json 2> $O/r2a_bench_c2_sah_opt.err
for f in $O/r2a_bench_c2_*.json; do echo $f; cut -c1-300 $f; echo; done
lap "kernel variants"
if ls vulkan-pathtracer_b200/variants/*.so > /dev/null 2>&1; then
  for c in c2 c4 c3_8th; do echo "== variants $c"; python tools/gpu_sweep.py one $c | tail -1; sh tools/gpu_variants.sh $c; done 2>&1 | tee $O/r2a_variants.txt
fi
lap "tile shares on one GPU"
for sh in 2 4; do
  python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu-baseline --partition tiles --tile-shares $sh > $O/r2a_bench_c3_shares$sh.json 2> $O/r2a_bench_c3_shares$sh.err
  python bench.py --no-cpu-baseline --partition tiles --tile-shares $sh > $O/r2a_bench_c2_shares$sh.json 2> $O/r2a_bench_c2_shares$sh.err
done
python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu-baseline > $O/r2a_bench_c3_default.json 2> /dev/null
for f in $O/r2a_bench_c3_*.json $O/r2a_bench_c2_shares*.json; do echo $f; cut -c1-200 $f; echo; done
lap "parity of the variants that are not bit-identical by construction"
for v in skipdead occl occlskip sincos; do
  f=$PWD/vulkan-pathtracer_b200/variants/libb200pt_$v.so
  [ -f $f ] && { echo "== parity with $v"; B200PT_LIB=$f timeout 900 python -m pytest tests/test_gpu_parity.py -q -k "not c3 and not c4 and not c5" 2>&1 | tail -5; } | tee -a $O/r2a_variant_parity.txt
done
lap "done"
