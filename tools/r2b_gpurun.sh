#!/bin/bash
# Second gpurun of round 2: full suite on the new defaults, A/B of the K-trace variants built by tools/build_variant.sh,
# round evidence (bench lines C1..C5, reference arm, ncu launch list + one --set full capture), smoke.
#   gpurun --timeout 2400 -- 'bash tools/r2b_gpurun.sh > gpurun_out/r2b.log 2>&1'
O=gpurun_out; mkdir -p $O
T0=$(date +%s); lap() { echo "#### $(( $(date +%s) - T0 )) s: $*"; }
lap "suite"
python -m pytest tests -m gpu -x -q > $O/r2b_pytest_gpu.log 2>&1; tail -8 $O/r2b_pytest_gpu.log
lap "smoke"
python __graft_entry__.py smoke 2>&1 | tail -2
lap "variants"
for c in c2 c4 c3_8th; do echo "== variants $c"; python tools/gpu_sweep.py one $c | tail -1; sh tools/gpu_variants.sh $c; done 2>&1 | tee $O/r2b_variants.txt
lap "evidence"
bash tools/evidence.sh r2b > $O/r2b_evidence.log 2>&1; tail -30 $O/r2b_evidence.log
lap "done"
