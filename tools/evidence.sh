#!/bin/bash
# Round evidence pass (run under gpurun from the repo root): bench lines for C1..C5, the reference arm, the ncu launch list
# and one `--set full` capture of the hot kernels.  Outputs land in gpurun_out/; tools/collect_evidence.py <tag> turns them
# into the tracked summaries under profiles/.      usage: bash tools/evidence.sh <tag>
TAG=${1:-r2}
O=gpurun_out; mkdir -p $O
python bench.py > $O/${TAG}_bench_c2.json 2> $O/${TAG}_bench_c2.err
python bench.py --impl reference --steps 3 --warmup 1 > $O/${TAG}_bench_ref_c2.json 2> $O/${TAG}_bench_ref_c2.err
python bench.py --workload c1 --steps 100 --no-strong > $O/${TAG}_bench_c1.json 2> /dev/null
python bench.py --workload c4 --steps 20 --no-cpu-baseline > $O/${TAG}_bench_c4.json 2> /dev/null
python bench.py --workload c3 --steps 3 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_c3.json 2> $O/${TAG}_bench_c3.err
python bench.py --workload c5 --steps 5 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_c5.json 2> $O/${TAG}_bench_c5.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches_c2.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-strong > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_trace_persistent|k_shade|k_raygen' -c 10 -f -o $O/${TAG}_full \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-strong > $O/${TAG}_ncu_full.log 2>&1
ls -la $O | tail -20
for f in $O/${TAG}_bench_*.json; do echo $f; cut -c1-400 $f; echo; done
