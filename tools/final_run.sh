#!/bin/bash
# Round-end pass (run under gpurun from the repo root): GPU suite, smoke, then tools/evidence.sh.   usage: bash tools/final_run.sh <tag>
TAG=${1:-r2}
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; tail -4 $O/${TAG}_pytest_gpu.log
python __graft_entry__.py smoke 2>&1 | tail -1
bash tools/evidence.sh $TAG > $O/${TAG}_evidence.log 2>&1; tail -14 $O/${TAG}_evidence.log | cut -c1-300
