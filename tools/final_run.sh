O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/r2h_pytest_gpu.log 2>&1; tail -4 $O/r2h_pytest_gpu.log
python __graft_entry__.py smoke 2>&1 | tail -1
bash tools/evidence.sh r2h > $O/r2h_evidence.log 2>&1; tail -14 $O/r2h_evidence.log | cut -c1-300
