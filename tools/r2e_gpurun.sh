#!/bin/bash
# Fourth gpurun of round 2: parity on the templated kernels (flat K-trace, atlas-only K-shade), K-shade variants, HW bilinear A/B
O=gpurun_out; mkdir -p $O
T0=$(date +%s); lap() { echo "#### $(( $(date +%s) - T0 )) s: $*"; }
lap "suite (without the config-size C3/C4/C5 samples)"
python -m pytest tests -m gpu -x -q -k "not config_c3 and not config_c4 and not config_c5" > $O/r2e_pytest_gpu.log 2>&1; tail -6 $O/r2e_pytest_gpu.log
grep "hw bilinear" $O/r2e_pytest_gpu.log
python -m pytest tests/test_gpu_parity.py -m gpu -q -s -k hw_bilinear 2>&1 | grep -E "hw bilinear|passed|failed"
lap "variants"
for c in c2 c4 c3_8th; do echo "== variants $c"; python tools/gpu_sweep.py one $c | tail -1; sh tools/gpu_variants.sh $c; done 2>&1 | tee $O/r2e_variants.txt
lap "C5: default / variants / HW bilinear"
( echo "== c5 default"; python tools/gpu_sweep.py one c5 | tail -1
  echo "== c5 PT_FLAG_HW_BILINEAR"; SWEEP_FLAGS=2048 python tools/gpu_sweep.py one c5 | tail -1
  for v in shade5m noatlasonly; do echo "== c5 $v"; B200PT_LIB=$PWD/vulkan-pathtracer_b200/variants/libb200pt_$v.so python tools/gpu_sweep.py one c5 | tail -1; done ) 2>&1 | tee $O/r2e_c5.txt
lap "ncu: K-shade on C5, default and HW bilinear"
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors_srcunit_tex_op_read.sum,l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum,l1tex__t_sectors_pipe_tex_mem_texture.sum,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active
SWEEP_FRAMES=1 ncu --metrics $M --clock-control none -k regex:k_shade --launch-skip 243 -c 20 --csv --log-file $O/r2e_ncu_shade_c5_default.csv python tools/gpu_sweep.py one c5 > /dev/null 2>&1
SWEEP_FRAMES=1 SWEEP_FLAGS=2048 ncu --metrics $M --clock-control none -k regex:k_shade --launch-skip 243 -c 20 --csv --log-file $O/r2e_ncu_shade_c5_hwtex.csv python tools/gpu_sweep.py one c5 > /dev/null 2>&1
lap "ncu --set full: K-shade<multi> on C4 (source-level)"
SWEEP_FRAMES=1 ncu --set full --clock-control none --import-source on -k regex:k_shade --launch-skip 41 -c 4 -f -o $O/r2e_shade_c4 python tools/gpu_sweep.py one c4 > $O/r2e_ncu_c4.log 2>&1
lap "done"
