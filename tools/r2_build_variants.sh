#!/bin/sh
# Run HERE (no GPU needed) before tools/r2_first_gpurun.sh: A/B kernel variants for tools/gpu_variants.sh.
set -e
cd "$(dirname "$0")/.."
for k in 4 8 16; do sh tools/build_variant.sh refill$k -DPT_TRACE_REFILL_MIN=$k; done
sh tools/build_variant.sh latebary -DPT_TRACE_LATE_BARY
sh tools/build_variant.sh skipdead -DPT_SKIP_DEAD_LAST        # zero-throughput last segments are not traced (11 % of C2's bounce rays)
sh tools/build_variant.sh occlskip -DPT_SKIP_DEAD_LAST -DPT_OCCLUSION_LAST
sh tools/build_variant.sh occl -DPT_OCCLUSION_LAST        # last segments as occlusion queries; run the parity suite with B200PT_LIB set to it
sh tools/build_variant.sh prefetch -DPT_TRACE_PREFETCH
sh tools/build_variant.sh sincos -DPT_SHADE_SINCOS     # needs the parity tests: B200PT_LIB=.../variants/libb200pt_sincos.so pytest -m gpu
for k in 8 12; do sh tools/build_variant.sh smem$k -DPT_TRACE_SMEM_STACK=$k; done
ls -la vulkan-pathtracer_b200/variants/
