#!/bin/sh
# runs tools/gpu_sweep.py modes-lite for every variant library
for f in vulkan-pathtracer_b200/variants/*.so; do
  echo "== $f"; B200PT_LIB=$PWD/$f python tools/gpu_sweep.py one "$@" | tail -1
done
