#!/bin/bash
# Lab build of the render core: the Tuning struct reads RT_* environment overrides (sweeps, A/B runs).
# -> simple_raytracer_b200/lib/librt_b200_lab.so (load it with RT_B200_LIB=...); the product library is untouched.
set -e
cd "$(dirname "$0")/.."
mkdir -p build_lab
FLAGS="-std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC,-ffp-contract=off -Iinclude -Isimple_raytracer_b200/csrc --expt-relaxed-constexpr -cudart static -DRT_TUNING_ENV $EXTRA"
ls simple_raytracer_b200/csrc/*.cu | xargs -P 8 -I{} sh -c "/usr/local/cuda/bin/nvcc $FLAGS -c -o build_lab/\$(basename {} .cu).o {}"
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -cudart static -shared -o simple_raytracer_b200/lib/librt_b200_lab.so build_lab/*.o -ldl
ls -la simple_raytracer_b200/lib/librt_b200_lab.so
