#!/usr/bin/env bash
# TEST INFRASTRUCTURE. Builds oracle/_ref/libref_oracle.so from the reference sources WHERE THEY
# LIE under /root/reference (nothing is copied into the repo; outputs only into oracle/_ref/, which
# is git-ignored but travels to the GPU box). Recipe follows SURVEY.md Appendix B:
#   - RayTracing/BVH.cpp compiled verbatim (Linux pch stand-in, two empty Assimp stubs, a case shim)
#   - RayTracing/main_raytracing.cu compiled as host C++ after TWO mechanical edits (the <<<>>> launch; the jitter draw order)
#   - cuRAND XORWOW from the toolkit's own header, host-side
# The reference's own build system (MSBuild/vcpkg) is not run.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
R="${REFERENCE_ROOT:-/root/reference}"
OUT="$HERE/_ref"
if [ ! -d "$R/RayTracing" ]; then echo "build_ref: $R not present - keeping prebuilt $OUT" >&2; exit 0; fi
mkdir -p "$OUT"
CUDA_INC="${CUDA_HOME:-/usr/local/cuda}/include"
INC="-I$R -I$R/include -I$R/Common -I$CUDA_INC -DGLM_ENABLE_EXPERIMENTAL -I$HERE/ref_shim"
# -ffp-contract=off: the oracle's arithmetic is plain IEEE mul/add (no FMA) - DESIGN.md "contraction policy"
CXXFLAGS="-std=c++20 -O2 -w -fPIC -fopenmp -ffp-contract=off"
g++ $CXXFLAGS $INC -include "$HERE/ref_shim/pch_linux.h" -c "$R/RayTracing/BVH.cpp" -o "$OUT/bvh_ref.o"
# Edit 2 pins an evaluation order C++ leaves unspecified to the one nvcc emits for the device (x jitter drawn first).
JITTER='s/vec2(rng.GetFloat01(), rng.GetFloat01())/ref_jitter_device_order<vec2>(rng)/'
sed -e 's/raytracing_kernel_main<<<Dg, Db>>>(render_data);/(void)Dg; (void)render_data;/' -e "$JITTER" \
    "$R/RayTracing/main_raytracing.cu" > "$OUT/rt_host.inc"
grep -q 'ref_jitter_device_order<vec2>(rng)' "$OUT/rt_host.inc"
g++ $CXXFLAGS $INC -I"$R/RayTracing" -DRT_HOST_SRC="\"$OUT/rt_host.inc\"" \
    -c "$HERE/ref_shim/ref_api.cpp" -o "$OUT/ref_api.o"
# CPU-baseline variant (bench.py only): constants -> variables, GetRayHit counter, -O3 -march=native is NOT used
# (the .so is built in the container and must run on the GPU box's host CPU): -O3 only.
sed -e 's/raytracing_kernel_main<<<Dg, Db>>>(render_data);/(void)Dg; (void)render_data;/' -e "$JITTER" \
    -e 's/constexpr int num_bounces = 6;/const int num_bounces = refbench_bounces;/' \
    -e 's/constexpr int sample_count = 5;/const int sample_count = refbench_spp;/' \
    -e 's/result.distance = max_distance;/result.distance = max_distance; refbench_rays++;/' \
    -e 's/extern "C" void raytracing_process/extern "C" void refbench_raytracing_process/' \
    -e 's/^#include "Math.h"//' -e 's/^#include "GPUScene.h"//' \
    "$R/RayTracing/main_raytracing.cu" > "$OUT/rt_bench.inc"
grep -q refbench_bounces "$OUT/rt_bench.inc" && grep -q refbench_spp "$OUT/rt_bench.inc" && grep -q 'refbench_rays++' "$OUT/rt_bench.inc"
g++ ${CXXFLAGS/-O2/-O3} $INC -I"$R/RayTracing" -DRT_BENCH_SRC="\"$OUT/rt_bench.inc\"" \
    -c "$HERE/ref_shim/ref_bench.cpp" -o "$OUT/ref_bench.o"
# the reference's DDS container parser, verbatim TU + a Windows.h stand-in (types / SAL / HRESULT names only): pins csrc/dds.cpp ParseDds
g++ $CXXFLAGS '-D__declspec(x)=__attribute__((weak))' -include "$HERE/ref_shim/winshim/Windows.h" -I"$HERE/ref_shim/winshim" -I"$R" -I"$R/utils/image" \
    -c "$R/utils/image/DDSLoader.cpp" -o "$OUT/dds_ref.o"
g++ $CXXFLAGS '-D__declspec(x)=__attribute__((weak))' -include "$HERE/ref_shim/winshim/Windows.h" -I"$HERE/ref_shim/winshim" -I"$R" -I"$R/utils/image" \
    -c "$HERE/ref_shim/ref_dds.cpp" -o "$OUT/ref_dds.o"
g++ -shared -fopenmp -o "$OUT/libref_oracle.so" "$OUT/ref_api.o" "$OUT/ref_bench.o" "$OUT/bvh_ref.o" "$OUT/dds_ref.o" "$OUT/ref_dds.o"
rm -f "$OUT/rt_host.inc" "$OUT/rt_bench.inc" "$OUT"/*.o
echo "built $OUT/libref_oracle.so"
# The reference's own CUDA kernels, UNMODIFIED, for sm_100a (BASELINE.json: "where the reference's own CUDA kernel builds
# for sm_100, its numbers are reported alongside"). Default nvcc flags as in RayTracing.vcxproj (no fast-math, fmad on).
# Exports the reference's extern "C" raytracing_process / init_rng; driven headless by tests/bench through ctypes.
NVCC="${CUDA_HOME:-/usr/local/cuda}/bin/nvcc"
if [ -x "$NVCC" ]; then
  "$NVCC" -std=c++20 -O3 -w -gencode arch=compute_100a,code=sm_100a -I"$R" -I"$R/include" -I"$R/Common" -I"$R/RayTracing" \
      -DGLM_ENABLE_EXPERIMENTAL -Xcompiler -fPIC -shared -cudart static \
      "$R/RayTracing/main_raytracing.cu" "$R/RayTracing/Random.cu" -o "$OUT/libref_gpu.so"
  echo "built $OUT/libref_gpu.so"
fi
