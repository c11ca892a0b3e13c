#!/bin/bash
# TEST INFRASTRUCTURE.  Builds oracle/_ref/libmcpt_ref.so from the reference's OWN host sources
# where they lie under /root/reference (read-only); nothing is copied into the repository history
# (oracle/_ref/ is git-ignored, but it travels to the GPU box with the working-tree snapshot).
#
# The reference's build system is a Visual Studio .sln, so this is a hand-written recipe:
#   * BVH/hlbvh.cpp spells its constructors `HLBVH<CPU>::HLBVH<CPU>(` (accepted by MSVC only); a
#     one-line sed writes a patched copy to oracle/_ref/src/hlbvh_patched.cpp.
#   * gnu++17 (not c++17): `.x/.y/.z` on cl_float4 need the anonymous-struct extension.
#   * -ffp-contract=off: no FMA contraction, so float results equal an SSE2/MSVC build.
#   * libOpenCL.so.1 (Khronos ICD loader shipped with the CUDA toolkit) only satisfies the cl.hpp
#     wrappers' release/retain symbols; no OpenCL platform is ever opened.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${MCPT_REFERENCE_DIR:-/root/reference}/MonteCarloPathTracing"
OUT="$HERE/_ref"
if [ ! -d "$REF" ]; then
  echo "build_ref: $REF not present (GPU box) - keeping prebuilt $OUT/libmcpt_ref.so" >&2
  exit 0
fi
mkdir -p "$OUT/src"
sed 's/HLBVH<CPU>::HLBVH<CPU>(/HLBVH<CPU>::HLBVH(/' "$REF/BVH/hlbvh.cpp" > "$OUT/src/hlbvh_patched.cpp"
OCL=""
for c in /usr/local/cuda/targets/x86_64-linux/lib/libOpenCL.so.1 /usr/local/cuda/lib64/libOpenCL.so.1; do
  [ -e "$c" ] && OCL="$c" && break
done
CXXFLAGS="-std=gnu++17 -O2 -fPIC -ffp-contract=off -w -I$REF -I$REF/BVH -I$OUT/src -DCL_TARGET_OPENCL_VERSION=120"
g++ $CXXFLAGS -c "$HERE/ref_shim.cpp" -o "$OUT/ref_shim.o"
g++ $CXXFLAGS -c "$REF/tiny_obj_loader.cpp" -o "$OUT/tiny_obj_loader.o"
g++ $CXXFLAGS -c "$REF/stb_image_write.cpp" -o "$OUT/stb_image_write.o"
g++ -shared -o "$OUT/libmcpt_ref.so" "$OUT/ref_shim.o" "$OUT/tiny_obj_loader.o" "$OUT/stb_image_write.o" $OCL
echo "build_ref: wrote $OUT/libmcpt_ref.so"

# ---- the reference's OpenCL kernels compiled as C++ (oracle/ref_kernels.cpp + oracle/clc_shim.hpp).
# Adapted copies go to oracle/_ref/src/ (never into the repository).  The adaptations are purely syntactic:
#   1. `(floatN)(a, b, ...)` / `(int2)(a, b)` vector literals -> make_floatN(a, b, ...) / make_int2(a, b): C++ reads the
#      OpenCL form as a cast of a comma expression;
#   2. objdef.h: `#ifdef __cplusplus` -> `#if 0`, `#ifndef __cplusplus` -> `#if 1`, i.e. the branch an OpenCL compiler
#      sees (record definitions without the host namespace, EPSILON = 1e-5f, the device functions);
#   3. `a.s012 = b.s012;` -> `a.s012 = make_float3(b.s012);` (shade.cl:185): the swizzle view keeps its trivial copy
#      assignment so that the records stay trivially copyable; routing the right side through float3 selects the
#      three-lane assignment OpenCL performs.
CLSED='s/(float\(2\|3\|4\|16\))(/make_float\1(/g; s/(int2)(/make_int2(/g; s/\.s012 = \(.*\)\.s012;/.s012 = make_float3(\1.s012);/'
sed "$CLSED; s/#ifdef __cplusplus/#if 0/; s/#ifndef __cplusplus/#if 1/" "$REF/objdef.h" > "$OUT/src/objdef_cl.h"
for k in rayGenerator intersect shade history testkernel EPO; do
  sed "$CLSED" "$REF/kernels/$k.cl" > "$OUT/src/${k}_cl.inc"
done
g++ -std=gnu++17 -O2 -fPIC -fopenmp -ffp-contract=off -w -I"$HERE" -I"$OUT/src" -shared -o "$OUT/libmcpt_refcl.so" "$HERE/ref_kernels.cpp"
echo "build_ref: wrote $OUT/libmcpt_refcl.so"

# ---- the drop-in proof: integration/OpenCLApp_mcpt.cpp compiled against the reference's OWN
# config.cpp, thirdpartywrapper.cpp, auxiliary.cpp, BVH/*.cpp, tinyobj and stb, linked with libmcpt_cuda.so, driven by a
# headless main that loops update().  tests/test_gpu_dropin.py runs it on the GPU box (prebuilt binary travels there).
ROOT="$(cd "$HERE/.." && pwd)"
PKG="$ROOT/montecarlopathtracing_b200"
if [ -e "$PKG/libmcpt_cuda.so" ]; then
  printf '#include <cmath>\n#include <cfloat>\n#include <cstdint>\n#include <cstring>\n#include <deque>\n#include <functional>\n#include <queue>\n#include <unordered_set>\n#include <iostream>\n#undef M_PI\n' > "$OUT/src/prelude.h"
  DROPFLAGS="$CXXFLAGS -I$ROOT/include -include $OUT/src/prelude.h"
  OBJS=""
  for f in config thirdpartywrapper auxiliary BVH/bvh BVH/treeletBVH; do
    o="$OUT/dropin_$(basename $f).o"
    g++ $DROPFLAGS -c "$REF/$f.cpp" -o "$o"
    OBJS="$OBJS $o"
  done
  g++ $DROPFLAGS -c "$OUT/src/hlbvh_patched.cpp" -o "$OUT/dropin_hlbvh.o"
  for f in OpenCLApp_mcpt headless_main oclbasic_stub; do
    g++ $DROPFLAGS -c "$ROOT/integration/$f.cpp" -o "$OUT/dropin_$f.o"
    OBJS="$OBJS $OUT/dropin_$f.o"
  done
  g++ -o "$OUT/mcpt_dropin" $OBJS "$OUT/dropin_hlbvh.o" "$OUT/tiny_obj_loader.o" "$OUT/stb_image_write.o" $OCL \
      -L"$PKG" -lmcpt_cuda -Wl,-rpath,'$ORIGIN/../../montecarlopathtracing_b200'
  echo "build_ref: wrote $OUT/mcpt_dropin"
fi
