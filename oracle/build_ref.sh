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
