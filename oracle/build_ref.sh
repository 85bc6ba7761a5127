#!/usr/bin/env bash
# Compiles the reference's OWN sources (where they lie under $REF, unmodified) plus
# oracle/ref_harness.cpp into oracle/_ref/ref_harness.  TEST INFRASTRUCTURE ONLY.
# Needs third_party/build_deps.sh to have produced third_party/_deps/lib/*.a first.
#
#   -include rng_shim.h      per-(pixel, sample) counter-based streams (see rng_shim.h)
#   -fno-access-control      lets the harness own the pixel loop / read private state
#   -msse4.2 (no -march=native): the binary must also run on the GPU box's host CPU;
#       Embree picks its SSE4.2 / AVX kernels at run time.  No FMA contraction on x86-64
#       without -mfma, which is what the parity tests assume.
set -euo pipefail
REF=${REF:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
OUT=$HERE/_ref
[ -d "$REF" ] || { echo "build_ref.sh: $REF not present; keeping prebuilt $OUT"; exit 0; }
DEPS=$(cd "$HERE/../third_party" && pwd)/_deps
"$HERE/../third_party/build_deps.sh" > /dev/null
mkdir -p "$OUT/obj"
CXXFLAGS="-O2 -msse4.2 -std=c++14 -w -DGLM_FORCE_RADIANS -DGLM_SWIZZLE -fno-access-control \
 -include $HERE/rng_shim.h -include stdexcept \
 -I$REF -I$REF/submodules/glm -I$REF/submodules/embree/include -I$REF/submodules/assimp/include -I$DEPS/include"
SRCS="Technique PT BPT UPG Scene Intersector RayIsect Geometry AreaLights BSDF Cameras Beta Materials loader \
 threadpool ImageView utility system_utils statistics runtime_assert source_location"
pids=()
for s in $SRCS; do
  if [ ! -f "$OUT/obj/$s.o" ] || [ "$REF/$s.cpp" -nt "$OUT/obj/$s.o" ] || [ "$HERE/rng_shim.h" -nt "$OUT/obj/$s.o" ]; then
    g++ $CXXFLAGS -c "$REF/$s.cpp" -o "$OUT/obj/$s.o" &
    pids+=($!)
  fi
done
for p in "${pids[@]:-}"; do [ -n "$p" ] && wait "$p"; done
g++ $CXXFLAGS -c "$HERE/ref_harness.cpp" -o "$OUT/obj/ref_harness.o"
OBJS=""; for s in $SRCS ref_harness; do OBJS="$OBJS $OUT/obj/$s.o"; done
g++ -o "$OUT/ref_harness" $OBJS -L"$DEPS/lib" -lassimp -lembree -lembree_avx -lembree_sse42 -lsimd -llexers -lsys \
  -lz -lpthread -ldl
echo "built $OUT/ref_harness"

# ---- the same objects on a correctly rounded float libm (oracle/cr_libm.cpp interposes sinf, cosf, powf, ...): a diagnostic
# oracle that tells how much of a per-path disagreement is the last bit of the host's libm (tests/test_large_scene_gpu.py)
g++ -O2 -msse4.2 -fno-builtin -c "$HERE/cr_libm.cpp" -o "$OUT/obj/cr_libm.o"
g++ -o "$OUT/ref_harness_crlibm" "$OUT/obj/cr_libm.o" $OBJS -L"$DEPS/lib" -lassimp -lembree -lembree_avx -lembree_sse42 -lsimd -llexers -lsys \
  -lz -lpthread -ldl
echo "built $OUT/ref_harness_crlibm"

# ---- timing-only build, as the stock reference compiles (Makefile:27,30-37): no RNG shim (std::mt19937), -O2 with AVX2 + FMA
# (-march=x86-64-v3 instead of -march=native so that the binary runs on any AVX2 host: the GPU box is not this machine), linked
# with the AVX2 Embree.  bench.py uses it for cpu_baseline / --impl reference when the host CPU has AVX2, and says which build ran.
if [ -f "$DEPS/lib_avx2/libembree_avx2.a" ] || [ "${B2PT_STOCK:-1}" = 1 ]; then
  DEPS_AVX2=1 "$HERE/../third_party/build_deps.sh" > /dev/null
  mkdir -p "$OUT/obj_stock"
  SFLAGS="-O2 -march=x86-64-v3 -std=c++14 -w -DGLM_FORCE_RADIANS -DGLM_SWIZZLE -fno-access-control -DB2PT_STOCK_TIMING -include stdexcept \
   -I$REF -I$REF/submodules/glm -I$REF/submodules/embree/include -I$REF/submodules/assimp/include -I$DEPS/include"
  pids=()
  for s in $SRCS; do
    if [ ! -f "$OUT/obj_stock/$s.o" ] || [ "$REF/$s.cpp" -nt "$OUT/obj_stock/$s.o" ]; then
      g++ $SFLAGS -c "$REF/$s.cpp" -o "$OUT/obj_stock/$s.o" &
      pids+=($!)
    fi
  done
  for p in "${pids[@]:-}"; do [ -n "$p" ] && wait "$p"; done
  g++ $SFLAGS -c "$HERE/ref_harness.cpp" -o "$OUT/obj_stock/ref_harness.o"
  OBJS=""; for s in $SRCS ref_harness; do OBJS="$OBJS $OUT/obj_stock/$s.o"; done
  g++ -o "$OUT/ref_harness_stock" $OBJS -L"$DEPS/lib_avx2" -L"$DEPS/lib" -lassimp -lembree -lembree_avx2 -lembree_avx -lembree_sse42 \
    -lsimd -llexers -lsys -lz -lpthread -ldl
  echo "built $OUT/ref_harness_stock"
fi
