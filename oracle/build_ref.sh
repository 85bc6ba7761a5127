#!/usr/bin/env bash
# Compiles the reference's OWN sources (where they lie under $REF, unmodified) plus
# oracle/ref_harness.cpp into oracle/_ref/ref_harness.  TEST INFRASTRUCTURE ONLY.
# Needs oracle/build_ref_deps.sh to have produced oracle/_ref/lib/*.a first.
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
"$HERE/build_ref_deps.sh" > /dev/null
mkdir -p "$OUT/obj"
CXXFLAGS="-O2 -msse4.2 -std=c++14 -w -DGLM_FORCE_RADIANS -DGLM_SWIZZLE -fno-access-control \
 -include $HERE/rng_shim.h -include stdexcept \
 -I$REF -I$REF/submodules/glm -I$REF/submodules/embree/include -I$REF/submodules/assimp/include -I$OUT/include"
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
g++ -o "$OUT/ref_harness" $OBJS -L"$OUT/lib" -lassimp -lembree -lembree_avx -lembree_sse42 -lsimd -llexers -lsys \
  -lz -lpthread -ldl
echo "built $OUT/ref_harness"
