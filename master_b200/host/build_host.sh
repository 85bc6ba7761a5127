#!/usr/bin/env bash
# Builds master_b200/host/_build/b2pt_master: the reference's own classes (compiled from the sources where they
# lie under $REF, unmodified, WITHOUT the oracle's RNG shim) + GpuTechnique + libb2pt.so.  This is the drop-in
# as a maintainer of the reference would link it.  Needs $REF at BUILD time only; the binary travels to the
# GPU box (master_b200/host/_build/ is git-ignored, not gpurun-ignored).
set -euo pipefail
REF=${REF:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
ROOT=$(cd "$HERE/../.." && pwd)
OUT=$HERE/_build
[ -d "$REF" ] || { echo "build_host.sh: $REF not present; keeping prebuilt $OUT"; exit 0; }
# the reference's vendored Embree (static libs) lives in third_party/_deps, a neutral directory: nothing here reads oracle/
[ -f "$ROOT/third_party/_deps/lib/libembree.a" ] || bash "$ROOT/third_party/build_deps.sh" > /dev/null
[ -f "$ROOT/master_b200/libb2pt.so" ] || { echo "build libb2pt.so first (python -m master_b200.build)"; exit 1; }
mkdir -p "$OUT/obj"
CXXFLAGS="-O2 -msse4.2 -std=c++14 -w -DGLM_FORCE_RADIANS -DGLM_SWIZZLE -fno-access-control -include stdexcept \
 -I$REF -I$REF/submodules/glm -I$REF/submodules/embree/include -I$REF/submodules/assimp/include -I$ROOT/third_party/_deps/include"
# the reference classes GpuTechnique and b2s_io.hpp touch; Embree is still linked because haste::Scene owns an RTCScene
SRCS="Technique Scene Intersector RayIsect Geometry AreaLights BSDF Cameras Materials threadpool ImageView utility \
 system_utils statistics runtime_assert source_location"
pids=()
for s in $SRCS; do
  if [ ! -f "$OUT/obj/$s.o" ] || [ "$REF/$s.cpp" -nt "$OUT/obj/$s.o" ]; then
    g++ $CXXFLAGS -c "$REF/$s.cpp" -o "$OUT/obj/$s.o" &
    pids+=($!)
  fi
done
for p in "${pids[@]:-}"; do [ -n "$p" ] && wait "$p"; done
g++ $CXXFLAGS -c "$HERE/GpuTechnique.cpp" -o "$OUT/obj/GpuTechnique.o"
g++ $CXXFLAGS -c "$HERE/host_main.cpp" -o "$OUT/obj/host_main.o"
OBJS=""; for s in $SRCS GpuTechnique host_main; do OBJS="$OBJS $OUT/obj/$s.o"; done
g++ -o "$OUT/b2pt_master" $OBJS -L"$ROOT/third_party/_deps/lib" -lembree -lembree_avx -lembree_sse42 -lsimd -llexers -lsys \
  -L"$ROOT/master_b200" -lb2pt -Wl,-rpath,'$ORIGIN/../..' -L/usr/local/cuda/lib64 -lcudart -lz -lpthread -ldl
echo "built $OUT/b2pt_master"
