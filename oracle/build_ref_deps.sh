#!/usr/bin/env bash
# Builds the reference's two vendored third-party libraries (Embree 2.12.0 and the
# Assimp 3.2 fork) out-of-tree, from the sources where they lie under $REF, into
# oracle/_ref/.  TEST INFRASTRUCTURE ONLY: nothing under oracle/ is linked into the
# product library.  Nothing is written to $REF; Embree is copied to a scratch dir
# because two one-line fixes are needed for cmake 4 / g++ 13.
#
# Flags follow the reference's own recipes: submodules/embree.makefile:2-10 and
# submodules/assimp.makefile:2-44.
set -euo pipefail
REF=${REF:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
OUT=$HERE/_ref
SCRATCH=${SCRATCH:-/tmp/b2pt_ref_build}
JOBS=${JOBS:-8}
mkdir -p "$OUT/lib" "$OUT/include" "$SCRATCH"

if [ ! -f "$OUT/lib/libembree.a" ]; then
  rm -rf "$SCRATCH/embree-src" "$SCRATCH/embree-build"
  cp -r "$REF/submodules/embree" "$SCRATCH/embree-src"
  chmod -R u+w "$SCRATCH/embree-src"
  # (1) cmake 4 rejects OLD policies
  sed -i 's/^\(\s*\)cmake_policy(SET CMP0042 OLD)/\1#cmake_policy(SET CMP0042 OLD)/I' "$SCRATCH/embree-src/CMakeLists.txt"
  # (2) g++ 13: __m128i -> __m128 needs an explicit cast
  sed -i '121s/.*/    return _mm_castsi128_ps(_mm_shuffle_epi32(_mm_castps_si128(a), _MM_SHUFFLE(i3, i2, i1, i0)));/' "$SCRATCH/embree-src/common/simd/vboolf4_sse2.h"
  mkdir -p "$SCRATCH/embree-build"
  ( cd "$SCRATCH/embree-build" && cmake "$SCRATCH/embree-src" \
      -DCMAKE_POLICY_VERSION_MINIMUM=3.5 -DCMAKE_BUILD_TYPE=Release \
      -DEMBREE_STATIC_LIB=ON -DEMBREE_TUTORIALS=OFF -DEMBREE_RAY_MASK=ON \
      -DEMBREE_TASKING_SYSTEM=INTERNAL -DEMBREE_GEOMETRY_LINES=OFF \
      -DEMBREE_GEOMETRY_HAIR=OFF -DEMBREE_GEOMETRY_SUBDIV=OFF -DEMBREE_GEOMETRY_USER=OFF \
      -DEMBREE_ISPC_SUPPORT=OFF -DEMBREE_MAX_ISA=AVX > cmake.log 2>&1 \
    && make embree -j"$JOBS" > make.log 2>&1 )
  cp "$SCRATCH"/embree-build/*.a "$OUT/lib/"
fi

if [ ! -f "$OUT/lib/libassimp.a" ]; then
  rm -rf "$SCRATCH/assimp-build"; mkdir -p "$SCRATCH/assimp-build"
  FLAGS="-DBUILD_SHARED_LIBS=OFF -DASSIMP_NO_EXPORT=ON -DASSIMP_BUILD_ASSIMP_TOOLS=OFF -DASSIMP_BUILD_TESTS=OFF"
  for imp in 3DS AC ASE ASSBIN ASSXML B3D BVH COLLADA DXF CSM HMP LWO LWS MD2 MD3 MD5 MDC MDL NFF NDO OFF OGRE \
             OPENGEX PLY MS3D COB IFC XGL FBX Q3D Q3BSP RAW SIB SMD STL TERRAGEN 3D X GLTF; do
    FLAGS="$FLAGS -DASSIMP_BUILD_${imp}_IMPORTER=FALSE"
  done
  ( cd "$SCRATCH/assimp-build" && cmake "$REF/submodules/assimp" -DCMAKE_POLICY_VERSION_MINIMUM=3.5 \
      -DCMAKE_BUILD_TYPE=Release $FLAGS > cmake.log 2>&1 && make -j"$JOBS" > make.log 2>&1 )
  cp "$SCRATCH/assimp-build/code/libassimp.a" "$OUT/lib/"
  # this Assimp revision keeps config.h in its source tree; newer ones generate it
  if [ -f "$SCRATCH/assimp-build/include/assimp/config.h" ]; then
    mkdir -p "$OUT/include/assimp"
    cp "$SCRATCH/assimp-build/include/assimp/config.h" "$OUT/include/assimp/"
  fi
fi
ls -la "$OUT/lib"
