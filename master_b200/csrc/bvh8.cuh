// bvh8.cuh — compressed 8-wide BVH: node layout and single-ray traversal.
//
// The layout is the compressed wide BVH of Ylitie, Karras and Laine ("Efficient Incoherent Ray Traversal on GPUs
// Through Compressed Wide BVHs", HPG 2017), restated from the paper: child boxes are quantised to 8 bits per plane
// on a per-node power-of-two grid, eight children share one 80-byte node (five 128-bit loads instead of the seven
// loads a 4-wide uncompressed node takes), leaves of up to three triangles hang directly off the inner node, and the
// child slots are assigned at build time so that visiting them in the order of (slot XOR ray-octant) is roughly front
// to back — no per-node sort.
//
//   float4 0   origin.x  origin.y  origin.z  | ex | ey | ez | imask |     2^e = grid step per axis; imask bit s: slot s is an inner node
//   float4 1   child_base  tri_base  meta[0..3]  meta[4..7]              inner child of slot s = child_base + popc(imask & below(s))
//   float4 2   qlo.x[0..7]  qlo.y[0..7]                                  meta[s] of a leaf: (unary triangle count) << 5 | offset from tri_base
//   float4 3   qlo.z[0..7]  qhi.x[0..7]
//   float4 4   qhi.y[0..7]  qhi.z[0..7]
//
// Triangles are stored so that the leaves of one node are contiguous (tri_base .. tri_base + 24), in the three float4
// streams of scene.cuh.  The triangle test and the hit finalisation are the Embree-exact ones of trace.cuh: the tree
// changes which triangles are tested, never what a test returns.
#pragma once
#include "scene.cuh"

namespace b2pt {

constexpr int kNode8Float4s = 5;
constexpr int kStack8 = 48;  // uint2 entries; a ray pushes at most one node group and one triangle group per level

__device__ __forceinline__ float byte_to_float(uint32_t word, uint32_t k) {
  // exact u8 -> float without the conversion pipe: place the byte in the low mantissa bits of 2^23, subtract 2^23
  return __uint_as_float(__byte_perm(word, 0x4B000000u, 0x7540u | k)) - 8388608.0f;
}

// 2^15 + byte, exactly: the byte lands in mantissa bits 8..15 of 2^15 (whose ulp is 2^-8).
// `base` is 0x47000000 handed in as a RUN-TIME value (DeviceScene::prmt_base, a kernel parameter): with both the base and the selector
// known at compile time ptxas keeps the base as the immediate of PRMT and moves the selector into a register before EVERY permute
// (UMOV + IMAD.U32: 48 extra instructions per node, a sixth of the node test — cuobjdump, round 2); with the base in the constant
// bank the selector is the immediate and the permute is one instruction.
__device__ __forceinline__ float byte_to_float15(uint32_t word, uint32_t k, uint32_t base) {
  return __uint_as_float(__byte_perm(word, base, 0x7604u | (k << 4)));
}

}  // namespace b2pt
