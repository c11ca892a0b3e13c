// Device memory layout of the derived traversal structure (shared by the host builder and the kernels).
#pragma once
#include <cstdint>

namespace mcpt {

// Inner node: both children's boxes + child references.  ref >= 0: inner node index; ref < 0: ~triangle.
struct alignas(16) DNode {
	float lminx, lminy, lminz, lmaxx;
	float lmaxy, lmaxz, rminx, rminy;
	float rminz, rmaxx, rmaxy, rmaxz;
	int32_t left, right, pad0, pad1;
};
static_assert(sizeof(DNode) == 64, "DNode");

// Leaf record: the leaf's box (the exact floats of the reference leaf node), the canonical rank of the
// leaf (its position in the reference's left-first DFS leaf order) and everything the exact triangle test
// needs, 6 x 16 B.
struct alignas(16) DLeaf {
	float bminx, bminy, bminz; int32_t rank;
	float bmaxx, bmaxy, bmaxz; int32_t material;
	float v0x, v0y, v0z, nx;
	float e1x, e1y, e1z, ny;       // e1 = -(v1 - v0)
	float e2x, e2y, e2z, nz;       // e2 = -(v2 - v0)
	float cr0, cr1, cr2, pad;      // ray-independent 2x2 determinants of (e1, e2)
};
static_assert(sizeof(DLeaf) == 96, "DLeaf");

}  // namespace mcpt
