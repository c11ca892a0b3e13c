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

// Quantised inner node of the fast walk: the two child boxes of a DNode on a 15-bit grid spanning the scene
// (plane = frame.lo[a] + q * frame.step[a]), rounded OUTWARDS by QNODE_MARGIN extra grid steps, 32 B = ONE
// 256-bit load per node visit.  The walk is bound by L1 wavefronts (one per lane per load instruction on
// divergent rays), so halving the node halves the dominant cost.  Only inner tests use it - they need to be
// conservative, not exact; leaf records keep the reference's floats.
//   planes[0..2] = left child x, y, z: low half = min plane, high half = max plane; planes[3..5] = right child.
struct alignas(32) QNode {
	uint32_t planes[6];
	int32_t left, right;            // as in DNode
};
static_assert(sizeof(QNode) == 32, "QNode");
constexpr int QNODE_GRID = 32767;   // q in [0, 32767]: (q << 8) fits the mantissa of the magic float 2^15 + q
constexpr int QNODE_MARGIN = 2;     // grid steps of outward margin on top of outward rounding (error budget: mcpt_traverse.cuh)
struct QFrame { float lo[3], step[3], extent; };

// 8-wide compressed node (after Ylitie, Karras, Laine: "Efficient Incoherent Ray Traversal on GPUs Through
// Compressed Wide BVHs", 2017): child boxes are quantised to 8 bits relative to the node origin with a
// power-of-two scale per axis, rounded OUTWARDS, so a quantised box always contains the child's box (the
// traversal only needs inner tests to be conservative).  80 bytes = 5 x 16 B.
//   meta[slot]: 0 = empty; inner child: (1 << 5) | (24 + slot); leaf slot with k = 1..3 triangles:
//   ((1 << k) - 1) << 5 | offset, offset = index of its first record relative to leaf_base (< 24).
struct alignas(16) WNode {
	float px, py, pz;
	uint8_t ex, ey, ez, imask;      // biased exponents of the per-axis scale; bit s of imask = slot s is an inner node
	uint32_t child_base;            // index of the first inner child (inner children are consecutive, slot order)
	uint32_t leaf_base;             // index of the first leaf record of this node in the wide-order leaf array
	uint8_t meta[8];
	uint8_t qlox[8], qloy[8], qloz[8];
	uint8_t qhix[8], qhiy[8], qhiz[8];
};
static_assert(sizeof(WNode) == 80, "WNode");

}  // namespace mcpt
