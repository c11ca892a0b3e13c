// EXPERIMENTAL: layout and builder of the 8-wide compressed nodes of MCPT_TRAVERSAL_WIDE8 (measured slower).
#pragma once
#include <cstdint>
#include <vector>

#include "../accel_layout.hpp"

namespace mcpt {

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

// Collapses the binary tree into 8-wide compressed nodes (root = 0, breadth-first).  out_leaves is the leaf
// array reordered so that every wide node's triangles are consecutive; DLeaf::pad of each record holds the
// ORIGINAL triangle index (bit pattern of an int).
void build_wide_bvh8(const std::vector<DNode>& bvh2, const DLeaf* leaves, int ntris, std::vector<WNode>& out_nodes,
                     std::vector<DLeaf>& out_leaves);

}  // namespace mcpt
