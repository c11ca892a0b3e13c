#pragma once
#include <vector>

#include "accel_layout.hpp"

namespace mcpt {
// Binned-SAH binary BVH over the leaf boxes in `leaves`; out_nodes gets ntris-1 nodes in preorder, root = 0.
void build_sah_bvh2(const DLeaf* leaves, int ntris, std::vector<DNode>& out_nodes);
// Collapses the binary tree into 8-wide compressed nodes (root = 0, breadth-first).  out_leaves is the leaf
// array reordered so that every wide node's triangles are consecutive; DLeaf::pad of each record holds the
// ORIGINAL triangle index (bit pattern of an int).
void build_wide_bvh8(const std::vector<DNode>& bvh2, const DLeaf* leaves, int ntris, std::vector<WNode>& out_nodes,
                     std::vector<DLeaf>& out_leaves);
}
