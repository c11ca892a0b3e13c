#pragma once
#include <vector>

#include "accel_layout.hpp"

namespace mcpt {
// Identifies the algorithm behind build_sah_bvh2 (bins, split rules, layout): layouts cached on disk (mcpt_derive_scene
// blobs) are only accepted from the same builder.  Bump when the builder's output may change.
constexpr unsigned FAST_TREE_BUILDER_ID = 0x00010010u;   // revision 1, 16 bins
// Binned-SAH binary BVH over the leaf boxes in `leaves`; out_nodes gets ntris-1 nodes in preorder, root = 0.
void build_sah_bvh2(const DLeaf* leaves, int ntris, std::vector<DNode>& out_nodes);
// Collapses the binary tree into 8-wide compressed nodes (root = 0, breadth-first).  out_leaves is the leaf
// array reordered so that every wide node's triangles are consecutive; DLeaf::pad of each record holds the
// ORIGINAL triangle index (bit pattern of an int).
void build_wide_bvh8(const std::vector<DNode>& bvh2, const DLeaf* leaves, int ntris, std::vector<WNode>& out_nodes,
                     std::vector<DLeaf>& out_leaves);
// Quantises the binary tree's child boxes onto a 15-bit scene grid (QNode); returns false when the scene
// cannot be framed (non-finite or empty boxes), in which case the caller keeps the float nodes only.
// Every quantised box is checked, in double arithmetic, to contain its child box plus QNODE_MARGIN steps.
bool build_qnodes(const std::vector<DNode>& bvh2, std::vector<QNode>& out_nodes, QFrame& frame);
}
