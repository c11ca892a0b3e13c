#pragma once
#include <vector>

#include "accel_layout.hpp"

namespace mcpt {
// Identifies the algorithm behind build_sah_bvh2 (bins, split rules, layout): layouts cached on disk (mcpt_derive_scene
// blobs) are only accepted from the same builder.  Bump when the builder's output may change.
constexpr unsigned FAST_TREE_BUILDER_ID = 0x00010010u;   // revision 1, 16 bins
// Binned-SAH binary BVH over the leaf boxes in `leaves`; out_nodes gets ntris-1 nodes in preorder, root = 0.
void build_sah_bvh2(const DLeaf* leaves, int ntris, std::vector<DNode>& out_nodes);
}
