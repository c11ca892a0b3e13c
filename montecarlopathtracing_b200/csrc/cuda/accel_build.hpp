#pragma once
#include <vector>

#include "accel_layout.hpp"

namespace mcpt {
// Binned-SAH binary BVH over the leaf boxes in `leaves`; out_nodes gets ntris-1 nodes in preorder, root = 0.
void build_sah_bvh2(const DLeaf* leaves, int ntris, std::vector<DNode>& out_nodes);
}
