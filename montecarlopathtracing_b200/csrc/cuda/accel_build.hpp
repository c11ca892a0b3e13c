#pragma once
#include <vector>

#include "accel_layout.hpp"

namespace mcpt {
// Identifies the algorithm behind build_sah_bvh2 (bins, split rules, layout): layouts cached on disk (mcpt_derive_scene
// blobs) are only accepted from the same builder.  Bump when the builder's output may change.
constexpr unsigned FAST_TREE_BUILDER_ID = 0x00020010u;   // revision 2 (zeros of the leaf boxes made positive), 16 bins
// Binned-SAH binary BVH over the leaf boxes in `leaves`; out_nodes gets ntris-1 nodes in preorder, root = 0.
void build_sah_bvh2(const DLeaf* leaves, int ntris, std::vector<DNode>& out_nodes);
// The same tree, byte for byte, built on `device` (sah_build.cu): d_leaves = the records in device memory, h_leaves = the
// same records on the host, out = ntris - 1 nodes (ntris >= 2).  Returns MCPT_OK or an MCPT_ERR_* code.
// *depth (optional) = levels of inner nodes (root = 1): what the walk's stack has to hold.
int build_sah_bvh2_device(int device, const DLeaf* d_leaves, const DLeaf* h_leaves, int ntris, DNode* out, float* build_ms, int* depth = nullptr);
}
