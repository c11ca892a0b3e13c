// Host-side build of the traversal acceleration structure used by MCPT_TRAVERSAL_FAST.
//
// The reference's LBVH (BVH/hlbvh.cpp) stays the canonical structure: it defines the leaf order that
// the accept rule is replayed in and is what MCPT_TRAVERSAL_REFERENCE / _CANONICAL walk.  Because the
// fast traversal only has to be CONSERVATIVE over inner nodes (see mcpt_traverse.cuh), it is free to
// use a better tree over the same leaves: a binned-SAH binary BVH, one triangle per leaf, nodes laid
// out in depth-first preorder (a subtree with n leaves owns n-1 consecutive node slots, which also makes
// the build embarrassingly parallel without atomics).
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <future>
#include <thread>
#include <vector>

#include "accel_build.hpp"

namespace mcpt {
namespace {

struct Box {
	float lo[3], hi[3];
	void reset() { lo[0] = lo[1] = lo[2] = FLT_MAX; hi[0] = hi[1] = hi[2] = -FLT_MAX; }
	void grow(const Box& b) {
		for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], b.lo[k]); hi[k] = std::max(hi[k], b.hi[k]); }
	}
	void grow_point(const float* p) {
		for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], p[k]); hi[k] = std::max(hi[k], p[k]); }
	}
	float half_area() const {
		float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
		return dx * dy + dy * dz + dz * dx;
	}
};

#ifndef MCPT_SAH_BINS
#define MCPT_SAH_BINS 16
#endif
constexpr int BINS = MCPT_SAH_BINS;
constexpr int MAX_SAH_DEPTH = 40;        // below this depth: median splits, so the traversal stack (64) cannot overflow

struct Builder {
	const DLeaf* leaves;
	std::vector<Box> tri_box;
	std::vector<float> centroid;         // 3 per triangle
	std::vector<int> idx;
	DNode* nodes;
	int max_parallel_depth;

	// builds the subtree over idx[lo, hi); inner nodes occupy slots [base, base + (hi-lo) - 1).
	// returns the child reference (inner index or ~triangle) and its box.
	int build(int lo, int hi, int base, int depth, Box& out_box) {
		const int n = hi - lo;
		if (n == 1) {
			out_box = tri_box[idx[lo]];
			return ~idx[lo];
		}
		Box cb;
		cb.reset();
		for (int i = lo; i < hi; ++i) cb.grow_point(&centroid[3 * size_t(idx[i])]);
		int mid = -1;
		if (depth < MAX_SAH_DEPTH && n > 2) {
			float best_cost = FLT_MAX;
			int best_axis = -1, best_bin = -1;
			for (int axis = 0; axis < 3; ++axis) {
				const float ext = cb.hi[axis] - cb.lo[axis];
				if (!(ext > 0.0f)) continue;
				Box bb[BINS];
				int cnt[BINS];
				for (int b = 0; b < BINS; ++b) { bb[b].reset(); cnt[b] = 0; }
				const float scale = float(BINS) / ext;
				for (int i = lo; i < hi; ++i) {
					int t = idx[i];
					int b = std::min(BINS - 1, std::max(0, int((centroid[3 * size_t(t) + axis] - cb.lo[axis]) * scale)));
					bb[b].grow(tri_box[t]);
					++cnt[b];
				}
				float right_area[BINS];
				int right_cnt[BINS];
				Box acc;
				acc.reset();
				int c = 0;
				for (int b = BINS - 1; b > 0; --b) {
					acc.grow(bb[b]);
					c += cnt[b];
					right_area[b] = c ? acc.half_area() : 0.0f;
					right_cnt[b] = c;
				}
				acc.reset();
				c = 0;
				for (int b = 0; b < BINS - 1; ++b) {
					acc.grow(bb[b]);
					c += cnt[b];
					if (c == 0 || right_cnt[b + 1] == 0) continue;
					float cost = acc.half_area() * float(c) + right_area[b + 1] * float(right_cnt[b + 1]);
					if (cost < best_cost) { best_cost = cost; best_axis = axis; best_bin = b; }
				}
			}
			if (best_axis >= 0) {
				const float ext = cb.hi[best_axis] - cb.lo[best_axis];
				const float scale = float(BINS) / ext;
				const float lo_c = cb.lo[best_axis];
				int* first = idx.data() + lo;
				int* last = idx.data() + hi;
				int* m = std::partition(first, last, [&](int t) {
					int b = std::min(BINS - 1, std::max(0, int((centroid[3 * size_t(t) + best_axis] - lo_c) * scale)));
					return b <= best_bin;
				});
				mid = int(m - idx.data());
			}
		}
		if (mid <= lo || mid >= hi) {       // median split along the widest centroid axis (also the deep-tree guard)
			int axis = 0;
			float e0 = cb.hi[0] - cb.lo[0], e1 = cb.hi[1] - cb.lo[1], e2 = cb.hi[2] - cb.lo[2];
			if (e1 > e0 && e1 >= e2) axis = 1;
			else if (e2 > e0 && e2 > e1) axis = 2;
			mid = lo + n / 2;
			std::nth_element(idx.begin() + lo, idx.begin() + mid, idx.begin() + hi, [&](int a, int b) {
				float ca = centroid[3 * size_t(a) + axis], cb2 = centroid[3 * size_t(b) + axis];
				return ca < cb2 || (ca == cb2 && a < b);
			});
		}
		const int nl = mid - lo;
		Box bl, br;
		int rl, rr;
		if (depth < max_parallel_depth && n > 16384) {
			auto fut = std::async(std::launch::async, [&] { return build(lo, mid, base + 1, depth + 1, bl); });
			rr = build(mid, hi, base + nl, depth + 1, br);
			rl = fut.get();
		} else {
			rl = build(lo, mid, base + 1, depth + 1, bl);
			rr = build(mid, hi, base + nl, depth + 1, br);
		}
		DNode& d = nodes[base];
		d.lminx = bl.lo[0]; d.lminy = bl.lo[1]; d.lminz = bl.lo[2]; d.lmaxx = bl.hi[0]; d.lmaxy = bl.hi[1]; d.lmaxz = bl.hi[2];
		d.rminx = br.lo[0]; d.rminy = br.lo[1]; d.rminz = br.lo[2]; d.rmaxx = br.hi[0]; d.rmaxy = br.hi[1]; d.rmaxz = br.hi[2];
		d.left = rl;
		d.right = rr;
		d.pad0 = d.pad1 = 0;
		out_box = bl;
		out_box.grow(br);
		return base;
	}
};

}  // namespace

void build_sah_bvh2(const DLeaf* leaves, int ntris, std::vector<DNode>& out_nodes) {
	out_nodes.assign(size_t(std::max(1, ntris - 1)), DNode{});
	if (ntris < 2) return;
	Builder b;
	b.leaves = leaves;
	b.tri_box.resize(size_t(ntris));
	b.centroid.resize(3 * size_t(ntris));
	b.idx.resize(size_t(ntris));
	for (int i = 0; i < ntris; ++i) {
		const DLeaf& l = leaves[i];
		Box& x = b.tri_box[size_t(i)];
		// + 0.0f: a negative zero becomes a positive one, so that min / max never have to choose between the two (the
		// device builder, sah_build.cu, reduces in another order and must arrive at the same bits)
		x.lo[0] = l.bminx + 0.0f; x.lo[1] = l.bminy + 0.0f; x.lo[2] = l.bminz + 0.0f;
		x.hi[0] = l.bmaxx + 0.0f; x.hi[1] = l.bmaxy + 0.0f; x.hi[2] = l.bmaxz + 0.0f;
		for (int k = 0; k < 3; ++k) b.centroid[3 * size_t(i) + k] = 0.5f * (x.lo[k] + x.hi[k]);
		b.idx[size_t(i)] = i;
	}
	b.nodes = out_nodes.data();
	unsigned hw = std::thread::hardware_concurrency();
	b.max_parallel_depth = hw >= 16 ? 4 : hw >= 8 ? 3 : hw >= 4 ? 2 : hw >= 2 ? 1 : 0;
	Box root;
	b.build(0, ntris, 0, 0, root);
}

}  // namespace mcpt
