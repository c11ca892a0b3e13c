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
		x.lo[0] = l.bminx; x.lo[1] = l.bminy; x.lo[2] = l.bminz;
		x.hi[0] = l.bmaxx; x.hi[1] = l.bmaxy; x.hi[2] = l.bmaxz;
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

// ================================================================== 8-wide collapse
namespace mcpt {
namespace {

struct Ref {            // a child reference into the binary tree
	int ref;            // >= 0 inner node, < 0 ~triangle
	Box box;
	int count;          // triangles below
};

struct WideBuilder {
	const std::vector<DNode>& b2;
	const DLeaf* leaves;
	std::vector<int> count;     // triangles below each binary inner node

	WideBuilder(const std::vector<DNode>& n, const DLeaf* l) : b2(n), leaves(l), count(n.size(), 0) {}

	int fill_counts(int node) {
		// iterative post-order (the tree can be ~40+ levels deep but recursion is fine at that depth)
		const DNode& d = b2[size_t(node)];
		int c = (d.left < 0 ? 1 : fill_counts(d.left)) + (d.right < 0 ? 1 : fill_counts(d.right));
		count[size_t(node)] = c;
		return c;
	}
	Ref child(const DNode& d, bool left) const {
		Ref r;
		r.ref = left ? d.left : d.right;
		if (left) { r.box.lo[0] = d.lminx; r.box.lo[1] = d.lminy; r.box.lo[2] = d.lminz; r.box.hi[0] = d.lmaxx; r.box.hi[1] = d.lmaxy; r.box.hi[2] = d.lmaxz; }
		else { r.box.lo[0] = d.rminx; r.box.lo[1] = d.rminy; r.box.lo[2] = d.rminz; r.box.hi[0] = d.rmaxx; r.box.hi[1] = d.rmaxy; r.box.hi[2] = d.rmaxz; }
		r.count = r.ref < 0 ? 1 : count[size_t(r.ref)];
		return r;
	}
	void collect(int ref, std::vector<int>& tris) const {
		if (ref < 0) { tris.push_back(~ref); return; }
		collect(b2[size_t(ref)].left, tris);
		collect(b2[size_t(ref)].right, tris);
	}
};

constexpr int LEAF_SLOT_TRIS = 3;

}  // namespace

void build_wide_bvh8(const std::vector<DNode>& bvh2, const DLeaf* leaves, int ntris, std::vector<WNode>& out_nodes,
                     std::vector<DLeaf>& out_leaves) {
	out_nodes.clear();
	out_leaves.clear();
	if (ntris < 2 || bvh2.empty()) { out_nodes.assign(1, WNode{}); return; }
	WideBuilder wb(bvh2, leaves);
	wb.fill_counts(0);
	out_leaves.reserve(size_t(ntris));
	out_nodes.reserve(size_t(ntris) / 4 + 16);
	out_nodes.push_back(WNode{});
	std::vector<int> work;          // binary node behind each wide node, index-aligned with out_nodes
	work.push_back(0);
	std::vector<int> tris;
	for (size_t wi = 0; wi < out_nodes.size(); ++wi) {
		const DNode& root = bvh2[size_t(work[wi])];
		std::vector<Ref> ch;
		ch.push_back(wb.child(root, true));
		ch.push_back(wb.child(root, false));
		// greedy: open the largest-area child that is still worth opening (more triangles than a leaf slot holds)
		while (ch.size() < 8) {
			int best = -1;
			float best_area = -1.0f;
			for (size_t k = 0; k < ch.size(); ++k)
				if (ch[k].ref >= 0 && ch[k].count > LEAF_SLOT_TRIS && ch[k].box.half_area() > best_area) {
					best_area = ch[k].box.half_area();
					best = int(k);
				}
			if (best < 0) break;
			const DNode& d = bvh2[size_t(ch[size_t(best)].ref)];
			ch[size_t(best)] = wb.child(d, true);
			ch.push_back(wb.child(d, false));
		}
		Box nb;
		nb.reset();
		for (const Ref& r : ch) nb.grow(r.box);
		// ---- slot assignment: child k goes to the free slot whose octant direction best matches its offset
		// from the node centre (greedy on the largest score), so that slot ^ ray octant approximates distance order
		int slot_of[8];
		bool slot_used[8] = {false, false, false, false, false, false, false, false};
		bool child_done[8] = {false, false, false, false, false, false, false, false};
		const float cx = 0.5f * (nb.lo[0] + nb.hi[0]), cy = 0.5f * (nb.lo[1] + nb.hi[1]), cz = 0.5f * (nb.lo[2] + nb.hi[2]);
		for (size_t it = 0; it < ch.size(); ++it) {
			float best = -FLT_MAX;
			int bk = -1, bs = -1;
			for (size_t k = 0; k < ch.size(); ++k) {
				if (child_done[k]) continue;
				const Box& b = ch[k].box;
				const float ox = 0.5f * (b.lo[0] + b.hi[0]) - cx, oy = 0.5f * (b.lo[1] + b.hi[1]) - cy, oz = 0.5f * (b.lo[2] + b.hi[2]) - cz;
				for (int s = 0; s < 8; ++s) {
					if (slot_used[s]) continue;
					float score = ((s & 1) ? ox : -ox) + ((s & 2) ? oy : -oy) + ((s & 4) ? oz : -oz);
					if (score > best) { best = score; bk = int(k); bs = s; }
				}
			}
			slot_of[bk] = bs;
			slot_used[bs] = true;
			child_done[bk] = true;
		}
		// ---- quantisation frame
		WNode w{};
		w.px = nb.lo[0]; w.py = nb.lo[1]; w.pz = nb.lo[2];
		int ebyte[3];
		double scale[3];
		for (int a = 0; a < 3; ++a) {
			double ext = double(nb.hi[a]) - double(nb.lo[a]);
			int e = -126;
			if (ext > 0.0) {
				e = int(std::ceil(std::log2(ext / 255.0)));
				while (std::ldexp(255.0, e) < ext) ++e;           // guard against log2 rounding
				if (e < -126) e = -126;
				if (e > 126) e = 126;
			}
			ebyte[a] = e + 127;
			scale[a] = std::ldexp(1.0, e);
		}
		w.ex = uint8_t(ebyte[0]); w.ey = uint8_t(ebyte[1]); w.ez = uint8_t(ebyte[2]);
		w.child_base = uint32_t(out_nodes.size());
		w.leaf_base = uint32_t(out_leaves.size());
		uint8_t* qlo[3] = {w.qlox, w.qloy, w.qloz};
		uint8_t* qhi[3] = {w.qhix, w.qhiy, w.qhiz};
		const float p[3] = {w.px, w.py, w.pz};
		int order[8];
		for (int s = 0; s < 8; ++s) order[s] = -1;
		for (size_t k = 0; k < ch.size(); ++k) order[slot_of[k]] = int(k);
		int leaf_off = 0;
		for (int s = 0; s < 8; ++s) {
			if (order[s] < 0) continue;
			const Ref& r = ch[size_t(order[s])];
			for (int a = 0; a < 3; ++a) {
				// outward rounding, checked in exact (double) arithmetic: p + q * scale is exact in double
				double lo = std::floor((double(r.box.lo[a]) - double(p[a])) / scale[a]);
				double hi = std::ceil((double(r.box.hi[a]) - double(p[a])) / scale[a]);
				while (lo > 0.0 && double(p[a]) + lo * scale[a] > double(r.box.lo[a])) lo -= 1.0;
				while (hi < 255.0 && double(p[a]) + hi * scale[a] < double(r.box.hi[a])) hi += 1.0;
				if (lo < 0.0) lo = 0.0;
				if (hi > 255.0) hi = 255.0;
				qlo[a][s] = uint8_t(lo);
				qhi[a][s] = uint8_t(hi);
			}
			if (r.ref >= 0 && r.count > LEAF_SLOT_TRIS) {
				w.imask = uint8_t(w.imask | (1u << s));
				w.meta[s] = uint8_t((1u << 5) | (24u + unsigned(s)));
				out_nodes.push_back(WNode{});
				work.push_back(r.ref);
			} else {
				tris.clear();
				wb.collect(r.ref, tris);
				const unsigned k = unsigned(tris.size());       // 1..3
				w.meta[s] = uint8_t((((1u << k) - 1u) << 5) | unsigned(leaf_off));
				for (int t : tris) {
					DLeaf rec = leaves[t];
					std::memcpy(&rec.pad, &t, sizeof(int));
					out_leaves.push_back(rec);
				}
				leaf_off += int(k);
			}
		}
		out_nodes[wi] = w;
	}
}

// ------------------------------------------------------------------ 15-bit scene-grid nodes
bool build_qnodes(const std::vector<DNode>& bvh2, std::vector<QNode>& out, QFrame& fr) {
	out.clear();
	if (bvh2.empty()) return false;
	// frame = box of the root's two children, padded so that margins never clip
	const DNode& r = bvh2[0];
	double lo[3] = {std::min(r.lminx, r.rminx), std::min(r.lminy, r.rminy), std::min(r.lminz, r.rminz)};
	double hi[3] = {std::max(r.lmaxx, r.rmaxx), std::max(r.lmaxy, r.rmaxy), std::max(r.lmaxz, r.rmaxz)};
	double ext = 0.0;
	for (int a = 0; a < 3; ++a) {
		if (!(lo[a] <= hi[a]) || !std::isfinite(lo[a]) || !std::isfinite(hi[a])) return false;
		ext = std::max(ext, hi[a] - lo[a]);
	}
	if (!(ext > 0.0) || !(ext < 1e30)) return false;
	// ONE step for all axes: the error budget of the device arithmetic is in units of (distance of the ray origin
	// from the frame) / step, so a finer grid on a flat axis would only shrink the set of rays that may use it
	const float step = float(ext / double(QNODE_GRID - 4 * (QNODE_MARGIN + 2)));
	if (!(step > 0.0f)) return false;
	for (int a = 0; a < 3; ++a) {
		fr.lo[a] = float(lo[a] - double(step) * 2 * (QNODE_MARGIN + 2));
		fr.step[a] = step;
		if (!std::isfinite(fr.lo[a])) return false;
	}
	fr.extent = float(ext);
	out.resize(bvh2.size());
	bool ok = true;
	auto qlo = [&](float p, int a) {          // largest grid plane <= p - MARGIN steps
		const double g = (double(p) - double(fr.lo[a])) / double(fr.step[a]);
		long q = long(std::floor(g)) - QNODE_MARGIN;
		while (q > 0 && double(fr.lo[a]) + double(q) * double(fr.step[a]) > double(p) - QNODE_MARGIN * double(fr.step[a])) --q;
		if (q < 0 || q > QNODE_GRID) ok = false;
		return uint32_t(std::min<long>(std::max<long>(q, 0), QNODE_GRID));
	};
	auto qhi = [&](float p, int a) {          // smallest grid plane >= p + MARGIN steps
		const double g = (double(p) - double(fr.lo[a])) / double(fr.step[a]);
		long q = long(std::ceil(g)) + QNODE_MARGIN;
		while (q < QNODE_GRID && double(fr.lo[a]) + double(q) * double(fr.step[a]) < double(p) + QNODE_MARGIN * double(fr.step[a])) ++q;
		if (q < 0 || q > QNODE_GRID) ok = false;
		return uint32_t(std::min<long>(std::max<long>(q, 0), QNODE_GRID));
	};
	for (size_t i = 0; i < bvh2.size(); ++i) {
		const DNode& n = bvh2[i];
		QNode& q = out[i];
		q.planes[0] = qlo(n.lminx, 0) | (qhi(n.lmaxx, 0) << 16);
		q.planes[1] = qlo(n.lminy, 1) | (qhi(n.lmaxy, 1) << 16);
		q.planes[2] = qlo(n.lminz, 2) | (qhi(n.lmaxz, 2) << 16);
		q.planes[3] = qlo(n.rminx, 0) | (qhi(n.rmaxx, 0) << 16);
		q.planes[4] = qlo(n.rminy, 1) | (qhi(n.rmaxy, 1) << 16);
		q.planes[5] = qlo(n.rminz, 2) | (qhi(n.rmaxz, 2) << 16);
		q.left = n.left;
		q.right = n.right;
	}
	if (!ok) out.clear();
	return ok;
}

}  // namespace mcpt
