// EXPERIMENTAL (MCPT_TRAVERSAL_WIDE8): collapse of the fast walk's binary tree into 8-wide compressed nodes.  Measured
// slower than the binary walk on B200; built lazily on first use (ensure_wide in mcpt_cuda.cu), never on
// the product path.
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#include "accel_wide.hpp"

namespace mcpt {
namespace {
struct Box {
	float lo[3], hi[3];
	void reset() { lo[0] = lo[1] = lo[2] = FLT_MAX; hi[0] = hi[1] = hi[2] = -FLT_MAX; }
	void grow(const Box& b) {
		for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], b.lo[k]); hi[k] = std::max(hi[k], b.hi[k]); }
	}
	float half_area() const {
		float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
		return dx * dy + dy * dz + dz * dx;
	}
};
}  // namespace
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

}  // namespace mcpt
