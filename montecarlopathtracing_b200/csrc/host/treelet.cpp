// CPU treelet restructuring ("bvhtype": "treelet"), the host-side counterpart of the reference's
// BVH::TreeletBVH<CPU> (BVH/treeletBVH.cpp:30-365): Karras-Aila 2013 treelets of 7 leaves, processed
// bottom-up, node slots reused.  The output must be memcmp-equal to the reference's, so its observable
// behaviour is reproduced including the parts that look accidental:
//   * per-node SAH values are normalised by the root area, but the dynamic programme adds the
//     UN-normalised Cinn * area of a subset (treeletBVH.cpp:194);
//   * the first leaf slot (index N-1) is classified as an inner node by `id > size/2`
//     (treeletBVH.cpp:303) and gets 2*SAH[node #triangleId] + Cinn*area/rootArea;
//   * treelet leaf j is tied to subset bit (n-1-j) for areas and reconstruction but to bit j for its
//     initial cost (treeletBVH.cpp:99-126 vs :175-177);
//   * the bottom-up schedule follows the ORIGINAL parent links while the working copy changes (:346-358);
//   * treelets that cannot grow to 7 leaves are optimised with the smaller size (:84-87).
// Heap order is observable (it decides which leaf gets which bit), so the same std::push_heap/pop_heap
// primitives are used.
#include <algorithm>
#include <cfloat>
#include <cstdint>
#include <stdexcept>
#include <vector>

#include "host.hpp"

namespace mcpt::host {
namespace {

constexpr float C_INNER = 1.2f;     // auxiliary.h:9-11
constexpr float C_LEAF = 0.0f;
constexpr float C_TRI = 1.0f;
constexpr int TREELET_LEAVES = 7;   // treeletBVH.cpp:15

struct Box4 { float lo[4], hi[4]; };

inline float area_of(const mcpt_float4& bmin, const mcpt_float4& bmax) {     // auxiliary.cpp:15-18
	float x = bmax.x - bmin.x, y = bmax.y - bmin.y, z = bmax.z - bmin.z;
	return 2.0f * (x * y + x * z + y * z);
}
inline float area_of(const Box4& b) {
	float x = b.hi[0] - b.lo[0], y = b.hi[1] - b.lo[1], z = b.hi[2] - b.lo[2];
	return 2.0f * (x * y + x * z + y * z);
}
inline int popcount7(int v) { int c = 0; while (v > 0) { c += v & 1; v >>= 1; } return c; }
inline int single_bit_index(int v) { int b = 0; while ((v >> b) != 1) ++b; return b; }

struct HeapEntry {
	int id;
	float value;
	bool operator<(const HeapEntry& o) const {          // max-heap on (value, id)
		if (value < o.value) return true;
		return value == o.value && id < o.id;
	}
};

struct Optimiser {
	std::vector<mcpt_bvh_node>& bvh;
	std::vector<float> sah;
	float root_area = 0.0f;

	explicit Optimiser(std::vector<mcpt_bvh_node>& nodes) : bvh(nodes) {}

	// treeletBVH.cpp:301-326, memoised DFS with -1 as "not computed"; `active` detects the non-terminating
	// recursion the reference would run into when the first leaf's triangle id names one of its ancestors.
	void sah_of(int id, std::vector<char>& active) {
		if (sah[size_t(id)] != -1.0f) return;
		if (active[size_t(id)]) throw std::runtime_error("treelet: SAH recursion does not terminate for this node array");
		active[size_t(id)] = 1;
		const mcpt_bvh_node& n = bvh[size_t(id)];
		if (size_t(id) > (bvh.size() >> 1)) {
			sah[size_t(id)] = (C_TRI + C_LEAF) * area_of(n.bbmin, n.bbmax) / root_area;
		} else {
			sah_of(n.left, active);
			sah_of(n.right, active);
			sah[size_t(id)] = sah[size_t(n.left)] + sah[size_t(n.right)] + C_INNER * area_of(n.bbmin, n.bbmax) / root_area;
		}
		active[size_t(id)] = 0;
	}

	void restructure(int root) {                         // treeletBVH.cpp:30-297
		std::vector<HeapEntry> heap;
		std::vector<int> free_slots;                     // inner nodes of the old treelet, in expansion order
		heap.push_back({root, sah[size_t(root)]});
		while (int(heap.size()) < TREELET_LEAVES) {
			HeapEntry top = heap.front();
			std::pop_heap(heap.begin(), heap.end());
			heap.pop_back();
			if (top.value < 0.0f) {                      // only leaves left: put it back un-heapified and stop
				heap.push_back({top.id, -1.0f});
				break;
			}
			const int l = bvh[size_t(top.id)].left, r = bvh[size_t(top.id)].right;
			if (l == r) {                                // a leaf stays in the treelet with a negative key
				heap.push_back({top.id, float(top.id) * -1.0f});
				std::push_heap(heap.begin(), heap.end());
				continue;
			}
			heap.push_back({l, sah[size_t(l)]});
			std::push_heap(heap.begin(), heap.end());
			heap.push_back({r, sah[size_t(r)]});
			std::push_heap(heap.begin(), heap.end());
			free_slots.push_back(top.id);
		}
		const int n = int(heap.size());
		if (n < 3) return;
		const int full = (1 << n) - 1;
		// ---- surface area of every subset; element j <-> bit (n-1-j)
		std::vector<float> area(size_t(1) << n, 0.0f);
		for (int s = 1; s <= full; ++s) {
			Box4 b{{FLT_MAX, FLT_MAX, FLT_MAX, FLT_MAX}, {-FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX}};
			for (int j = 0; j < n; ++j) {
				if (!((s >> (n - 1 - j)) & 1)) continue;
				const mcpt_bvh_node& e = bvh[size_t(heap[size_t(j)].id)];
				b.lo[0] = std::min(b.lo[0], e.bbmin.x); b.lo[1] = std::min(b.lo[1], e.bbmin.y); b.lo[2] = std::min(b.lo[2], e.bbmin.z);
				b.hi[0] = std::max(b.hi[0], e.bbmax.x); b.hi[1] = std::max(b.hi[1], e.bbmax.y); b.hi[2] = std::max(b.hi[2], e.bbmax.z);
			}
			area[size_t(s)] = area_of(b);
		}
		// ---- optimal partition of every subset, subsets in order of (size, value); element i <-> bit i here
		std::vector<float> cost(size_t(1) << n, 0.0f);
		std::vector<int> split(size_t(1) << n, 0);
		for (int i = 0; i < n; ++i) cost[size_t(1) << i] = sah[size_t(heap[size_t(i)].id)];
		std::vector<std::pair<int, int>> order;          // (popcount, subset)
		for (int s = 1; s <= full; ++s) order.push_back({popcount7(s), s});
		std::sort(order.begin(), order.end());
		for (const auto& [cnt, s] : order) {
			if (cnt < 2) continue;
			float best = FLT_MAX;
			int best_p = 0;
			const int delta = (s - 1) & s;
			int p = (-delta) & s;
			do {
				const float c = cost[size_t(p)] + cost[size_t(s ^ p)];
				if (c < best) { best = c; best_p = p; }
				p = (p - delta) & s;
			} while (p != 0);
			cost[size_t(s)] = C_INNER * area[size_t(s)] + best;
			split[size_t(s)] = best_p;
		}
		// ---- rewrite the topology breadth-first into the freed slots
		struct Pending { int whole, left_part, slot; };
		std::vector<Pending> level, next;
		size_t used = 0;
		level.push_back({full, split[size_t(full)], free_slots[used++]});
		auto leaf_of = [&](int code) { return heap[size_t(n - single_bit_index(code) - 1)].id; };
		while (!level.empty()) {
			next.clear();
			for (const Pending& it : level) {
				const int parts[2] = {it.left_part, it.whole ^ it.left_part};
				for (int side = 0; side < 2; ++side) {
					const int code = parts[side];
					int child;
					if (popcount7(code) == 1) {
						child = leaf_of(code);
					} else {
						child = free_slots[used++];
						next.push_back({code, split[size_t(code)], child});
					}
					if (side == 0) bvh[size_t(it.slot)].left = child;
					else bvh[size_t(it.slot)].right = child;
					bvh[size_t(child)].parent = it.slot;
				}
			}
			level.swap(next);
		}
		// ---- refit boxes and SAH values bottom-up (slots were handed out top-down)
		for (size_t k = free_slots.size(); k-- > 0;) {
			mcpt_bvh_node& nd = bvh[size_t(free_slots[k])];
			const mcpt_bvh_node& L = bvh[size_t(nd.left)];
			const mcpt_bvh_node& R = bvh[size_t(nd.right)];
			nd.bbmax = {std::max(L.bbmax.x, R.bbmax.x), std::max(L.bbmax.y, R.bbmax.y), std::max(L.bbmax.z, R.bbmax.z), std::max(L.bbmax.w, R.bbmax.w)};
			nd.bbmin = {std::min(L.bbmin.x, R.bbmin.x), std::min(L.bbmin.y, R.bbmin.y), std::min(L.bbmin.z, R.bbmin.z), std::min(L.bbmin.w, R.bbmin.w)};
			sah[size_t(free_slots[k])] = sah[size_t(nd.left)] + sah[size_t(nd.right)] + C_INNER * area_of(nd.bbmin, nd.bbmax) / root_area;
		}
	}
};

}  // namespace

std::vector<mcpt_bvh_node> optimize_treelet(const std::vector<mcpt_bvh_node>& original) {
	if (original.size() < 3) return original;
	std::vector<mcpt_bvh_node> work = original;
	Optimiser opt(work);
	opt.root_area = area_of(work[0].bbmin, work[0].bbmax);
	opt.sah.assign(work.size(), -1.0f);
	std::vector<char> active(work.size(), 0);
	opt.sah_of(0, active);
	std::vector<char> seen(work.size() >> 1, 0);
	for (size_t i = work.size() >> 1; i < work.size(); ++i) {      // every leaf, bottom-up along the ORIGINAL parents
		int p = original[i].parent;
		while (p != -1) {
			if (!seen[size_t(p)]) { seen[size_t(p)] = 1; break; }
			opt.restructure(p);
			p = original[size_t(p)].parent;
		}
	}
	return work;
}

}  // namespace mcpt::host
