// Host LBVH build.  Contract: the returned node array is memcmp-equal to what the reference's
// HLBVH<CPU>::build (BVH/hlbvh.cpp:92-200) produces for the same triangle list; tests compare it
// with the reference's own code compiled into oracle/_ref.
//
// What pins the bytes (hlbvh.cpp line numbers):
//   :97-106   per-triangle box = 4-lane min/max of the three float4 vertices (w = 0); centroid =
//             0.5f * (bbmin + bbmax) in float
//   :107-114  scene box over the CENTROIDS
//   :118-129  q = round((c - gmin) / gsize * 1024.0f) per axis, two float roundings, half away from 0
//   :12-26    1024 clamps to 1023; 30-bit Morton code z<<2 | y<<1 | x
//   :27-63    stable sort on the 30 bits (ties keep triangle order)
//   :138-188  range [l,r] splits after the last element sharing the first element's prefix beyond the
//             range's common prefix; identical codes split at the midpoint; internal node of the left
//             part gets index `split`, of the right part `split+1`; a one-element part becomes leaf
//             `N-1+position`
//   :189-193  leaf: left = right = triangle index, box = triangle box
//   :64-76    internal boxes = min/max of the children, bottom-up
// Different from the reference only in mechanics: 3x10-bit radix passes instead of 5x6, an explicit
// stack instead of a deque (node indices do not depend on visiting order) and a reverse sweep over
// the creation order instead of recursion for the refit.
#include <cfloat>
#include <cmath>
#include <cstring>

#include "host.hpp"

namespace mcpt::host {
namespace {

inline uint32_t spread10(uint32_t x) {
	if (x == 1024u) x = 1023u;
	x = (x | (x << 16)) & 0x030000FFu;
	x = (x | (x << 8)) & 0x0300F00Fu;
	x = (x | (x << 4)) & 0x030C30C3u;
	x = (x | (x << 2)) & 0x09249249u;
	return x;
}

inline uint32_t quantize(float c, float lo, float extent) {
	float d = c - lo;
	d /= extent;
	d *= 1024.0f;
	float r = std::round(d);
	// 0/0 on a flat axis is undefined behaviour in the reference; x86 builds of it get
	// 0x80000000 from cvttss2si, which the bit-spreading masks reduce to 0.  Pin that outcome.
	if (!(r >= 0.0f)) return 0u;
	return uint32_t(r);
}

struct Keyed { uint32_t code; int32_t tri; };

void stable_sort_codes(std::vector<Keyed>& a) {
	std::vector<Keyed> tmp(a.size());
	Keyed* src = a.data();
	Keyed* dst = tmp.data();
	const size_t n = a.size();
	for (int pass = 0; pass < 3; ++pass) {
		const int shift = pass * 10;
		uint32_t count[1025];
		std::memset(count, 0, sizeof(count));
		for (size_t i = 0; i < n; ++i) ++count[((src[i].code >> shift) & 1023u) + 1];
		for (int b = 0; b < 1024; ++b) count[b + 1] += count[b];
		for (size_t i = 0; i < n; ++i) dst[count[(src[i].code >> shift) & 1023u]++] = src[i];
		std::swap(src, dst);
	}
	if (src != a.data()) std::memcpy(a.data(), src, n * sizeof(Keyed));
}

inline int prefix_len(uint32_t a, uint32_t b) {
	uint32_t x = a ^ b;
	return x == 0 ? 32 : __builtin_clz(x);
}

inline size_t find_split(const std::vector<Keyed>& k, size_t left, size_t right) {
	const int target = prefix_len(k[left].code, k[right].code);
	if (target == 32) return (left + right) >> 1;
	do {
		size_t mid = (left + right) >> 1;
		if (prefix_len(k[left].code, k[mid].code) > target) left = mid;
		else right = mid;
	} while (right > left + 1);
	return left;
}

inline mcpt_float4 min4(const mcpt_float4& a, const mcpt_float4& b) {
	return {std::min(a.x, b.x), std::min(a.y, b.y), std::min(a.z, b.z), std::min(a.w, b.w)};
}
inline mcpt_float4 max4(const mcpt_float4& a, const mcpt_float4& b) {
	return {std::max(a.x, b.x), std::max(a.y, b.y), std::max(a.z, b.z), std::max(a.w, b.w)};
}

}  // namespace

std::vector<mcpt_bvh_node> build_hlbvh(const std::vector<mcpt_triangle>& tris) {
	const size_t n = tris.size();
	if (n == 0) throw std::runtime_error("build_hlbvh: empty triangle list");
	std::vector<mcpt_float4> lo(n), hi(n), cen(n);
	mcpt_float4 glo{FLT_MAX, FLT_MAX, FLT_MAX, FLT_MAX}, ghi{-FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX};
	for (size_t i = 0; i < n; ++i) {
		const auto& t = tris[i];
		lo[i] = min4(min4(t.v[0], t.v[1]), t.v[2]);
		hi[i] = max4(max4(t.v[0], t.v[1]), t.v[2]);
		cen[i] = {0.5f * (lo[i].x + hi[i].x), 0.5f * (lo[i].y + hi[i].y), 0.5f * (lo[i].z + hi[i].z),
		          0.5f * (lo[i].w + hi[i].w)};
		glo = min4(glo, cen[i]);
		ghi = max4(ghi, cen[i]);
	}
	const float ex = ghi.x - glo.x, ey = ghi.y - glo.y, ez = ghi.z - glo.z;
	std::vector<Keyed> keys(n);
	for (size_t i = 0; i < n; ++i) {
		uint32_t qx = quantize(cen[i].x, glo.x, ex);
		uint32_t qy = quantize(cen[i].y, glo.y, ey);
		uint32_t qz = quantize(cen[i].z, glo.z, ez);
		keys[i] = {(spread10(qz) << 2) | (spread10(qy) << 1) | spread10(qx), int32_t(i)};
	}
	stable_sort_codes(keys);

	std::vector<mcpt_bvh_node> nodes(2 * n - 1);
	std::memset(nodes.data(), 0, nodes.size() * sizeof(mcpt_bvh_node));
	nodes[0].parent = -1;
	struct Range { uint32_t l, r, node; };
	std::vector<Range> todo;
	std::vector<uint32_t> created;          // internal nodes, parents before children
	created.reserve(n);
	if (n > 1) todo.push_back({0u, uint32_t(n - 1), 0u});
	while (!todo.empty()) {
		Range g = todo.back();
		todo.pop_back();
		created.push_back(g.node);
		size_t s = find_split(keys, g.l, g.r);
		size_t li = (s != g.l) ? s : s + n - 1;
		size_t ri = (s + 1 != g.r) ? s + 1 : s + n;
		nodes[g.node].left = int32_t(li);
		nodes[g.node].right = int32_t(ri);
		nodes[li].parent = int32_t(g.node);
		nodes[ri].parent = int32_t(g.node);
		if (li == s) todo.push_back({g.l, uint32_t(s), uint32_t(s)});
		if (ri == s + 1) todo.push_back({uint32_t(s + 1), g.r, uint32_t(s + 1)});
	}
	for (size_t i = 0; i < n; ++i) {
		auto& leaf = nodes[n - 1 + i];
		leaf.left = leaf.right = keys[i].tri;
		leaf.bbmin = lo[size_t(keys[i].tri)];
		leaf.bbmax = hi[size_t(keys[i].tri)];
	}
	for (size_t k = created.size(); k-- > 0;) {
		auto& nd = nodes[created[k]];
		nd.bbmin = min4(nodes[size_t(nd.left)].bbmin, nodes[size_t(nd.right)].bbmin);
		nd.bbmax = max4(nodes[size_t(nd.left)].bbmax, nodes[size_t(nd.right)].bbmax);
	}
	return nodes;
}

// bvhtest.cpp:104-115 — SAH with Cinn = 1.2, Ctri = 1 (auxiliary.h:9-11), normalised by root area.
double sah_cost(const std::vector<mcpt_bvh_node>& nodes) {
	auto area = [](const mcpt_bvh_node& b) {
		float dx = b.bbmax.x - b.bbmin.x, dy = b.bbmax.y - b.bbmin.y, dz = b.bbmax.z - b.bbmin.z;
		return 2.0f * (dx * dy + dx * dz + dy * dz);
	};
	const double root = area(nodes[0]);
	double cost = 0.0;
	for (const auto& nd : nodes) cost += (nd.left == nd.right ? 1.0 : 1.2) * area(nd);
	return cost / root;
}

}  // namespace mcpt::host
