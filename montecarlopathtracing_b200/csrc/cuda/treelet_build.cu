// Treelet restructuring on the device ("bvhtype": "treeletGPU").
// Contract: the node array is memcmp-equal to the host optimiser (csrc/host/treelet.cpp), which is
// memcmp-equal to the reference's BVH::TreeletBVH<CPU> (BVH/treeletBVH.cpp:30-365).  The reference's own
// device kernel (kernels/treeletBVH.cl:230-531) relies on lock-step warps and unordered flag updates and is
// not deterministic, so the bar is the CPU result (SURVEY f-3) - including its accidental behaviours, which
// are listed in treelet.cpp and reproduced here.
//
// Why a parallel schedule can give the sequential answer.  The reference walks the leaves in index order and
// climbs the ORIGINAL parent links; the first arrival at a node only marks it, the second restructures it
// (treeletBVH.cpp:346-358).  That is a post-order: a node is restructured after both child subtrees are
// finished.  restructure(p) reads and writes only slots of subtree(p) - it re-uses the inner slots of its
// treelet, so the SET of slots below p never changes - hence restructurings in disjoint subtrees commute, and
// any schedule that respects "children before parent" produces the same bytes.  Here: one warp per leaf, an
// atomic arrival counter per inner node (Karras & Aila 2013), the second arriver restructures and climbs on.
//
// One warp per treelet:
//   lane 0   grows the treelet with a max-heap on (SAH, id); heap ORDER is observable (it decides which leaf
//            gets which subset bit), so push/pop are libstdc++'s __push_heap/__adjust_heap step for step;
//   32 lanes surface area of the 2^n subsets, then the subset DP level by level (popcount 2..n).  Up to
//            popcount 5 a lane owns whole subsets and scans their partitions in the reference's order
//            (p = (p - delta) & s, first strict minimum wins); for popcount 6 and 7 the 31 / 63 partitions
//            of a subset are spread over the lanes - partition j is pdep(2j, s) - and combined by
//            (cost, j) lexicographic minimum, which is the same "first strict minimum";
//   lane 0   writes the new topology breadth-first into the freed slots, refits boxes and SAH bottom-up
//            with the host's operand order.
// All node / SAH traffic of the second pass uses ld.global.cg + __threadfence around the arrival counter:
// an SM's L1 may hold lines another SM's warp has rewritten since.
#include <cuda_runtime.h>

#ifndef MCPT_CHECKED
#define MCPT_CHECKED 0
#endif

#include <cfloat>
#include <cstdint>
#include <string>
#include <vector>

#include "mcpt.h"

namespace mcpt {
void set_last_error(const std::string& s);     // mcpt_cuda.cu
namespace {

constexpr float C_INNER = 1.2f;      // auxiliary.h:9-11
constexpr float C_LEAF = 0.0f;
constexpr float C_TRI = 1.0f;
constexpr int TL = 7;                // treeletBVH.cpp:15
constexpr int WARPS = 4;             // warps per block

__device__ __forceinline__ float hmin(float a, float b) { return (b < a) ? b : a; }     // std::min(a, b)
__device__ __forceinline__ float hmax(float a, float b) { return (a < b) ? b : a; }     // std::max(a, b)

__device__ __forceinline__ float area3(float lx, float ly, float lz, float hx, float hy, float hz) {   // auxiliary.cpp:15-18
	const float x = hx - lx, y = hy - ly, z = hz - lz;
	return 2.0f * (x * y + x * z + y * z);
}

struct NodeView {                     // 64-byte node seen as four 16-byte words
	float4 lo, hi;
	int4 link;                        // parent, left, right, pad
};
__device__ __forceinline__ float4 ld_lo(const mcpt_bvh_node* n, int i) { return __ldcg(reinterpret_cast<const float4*>(n + i)); }
__device__ __forceinline__ float4 ld_hi(const mcpt_bvh_node* n, int i) { return __ldcg(reinterpret_cast<const float4*>(n + i) + 1); }
__device__ __forceinline__ int4 ld_link(const mcpt_bvh_node* n, int i) { return __ldcg(reinterpret_cast<const int4*>(n + i) + 3); }

// ---------------------------------------------------------------------------------------------------------
// Pass 1: per-node SAH of the input tree (treeletBVH.cpp:301-326), normalised by the root area.
// Leaf slot N-1 is the reference's accident: `id > size/2` classifies it as an inner node whose two
// children are both "node #triangleId".  It does not start a climb here; k_sah_first_leaf finishes its chain.
__global__ void k_sah_init(const mcpt_bvh_node* __restrict__ nodes, int nleaves, float root_area, float* sah, int* arrive) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= nleaves || i == 0) return;
	int id = nleaves - 1 + i;
	{
		const float4 lo = ld_lo(nodes, id), hi = ld_hi(nodes, id);
		sah[id] = (C_TRI + C_LEAF) * area3(lo.x, lo.y, lo.z, hi.x, hi.y, hi.z) / root_area;
	}
	int p = ld_link(nodes, id).x;
	while (p >= 0) {
		__threadfence();
		if (atomicAdd(arrive + p, 1) == 0) return;
		__threadfence();
		const int4 lk = ld_link(nodes, p);
		const float4 lo = ld_lo(nodes, p), hi = ld_hi(nodes, p);
		const float sl = __ldcg(sah + lk.y), sr = __ldcg(sah + lk.z);
		sah[p] = sl + sr + C_INNER * area3(lo.x, lo.y, lo.z, hi.x, hi.y, hi.z) / root_area;
		p = lk.x;
	}
}

// status: 0 ok, 1 = the reference's recursion would not terminate (the first leaf's triangle id names the
// leaf itself or one of its ancestors, whose value needs the leaf's).
__global__ void k_sah_first_leaf(const mcpt_bvh_node* __restrict__ nodes, int nleaves, float root_area, float* sah, int* arrive, int* status) {
	const int first = nleaves - 1;
	const int4 lk0 = ld_link(nodes, first);
	const int t = lk0.y;
	bool bad = (t < 0 || t >= 2 * nleaves - 1 || t == first);
	int steps = 0;                           // a parent chain longer than the node count is a cycle: malformed input
	for (int a = lk0.x; a >= 0 && !bad; a = ld_link(nodes, a).x) bad = (a == t) || (++steps > 2 * nleaves);
	if (bad) { *status = steps > 2 * nleaves ? 3 : 1; return; }
	{
		const float4 lo = ld_lo(nodes, first), hi = ld_hi(nodes, first);
		const float st = __ldcg(sah + t);
		sah[first] = st + st + C_INNER * area3(lo.x, lo.y, lo.z, hi.x, hi.y, hi.z) / root_area;
	}
	for (int p = lk0.x; p >= 0;) {          // every sibling subtree on this chain is complete (chain length checked above)
		const int4 lk = ld_link(nodes, p);
		const float4 lo = ld_lo(nodes, p), hi = ld_hi(nodes, p);
		const float sl = __ldcg(sah + lk.y), sr = __ldcg(sah + lk.z);
		sah[p] = sl + sr + C_INNER * area3(lo.x, lo.y, lo.z, hi.x, hi.y, hi.z) / root_area;
		__threadfence();
		p = lk.x;
	}
	(void)arrive;
}

// ---------------------------------------------------------------------------------------------------------
// Pass 2: the restructuring itself.
struct Ent { int id; float v; };
__device__ __forceinline__ bool ent_less(const Ent& a, const Ent& b) {      // max-heap on (value, id), treeletBVH.cpp:35-41
	if (a.v < b.v) return true;
	return a.v == b.v && a.id < b.id;
}
// libstdc++ std::push_heap over h[0, len): the new element sits at h[len-1]
__device__ void heap_push(Ent* h, int len) {
	const Ent value = h[len - 1];
	int hole = len - 1, parent = (hole - 1) / 2;
	while (hole > 0 && ent_less(h[parent], value)) { h[hole] = h[parent]; hole = parent; parent = (hole - 1) / 2; }
	h[hole] = value;
}
// libstdc++ std::pop_heap over h[0, len): the old top ends at h[len-1], h[0, len-1) is a heap again
__device__ void heap_pop(Ent* h, int len) {
	if (len <= 1) return;
	const Ent value = h[len - 1];
	h[len - 1] = h[0];
	len -= 1;
	int hole = 0, child = 0;
	while (child < (len - 1) / 2) {
		child = 2 * (child + 1);
		if (ent_less(h[child], h[child - 1])) --child;
		h[hole] = h[child];
		hole = child;
	}
	if ((len & 1) == 0 && child == (len - 2) / 2) {
		child = 2 * (child + 1);
		h[hole] = h[child - 1];
		hole = child - 1;
	}
	int parent = (hole - 1) / 2;
	while (hole > 0 && ent_less(h[parent], value)) { h[hole] = h[parent]; hole = parent; parent = (hole - 1) / 2; }
	h[hole] = value;
}

struct WarpScratch {
	Ent heap[8];
	int slots[8];                 // inner nodes of the old treelet, in expansion order (free_slots)
	int new_left[8], new_right[8];
	float box[TL][6];
	float area[1 << TL];
	float cost[1 << TL];
	unsigned char split[1 << TL];
	int n;
};

// partition number j (1-based) of subset s in the reference's enumeration order p = (p - delta) & s:
// bit r >= 1 of s (by rank) carries bit r-1 of j, the lowest bit of s stays on the other side
__device__ __forceinline__ int partition_of(int j, int s) {
	int p = 0, r = 0;
#pragma unroll
	for (int b = 0; b < TL; ++b) {
		if ((s >> b) & 1) {
			if (r > 0 && ((j >> (r - 1)) & 1)) p |= 1 << b;
			++r;
		}
	}
	return p;
}

__device__ void restructure(mcpt_bvh_node* nodes, float* sah, float root_area, int root, WarpScratch& w, int lane, int* status) {
	// ---- grow the treelet (treeletBVH.cpp:44-92)
	if (lane == 0) {
		int hn = 1, nfree = 0;
		w.heap[0] = Ent{root, __ldcg(sah + root)};
		while (hn < TL) {
			const Ent top = w.heap[0];
			heap_pop(w.heap, hn);
			--hn;
			if (top.v < 0.0f) {                      // only leaves left: back in, un-heapified
				w.heap[hn++] = Ent{top.id, -1.0f};
				break;
			}
			const int4 lk = ld_link(nodes, top.id);
			if (lk.y == lk.z) {                      // a leaf stays with a negative key
				w.heap[hn++] = Ent{top.id, float(top.id) * -1.0f};
				heap_push(w.heap, hn);
				continue;
			}
			w.heap[hn++] = Ent{lk.y, __ldcg(sah + lk.y)};
			heap_push(w.heap, hn);
			w.heap[hn++] = Ent{lk.z, __ldcg(sah + lk.z)};
			heap_push(w.heap, hn);
			w.slots[nfree++] = top.id;
		}
		w.n = hn;
	}
	__syncwarp();
	const int n = w.n;
	if (n < 3) return;
	const int full = (1 << n) - 1;
	if (lane < n) {
		const int id = w.heap[lane].id;
		const float4 lo = ld_lo(nodes, id), hi = ld_hi(nodes, id);
		w.box[lane][0] = lo.x; w.box[lane][1] = lo.y; w.box[lane][2] = lo.z;
		w.box[lane][3] = hi.x; w.box[lane][4] = hi.y; w.box[lane][5] = hi.z;
		w.cost[1 << lane] = __ldcg(sah + id);        // element i <-> bit i for the initial costs (:175-177)
	}
	__syncwarp();
	// ---- surface area of every subset; element j <-> bit (n-1-j) (:99-126)
	for (int s = lane; s <= full; s += 32) {
		if (s == 0) continue;
		float lx = FLT_MAX, ly = FLT_MAX, lz = FLT_MAX, hx = -FLT_MAX, hy = -FLT_MAX, hz = -FLT_MAX;
		for (int j = 0; j < n; ++j) {
			if (!((s >> (n - 1 - j)) & 1)) continue;
			lx = hmin(lx, w.box[j][0]); ly = hmin(ly, w.box[j][1]); lz = hmin(lz, w.box[j][2]);
			hx = hmax(hx, w.box[j][3]); hy = hmax(hy, w.box[j][4]); hz = hmax(hz, w.box[j][5]);
		}
		w.area[s] = area3(lx, ly, lz, hx, hy, hz);
	}
	__syncwarp();
	// ---- optimal partition of every subset, by popcount (:128-200)
	for (int k = 2; k <= n; ++k) {
		if (k <= 5) {
			for (int s = lane; s <= full; s += 32) {
				if (__popc(s) != k) continue;
				float best = FLT_MAX;
				int best_p = 0;
				const int delta = (s - 1) & s;
				int p = (-delta) & s;
				do {
					const float c = w.cost[p] + w.cost[s ^ p];
					if (c < best) { best = c; best_p = p; }
					p = (p - delta) & s;
				} while (p != 0);
				w.cost[s] = C_INNER * w.area[s] + best;
				w.split[s] = (unsigned char)best_p;
			}
		} else {
			const int nsub = (k == n) ? 1 : n;       // popcount n: the full set; popcount n-1 (n = 7, k = 6): drop one bit
			const int parts = (1 << (k - 1)) - 1;
			for (int b = 0; b < nsub; ++b) {
				const int s = (k == n) ? full : (full ^ (1 << b));
				float best = FLT_MAX;
				int best_j = 0x7fffffff, best_p = 0;
				for (int j = lane + 1; j <= parts; j += 32) {
					const int p = partition_of(j, s);
					const float c = w.cost[p] + w.cost[s ^ p];
					if (c < best) { best = c; best_j = j; best_p = p; }
				}
#pragma unroll
				for (int o = 16; o > 0; o >>= 1) {
					const float oc = __shfl_xor_sync(0xffffffffu, best, o);
					const int oj = __shfl_xor_sync(0xffffffffu, best_j, o);
					const int op = __shfl_xor_sync(0xffffffffu, best_p, o);
					if (oc < best || (oc == best && oj < best_j)) { best = oc; best_j = oj; best_p = op; }
				}
				if (lane == 0) {
					w.cost[s] = C_INNER * w.area[s] + best;
					w.split[s] = (unsigned char)best_p;
				}
			}
		}
		__syncwarp();
	}
	// ---- new topology breadth-first into the freed slots, then boxes and SAH bottom-up (:202-297)
	if (lane == 0) {
		const int nfree = n - 1;
		int q_code[8], q_pos[8];
		int qh = 0, qt = 0, used = 0;
		bool ok = true;
		q_code[qt] = full; q_pos[qt] = used++; ++qt;
		while (qh < qt && ok) {
			const int whole = q_code[qh], pos = q_pos[qh];
			++qh;
			const int lp = w.split[whole];
			const int slot = w.slots[pos];
			for (int side = 0; side < 2; ++side) {
				const int code = side == 0 ? lp : (whole ^ lp);
				int child;
				if (__popc(code) == 1) {
					child = w.heap[n - (__ffs(code) - 1) - 1].id;
				} else {
					if (used >= nfree || code == 0) { ok = false; break; }      // NaN costs: the reference runs off its vector here
					q_code[qt] = code; q_pos[qt] = used; ++qt;
					child = w.slots[used++];
				}
				if (side == 0) w.new_left[pos] = child; else w.new_right[pos] = child;
				int* cl = reinterpret_cast<int*>(nodes + child);
				cl[12] = slot;                                                  // parent
			}
			int* sl = reinterpret_cast<int*>(nodes + slot);
			sl[13] = w.new_left[pos];
			sl[14] = w.new_right[pos];
		}
		if (!ok) { atomicExch(status, 2); return; }
		for (int k = nfree - 1; k >= 0; --k) {
			const int slot = w.slots[k], l = w.new_left[k], r = w.new_right[k];
			const float4 Llo = ld_lo(nodes, l), Lhi = ld_hi(nodes, l), Rlo = ld_lo(nodes, r), Rhi = ld_hi(nodes, r);
			const float4 hi = make_float4(hmax(Lhi.x, Rhi.x), hmax(Lhi.y, Rhi.y), hmax(Lhi.z, Rhi.z), hmax(Lhi.w, Rhi.w));
			const float4 lo = make_float4(hmin(Llo.x, Rlo.x), hmin(Llo.y, Rlo.y), hmin(Llo.z, Rlo.z), hmin(Llo.w, Rlo.w));
			float4* dst = reinterpret_cast<float4*>(nodes + slot);
			dst[0] = lo;
			dst[1] = hi;
			sah[slot] = __ldcg(sah + l) + __ldcg(sah + r) + C_INNER * area3(lo.x, lo.y, lo.z, hi.x, hi.y, hi.z) / root_area;
		}
	}
}

__global__ void __launch_bounds__(32 * WARPS) k_treelet(mcpt_bvh_node* nodes, const int* __restrict__ orig_parent, int nleaves,
                                                         float root_area, float* sah, int* arrive, int* status) {
	__shared__ WarpScratch scratch[WARPS];
	const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
	const int leaf = blockIdx.x * WARPS + wib;
	if (leaf >= nleaves) return;
	WarpScratch& w = scratch[wib];
	int p = orig_parent[nleaves - 1 + leaf];
	while (p >= 0) {
		int prev = 0;
		if (lane == 0) {
			__threadfence();
			prev = atomicAdd(arrive + p, 1);
			__threadfence();
		}
		prev = __shfl_sync(0xffffffffu, prev, 0);
#if MCPT_CHECKED
		// protocol check (checked build): an inner node is reached exactly twice, through in-range parent links
		if (prev < 0 || prev > 1 || p >= nleaves - 1) { if (lane == 0) atomicExch(status, 7); return; }
#endif
		if (prev == 0) return;                       // first arrival only marks the node (:349-352)
		restructure(nodes, sah, root_area, p, w, lane, status);
		__syncwarp();
		p = orig_parent[p];
	}
}

__global__ void k_copy_parents(const mcpt_bvh_node* __restrict__ nodes, int nn, int* parent) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < nn) parent[i] = nodes[i].parent;
}

template <class T>
struct Scoped {
	T* p = nullptr;
	cudaError_t alloc(size_t n) { return cudaMalloc(reinterpret_cast<void**>(&p), (n ? n : 1) * sizeof(T)); }
	~Scoped() { if (p) cudaFree(p); }
};
inline int blocks(long n, int bs) { return int((n + bs - 1) / bs); }

}  // namespace
}  // namespace mcpt

using namespace mcpt;

#define TRY_(x)                                                                                         \
	do {                                                                                                \
		cudaError_t e_ = (x);                                                                           \
		if (e_ != cudaSuccess) {                                                                        \
			set_last_error(std::string(#x) + ": " + cudaGetErrorString(e_));                            \
			return MCPT_ERR_CUDA;                                                                       \
		}                                                                                               \
	} while (0)

extern "C" int mcpt_optimize_treelet(int device, const mcpt_bvh_node* nodes_in, int nnodes, mcpt_bvh_node* nodes_out, float* build_ms) {
	if (!nodes_in || !nodes_out || nnodes <= 0 || (nnodes & 1) == 0) {
		set_last_error("mcpt_optimize_treelet: expected a node array of odd length 2*ntris-1 and room for as many nodes");
		return MCPT_ERR_ARG;
	}
	if (build_ms) *build_ms = 0.0f;
	if (nnodes < 3) {                                    // treeletBVH.cpp: nothing to restructure
		if (nodes_out != nodes_in) nodes_out[0] = nodes_in[0];
		return nnodes;
	}
	const int nleaves = (nnodes + 1) / 2;
	// The kernels climb parent links and descend child links without bounds: make sure the array is ONE binary tree
	// in the reference's layout (inner nodes 0..N-2, leaves N-1..2N-2, consistent parent links) before launching.
	{
		std::vector<int> todo{0};
		std::vector<char> seen(size_t(nnodes), 0);
		size_t reached = 0;
		bool ok = nodes_in[0].parent == -1;
		while (ok && !todo.empty()) {
			const int i = todo.back();
			todo.pop_back();
			if (seen[size_t(i)]) { ok = false; break; }
			seen[size_t(i)] = 1;
			++reached;
			const mcpt_bvh_node& nd = nodes_in[i];
			if (i >= nleaves - 1) { ok = nd.left == nd.right; continue; }
			ok = nd.left >= 0 && nd.left < nnodes && nd.right >= 0 && nd.right < nnodes && nd.left != nd.right &&
			     nodes_in[nd.left].parent == i && nodes_in[nd.right].parent == i;
			if (ok) { todo.push_back(nd.right); todo.push_back(nd.left); }
		}
		if (!ok || reached != size_t(nnodes)) {
			set_last_error("mcpt_optimize_treelet: the node array is not a binary tree in the reference's layout (inner nodes first, consistent parent / child links)");
			return MCPT_ERR_ARG;
		}
	}
	TRY_(cudaSetDevice(device));
	cudaStream_t st;
	TRY_(cudaStreamCreate(&st));
	struct StreamGuard { cudaStream_t s; ~StreamGuard() { cudaStreamDestroy(s); } } guard{st};
	Scoped<mcpt_bvh_node> nodes;
	Scoped<float> sah;
	Scoped<int> arrive, parent, status;
	TRY_(nodes.alloc(size_t(nnodes))); TRY_(sah.alloc(size_t(nnodes))); TRY_(arrive.alloc(size_t(nleaves)));
	TRY_(parent.alloc(size_t(nnodes))); TRY_(status.alloc(1));
	TRY_(cudaMemcpyAsync(nodes.p, nodes_in, sizeof(mcpt_bvh_node) * size_t(nnodes), cudaMemcpyHostToDevice, st));
	cudaEvent_t e0, e1;
	TRY_(cudaEventCreate(&e0)); TRY_(cudaEventCreate(&e1));
	struct EventGuard { cudaEvent_t a, b; ~EventGuard() { cudaEventDestroy(a); cudaEventDestroy(b); } } eguard{e0, e1};
	TRY_(cudaEventRecord(e0, st));
	const mcpt_bvh_node& root = nodes_in[0];
	const float rx = root.bbmax.x - root.bbmin.x, ry = root.bbmax.y - root.bbmin.y, rz = root.bbmax.z - root.bbmin.z;
	const float root_area = 2.0f * (rx * ry + rx * rz + ry * rz);
	TRY_(cudaMemsetAsync(arrive.p, 0, sizeof(int) * size_t(nleaves), st));
	TRY_(cudaMemsetAsync(status.p, 0, sizeof(int), st));
	k_copy_parents<<<blocks(nnodes, 256), 256, 0, st>>>(nodes.p, nnodes, parent.p);
	k_sah_init<<<blocks(nleaves, 128), 128, 0, st>>>(nodes.p, nleaves, root_area, sah.p, arrive.p);
	k_sah_first_leaf<<<1, 1, 0, st>>>(nodes.p, nleaves, root_area, sah.p, arrive.p, status.p);
	int h_status = 0;
	TRY_(cudaMemcpyAsync(&h_status, status.p, sizeof(int), cudaMemcpyDeviceToHost, st));
	TRY_(cudaStreamSynchronize(st));
	if (h_status == 3) {
		set_last_error("mcpt_optimize_treelet: the parent links of the first leaf do not reach the root (malformed node array)");
		return MCPT_ERR_ARG;
	}
	if (h_status == 1) {                                 // same condition and wording as csrc/host/treelet.cpp
		set_last_error("treelet: SAH recursion does not terminate for this node array");
		return MCPT_ERR_STATE;
	}
	TRY_(cudaMemsetAsync(arrive.p, 0, sizeof(int) * size_t(nleaves), st));
	k_treelet<<<blocks(nleaves, WARPS), 32 * WARPS, 0, st>>>(nodes.p, parent.p, nleaves, root_area, sah.p, arrive.p, status.p);
	TRY_(cudaEventRecord(e1, st));
	TRY_(cudaMemcpyAsync(&h_status, status.p, sizeof(int), cudaMemcpyDeviceToHost, st));
	TRY_(cudaMemcpyAsync(nodes_out, nodes.p, sizeof(mcpt_bvh_node) * size_t(nnodes), cudaMemcpyDeviceToHost, st));
	TRY_(cudaStreamSynchronize(st));
	TRY_(cudaGetLastError());
#if MCPT_CHECKED
	if (h_status == 7) {
		set_last_error("treelet (checked build): arrival protocol violated (a node reached a third time or a parent link out of range)");
		return MCPT_ERR_STATE;
	}
	if (h_status == 0 && nleaves > 1) {          // every inner node must have been reached exactly twice
		std::vector<int> h_arrive(size_t(nleaves) - 1);
		TRY_(cudaMemcpy(h_arrive.data(), arrive.p, sizeof(int) * h_arrive.size(), cudaMemcpyDeviceToHost));
		for (size_t i = 0; i < h_arrive.size(); ++i)
			if (h_arrive[i] != 2) {
				set_last_error("treelet (checked build): inner node " + std::to_string(i) + " was reached " + std::to_string(h_arrive[i]) + " times");
				return MCPT_ERR_STATE;
			}
	}
#endif
	if (h_status != 0) {
		set_last_error("treelet: a subset partition came out empty (non-finite SAH costs); the reference indexes past its slot list here");
		return MCPT_ERR_STATE;
	}
	if (build_ms) TRY_(cudaEventElapsedTime(build_ms, e0, e1));
	return nnodes;
}
