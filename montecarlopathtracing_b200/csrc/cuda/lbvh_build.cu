// LBVH build on the device.  Contract: the node array is memcmp-equal to the host
// build (csrc/host/lbvh.cpp), which is memcmp-equal to the reference's HLBVH<CPU>::build
// (BVH/hlbvh.cpp:92-200).  What pins the bytes is listed in lbvh.cpp; the device version keeps every
// rounding step (IEEE subtract / divide / multiply / round-half-away, -fmad=false) and replaces only
// order-free mechanics:
//   * triangle boxes, centroids and Morton codes: one thread per triangle; the scene box over the centroids
//     is a min/max reduction (exact in any order);
//   * the stable sort on the 30-bit code: cub::DeviceRadixSort::SortPairs (LSD radix sort, stable);
//   * the hierarchy: the reference splits ranges top-down (hlbvh.cpp:138-188) and names the internal node of
//     a range after its split position, so the node set does not depend on processing order: one kernel per
//     LEVEL splits all ranges of that level in parallel (binary search per range) and appends the child
//     ranges to the next level's queue.  Ranges of identical codes split at their midpoint like the
//     reference (Karras' index-augmented keys would split them differently);
//   * boxes: levels in reverse order, parents = min/max of the children with the host's operand order.
#include <cuda_runtime.h>

#include <algorithm>
#include <cfloat>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include <cub/device/device_radix_sort.cuh>

#include "mcpt.h"

namespace mcpt {
void set_last_error(const std::string& s);     // mcpt_cuda.cu
namespace {

__device__ __forceinline__ float hmin(float a, float b) { return (b < a) ? b : a; }     // std::min(a, b)
__device__ __forceinline__ float hmax(float a, float b) { return (a < b) ? b : a; }     // std::max(a, b)
__device__ __forceinline__ float4 min4(float4 a, float4 b) { return make_float4(hmin(a.x, b.x), hmin(a.y, b.y), hmin(a.z, b.z), hmin(a.w, b.w)); }
__device__ __forceinline__ float4 max4(float4 a, float4 b) { return make_float4(hmax(a.x, b.x), hmax(a.y, b.y), hmax(a.z, b.z), hmax(a.w, b.w)); }

// order-preserving float <-> int map for atomicMin/atomicMax
__device__ __forceinline__ int f2ord(float f) { int b = __float_as_int(f); return b >= 0 ? b : b ^ 0x7fffffff; }
__device__ __forceinline__ float ord2f(int o) { return __int_as_float(o >= 0 ? o : o ^ 0x7fffffff); }

__global__ void k_tri_boxes(const mcpt_triangle* __restrict__ tris, int n, float4* lo, float4* hi, float4* cen, int* gbox) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	float4 c = make_float4(FLT_MAX, FLT_MAX, FLT_MAX, FLT_MAX), cmax = make_float4(-FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX);
	if (i < n) {
		const float4* v = reinterpret_cast<const float4*>(tris + i);
		const float4 v0 = v[0], v1 = v[1], v2 = v[2];
		const float4 l = min4(min4(v0, v1), v2), h = max4(max4(v0, v1), v2);
		c = make_float4(0.5f * (l.x + h.x), 0.5f * (l.y + h.y), 0.5f * (l.z + h.z), 0.5f * (l.w + h.w));
		cmax = c;
		lo[i] = l; hi[i] = h; cen[i] = c;
	}
	// warp, then one atomic per warp and component (xyz only: the w lane never enters a Morton code)
	for (int o = 16; o > 0; o >>= 1) {
		c.x = fminf(c.x, __shfl_xor_sync(0xffffffffu, c.x, o)); c.y = fminf(c.y, __shfl_xor_sync(0xffffffffu, c.y, o)); c.z = fminf(c.z, __shfl_xor_sync(0xffffffffu, c.z, o));
		cmax.x = fmaxf(cmax.x, __shfl_xor_sync(0xffffffffu, cmax.x, o)); cmax.y = fmaxf(cmax.y, __shfl_xor_sync(0xffffffffu, cmax.y, o)); cmax.z = fmaxf(cmax.z, __shfl_xor_sync(0xffffffffu, cmax.z, o));
	}
	if ((threadIdx.x & 31) == 0) {
		atomicMin(gbox + 0, f2ord(c.x)); atomicMin(gbox + 1, f2ord(c.y)); atomicMin(gbox + 2, f2ord(c.z));
		atomicMax(gbox + 3, f2ord(cmax.x)); atomicMax(gbox + 4, f2ord(cmax.y)); atomicMax(gbox + 5, f2ord(cmax.z));
	}
}

__device__ __forceinline__ uint32_t spread10(uint32_t x) {      // hlbvh.cpp:12-26
	if (x == 1024u) x = 1023u;
	x = (x | (x << 16)) & 0x030000FFu;
	x = (x | (x << 8)) & 0x0300F00Fu;
	x = (x | (x << 4)) & 0x030C30C3u;
	x = (x | (x << 2)) & 0x09249249u;
	return x;
}
__device__ __forceinline__ uint32_t quantize(float c, float lo, float extent) {     // hlbvh.cpp:118-129
	float d = c - lo;
	d = d / extent;
	d = d * 1024.0f;
	const float r = roundf(d);               // half away from zero, like std::round
	if (!(r >= 0.0f)) return 0u;             // 0/0 on a flat axis: pinned to 0 like the host build
	return uint32_t(r);
}

__global__ void k_morton(const float4* __restrict__ cen, int n, const int* __restrict__ gbox, uint32_t* codes, int32_t* ids) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const float lx = ord2f(gbox[0]), ly = ord2f(gbox[1]), lz = ord2f(gbox[2]);
	const float ex = ord2f(gbox[3]) - lx, ey = ord2f(gbox[4]) - ly, ez = ord2f(gbox[5]) - lz;
	const float4 c = cen[i];
	codes[i] = (spread10(quantize(c.z, lz, ez)) << 2) | (spread10(quantize(c.y, ly, ey)) << 1) | spread10(quantize(c.x, lx, ex));
	ids[i] = i;
}

struct Range { uint32_t l, r, node; };

__device__ __forceinline__ int prefix_len(uint32_t a, uint32_t b) { const uint32_t x = a ^ b; return x == 0 ? 32 : __clz(int(x)); }

// One level of hlbvh.cpp:138-188: every range of `in` is split; child ranges with more than one key go to `out`.
__global__ void k_split_level(const uint32_t* __restrict__ codes, int n, const Range* __restrict__ in, uint32_t nin, Range* out, uint32_t* nout,
                              mcpt_bvh_node* nodes) {
	const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
	if (k >= nin) return;
	const Range g = in[k];
	uint32_t left = g.l, right = g.r;
	const int target = prefix_len(codes[left], codes[right]);
	uint32_t s;
	if (target == 32) s = (left + right) >> 1;
	else {
		const uint32_t first = codes[g.l];
		do {
			const uint32_t mid = (left + right) >> 1;
			if (prefix_len(first, codes[mid]) > target) left = mid; else right = mid;
		} while (right > left + 1);
		s = left;
	}
	const uint32_t li = (s != g.l) ? s : s + uint32_t(n) - 1u;
	const uint32_t ri = (s + 1 != g.r) ? s + 1 : s + uint32_t(n);
	nodes[g.node].left = int32_t(li);
	nodes[g.node].right = int32_t(ri);
	nodes[li].parent = int32_t(g.node);
	nodes[ri].parent = int32_t(g.node);
	const bool pl = li == s, pr = ri == s + 1;
	if (pl || pr) {
		uint32_t pos = atomicAdd(nout, uint32_t(pl) + uint32_t(pr));
		if (pl) out[pos++] = Range{g.l, s, s};
		if (pr) out[pos] = Range{s + 1, g.r, s + 1};
	}
}

__global__ void k_leaves(const int32_t* __restrict__ ids, const float4* __restrict__ lo, const float4* __restrict__ hi, int n, mcpt_bvh_node* nodes) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const int t = ids[i];
	mcpt_bvh_node& leaf = nodes[n - 1 + i];
	leaf.left = leaf.right = t;
	*reinterpret_cast<float4*>(&leaf.bbmin) = lo[t];
	*reinterpret_cast<float4*>(&leaf.bbmax) = hi[t];
}

__global__ void k_refit_level(const Range* __restrict__ level, uint32_t cnt, mcpt_bvh_node* nodes) {
	const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
	if (k >= cnt) return;
	mcpt_bvh_node& nd = nodes[level[k].node];
	const mcpt_bvh_node& L = nodes[nd.left];
	const mcpt_bvh_node& R = nodes[nd.right];
	*reinterpret_cast<float4*>(&nd.bbmin) = min4(*reinterpret_cast<const float4*>(&L.bbmin), *reinterpret_cast<const float4*>(&R.bbmin));
	*reinterpret_cast<float4*>(&nd.bbmax) = max4(*reinterpret_cast<const float4*>(&L.bbmax), *reinterpret_cast<const float4*>(&R.bbmax));
}

template <class T>
struct Scoped {
	T* p = nullptr;
	cudaError_t alloc(size_t n) { return cudaMalloc(reinterpret_cast<void**>(&p), std::max<size_t>(1, n) * sizeof(T)); }
	~Scoped() { if (p) cudaFree(p); }
};

inline int blocks(long n, int bs) { return int((n + bs - 1) / bs); }

}  // namespace
}  // namespace mcpt

using namespace mcpt;

#define CUB_(x)                                                                                         \
	do {                                                                                                \
		cudaError_t e_ = (x);                                                                           \
		if (e_ != cudaSuccess) {                                                                        \
			set_last_error(std::string(#x) + ": " + cudaGetErrorString(e_));                            \
			return MCPT_ERR_CUDA;                                                                       \
		}                                                                                               \
	} while (0)

extern "C" int mcpt_build_lbvh(int device, const mcpt_triangle* tris, int ntris, mcpt_bvh_node* nodes_out, float* build_ms) {
	if (!tris || !nodes_out || ntris <= 0) {
		set_last_error("mcpt_build_lbvh: expected triangles, ntris > 0 and room for 2*ntris-1 nodes");
		return MCPT_ERR_ARG;
	}
	CUB_(cudaSetDevice(device));
	const int n = ntris;
	const size_t nn = size_t(2) * n - 1;
	cudaStream_t st;
	CUB_(cudaStreamCreate(&st));
	struct StreamGuard { cudaStream_t s; ~StreamGuard() { cudaStreamDestroy(s); } } guard{st};
	Scoped<mcpt_triangle> d_tris;
	Scoped<float4> lo, hi, cen;
	Scoped<int> gbox;
	Scoped<uint32_t> codes_a, codes_b, counts;
	Scoped<int32_t> ids_a, ids_b;
	Scoped<mcpt_bvh_node> nodes;
	Scoped<Range> ranges;                 // all levels back to back: n-1 internal nodes in total
	Scoped<unsigned char> tmp;
	CUB_(d_tris.alloc(n)); CUB_(lo.alloc(n)); CUB_(hi.alloc(n)); CUB_(cen.alloc(n)); CUB_(gbox.alloc(6));
	CUB_(codes_a.alloc(n)); CUB_(codes_b.alloc(n)); CUB_(ids_a.alloc(n)); CUB_(ids_b.alloc(n));
	CUB_(nodes.alloc(nn)); CUB_(ranges.alloc(size_t(n))); CUB_(counts.alloc(130));
	size_t tmp_bytes = 0;
	CUB_(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, codes_a.p, codes_b.p, ids_a.p, ids_b.p, n, 0, 30, st));
	CUB_(tmp.alloc(tmp_bytes));
	CUB_(cudaMemcpyAsync(d_tris.p, tris, sizeof(mcpt_triangle) * n, cudaMemcpyHostToDevice, st));
	cudaEvent_t e0, e1;
	CUB_(cudaEventCreate(&e0)); CUB_(cudaEventCreate(&e1));
	CUB_(cudaEventRecord(e0, st));
	const int init_box[6] = {0x7fffffff, 0x7fffffff, 0x7fffffff, int(0x80000000), int(0x80000000), int(0x80000000)};
	CUB_(cudaMemcpyAsync(gbox.p, init_box, sizeof(init_box), cudaMemcpyHostToDevice, st));
	CUB_(cudaMemsetAsync(nodes.p, 0, sizeof(mcpt_bvh_node) * nn, st));
	CUB_(cudaMemsetAsync(counts.p, 0, sizeof(uint32_t) * 130, st));
	k_tri_boxes<<<blocks(n, 256), 256, 0, st>>>(d_tris.p, n, lo.p, hi.p, cen.p, gbox.p);
	k_morton<<<blocks(n, 256), 256, 0, st>>>(cen.p, n, gbox.p, codes_a.p, ids_a.p);
	CUB_(cub::DeviceRadixSort::SortPairs(tmp.p, tmp_bytes, codes_a.p, codes_b.p, ids_a.p, ids_b.p, n, 0, 30, st));
	k_leaves<<<blocks(n, 256), 256, 0, st>>>(ids_b.p, lo.p, hi.p, n, nodes.p);
	// ---- hierarchy, one launch per level; the level sizes come back to the host (4 bytes per level)
	std::vector<uint32_t> level_off{0u}, level_cnt;
	if (n > 1) {
		const Range root{0u, uint32_t(n - 1), 0u};
		CUB_(cudaMemcpyAsync(ranges.p, &root, sizeof(Range), cudaMemcpyHostToDevice, st));
		uint32_t cnt = 1, off = 0;
		int level = 0;
		while (cnt > 0) {
			if (level >= 128) { set_last_error("mcpt_build_lbvh: hierarchy deeper than 128 levels"); return MCPT_ERR_STATE; }
			level_cnt.push_back(cnt);
			k_split_level<<<blocks(cnt, 128), 128, 0, st>>>(codes_b.p, n, ranges.p + off, cnt, ranges.p + off + cnt, counts.p + level, nodes.p);
			uint32_t next = 0;
			CUB_(cudaMemcpyAsync(&next, counts.p + level, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
			CUB_(cudaStreamSynchronize(st));
			off += cnt;
			level_off.push_back(off);
			cnt = next;
			++level;
		}
		for (int l = int(level_cnt.size()) - 1; l >= 0; --l)
			k_refit_level<<<blocks(level_cnt[size_t(l)], 128), 128, 0, st>>>(ranges.p + level_off[size_t(l)], level_cnt[size_t(l)], nodes.p);
	}
	const int32_t minus1 = -1;
	CUB_(cudaMemcpyAsync(&nodes.p[0].parent, &minus1, sizeof(int32_t), cudaMemcpyHostToDevice, st));
	CUB_(cudaEventRecord(e1, st));
	CUB_(cudaMemcpyAsync(nodes_out, nodes.p, sizeof(mcpt_bvh_node) * nn, cudaMemcpyDeviceToHost, st));
	CUB_(cudaStreamSynchronize(st));
	CUB_(cudaGetLastError());
	if (build_ms) CUB_(cudaEventElapsedTime(build_ms, e0, e1));
	cudaEventDestroy(e0);
	cudaEventDestroy(e1);
	return int(nn);
}
