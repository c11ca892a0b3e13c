// Binned-SAH tree of the fast walk, built on the device (SURVEY §8 f-2 "next": what mcpt_upload_scene derives for a NEW
// scene; the reference derives nothing - its kernels walk the uploaded LBVH, scenebuild.cpp:97-100).
// Contract: the node array is memcmp-equal to the host builder's (accel_build.cpp::build_sah_bvh2), which stays the
// definition (mcpt_derive_scene needs no GPU).  What makes that possible: a node's split depends only on the SET of
// triangles below it - centroid bounds and per-bin boxes are min / max reductions (exact in any order), bin counts are
// integers, the sweep over 16 bins x 3 axes is evaluated with the host's float operations in the host's order - and a
// subtree with n leaves owns the n - 1 consecutive slots after its root (preorder), so slots need no allocation.
//   * tasks of more than SMALL triangles: one level at a time, several blocks per task (chunks of CHUNK triangles):
//     centroid bounds -> bins (shared-memory bins per block, flushed with atomics) -> split + node record -> partition
//     (warp-aggregated atomics; the order inside a side is irrelevant);
//   * tasks of at most SMALL triangles: one block builds the whole subtree out of shared memory (explicit stack of
//     subtasks, block-wide reductions and a scan-based partition).
// The host's median fallback (no split found, two triangles, depth >= MAX_SAH_DEPTH: the n/2 smallest by (centroid on
// the widest axis, index)) is a sort in the small phase; for a large task - degenerate input - it runs on the host.
#include <cuda_runtime.h>

#include <algorithm>
#include <cfloat>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "mcpt.h"
#include "accel_build.hpp"

namespace mcpt {
void set_last_error(const std::string& s);     // mcpt_cuda.cu
namespace {

constexpr int BINS = 16;                 // accel_build.cpp
constexpr int MAX_SAH_DEPTH = 40;        // accel_build.cpp
constexpr int SMALL = 1024;              // triangles a single block takes over
constexpr int CHUNK = 2048;              // triangles per block in the level-wise phase
constexpr int BT = 256;                  // threads per block

struct Task { int lo, hi, base, depth; };
// per large task: centroid bounds (ordered ints), bins, the chosen split
struct TaskState {
	int cb[6];                           // centroid min xyz, max xyz as ordered ints
	int cnt[3][BINS];
	int box[3][BINS][6];                 // lo xyz (min), hi xyz (max) as ordered ints
	int axis, bin, nl;                   // axis < 0: no split (median fallback on the host)
	int lcount, rcount;                  // partition cursors
};

__device__ __forceinline__ int f2ord(float f) { int b = __float_as_int(f); return b >= 0 ? b : b ^ 0x7fffffff; }
__device__ __forceinline__ float ord2f(int o) { return __int_as_float(o >= 0 ? o : o ^ 0x7fffffff); }
constexpr int ORD_MAX = 0x7f7fffff;                   // f2ord(FLT_MAX)
constexpr int ORD_MIN = int(0x80800000u);             // f2ord(-FLT_MAX) = 0xff7fffff ^ 0x7fffffff

// the leaf's box and centroid as the host builder sees them (accel_build.cpp: tri_box from the DLeaf record with
// zeros made positive, centroid = 0.5f * (lo + hi))
struct TriBox { float lo[3], hi[3], c[3]; };
__device__ __forceinline__ TriBox load_tri(const DLeaf* __restrict__ leaves, int t) {
	const float4* p = reinterpret_cast<const float4*>(leaves + t);
	const float4 a = __ldg(p), b = __ldg(p + 1);
	TriBox r;
	r.lo[0] = a.x + 0.0f; r.lo[1] = a.y + 0.0f; r.lo[2] = a.z + 0.0f;
	r.hi[0] = b.x + 0.0f; r.hi[1] = b.y + 0.0f; r.hi[2] = b.z + 0.0f;
#pragma unroll
	for (int k = 0; k < 3; ++k) r.c[k] = 0.5f * (r.lo[k] + r.hi[k]);
	return r;
}
__device__ __forceinline__ int bin_of(float c, float lo, float scale) {
	return min(BINS - 1, max(0, int((c - lo) * scale)));
}
__device__ __forceinline__ float half_area(const float* lo, const float* hi) {
	const float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
	return dx * dy + dy * dz + dz * dx;
}

// The host's sweep (accel_build.cpp:76-103) over the bins of the three axes.  cnt / box are the bins (box as floats:
// lo xyz, hi xyz; empty bins hold +FLT_MAX / -FLT_MAX like Box::reset).  Returns the axis (or -1) and the bin.
struct Split { int axis, bin; };
template <class Cnt, class BoxF>
__device__ Split sweep_bins(const float* cb_lo, const float* cb_hi, Cnt cnt, BoxF box) {
	float best_cost = FLT_MAX;
	Split s{-1, -1};
	for (int axis = 0; axis < 3; ++axis) {
		const float ext = cb_hi[axis] - cb_lo[axis];
		if (!(ext > 0.0f)) continue;
		float right_area[BINS];
		int right_cnt[BINS];
		float alo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, ahi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
		int c = 0;
		for (int b = BINS - 1; b > 0; --b) {
			for (int k = 0; k < 3; ++k) { alo[k] = fminf(alo[k], box(axis, b, k)); ahi[k] = fmaxf(ahi[k], box(axis, b, 3 + k)); }
			c += cnt(axis, b);
			right_area[b] = c ? half_area(alo, ahi) : 0.0f;
			right_cnt[b] = c;
		}
		for (int k = 0; k < 3; ++k) { alo[k] = FLT_MAX; ahi[k] = -FLT_MAX; }
		c = 0;
		for (int b = 0; b < BINS - 1; ++b) {
			for (int k = 0; k < 3; ++k) { alo[k] = fminf(alo[k], box(axis, b, k)); ahi[k] = fmaxf(ahi[k], box(axis, b, 3 + k)); }
			c += cnt(axis, b);
			if (c == 0 || right_cnt[b + 1] == 0) continue;
			const float cost = half_area(alo, ahi) * float(c) + right_area[b + 1] * float(right_cnt[b + 1]);
			if (cost < best_cost) { best_cost = cost; s.axis = axis; s.bin = b; }
		}
	}
	return s;
}

// ------------------------------------------------------------------------------------------------ level-wise phase
// chunk c covers idx[task.lo + off, ...) of task chunk_task[c]
__global__ void __launch_bounds__(BT) k_bounds(const DLeaf* __restrict__ leaves, const int* __restrict__ idx, const Task* __restrict__ tasks,
                                                const int* __restrict__ chunk_task, const int* __restrict__ chunk_off, TaskState* st) {
	const int t = chunk_task[blockIdx.x];
	const Task tk = tasks[t];
	const int begin = tk.lo + chunk_off[blockIdx.x], end = min(tk.hi, begin + CHUNK);
	float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
	for (int i = begin + threadIdx.x; i < end; i += BT) {
		const TriBox b = load_tri(leaves, idx[i]);
		for (int k = 0; k < 3; ++k) { lo[k] = fminf(lo[k], b.c[k]); hi[k] = fmaxf(hi[k], b.c[k]); }
	}
	for (int o = 16; o > 0; o >>= 1)
		for (int k = 0; k < 3; ++k) {
			lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
			hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
		}
	if ((threadIdx.x & 31) == 0)
		for (int k = 0; k < 3; ++k) { atomicMin(&st[t].cb[k], f2ord(lo[k])); atomicMax(&st[t].cb[3 + k], f2ord(hi[k])); }
}

__global__ void __launch_bounds__(BT) k_bins(const DLeaf* __restrict__ leaves, const int* __restrict__ idx, const Task* __restrict__ tasks,
                                              const int* __restrict__ chunk_task, const int* __restrict__ chunk_off, TaskState* st) {
	__shared__ int s_cnt[3][BINS];
	__shared__ int s_box[3][BINS][6];
	const int t = chunk_task[blockIdx.x];
	const Task tk = tasks[t];
	for (int i = threadIdx.x; i < 3 * BINS; i += BT) {
		(&s_cnt[0][0])[i] = 0;
		for (int k = 0; k < 3; ++k) { (&s_box[0][0][0])[i * 6 + k] = ORD_MAX; (&s_box[0][0][0])[i * 6 + 3 + k] = ORD_MIN; }
	}
	__syncthreads();
	float cl[3], scale[3];
	bool use[3];
	for (int a = 0; a < 3; ++a) {
		cl[a] = ord2f(st[t].cb[a]);
		const float ext = ord2f(st[t].cb[3 + a]) - cl[a];
		use[a] = ext > 0.0f;
		scale[a] = float(BINS) / ext;
	}
	const int begin = tk.lo + chunk_off[blockIdx.x], end = min(tk.hi, begin + CHUNK);
	for (int i = begin + threadIdx.x; i < end; i += BT) {
		const TriBox b = load_tri(leaves, idx[i]);
		for (int a = 0; a < 3; ++a) {
			if (!use[a]) continue;
			const int bn = bin_of(b.c[a], cl[a], scale[a]);
			atomicAdd(&s_cnt[a][bn], 1);
			for (int k = 0; k < 3; ++k) { atomicMin(&s_box[a][bn][k], f2ord(b.lo[k])); atomicMax(&s_box[a][bn][3 + k], f2ord(b.hi[k])); }
		}
	}
	__syncthreads();
	for (int i = threadIdx.x; i < 3 * BINS; i += BT) {
		const int a = i / BINS, bn = i % BINS;
		if (s_cnt[a][bn] == 0) continue;
		atomicAdd(&st[t].cnt[a][bn], s_cnt[a][bn]);
		for (int k = 0; k < 3; ++k) { atomicMin(&st[t].box[a][bn][k], s_box[a][bn][k]); atomicMax(&st[t].box[a][bn][3 + k], s_box[a][bn][3 + k]); }
	}
}

struct Lists {
	int n_large, n_small, n_fallback;    // entries appended by k_split
	int max_depth;
};

__device__ __forceinline__ void write_node(DNode* nodes, int base, const float* llo, const float* lhi, const float* rlo, const float* rhi,
                                           int left, int right) {
	DNode d;
	d.lminx = llo[0]; d.lminy = llo[1]; d.lminz = llo[2]; d.lmaxx = lhi[0]; d.lmaxy = lhi[1]; d.lmaxz = lhi[2];
	d.rminx = rlo[0]; d.rminy = rlo[1]; d.rminz = rlo[2]; d.rmaxx = rhi[0]; d.rmaxy = rhi[1]; d.rmaxz = rhi[2];
	d.left = left; d.right = right; d.pad0 = 0; d.pad1 = 0;
	nodes[base] = d;
}

// one thread per task of the level: the sweep, the node record, the children
__global__ void k_split(const Task* __restrict__ tasks, int ntasks, TaskState* st, DNode* nodes, Task* next_large, Task* small_list,
                        int* fallback_list, Lists* lists) {
	const int t = blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= ntasks) return;
	const Task tk = tasks[t];
	TaskState& s = st[t];
	float cl[3], ch[3];
	for (int k = 0; k < 3; ++k) { cl[k] = ord2f(s.cb[k]); ch[k] = ord2f(s.cb[3 + k]); }
	Split sp{-1, -1};
	if (tk.depth < MAX_SAH_DEPTH)
		sp = sweep_bins(cl, ch, [&](int a, int b) { return s.cnt[a][b]; }, [&](int a, int b, int k) { return ord2f(s.box[a][b][k]); });
	s.axis = sp.axis; s.bin = sp.bin; s.lcount = 0; s.rcount = 0;
	atomicMax(&lists->max_depth, tk.depth + 1);
	if (sp.axis < 0) {
		s.nl = 0;
		fallback_list[atomicAdd(&lists->n_fallback, 1)] = t;
		return;
	}
	int nl = 0;
	float llo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, lhi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
	float rlo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, rhi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
	for (int b = 0; b < BINS; ++b) {
		float* lo = b <= sp.bin ? llo : rlo;
		float* hi = b <= sp.bin ? lhi : rhi;
		if (b <= sp.bin) nl += s.cnt[sp.axis][b];
		for (int k = 0; k < 3; ++k) { lo[k] = fminf(lo[k], ord2f(s.box[sp.axis][b][k])); hi[k] = fmaxf(hi[k], ord2f(s.box[sp.axis][b][3 + k])); }
	}
	s.nl = nl;
	const int n = tk.hi - tk.lo, nr = n - nl;
	// single-triangle sides are linked by k_links once the partition has put their triangle in place
	write_node(nodes, tk.base, llo, lhi, rlo, rhi, tk.base + 1, tk.base + nl);
	const Task L{tk.lo, tk.lo + nl, tk.base + 1, tk.depth + 1}, R{tk.lo + nl, tk.hi, tk.base + nl, tk.depth + 1};
	if (nl > SMALL) next_large[atomicAdd(&lists->n_large, 1)] = L;
	else if (nl > 1) small_list[atomicAdd(&lists->n_small, 1)] = L;
	if (nr > SMALL) next_large[atomicAdd(&lists->n_large, 1)] = R;
	else if (nr > 1) small_list[atomicAdd(&lists->n_small, 1)] = R;
}

__global__ void __launch_bounds__(BT) k_partition(const DLeaf* __restrict__ leaves, const int* __restrict__ idx, int* __restrict__ out,
                                                   const Task* __restrict__ tasks, const int* __restrict__ chunk_task,
                                                   const int* __restrict__ chunk_off, TaskState* st) {
	const int t = chunk_task[blockIdx.x];
	const Task tk = tasks[t];
	TaskState& s = st[t];
	const int axis = s.axis, bin = s.bin, nl = s.nl;
	const int begin = tk.lo + chunk_off[blockIdx.x], end = min(tk.hi, begin + CHUNK);
	if (axis < 0) {                                  // host fallback: keep the range as it is
		for (int i = begin + threadIdx.x; i < end; i += BT) out[i] = idx[i];
		return;
	}
	const float cl = ord2f(s.cb[axis]);
	const float scale = float(BINS) / (ord2f(s.cb[3 + axis]) - cl);
	const unsigned lane = threadIdx.x & 31u;
	for (int i0 = begin; i0 < end; i0 += BT) {       // block-uniform trip count: every warp reaches the ballots
		const int i = i0 + int(threadIdx.x);
		const bool valid = i < end;
		int tri = 0;
		bool left = false;
		if (valid) {
			tri = idx[i];
			const TriBox b = load_tri(leaves, tri);
			left = bin_of(b.c[axis], cl, scale) <= bin;
		}
		const unsigned ml = __ballot_sync(0xffffffffu, valid && left), mr = __ballot_sync(0xffffffffu, valid && !left);
		int bl = 0, br = 0;
		if (lane == 0) {
			if (ml) bl = atomicAdd(&s.lcount, __popc(ml));
			if (mr) br = atomicAdd(&s.rcount, __popc(mr));
		}
		bl = __shfl_sync(0xffffffffu, bl, 0);
		br = __shfl_sync(0xffffffffu, br, 0);
		if (valid) {
			const unsigned below = (1u << lane) - 1u;
			out[left ? tk.lo + bl + __popc(ml & below) : tk.lo + nl + br + __popc(mr & below)] = tri;
		}
	}
}

__global__ void __launch_bounds__(BT) k_copy_back(int* __restrict__ idx, const int* __restrict__ tmp, const Task* __restrict__ tasks,
                                                   const int* __restrict__ chunk_task, const int* __restrict__ chunk_off) {
	const Task tk = tasks[chunk_task[blockIdx.x]];
	const int begin = tk.lo + chunk_off[blockIdx.x], end = min(tk.hi, begin + CHUNK);
	for (int i = begin + threadIdx.x; i < end; i += BT) idx[i] = tmp[i];
}

__global__ void k_links(const Task* __restrict__ tasks, int ntasks, const TaskState* __restrict__ st, const int* __restrict__ idx, DNode* nodes) {
	const int t = blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= ntasks || st[t].axis < 0) return;
	const Task tk = tasks[t];
	const int nl = st[t].nl;
	if (nl == 1) nodes[tk.base].left = ~idx[tk.lo];
	if (tk.hi - tk.lo - nl == 1) nodes[tk.base].right = ~idx[tk.lo + nl];
}

__global__ void k_init_state(TaskState* st, int ntasks) {
	const int t = blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= ntasks) return;
	TaskState& s = st[t];
	for (int k = 0; k < 3; ++k) { s.cb[k] = ORD_MAX; s.cb[3 + k] = ORD_MIN; }
	for (int a = 0; a < 3; ++a)
		for (int b = 0; b < BINS; ++b) {
			s.cnt[a][b] = 0;
			for (int k = 0; k < 3; ++k) { s.box[a][b][k] = ORD_MAX; s.box[a][b][3 + k] = ORD_MIN; }
		}
	s.axis = -1; s.bin = -1; s.nl = 0; s.lcount = 0; s.rcount = 0;
}

__global__ void k_iota_idx(int* idx, int n) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) idx[i] = i;
}

// ------------------------------------------------------------------------------------------------ small-task phase
// One block per task of at most SMALL triangles: the whole subtree from shared memory.
constexpr int SMALL_STACK = 96;          // pending subtasks (one per level at most; depth <= MAX_SAH_DEPTH + log2(SMALL) + slack)

__global__ void __launch_bounds__(BT) k_small(const DLeaf* __restrict__ leaves, const int* __restrict__ idx, const Task* __restrict__ tasks,
                                               DNode* nodes, Lists* lists, int* status) {
	__shared__ int s_idx[SMALL], s_tmp[SMALL];
	__shared__ unsigned long long s_key[SMALL];
	__shared__ int s_cnt[3][BINS];
	__shared__ int s_box[3][BINS][6];
	__shared__ int s_cb[6];
	__shared__ Task s_stack[SMALL_STACK];
	__shared__ Task s_cur;
	__shared__ int s_sp, s_axis, s_bin, s_nl, s_l, s_r;
	const Task root = tasks[blockIdx.x];
	const int n0 = root.hi - root.lo;
	for (int i = threadIdx.x; i < n0; i += BT) s_idx[i] = idx[root.lo + i];
	int max_depth = 0;
	if (threadIdx.x == 0) { s_stack[0] = Task{0, n0, root.base, root.depth}; s_sp = 1; }
	for (;;) {
		__syncthreads();
		if (s_sp == 0) break;
		__syncthreads();
		if (threadIdx.x == 0) s_cur = s_stack[--s_sp];
		if (threadIdx.x < 6) s_cb[threadIdx.x] = threadIdx.x < 3 ? ORD_MAX : ORD_MIN;
		for (int i = threadIdx.x; i < 3 * BINS; i += BT) {
			(&s_cnt[0][0])[i] = 0;
			for (int k = 0; k < 3; ++k) { (&s_box[0][0][0])[i * 6 + k] = ORD_MAX; (&s_box[0][0][0])[i * 6 + 3 + k] = ORD_MIN; }
		}
		__syncthreads();
		const Task tk = s_cur;
		const int n = tk.hi - tk.lo;
		max_depth = max(max_depth, tk.depth + 1);
		// ---- centroid bounds
		for (int i = tk.lo + threadIdx.x; i < tk.hi; i += BT) {
			const TriBox b = load_tri(leaves, s_idx[i]);
			for (int k = 0; k < 3; ++k) { atomicMin(&s_cb[k], f2ord(b.c[k])); atomicMax(&s_cb[3 + k], f2ord(b.c[k])); }
		}
		__syncthreads();
		float cl[3], ch[3];
		for (int k = 0; k < 3; ++k) { cl[k] = ord2f(s_cb[k]); ch[k] = ord2f(s_cb[3 + k]); }
		const bool try_sah = tk.depth < MAX_SAH_DEPTH && n > 2;
		if (try_sah) {
			float scale[3];
			bool use[3];
			for (int a = 0; a < 3; ++a) { const float ext = ch[a] - cl[a]; use[a] = ext > 0.0f; scale[a] = float(BINS) / ext; }
			for (int i = tk.lo + threadIdx.x; i < tk.hi; i += BT) {
				const TriBox b = load_tri(leaves, s_idx[i]);
				for (int a = 0; a < 3; ++a) {
					if (!use[a]) continue;
					const int bn = bin_of(b.c[a], cl[a], scale[a]);
					atomicAdd(&s_cnt[a][bn], 1);
					for (int k = 0; k < 3; ++k) { atomicMin(&s_box[a][bn][k], f2ord(b.lo[k])); atomicMax(&s_box[a][bn][3 + k], f2ord(b.hi[k])); }
				}
			}
			__syncthreads();
			if (threadIdx.x == 0) {
				const Split sp = sweep_bins(cl, ch, [&](int a, int b) { return s_cnt[a][b]; }, [&](int a, int b, int k) { return ord2f(s_box[a][b][k]); });
				s_axis = sp.axis; s_bin = sp.bin;
				int nl = 0;
				if (sp.axis >= 0) for (int b = 0; b <= sp.bin; ++b) nl += s_cnt[sp.axis][b];
				s_nl = nl;
			}
			__syncthreads();
		}
		const bool median = !try_sah || s_axis < 0;
		float llo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, lhi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
		float rlo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, rhi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
		int nl;
		if (!median) {
			// ---- partition by bin (order inside a side is irrelevant), boxes from the bins
			const int axis = s_axis, bin = s_bin;
			nl = s_nl;
			if (threadIdx.x == 0) { s_l = 0; s_r = 0; }
			__syncthreads();
			const float scale = float(BINS) / (ch[axis] - cl[axis]);
			for (int i = tk.lo + threadIdx.x; i < tk.hi; i += BT) {
				const int tri = s_idx[i];
				const TriBox b = load_tri(leaves, tri);
				const bool left = bin_of(b.c[axis], cl[axis], scale) <= bin;
				s_tmp[left ? tk.lo + atomicAdd(&s_l, 1) : tk.lo + nl + atomicAdd(&s_r, 1)] = tri;
			}
			for (int b = 0; b < BINS; ++b) {
				float* lo = b <= bin ? llo : rlo;
				float* hi = b <= bin ? lhi : rhi;
				for (int k = 0; k < 3; ++k) { lo[k] = fminf(lo[k], ord2f(s_box[axis][b][k])); hi[k] = fmaxf(hi[k], ord2f(s_box[axis][b][3 + k])); }
			}
			__syncthreads();
			for (int i = tk.lo + threadIdx.x; i < tk.hi; i += BT) s_idx[i] = s_tmp[i];
		} else {
			// ---- the host's median split (accel_build.cpp:118-129): widest centroid axis, the n / 2 smallest by (centroid, index)
			int axis = 0;
			const float e0 = ch[0] - cl[0], e1 = ch[1] - cl[1], e2 = ch[2] - cl[2];
			if (e1 > e0 && e1 >= e2) axis = 1;
			else if (e2 > e0 && e2 > e1) axis = 2;
			nl = n / 2;
			int p2 = 1;
			while (p2 < n) p2 <<= 1;
			for (int i = threadIdx.x; i < p2; i += BT) {
				unsigned long long key = ~0ull;
				if (i < n) {
					const int tri = s_idx[tk.lo + i];
					const TriBox b = load_tri(leaves, tri);
					key = (static_cast<unsigned long long>(unsigned(f2ord(b.c[axis])) ^ 0x80000000u) << 32) | unsigned(tri);
				}
				s_key[i] = key;
			}
			__syncthreads();
			for (int k = 2; k <= p2; k <<= 1)
				for (int j = k >> 1; j > 0; j >>= 1) {
					for (int i = threadIdx.x; i < p2; i += BT) {
						const int o = i ^ j;
						if (o > i) {
							const unsigned long long a = s_key[i], b = s_key[o];
							if (((i & k) == 0) == (a > b)) { s_key[i] = b; s_key[o] = a; }
						}
					}
					__syncthreads();
				}
			for (int i = threadIdx.x; i < n; i += BT) s_idx[tk.lo + i] = int(unsigned(s_key[i] & 0xffffffffull));
			// boxes of the two sides: bins 0 / 1 of axis 0 serve as accumulators (they were reset above; the SAH attempt, if
			// there was one, found no axis with extent and left them empty)
			if (threadIdx.x < 2 * 6) (&s_box[0][0][0])[threadIdx.x] = (threadIdx.x % 6) < 3 ? ORD_MAX : ORD_MIN;
			__syncthreads();
			for (int i = threadIdx.x; i < n; i += BT) {
				const TriBox b = load_tri(leaves, s_idx[tk.lo + i]);
				const int side = i < nl ? 0 : 1;
				for (int k = 0; k < 3; ++k) { atomicMin(&s_box[0][side][k], f2ord(b.lo[k])); atomicMax(&s_box[0][side][3 + k], f2ord(b.hi[k])); }
			}
			__syncthreads();
			for (int k = 0; k < 3; ++k) {
				llo[k] = ord2f(s_box[0][0][k]); lhi[k] = ord2f(s_box[0][0][3 + k]);
				rlo[k] = ord2f(s_box[0][1][k]); rhi[k] = ord2f(s_box[0][1][3 + k]);
			}
		}
		__syncthreads();
		if (threadIdx.x == 0) {
			const int nr = n - nl;
			if (nl < 1 || nr < 1) atomicExch(status, 1);            // cannot happen: both rules put something on each side
			const int left = nl == 1 ? ~s_idx[tk.lo] : tk.base + 1;
			const int right = nr == 1 ? ~s_idx[tk.lo + nl] : tk.base + nl;
			write_node(nodes, tk.base, llo, lhi, rlo, rhi, left, right);
			if (s_sp + 2 > SMALL_STACK) atomicExch(status, 2);
			else {
				if (nr > 1) s_stack[s_sp++] = Task{tk.lo + nl, tk.hi, tk.base + nl, tk.depth + 1};
				if (nl > 1) s_stack[s_sp++] = Task{tk.lo, tk.lo + nl, tk.base + 1, tk.depth + 1};
			}
		}
	}
	if (threadIdx.x == 0) atomicMax(&lists->max_depth, max_depth);
}

#define TRY_(x)                                                                                        \
	do {                                                                                               \
		cudaError_t e_ = (x);                                                                          \
		if (e_ != cudaSuccess) {                                                                       \
			set_last_error(std::string("fast-tree build: ") + #x + ": " + cudaGetErrorString(e_));     \
			return MCPT_ERR_CUDA;                                                                      \
		}                                                                                              \
	} while (0)

template <class T>
struct Scoped {
	T* p = nullptr;
	cudaError_t alloc(size_t n) { return cudaMalloc(reinterpret_cast<void**>(&p), std::max<size_t>(1, n) * sizeof(T)); }
	~Scoped() { if (p) cudaFree(p); }
};

inline int blocks(long n, int bs) { return int((n + bs - 1) / bs); }

// the host's median split for a LARGE task (degenerate input: more than SMALL triangles with one centroid, or a chain
// of lopsided splits deeper than MAX_SAH_DEPTH): sorted on the host, written back, node record included
int host_median(const DLeaf* h_leaves, int* d_idx, DNode* d_nodes, const Task& tk, std::vector<Task>& next_large, std::vector<Task>& new_small) {
	const int n = tk.hi - tk.lo;
	std::vector<int> ids(static_cast<size_t>(n), 0);
	TRY_(cudaMemcpy(ids.data(), d_idx + tk.lo, sizeof(int) * size_t(n), cudaMemcpyDeviceToHost));
	auto lo_of = [&](int t, int k) { const DLeaf& l = h_leaves[t]; return (k == 0 ? l.bminx : k == 1 ? l.bminy : l.bminz) + 0.0f; };
	auto hi_of = [&](int t, int k) { const DLeaf& l = h_leaves[t]; return (k == 0 ? l.bmaxx : k == 1 ? l.bmaxy : l.bmaxz) + 0.0f; };
	auto cen = [&](int t, int k) { return 0.5f * (lo_of(t, k) + hi_of(t, k)); };
	float cl[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, ch[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
	for (int t : ids) for (int k = 0; k < 3; ++k) { cl[k] = std::min(cl[k], cen(t, k)); ch[k] = std::max(ch[k], cen(t, k)); }
	int axis = 0;
	const float e0 = ch[0] - cl[0], e1 = ch[1] - cl[1], e2 = ch[2] - cl[2];
	if (e1 > e0 && e1 >= e2) axis = 1;
	else if (e2 > e0 && e2 > e1) axis = 2;
	const int nl = n / 2;
	std::nth_element(ids.begin(), ids.begin() + nl, ids.end(), [&](int a, int b) {
		const float ca = cen(a, axis), cb = cen(b, axis);
		return ca < cb || (ca == cb && a < b);
	});
	TRY_(cudaMemcpy(d_idx + tk.lo, ids.data(), sizeof(int) * size_t(n), cudaMemcpyHostToDevice));
	DNode d{};
	float b[2][6] = {{FLT_MAX, FLT_MAX, FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX}, {FLT_MAX, FLT_MAX, FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX}};
	for (int i = 0; i < n; ++i)
		for (int k = 0; k < 3; ++k) {
			float* x = b[i < nl ? 0 : 1];
			x[k] = std::min(x[k], lo_of(ids[size_t(i)], k));
			x[3 + k] = std::max(x[3 + k], hi_of(ids[size_t(i)], k));
		}
	d.lminx = b[0][0]; d.lminy = b[0][1]; d.lminz = b[0][2]; d.lmaxx = b[0][3]; d.lmaxy = b[0][4]; d.lmaxz = b[0][5];
	d.rminx = b[1][0]; d.rminy = b[1][1]; d.rminz = b[1][2]; d.rmaxx = b[1][3]; d.rmaxy = b[1][4]; d.rmaxz = b[1][5];
	const int nr = n - nl;
	d.left = nl == 1 ? ~ids[0] : tk.base + 1;
	d.right = nr == 1 ? ~ids[size_t(nl)] : tk.base + nl;
	TRY_(cudaMemcpy(d_nodes + tk.base, &d, sizeof(DNode), cudaMemcpyHostToDevice));
	const Task L{tk.lo, tk.lo + nl, tk.base + 1, tk.depth + 1}, R{tk.lo + nl, tk.hi, tk.base + nl, tk.depth + 1};
	for (const Task& c : {L, R}) {
		const int m = c.hi - c.lo;
		if (m > SMALL) next_large.push_back(c);
		else if (m > 1) new_small.push_back(c);
	}
	return MCPT_OK;
}

}  // namespace

// d_leaves: the DLeaf records on `device` (their boxes are all the builder reads); h_leaves: the same records on the
// host (median fallback of large tasks only); out: ntris - 1 nodes, memcmp-equal to build_sah_bvh2's.
int build_sah_bvh2_device(int device, const DLeaf* d_leaves, const DLeaf* h_leaves, int ntris, DNode* out, float* build_ms, int* depth) {
	if (ntris < 2) { set_last_error("fast-tree build: needs at least two triangles"); return MCPT_ERR_ARG; }
	TRY_(cudaSetDevice(device));
	cudaStream_t st = nullptr;       // the legacy default stream: the calls below are ordered and the copies blocking
	const size_t cap_large = size_t(ntris) / SMALL + 4, cap_small = size_t(ntris) / 2 + 4, cap_chunks = size_t(ntris) / CHUNK + cap_large + 4;
	Scoped<int> idx, tmp, chunk_task, chunk_off, fallback, status;
	Scoped<DNode> nodes;
	Scoped<Task> cur, next, small;
	Scoped<TaskState> state;
	Scoped<Lists> lists;
	TRY_(idx.alloc(size_t(ntris))); TRY_(tmp.alloc(size_t(ntris))); TRY_(nodes.alloc(size_t(ntris) - 1));
	TRY_(cur.alloc(cap_large)); TRY_(next.alloc(cap_large)); TRY_(small.alloc(cap_small)); TRY_(state.alloc(cap_large));
	TRY_(chunk_task.alloc(cap_chunks)); TRY_(chunk_off.alloc(cap_chunks)); TRY_(fallback.alloc(cap_large));
	TRY_(lists.alloc(1)); TRY_(status.alloc(1));
	cudaEvent_t e0, e1;
	TRY_(cudaEventCreate(&e0)); TRY_(cudaEventCreate(&e1));
	TRY_(cudaEventRecord(e0, st));
	TRY_(cudaMemsetAsync(lists.p, 0, sizeof(Lists), st));
	TRY_(cudaMemsetAsync(status.p, 0, sizeof(int), st));
	k_iota_idx<<<blocks(ntris, BT), BT, 0, st>>>(idx.p, ntris);
	std::vector<Task> level, new_small;
	int n_small = 0;
	if (ntris > SMALL) level.push_back(Task{0, ntris, 0, 0});
	else new_small.push_back(Task{0, ntris, 0, 0});
	std::vector<int> h_ct, h_co, h_fb;
	while (!level.empty()) {
		const int nt = int(level.size());
		if (size_t(nt) > cap_large) { set_last_error("fast-tree build: task list overflow"); return MCPT_ERR_STATE; }
		h_ct.clear(); h_co.clear();
		for (int t = 0; t < nt; ++t)
			for (int off = 0; off < level[size_t(t)].hi - level[size_t(t)].lo; off += CHUNK) { h_ct.push_back(t); h_co.push_back(off); }
		const int nc = int(h_ct.size());
		TRY_(cudaMemcpy(cur.p, level.data(), sizeof(Task) * size_t(nt), cudaMemcpyHostToDevice));
		TRY_(cudaMemcpy(chunk_task.p, h_ct.data(), sizeof(int) * size_t(nc), cudaMemcpyHostToDevice));
		TRY_(cudaMemcpy(chunk_off.p, h_co.data(), sizeof(int) * size_t(nc), cudaMemcpyHostToDevice));
		Lists h_lists{};
		TRY_(cudaMemcpy(&h_lists, lists.p, sizeof(Lists), cudaMemcpyDeviceToHost));
		h_lists.n_large = 0; h_lists.n_fallback = 0; h_lists.n_small = n_small;
		TRY_(cudaMemcpy(lists.p, &h_lists, sizeof(Lists), cudaMemcpyHostToDevice));
		k_init_state<<<blocks(nt, 128), 128, 0, st>>>(state.p, nt);
		k_bounds<<<nc, BT, 0, st>>>(d_leaves, idx.p, cur.p, chunk_task.p, chunk_off.p, state.p);
		k_bins<<<nc, BT, 0, st>>>(d_leaves, idx.p, cur.p, chunk_task.p, chunk_off.p, state.p);
		k_split<<<blocks(nt, 64), 64, 0, st>>>(cur.p, nt, state.p, nodes.p, next.p, small.p, fallback.p, lists.p);
		k_partition<<<nc, BT, 0, st>>>(d_leaves, idx.p, tmp.p, cur.p, chunk_task.p, chunk_off.p, state.p);
		k_copy_back<<<nc, BT, 0, st>>>(idx.p, tmp.p, cur.p, chunk_task.p, chunk_off.p);
		k_links<<<blocks(nt, 64), 64, 0, st>>>(cur.p, nt, state.p, idx.p, nodes.p);
		TRY_(cudaMemcpy(&h_lists, lists.p, sizeof(Lists), cudaMemcpyDeviceToHost));
		TRY_(cudaGetLastError());
		if (size_t(h_lists.n_large) > cap_large || size_t(h_lists.n_small) > cap_small) { set_last_error("fast-tree build: task list overflow"); return MCPT_ERR_STATE; }
		std::vector<Task> nextv(size_t(h_lists.n_large));
		if (h_lists.n_large) TRY_(cudaMemcpy(nextv.data(), next.p, sizeof(Task) * nextv.size(), cudaMemcpyDeviceToHost));
		n_small = h_lists.n_small;
		if (h_lists.n_fallback) {
			h_fb.resize(size_t(h_lists.n_fallback));
			TRY_(cudaMemcpy(h_fb.data(), fallback.p, sizeof(int) * h_fb.size(), cudaMemcpyDeviceToHost));
			for (int t : h_fb)
				if (int r = host_median(h_leaves, idx.p, nodes.p, level[size_t(t)], nextv, new_small)) return r;
		}
		level.swap(nextv);
	}
	if (!new_small.empty()) {        // tasks the host produced (median fallback, or a scene of at most SMALL triangles)
		if (size_t(n_small) + new_small.size() > cap_small) { set_last_error("fast-tree build: task list overflow"); return MCPT_ERR_STATE; }
		TRY_(cudaMemcpy(small.p + n_small, new_small.data(), sizeof(Task) * new_small.size(), cudaMemcpyHostToDevice));
		n_small += int(new_small.size());
	}
	if (n_small > 0) k_small<<<n_small, BT, 0, st>>>(d_leaves, idx.p, small.p, nodes.p, lists.p, status.p);
	TRY_(cudaEventRecord(e1, st));
	int h_status = 0;
	TRY_(cudaMemcpy(&h_status, status.p, sizeof(int), cudaMemcpyDeviceToHost));
	TRY_(cudaMemcpy(out, nodes.p, sizeof(DNode) * (size_t(ntris) - 1), cudaMemcpyDeviceToHost));
	TRY_(cudaGetLastError());
	if (build_ms) TRY_(cudaEventElapsedTime(build_ms, e0, e1));
	cudaEventDestroy(e0); cudaEventDestroy(e1);
	if (h_status != 0) { set_last_error("fast-tree build: internal error " + std::to_string(h_status)); return MCPT_ERR_STATE; }
	if (depth) {
		Lists h_lists{};
		TRY_(cudaMemcpy(&h_lists, lists.p, sizeof(Lists), cudaMemcpyDeviceToHost));
		*depth = h_lists.max_depth;
	}
	return MCPT_OK;
}

}  // namespace mcpt
