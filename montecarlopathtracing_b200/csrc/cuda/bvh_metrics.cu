// BVH quality metrics on the device: SAH, EPO and LCV of the reference's research harness
// (`testbvh` / `testall` modes, bvhtest.cpp:97-444, kernels/EPO.cl).
//   SAH  bvhtest.cpp:97-110   sum of Cinn*area over inner nodes + Ctri*area over leaves, over the root area.
//        Per-node terms on the device; the sum runs on the host in the reference's order (sequential, double), so
//        the value equals the reference's bit for bit - what EPO_GPU does for its own partial results.
//   EPO  kernels/EPO.cl:133-196 + bvhtest.cpp:283-321: for every leaf's triangle, the area it covers inside boxes of
//        nodes that are not its ancestors (clipped against the six box planes), weighted Cinn / Ctri; one thread per
//        leaf, the kernel's depth-first order (left child popped first) so the float sums are the kernel's; host
//        double sums as in EPO_GPU.  The "is this an ancestor" test is a lookup by depth instead of the kernel's
//        scan over the ancestor list (same predicate).  Point buffers hold 10 vertices (the kernel's 8 overflow
//        when six planes cut a triangle into a 9-gon).
//   LCV  bvhtest.cpp:324-444: standard deviation of the number of leaf BOXES a primary ray touches (no triangle
//        tests), rays from the reference's own generator (un-normalised directions, pixel centres).  Counts are
//        integers, so sum and sum of squares are exact in 64-bit atomics and the result equals the reference's.
// Arithmetic: -fmad=false, IEEE division and square root, std::min/std::max operand order of the host code.
#include <cuda_runtime.h>

#include <cfloat>
#include <cmath>
#include <cstdint>
#include <string>
#include <vector>

#include "mcpt.h"

namespace mcpt {
void set_last_error(const std::string& s);     // mcpt_cuda.cu
namespace {

constexpr float C_INNER = 1.2f;      // auxiliary.h:9-11
constexpr float C_TRI = 1.0f;
constexpr int STACK = 128;           // EPO.cl:150,164

__device__ __forceinline__ float hmin(float a, float b) { return (b < a) ? b : a; }     // std::min(a, b)
__device__ __forceinline__ float hmax(float a, float b) { return (a < b) ? b : a; }     // std::max(a, b)

__device__ __forceinline__ float area3(float4 lo, float4 hi) {                           // auxiliary.cpp:15-18
	const float x = hi.x - lo.x, y = hi.y - lo.y, z = hi.z - lo.z;
	return 2.0f * (x * y + x * z + y * z);
}
__device__ __forceinline__ float4 node_lo(const mcpt_bvh_node* n, int i) { return __ldg(reinterpret_cast<const float4*>(n + i)); }
__device__ __forceinline__ float4 node_hi(const mcpt_bvh_node* n, int i) { return __ldg(reinterpret_cast<const float4*>(n + i) + 1); }
__device__ __forceinline__ int4 node_link(const mcpt_bvh_node* n, int i) { return __ldg(reinterpret_cast<const int4*>(n + i) + 3); }

__global__ void k_sah_terms(const mcpt_bvh_node* __restrict__ nodes, int nn, float* terms) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= nn) return;
	terms[i] = (i < (nn >> 1) ? C_INNER : C_TRI) * area3(node_lo(nodes, i), node_hi(nodes, i));
}

// ---------------------------------------------------------------------------------------------------- EPO
struct V3 { float s[3]; };

__device__ __forceinline__ bool point_in_box(const float* p, float4 lo, float4 hi) {      // EPO.cl:29-40
	return p[0] >= lo.x && p[0] <= hi.x && p[1] >= lo.y && p[1] <= hi.y && p[2] >= lo.z && p[2] <= hi.z;
}
__device__ __forceinline__ float half_cross_len(const float* a, const float* b, const float* o) {
	const float e1x = a[0] - o[0], e1y = a[1] - o[1], e1z = a[2] - o[2];
	const float e2x = b[0] - o[0], e2y = b[1] - o[1], e2z = b[2] - o[2];
	const float cx = e1y * e2z - e1z * e2y, cy = e1z * e2x - e1x * e2z, cz = e1x * e2y - e1y * e2x;
	return sqrtf(cx * cx + cy * cy + cz * cz) * 0.5f;
}
__device__ void clip(V3* pts, int& size, int axis, float pos, int arg) {                  // roundTr, EPO.cl:43-89
	if (size == 0) return;
	V3 buf[10];
	const int n = size;
	unsigned inside = 0;
	for (int i = 0; i < n; ++i) {
		buf[i] = pts[i];
		const float c = buf[i].s[axis];
		if (arg > 0 ? (c >= pos) : (c <= pos)) inside |= 1u << i;
	}
	size = 0;
	for (int i = 0; i < n; ++i) {
		const int j = (i + 1 == n) ? 0 : i + 1;
		const bool in_i = (inside >> i) & 1u, in_j = (inside >> j) & 1u;
		if (!in_i && !in_j) continue;
		if (in_i && in_j) { pts[size++] = buf[i]; continue; }
		if (in_i) pts[size++] = buf[i];
		const float dx = buf[j].s[0] - buf[i].s[0], dy = buf[j].s[1] - buf[i].s[1], dz = buf[j].s[2] - buf[i].s[2];
		const float da = axis == 0 ? dx : (axis == 1 ? dy : dz);
		const float t = (pos - buf[i].s[axis]) / da;
		V3 q;
		q.s[0] = buf[i].s[0] + t * dx; q.s[1] = buf[i].s[1] + t * dy; q.s[2] = buf[i].s[2] + t * dz;
		pts[size++] = q;
	}
}
__device__ float intersect_area(const float* v0, const float* v1, const float* v2, float4 lo, float4 hi) {   // EPO.cl:102-130
	if (point_in_box(v0, lo, hi) && point_in_box(v1, lo, hi) && point_in_box(v2, lo, hi)) return half_cross_len(v1, v2, v0);
	// no separating-axis shortcut: interpolated vertices can round onto a plane the originals lie beyond, and the
	// float sums are compared bit for bit
	V3 pts[10];
	for (int k = 0; k < 3; ++k) { pts[0].s[k] = v0[k]; pts[1].s[k] = v1[k]; pts[2].s[k] = v2[k]; }
	int n = 3;
	clip(pts, n, 0, lo.x, 1); clip(pts, n, 1, lo.y, 1); clip(pts, n, 2, lo.z, 1);
	clip(pts, n, 0, hi.x, -1); clip(pts, n, 1, hi.y, -1); clip(pts, n, 2, hi.z, -1);
	float ans = 0.0f;                                                                      // pArea, EPO.cl:92-100
	if (n < 2) return ans;
	for (int i = 1; i < n - 1; ++i) ans += half_cross_len(pts[i].s, pts[i + 1].s, pts[0].s);
	return ans;
}

__global__ void __launch_bounds__(128) k_epo(const mcpt_bvh_node* __restrict__ nodes, const mcpt_triangle* __restrict__ tris,
                                             int prims, float* epo, float* area, int* status) {
	const int gid = blockIdx.x * blockDim.x + threadIdx.x;
	if (gid >= prims) return;
	const int me = gid + prims - 1;
	const int tri = node_link(nodes, me).y;
	float v0[3], v1[3], v2[3];
	{
		const float4* t = reinterpret_cast<const float4*>(tris + tri);
		const float4 a = __ldg(t), b = __ldg(t + 1), c = __ldg(t + 2);
		v0[0] = a.x; v0[1] = a.y; v0[2] = a.z; v1[0] = b.x; v1[1] = b.y; v1[2] = b.z; v2[0] = c.x; v2[1] = c.y; v2[2] = c.z;
	}
	// path[d] = the ancestor (or the leaf itself) at depth d; the kernel's list is the same set, leaf first
	int path[STACK];
	int depth = 0;
	for (int p = me; p != -1; p = node_link(nodes, p).x) {
		if (depth >= STACK) { atomicExch(status, 1); return; }
		path[depth++] = p;
	}
	for (int i = 0; i < depth / 2; ++i) { const int t = path[i]; path[i] = path[depth - 1 - i]; path[depth - 1 - i] = t; }
	int todo[STACK];
	unsigned char todo_depth[STACK];
	int ts = 0;
	float sum = 0.0f;
	todo[ts] = 0; todo_depth[ts] = 0; ++ts;
	while (ts > 0) {
		--ts;
		const int now = todo[ts];
		const int d = todo_depth[ts];
		bool descend = (d < depth && path[d] == now);
		if (!descend) {
			const float a = intersect_area(v0, v1, v2, node_lo(nodes, now), node_hi(nodes, now));
			if (a > 0) {
				sum += a * (now >= prims - 1 ? C_TRI : C_INNER);
				descend = true;
			}
		}
		if (descend) {
			const int4 lk = node_link(nodes, now);
			if (lk.y != lk.z) {
				if (ts + 2 > STACK || d + 1 > 255) { atomicExch(status, 1); return; }
				todo[ts] = lk.z; todo_depth[ts] = (unsigned char)(d + 1); ++ts;
				todo[ts] = lk.y; todo_depth[ts] = (unsigned char)(d + 1); ++ts;
			}
		}
	}
	epo[gid] = sum;
	area[gid] = half_cross_len(v1, v2, v0);
}

// ---------------------------------------------------------------------------------------------------- LCV
struct LcvCamera { float ox, oy, oz, ow; float4 front, horizontal, up; };   // front = distance * camera.direction

__global__ void __launch_bounds__(128) k_lcv(const mcpt_bvh_node* __restrict__ nodes, LcvCamera cam, int width, int height,
                                             unsigned long long* sums, int* status) {
	const long idx = long(blockIdx.x) * blockDim.x + threadIdx.x;
	unsigned long long c = 0;
	if (idx < long(width) * height) {
		const int i = int(idx / height), j = int(idx % height);                             // bvhtest.cpp:416-427
		const float t1 = (float(i) + 0.5f) / float(width) - 0.5f;
		const float t2 = (float(j) + 0.5f) / float(height) - 0.5f;
		const float dx = cam.front.x + t1 * cam.horizontal.x + t2 * cam.up.x;
		const float dy = cam.front.y + t1 * cam.horizontal.y + t2 * cam.up.y;
		const float dz = cam.front.z + t1 * cam.horizontal.z + t2 * cam.up.z;
		int stack[STACK];
		int sp = 0;
		int pt = 0;
		for (;;) {                                                                          // bvhtest.cpp:328-357
			const float4 lo = node_lo(nodes, pt), hi = node_hi(nodes, pt);
			const float ax = (lo.x - cam.ox) / dx, ay = (lo.y - cam.oy) / dy, az = (lo.z - cam.oz) / dz;   // intersectBox, :22-33
			const float bx = (hi.x - cam.ox) / dx, by = (hi.y - cam.oy) / dy, bz = (hi.z - cam.oz) / dz;
			const float tnear = hmax(hmax(hmin(ax, bx), hmin(ay, by)), hmin(az, bz));
			const float tfar = hmin(hmin(hmax(ax, bx), hmax(ay, by)), hmax(az, bz));
			bool pop = true;
			if (!(tfar < tnear || tfar < 0.001f)) {
				const int4 lk = node_link(nodes, pt);
				if (lk.y == lk.z) {
					++c;
				} else {
					if (sp >= STACK) { atomicExch(status, 1); break; }
					stack[sp++] = lk.z;
					pt = lk.y;
					pop = false;
				}
			}
			if (pop) {
				if (sp == 0) break;
				pt = stack[--sp];
			}
		}
	}
	// warp sums, one pair of atomics per warp
	unsigned long long c2 = c * c;
	for (int o = 16; o > 0; o >>= 1) {
		c += __shfl_xor_sync(0xffffffffu, c, o);
		c2 += __shfl_xor_sync(0xffffffffu, c2, o);
	}
	if ((threadIdx.x & 31) == 0 && c) { atomicAdd(sums, c); atomicAdd(sums + 1, c2); }
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

extern "C" int mcpt_bvh_metrics(int device, const mcpt_bvh_node* nodes, int nnodes, const mcpt_triangle* tris, int ntris,
                                const mcpt_camera* camera, int width, int height, mcpt_bvh_quality* out,
                                float* epo_per_leaf, float* area_per_leaf) {
	if (!nodes || !out || nnodes < 1 || (tris && nnodes != 2 * ntris - 1) || (camera && (width < 1 || height < 1))) {
		set_last_error("mcpt_bvh_metrics: expected 2*ntris-1 nodes, a result record and (with a camera) a resolution");
		return MCPT_ERR_ARG;
	}
	*out = mcpt_bvh_quality{};
	TRY_(cudaSetDevice(device));
	cudaStream_t st;
	TRY_(cudaStreamCreate(&st));
	struct StreamGuard { cudaStream_t s; ~StreamGuard() { cudaStreamDestroy(s); } } guard{st};
	cudaEvent_t ev[4];
	for (auto& e : ev) TRY_(cudaEventCreate(&e));
	struct EventGuard { cudaEvent_t* e; ~EventGuard() { for (int i = 0; i < 4; ++i) cudaEventDestroy(e[i]); } } eguard{ev};
	Scoped<mcpt_bvh_node> d_nodes;
	Scoped<mcpt_triangle> d_tris;
	Scoped<float> d_terms, d_epo, d_area;
	Scoped<unsigned long long> d_sums;
	Scoped<int> d_status;
	const size_t nn = size_t(nnodes);
	TRY_(d_nodes.alloc(nn)); TRY_(d_terms.alloc(nn)); TRY_(d_sums.alloc(2)); TRY_(d_status.alloc(1));
	TRY_(cudaMemcpyAsync(d_nodes.p, nodes, sizeof(mcpt_bvh_node) * nn, cudaMemcpyHostToDevice, st));
	TRY_(cudaMemsetAsync(d_sums.p, 0, 2 * sizeof(unsigned long long), st));
	TRY_(cudaMemsetAsync(d_status.p, 0, sizeof(int), st));
	if (tris) {
		TRY_(d_tris.alloc(size_t(ntris))); TRY_(d_epo.alloc(size_t(ntris))); TRY_(d_area.alloc(size_t(ntris)));
		TRY_(cudaMemcpyAsync(d_tris.p, tris, sizeof(mcpt_triangle) * size_t(ntris), cudaMemcpyHostToDevice, st));
	}
	// ---- SAH
	TRY_(cudaEventRecord(ev[0], st));
	k_sah_terms<<<blocks(nnodes, 256), 256, 0, st>>>(d_nodes.p, nnodes, d_terms.p);
	TRY_(cudaEventRecord(ev[1], st));
	// ---- EPO
	if (tris) k_epo<<<blocks(ntris, 128), 128, 0, st>>>(d_nodes.p, d_tris.p, ntris, d_epo.p, d_area.p, d_status.p);
	TRY_(cudaEventRecord(ev[2], st));
	// ---- LCV
	if (camera) {
		// bvhtest.cpp:420: `0.5f / tan(camera.arg / 2)` - with g++/glibc (the build that pins this) the unqualified
		// tan is the double one: float half-angle, double tangent and quotient, narrowed on assignment
		const float distance = float(0.5f / std::tan(double(camera->arg / 2)));
		LcvCamera c;
		c.ox = camera->center.x; c.oy = camera->center.y; c.oz = camera->center.z; c.ow = camera->center.w;
		c.front = make_float4(distance * camera->direction.x, distance * camera->direction.y, distance * camera->direction.z, 0.0f);
		c.horizontal = make_float4(camera->horizontal.x, camera->horizontal.y, camera->horizontal.z, 0.0f);
		c.up = make_float4(camera->up.x, camera->up.y, camera->up.z, 0.0f);
		k_lcv<<<blocks(long(width) * height, 128), 128, 0, st>>>(d_nodes.p, c, width, height, d_sums.p, d_status.p);
	}
	TRY_(cudaEventRecord(ev[3], st));
	std::vector<float> terms(nn), epo, area;
	unsigned long long sums[2] = {0, 0};
	int status = 0;
	TRY_(cudaMemcpyAsync(terms.data(), d_terms.p, sizeof(float) * nn, cudaMemcpyDeviceToHost, st));
	if (tris) {
		epo.resize(size_t(ntris)); area.resize(size_t(ntris));
		TRY_(cudaMemcpyAsync(epo.data(), d_epo.p, sizeof(float) * size_t(ntris), cudaMemcpyDeviceToHost, st));
		TRY_(cudaMemcpyAsync(area.data(), d_area.p, sizeof(float) * size_t(ntris), cudaMemcpyDeviceToHost, st));
	}
	TRY_(cudaMemcpyAsync(sums, d_sums.p, sizeof(sums), cudaMemcpyDeviceToHost, st));
	TRY_(cudaMemcpyAsync(&status, d_status.p, sizeof(int), cudaMemcpyDeviceToHost, st));
	TRY_(cudaStreamSynchronize(st));
	TRY_(cudaGetLastError());
	if (status != 0) {
		set_last_error("mcpt_bvh_metrics: tree deeper than the 128-entry stacks of kernels/EPO.cl (the reference overflows here)");
		return MCPT_ERR_STATE;
	}
	TRY_(cudaEventElapsedTime(&out->ms_sah, ev[0], ev[1]));
	TRY_(cudaEventElapsedTime(&out->ms_epo, ev[1], ev[2]));
	TRY_(cudaEventElapsedTime(&out->ms_lcv, ev[2], ev[3]));
	{                                                      // bvhtest.cpp:98-109, same order
		double s = 0.0f;
		for (size_t i = 0; i < (nn >> 1); ++i) s += terms[i];
		for (size_t i = (nn >> 1); i < nn; ++i) s += terms[i];
		const mcpt_bvh_node& r = nodes[0];
		const float x = r.bbmax.x - r.bbmin.x, y = r.bbmax.y - r.bbmin.y, z = r.bbmax.z - r.bbmin.z;
		s /= 2.0f * (x * y + x * z + y * z);
		out->sah_exact = s;
		out->sah = double(float(s));                       // the reference returns float
	}
	if (tris) {                                            // bvhtest.cpp:309-320
		double count = 0.0, a = 0.0;
		for (int i = 0; i < ntris; ++i) count += epo[size_t(i)];
		for (int i = 0; i < ntris; ++i) a += area[size_t(i)];
		out->epo_area = count;
		out->triangle_area = a;
		out->epo = double(float(count / a));
		if (epo_per_leaf) std::copy(epo.begin(), epo.end(), epo_per_leaf);
		if (area_per_leaf) std::copy(area.begin(), area.end(), area_per_leaf);
	}
	if (camera) {                                          // bvhtest.cpp:435-444
		const double n = double(size_t(width) * size_t(height));
		double en = double(sums[0]), en2 = double(sums[1]);
		en /= n;
		en2 /= n;
		out->lcv_rays = uint64_t(width) * uint64_t(height);
		out->lcv_mean_leaves = en;
		out->lcv = double(float(std::sqrt(en2 - en * en)));
	}
	return MCPT_OK;
}
