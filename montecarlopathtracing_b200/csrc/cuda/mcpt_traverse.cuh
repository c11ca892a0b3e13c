// BVH traversal on the device.
//
//  trace_reference : intersectObjects (objdef.h:240-275) as written — 64-byte BVHNode records, one
//                    box test per visited node, left-first DFS with a 64-entry stack, no pruning.
//  trace_canonical      : the same canonical left-first LEAF ORDER on a derived layout.  Why this is
//                    still the reference's answer:
//                    (1) the reference's box test is monotone (IEEE division, fmin/fmax are monotone
//                        and a parent box contains its children exactly, hlbvh.cpp:64-76), so without
//                        NaNs "every ancestor box passes" <=> "the leaf's own box passes".  Inner
//                        boxes may therefore be tested CONSERVATIVELY (reciprocal multiply + error
//                        slack, never rejecting what the reference accepts) as long as each leaf box
//                        gets the exact reference test before its triangle is touched.
//                    (2) candidates reach the accept rule `hit.t - t >= 1e-5` (objdef.h:213) in the
//                        same order, so ties and near-ties resolve identically (the cbox lamp is
//                        coplanar with the ceiling).
//                    (3) a subtree is skipped only when its entry distance exceeds the current hit by
//                        a margin far above the triangle solve's rounding error; such candidates
//                        could never satisfy the accept rule.
//                    Rays with a zero direction component (0/0 = NaN breaks (1)) take the reference
//                    walk instead.
#pragma once

#include "accel_layout.hpp"
#include "mcpt_device.cuh"

namespace mcpt {

struct WNode;   // experimental/accel_wide.hpp
struct SceneDev {
	const mcpt_bvh_node* nodes_ref;
	const mcpt_triangle* tris_ref;
	const DNode* nodes;          // reference topology, two-box layout (canonical walk)
	const DNode* fast_nodes;     // binary SAH tree over the same leaves (MCPT_TRAVERSAL_FAST)
	const WNode* wnodes;         // 8-wide compressed collapse of that tree (MCPT_TRAVERSAL_WIDE8)
	const DLeaf* wleaves;        // leaf records in wide-node order; DLeaf::pad = original triangle index
	const DLeaf* leaves;         // leaf records indexed by triangle
	const mcpt_material* materials;
	const float2* cs_table;
	int ntris, nnodes;
	int root_is_leaf;
	int opt_flags;               // development switches (MCPT_OPT): bit 1 = one-thread-per-ray kernels instead of persistent ones
	float max_abs;               // largest |coordinate| of the scene's bounding box (finite-arithmetic guard of the fast walk)
	float key_lo[3], key_scale[3];   // sort key of the bounce queues: cell = (origin - key_lo) * key_scale on a 32^3 grid
};

// ---- checked build (-DMCPT_CHECKED=1, tools/checked_run.sh): every index the walk, the queues and the resolve follow is
// compared with its bound before use; a violation is counted, its first occurrence recorded (code, value, bound) and
// the index replaced by a safe one, so the run finishes and mcpt_check_failures() reports it.  compute-sanitizer is
// closed on this pool; this is the bounds / protocol evidence that stands in for it.  The product build compiles the
// checks away.
#ifndef MCPT_CHECKED
#define MCPT_CHECKED 0
#endif
enum { CHK_NODE = 1, CHK_LEAF = 2, CHK_STACK = 3, CHK_HIT = 4, CHK_PERM = 5, CHK_QUEUE = 6, CHK_PATH = 7, CHK_MATERIAL = 8 };
#if MCPT_CHECKED
static __device__ unsigned long long g_check[4];      // failures, first code, first value, first bound
__device__ __forceinline__ bool check_failed(int code, long long value, long long bound) {
	if (atomicAdd(&g_check[0], 1ull) == 0ull) { g_check[1] = (unsigned long long)code; g_check[2] = (unsigned long long)value; g_check[3] = (unsigned long long)bound; }
	return true;
}
// index in [0, bound) or recorded and replaced by 0
#define MCPT_CHECK_INDEX(code, idx, bound) do { if ((long long)(idx) < 0 || (long long)(idx) >= (long long)(bound)) { check_failed(code, (long long)(idx), (long long)(bound)); (idx) = 0; } } while (0)
#else
#define MCPT_CHECK_INDEX(code, idx, bound) do { } while (0)
#endif

// 256-bit read-only loads (LDG.E.256, sm_100): a divergent warp costs the L1 one wavefront per lane PER
// LOAD INSTRUCTION, and this kernel is bound by exactly that, so a 64 B node is two loads instead of four.
#ifndef MCPT_LDG256
#define MCPT_LDG256 1
#endif
// development variants of the cache policy (A/B only): MCPT_LEAF_NOALLOC = leaf / triangle records bypass L1 allocation
// (each is read once per ray that reaches it), MCPT_NODE_EVICT_LAST = inner nodes are the last lines L1 gives up
#ifndef MCPT_LEAF_NOALLOC
#define MCPT_LEAF_NOALLOC 0
#endif
#ifndef MCPT_NODE_EVICT_LAST
#define MCPT_NODE_EVICT_LAST 0
#endif
__device__ __forceinline__ void ldg256_leaf(const float4* p, float4& lo, float4& hi);
__device__ __forceinline__ void ldg256_node(const float4* p, float4& lo, float4& hi);
__device__ __forceinline__ void ldg256(const float4* p, float4& lo, float4& hi) {
#if MCPT_LDG256
	asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
	             : "=f"(lo.x), "=f"(lo.y), "=f"(lo.z), "=f"(lo.w), "=f"(hi.x), "=f"(hi.y), "=f"(hi.z), "=f"(hi.w)
	             : "l"(p));
#else
	lo = __ldg(p); hi = __ldg(p + 1);
#endif
}
__device__ __forceinline__ void ldg256_leaf(const float4* p, float4& lo, float4& hi) {
#if MCPT_LEAF_NOALLOC
	asm volatile("ld.global.nc.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
	             : "=f"(lo.x), "=f"(lo.y), "=f"(lo.z), "=f"(lo.w), "=f"(hi.x), "=f"(hi.y), "=f"(hi.z), "=f"(hi.w)
	             : "l"(p));
#else
	ldg256(p, lo, hi);
#endif
}
__device__ __forceinline__ void ldg256_node(const float4* p, float4& lo, float4& hi) {
#if MCPT_NODE_EVICT_LAST
	asm volatile("ld.global.nc.L1::evict_last.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
	             : "=f"(lo.x), "=f"(lo.y), "=f"(lo.z), "=f"(lo.w), "=f"(hi.x), "=f"(hi.y), "=f"(hi.z), "=f"(hi.w)
	             : "l"(p));
#else
	ldg256(p, lo, hi);
#endif
}

struct HitRegs {
	float t;
	int32_t closest;        // triangle whose data sits in the hit, -1 = none
	int32_t last_touched;   // reference Hit.triangleID semantics (F7); reference walk only
	bool any;
};

constexpr int STACK_SIZE = 64;       // objdef.h:247
// pruning margin: skip a box when tnear > t_hit * (1 + 2^-7) + 2^-7
constexpr float PRUNE_REL = 1.0f + 0.0078125f;
constexpr float PRUNE_ABS = 0.0078125f;
constexpr float SLAB_SLACK = 9.5367431640625e-7f;   // 2^-20, >= 4x the reciprocal-multiply error bound

__device__ __forceinline__ TriTest leaf_triangle(const DLeaf* __restrict__ leaves, int tri, const RayRegs& ray, float tmin) {
	const float4* p = reinterpret_cast<const float4*>(leaves + tri);
	float4 q2 = __ldg(p + 2), q3 = __ldg(p + 3), q4 = __ldg(p + 4), q5 = __ldg(p + 5);
	return intersect_triangle(ray, q2.x, q2.y, q2.z, q3.x, q3.y, q3.z, q4.x, q4.y, q4.z, q5.x, q5.y, q5.z, q2.w, q3.w,
	                          q4.w, tmin);
}

// objdef.h:209-219 applied to one leaf in canonical order
__device__ __forceinline__ void test_leaf_sequential(const DLeaf* __restrict__ leaves, int tri, const RayRegs& ray, float tmin, HitRegs& hit) {
	TriTest r = leaf_triangle(leaves, tri, ray, tmin);
	if (r.valid) {
		hit.any = true;
		hit.last_touched = tri;
		if (hit.t - r.t >= KERNEL_EPSILON) {       // objdef.h:213
			hit.t = r.t;
			hit.closest = tri;
		}
	}
}

// Reference walk over the reference records (the triangle arithmetic reads the derived record, whose
// contents are the same floats).
template <bool COUNT>
__device__ __forceinline__ void trace_reference(const SceneDev& sc, const RayRegs& ray, float tmin, HitRegs& hit,
                                                unsigned long long* cnt_nodes, unsigned long long* cnt_tris) {
	int stack[STACK_SIZE];
	int sp = 1;
	stack[0] = 0;
	unsigned ln = 0, lt = 0;
	while (sp > 0) {
		int ni = stack[--sp];
		for (;;) {
			const float4* p = reinterpret_cast<const float4*>(sc.nodes_ref + ni);
			float4 bmin = __ldg(p), bmax = __ldg(p + 1);
			int4 link = __ldg(reinterpret_cast<const int4*>(p + 3));   // parent, left, right, pad
			if (COUNT) ++ln;
			float tn;
			if (!intersect_aabb_exact(bmin.x, bmin.y, bmin.z, bmax.x, bmax.y, bmax.z, ray, tmin, tn)) break;
			if (link.y == link.z) {
				if (COUNT) ++lt;
				test_leaf_sequential(sc.leaves, link.y, ray, tmin, hit);
				break;
			}
			if (sp < STACK_SIZE) stack[sp++] = link.z;
			ni = link.y;
		}
	}
	if (COUNT) { atomicAdd(cnt_nodes, (unsigned long long)ln); atomicAdd(cnt_tris, (unsigned long long)lt); }
}

// Conservative slab test with a precomputed reciprocal direction.  Returns false only when the
// reference's exact test is certain to reject (or the box lies beyond the pruning distance).
__device__ __forceinline__ bool slab_conservative(float bminx, float bminy, float bminz, float bmaxx, float bmaxy,
                                                  float bmaxz, const RayRegs& ray, float rx, float ry, float rz,
                                                  float tmin, float tprune) {
	float o1x = (bminx - ray.origin.x) * rx, o2x = (bmaxx - ray.origin.x) * rx;
	float o1y = (bminy - ray.origin.y) * ry, o2y = (bmaxy - ray.origin.y) * ry;
	float o1z = (bminz - ray.origin.z) * rz, o2z = (bmaxz - ray.origin.z) * rz;
	float tnear = fmaxf(fmaxf(fminf(o1x, o2x), fminf(o1y, o2y)), fminf(o1z, o2z));
	float tfar = fminf(fminf(fmaxf(o1x, o2x), fmaxf(o1y, o2y)), fmaxf(o1z, o2z));
	float slack = SLAB_SLACK * fmaxf(fmaxf(fabsf(tnear), fabsf(tfar)), tmin);
	float lo = tnear - slack;
	// every comparison is false on NaN => "not rejected"
	bool reject = (tfar < lo) || (tfar < tmin - slack) || (lo > tprune);
	return !reject;
}

__device__ __forceinline__ bool leaf_box_exact(float bminx, float bminy, float bminz, float bmaxx, float bmaxy,
                                               float bmaxz, const RayRegs& ray, float tmin, float tprune) {
	float tn;
	if (!intersect_aabb_exact(bminx, bminy, bminz, bmaxx, bmaxy, bmaxz, ray, tmin, tn)) return false;
	return !(tn > tprune);
}

template <bool COUNT>
__device__ __forceinline__ void trace_canonical(const SceneDev& sc, const RayRegs& ray, float tmin, HitRegs& hit,
                                           unsigned long long* cnt_nodes, unsigned long long* cnt_tris) {
	// 0/0 = NaN breaks the monotonicity argument: take the reference walk (measure-zero ray set, e.g.
	// the centre pixel column of an axis-aligned camera).
	if (ray.direction.x == 0.0f || ray.direction.y == 0.0f || ray.direction.z == 0.0f || sc.root_is_leaf) {
		trace_reference<COUNT>(sc, ray, tmin, hit, cnt_nodes, cnt_tris);
		return;
	}
	const float rx = 1.0f / ray.direction.x, ry = 1.0f / ray.direction.y, rz = 1.0f / ray.direction.z;
	int stack[STACK_SIZE];     // >= 0: inner node to visit; < 0: ~triangle whose leaf box already passed
	int sp = 0;
	int ni = 0;
	unsigned ln = 0, lt = 0;
	for (;;) {
		const float4* p = reinterpret_cast<const float4*>(sc.nodes + ni);
		float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
		int4 link = __ldg(reinterpret_cast<const int4*>(p + 3));
		if (COUNT) ++ln;
		const int left = link.x, right = link.y;
		float tprune = hit.t * PRUNE_REL + PRUNE_ABS;
		int next = -1;                 // inner node to descend into
		// ---- left child first (canonical order)
		if (left < 0) {
			if (leaf_box_exact(a.x, a.y, a.z, a.w, b.x, b.y, ray, tmin, tprune)) {
				if (COUNT) ++lt;
				test_leaf_sequential(sc.leaves, ~left, ray, tmin, hit);
				tprune = hit.t * PRUNE_REL + PRUNE_ABS;
			}
		} else if (slab_conservative(a.x, a.y, a.z, a.w, b.x, b.y, ray, rx, ry, rz, tmin, tprune)) {
			next = left;
		}
		// ---- right child
		if (right < 0) {
			if (leaf_box_exact(b.z, b.w, c.x, c.y, c.z, c.w, ray, tmin, tprune)) {
				if (next < 0) {
					if (COUNT) ++lt;
					test_leaf_sequential(sc.leaves, ~right, ray, tmin, hit);
				} else if (sp < STACK_SIZE) {
					stack[sp++] = right;      // the left subtree comes first: defer the triangle
				}
			}
		} else if (slab_conservative(b.z, b.w, c.x, c.y, c.z, c.w, ray, rx, ry, rz, tmin, tprune)) {
			if (next < 0) next = right;
			else if (sp < STACK_SIZE) stack[sp++] = right;
		}
		// ---- pop until an inner node turns up
		while (next < 0) {
			if (sp == 0) { next = -2; break; }
			int e = stack[--sp];
			if (e >= 0) { next = e; break; }
			if (COUNT) ++lt;
			test_leaf_sequential(sc.leaves, ~e, ray, tmin, hit);
			next = -1;
		}
		if (next == -2) break;
		ni = next;
	}
	if (COUNT) { atomicAdd(cnt_nodes, (unsigned long long)ln); atomicAdd(cnt_tris, (unsigned long long)lt); }
}

// ------------------------------------------------------------------ certified-fast leaf tests
// The reference's leaf decisions (six IEEE divisions for the box, nine for the triangle) are reproduced
// WITHOUT divisions whenever the outcome is certain: the same quantities are evaluated with a reciprocal
// multiply, an a-priori error bound is attached to each (|computed - reference's float value| <= bound),
// and a comparison is taken only if it holds by more than the bound.  Anything closer than that - and any
// NaN/Inf, which fails every "certain" test - is re-evaluated with the exact reference arithmetic
// (intersect_aabb_exact / intersect_triangle).  Only the decisions and the final `t` enter the result, and
// `t` of an accepted candidate is always computed with the reference's exact operations.
//
// Box: v~ = (bb - o) * RN(1/d) vs the reference's RN(RN(bb - o) / d): relative difference <= 3 * 2^-24;
//      bound used 2^-21 |v~|.  accept <=> n_i <= f_j for all i != j and f_j >= tmin (same-axis n_i <= f_i holds
//      identically in both evaluations, which also covers flat boxes where n_i == f_i).
// Triangle: b~ = (A . num_b) * rcp(det) vs the reference's sum of RN(A_k * RN(num_k / det)): both are within
//      a few 2^-24 of sum |A_k num_k / det| of the real value; bound used 2^-19 * sum |A_k num_k| |rcp(det)| + 1e-37.
constexpr float BOX_REL = 4.76837158203125e-7f;       // 2^-21
constexpr float TRI_REL = 1.9073486328125e-6f;        // 2^-19
constexpr float TRI_ABS = 1e-37f;

__device__ __forceinline__ float rcp_approx(float x) {
	float r;
	asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
	return r;
}

// 1 = box passes (and is not beyond tprune), 0 = rejected.  `fastok` is false for rays whose reciprocal
// direction is not finite-safe (|d| < 1e-30): they always take the exact test.
__device__ __forceinline__ bool leaf_box_certified(const float4& q0, const float4& q1, const RayRegs& ray, float rx, float ry,
                                                   float rz, bool fastok, float tmin, float tprune) {
	if (fastok) {
		float l, h;
		l = (q0.x - ray.origin.x) * rx; h = (q1.x - ray.origin.x) * rx;
		const float nx = fminf(l, h), fx = fmaxf(l, h);
		l = (q0.y - ray.origin.y) * ry; h = (q1.y - ray.origin.y) * ry;
		const float ny = fminf(l, h), fy = fmaxf(l, h);
		l = (q0.z - ray.origin.z) * rz; h = (q1.z - ray.origin.z) * rz;
		const float nz = fminf(l, h), fz = fmaxf(l, h);
		// upper bounds of the reference's n_i, lower bounds of its f_j
		const float Nx = __fmaf_rn(BOX_REL, fabsf(nx), nx), Ny = __fmaf_rn(BOX_REL, fabsf(ny), ny), Nz = __fmaf_rn(BOX_REL, fabsf(nz), nz);
		const float Fx = __fmaf_rn(-BOX_REL, fabsf(fx), fx), Fy = __fmaf_rn(-BOX_REL, fabsf(fy), fy), Fz = __fmaf_rn(-BOX_REL, fabsf(fz), fz);
		const bool sure_accept = Nx <= Fy && Nx <= Fz && Ny <= Fx && Ny <= Fz && Nz <= Fx && Nz <= Fy && Fx >= tmin && Fy >= tmin && Fz >= tmin;
		// lower bounds of n_i, upper bounds of f_j
		const float Lx = __fmaf_rn(-BOX_REL, fabsf(nx), nx), Ly = __fmaf_rn(-BOX_REL, fabsf(ny), ny), Lz = __fmaf_rn(-BOX_REL, fabsf(nz), nz);
		if (sure_accept) return !(fmaxf(fmaxf(Lx, Ly), Lz) > tprune);     // L <= true tnear: pruning stays safe
		const float Ux = __fmaf_rn(BOX_REL, fabsf(fx), fx), Uy = __fmaf_rn(BOX_REL, fabsf(fy), fy), Uz = __fmaf_rn(BOX_REL, fabsf(fz), fz);
		const bool sure_reject = Lx > Uy || Lx > Uz || Ly > Ux || Ly > Uz || Lz > Ux || Lz > Uy || Ux < tmin || Uy < tmin || Uz < tmin;
		if (sure_reject) return false;
	}
	float tn;
	if (!intersect_aabb_exact(q0.x, q0.y, q0.z, q1.x, q1.y, q1.z, ray, tmin, tn)) return false;
	return !(tn > tprune);
}

// Exact outcome of the reference triangle test for candidates that can matter: returns valid + the
// reference's t.  Candidates certainly farther than `tfar_skip` (best + 2 * WINDOW) are reported invalid:
// they can be neither near candidates nor band members.
__device__ __forceinline__ TriTest leaf_triangle_certified(const float4* __restrict__ p, const RayRegs& ray, bool fastok, float tmin,
                                                           float tfar_skip, int& orig_tri) {
	float4 q2, q3, q4, q5;
	ldg256_leaf(p + 2, q2, q3);
	ldg256_leaf(p + 4, q4, q5);
	orig_tri = __float_as_int(q5.w);
	TriTest out{0.0f, false};
	const float a1 = ray.direction.x, a2 = ray.direction.y, a3 = ray.direction.z;
	if (fabsf(q2.w * a1 + q3.w * a2 + q4.w * a3) < KERNEL_EPSILON) return out;      // objdef.h:186 (exact: no division)
	const float e1x = q3.x, e1y = q3.y, e1z = q3.z, e2x = q4.x, e2y = q4.y, e2z = q4.z;
	const float T2 = a2 * e2z - e2y * a3;
	const float T3 = a2 * e1z - e1y * a3;
	const float det = a1 * q5.x - e1x * T2 + e2x * T3;
	if (fabsf(det) < KERNEL_EPSILON) return out;                                      // objdef.h:152 (exact)
	const float Ax = q2.x - ray.origin.x, Ay = q2.y - ray.origin.y, Az = q2.z - ray.origin.z;
	bool ambiguous = !fastok;
	if (fastok) {
		const float rd = rcp_approx(det), ard = fabsf(rd);
		// b
		float p1 = Ax * -T2, p2 = Ay * (a1 * e2z - e2x * a3), p3 = Az * -(a1 * e2y - e2x * a2);
		const float bt = (p1 + p2 + p3) * rd;
		const float bb = __fmaf_rn(TRI_REL, (fabsf(p1) + fabsf(p2) + fabsf(p3)) * ard, TRI_ABS);
		if (bt < -bb) return out;
		ambiguous |= !(bt > bb);
		// c
		p1 = Ax * T3; p2 = Ay * -(a1 * e1z - e1x * a3); p3 = Az * (a1 * e1y - e1x * a2);
		const float ct = (p1 + p2 + p3) * rd;
		const float cb = __fmaf_rn(TRI_REL, (fabsf(p1) + fabsf(p2) + fabsf(p3)) * ard, TRI_ABS);
		if (ct < -cb) return out;
		ambiguous |= !(ct > cb);
		const float sum = bt + ct, sb = bb + cb + BOX_REL;
		if (sum - 1.0f > sb) return out;
		ambiguous |= !(1.0f - sum > sb);
		// t
		p1 = Ax * q5.x; p2 = Ay * -q5.y; p3 = Az * q5.z;
		const float tt = (p1 + p2 + p3) * rd;
		const float tb = __fmaf_rn(TRI_REL, (fabsf(p1) + fabsf(p2) + fabsf(p3)) * ard, TRI_ABS);
		if (tt + tb < tmin) return out;              // reference t <= tmin for sure
		if (tt - tb > tfar_skip) return out;         // certainly far: irrelevant to the near set
	}
	if (ambiguous)
		return intersect_triangle(ray, q2.x, q2.y, q2.z, e1x, e1y, e1z, e2x, e2y, e2z, q5.x, q5.y, q5.z, q2.w, q3.w, q4.w, tmin);
	// b, c and b + c certainly pass: the reference's t, with its exact operations (objdef.h:203-207)
	const float s0 = fdiv(q5.x, det), s4 = fdiv(-q5.y, det), s8 = fdiv(q5.z, det);
	const float t = Ax * s0 + Ay * s4 + Az * s8;
	if (t <= tmin) return out;
	out.t = t;
	out.valid = true;
	return out;
}

// ------------------------------------------------------------------ ordered traversal + canonical replay
//
// trace_fast visits the tree NEAR CHILD FIRST and prunes against the closest candidate found so far, so
// candidates no longer arrive in the reference's order.  The reference's answer is recovered exactly:
//   * A candidate is a triangle whose leaf box passes the exact reference box test and whose exact
//     triangle test returns true (the same set the reference evaluates, see (1) above), minus boxes whose
//     entry distance exceeds the best candidate by the pruning margin.
//   * Let T = min t over candidates, e = 1e-5f (objdef.h:16), w = WINDOW (2e-5f), b = BAND (9.9e-6f <= w - e).
//     NEAR = {t - T <= w}, FAR = the rest.  (t - T is exact in float near T by Sterbenz.)
//     If no near candidate has t - T > b, then for every far f and near c: f - c > w - b >= e, so in the
//     reference's sequential rule `hit.t - t >= e` (a) a near candidate arriving while hit.t is FLT_MAX or
//     a far value is always accepted, (b) once hit.t is near no far candidate is accepted (hit.t - f < 0).
//     Hence the reference's final (t, triangle) equals the rule replayed over the NEAR candidates alone in
//     canonical order (rank = position of the leaf in the reference's left-first DFS).
//   * Near candidates are kept in a 3-slot list.  Overflow or a non-empty band (both need >= 3 surfaces
//     within 2e-5 of each other along the ray) fall back to trace_canonical for that ray.
constexpr float WINDOW = 2.0e-5f;
constexpr float BAND = 9.9e-6f;
constexpr int LEAF_FLAG = int(0x80000000);

struct CandList {
	float t0, t1, t2;
	int tri0, tri1, tri2;
	int rank0, rank1, rank2;
	int n;
	bool overflow;
};

__device__ __forceinline__ void cand_insert(CandList& c, float best, float t, int tri, int rank) {
	// drop entries that are now far from the (possibly new) best: best only decreases, so they stay far
	bool k0 = c.n > 0 && !(c.t0 - best > WINDOW);
	bool k1 = c.n > 1 && !(c.t1 - best > WINDOW);
	bool k2 = c.n > 2 && !(c.t2 - best > WINDOW);
	if (!k0) {
		if (k1) { c.t0 = c.t1; c.tri0 = c.tri1; c.rank0 = c.rank1; k0 = true; k1 = false; }
		else if (k2) { c.t0 = c.t2; c.tri0 = c.tri2; c.rank0 = c.rank2; k0 = true; k2 = false; }
	}
	if (!k1 && k2) { c.t1 = c.t2; c.tri1 = c.tri2; c.rank1 = c.rank2; k1 = true; k2 = false; }
	c.n = int(k0) + int(k1) + int(k2);
	if (c.n == 0) { c.t0 = t; c.tri0 = tri; c.rank0 = rank; c.n = 1; }
	else if (c.n == 1) { c.t1 = t; c.tri1 = tri; c.rank1 = rank; c.n = 2; }
	else if (c.n == 2) { c.t2 = t; c.tri2 = tri; c.rank2 = rank; c.n = 3; }
	else c.overflow = true;
}

// Replays the accept rule over the near candidates in canonical order.  Returns false when the band
// condition fails (caller falls back).
__device__ __forceinline__ bool cand_resolve(const CandList& c, float best, HitRegs& hit) {
	float t[3] = {c.t0, c.t1, c.t2};
	int tri[3] = {c.tri0, c.tri1, c.tri2};
	int rank[3] = {c.rank0, c.rank1, c.rank2};
	bool ok = true;
#pragma unroll
	for (int k = 0; k < 3; ++k) {
		bool used = k < c.n && !(t[k] - best > WINDOW);
		if (!used) rank[k] = 0x7fffffff;
		else if (t[k] - best > BAND) ok = false;
	}
	// 3-element sorting network on rank
#define MCPT_CSWAP(i, j) if (rank[j] < rank[i]) { float tt = t[i]; t[i] = t[j]; t[j] = tt; int a = tri[i]; tri[i] = tri[j]; tri[j] = a; a = rank[i]; rank[i] = rank[j]; rank[j] = a; }
	MCPT_CSWAP(0, 1) MCPT_CSWAP(1, 2) MCPT_CSWAP(0, 1)
#undef MCPT_CSWAP
#pragma unroll
	for (int k = 0; k < 3; ++k) {
		if (rank[k] != 0x7fffffff && hit.t - t[k] >= KERNEL_EPSILON) { hit.t = t[k]; hit.closest = tri[k]; }
	}
	return ok;
}

// State of one ordered binary walk, so that the persistent kernels can advance many rays round by round
// (a "round" = descend inner nodes until a leaf turns up, then test the leaves at hand).
struct Walk2 {
	int stack[STACK_SIZE];     // >= 0 inner node; LEAF_FLAG | triangle for a leaf
	int top;                   // MCPT_TOP_CACHE: newest stack entry (0 = none)
	int sp, cur;
	float best;
	float rx, ry, rz;
	float c1x, c1y, c1z, c2x, c2y, c2z;   // plane distance = fma(plane, r, c): c1 for bmin planes, c2 for bmax planes
	CandList cl;
	bool fastok, slow;         // slow: zero direction component or single-triangle scene -> reference walk
	unsigned ln, lt, lr;       // instrumentation
};
constexpr int WALK_DONE = 0x7fffffff;

__device__ __forceinline__ void walk2_init(Walk2& w, const SceneDev& sc, const RayRegs& ray) {
	w.slow = ray.direction.x == 0.0f || ray.direction.y == 0.0f || ray.direction.z == 0.0f || sc.root_is_leaf;
	w.rx = 1.0f / ray.direction.x; w.ry = 1.0f / ray.direction.y; w.rz = 1.0f / ray.direction.z;
	w.fastok = fabsf(ray.direction.x) > 1e-30f && fabsf(ray.direction.y) > 1e-30f && fabsf(ray.direction.z) > 1e-30f;
	{
		// -o*r with the rounding error of the fused form (<= 2.1 * 2^-24 |o*r| per plane, see walk2_round) folded
		// in: bmin planes are the near ones when r > 0, so c1 moves them down and c2 moves the bmax planes up by
		// S = 2^-22 |o*r|; for r < 0 the roles (and the signs) swap.  The slack is per axis: an axis the ray is
		// almost parallel to gets a huge S, but only its own (already huge) plane distances see it.
		const float ox = ray.origin.x * w.rx, oy = ray.origin.y * w.ry, oz = ray.origin.z * w.rz;
		const float sx = copysignf(0x1p-22f * fabsf(ox), w.rx), sy = copysignf(0x1p-22f * fabsf(oy), w.ry),
		            sz = copysignf(0x1p-22f * fabsf(oz), w.rz);
		w.c1x = -ox - sx; w.c2x = -ox + sx;
		w.c1y = -oy - sy; w.c2y = -oy + sy;
		w.c1z = -oz - sz; w.c2z = -oz + sz;
		// everything must stay finite (no Inf - Inf): tiny direction components or far-away origins walk the reference way
		const float big = fmaxf(fmaxf(fabsf(ox), fabsf(oy)), fabsf(oz)) +
		                  sc.max_abs * fmaxf(fmaxf(fabsf(w.rx), fabsf(w.ry)), fabsf(w.rz));
		w.slow = w.slow || !w.fastok || !(big < 1e37f);
	}
	w.sp = 0;
	w.top = 0;
	w.cur = w.slow ? WALK_DONE : 0;
	w.best = FLT_MAX;
	w.cl.n = 0; w.cl.overflow = false;
	w.cl.t0 = w.cl.t1 = w.cl.t2 = 0.f; w.cl.tri0 = w.cl.tri1 = w.cl.tri2 = 0; w.cl.rank0 = w.cl.rank1 = w.cl.rank2 = 0;
	w.ln = w.lt = w.lr = 0;
}

__device__ __forceinline__ bool walk2_finished(const Walk2& w) { return w.cur == WALK_DONE; }

// The walk's stack lives in local memory: every push / pop is an LSU access whose lines also compete for L1 and,
// written through, for L2 (ncu r1n: 26% of the kernel's L1 wavefronts, 16% of its L2 requests).  MCPT_SHARED_STACK
// = N keeps the first N entries of every lane's stack in shared memory instead ([entry][thread], conflict-free,
// one wavefront per access); deeper entries stay in local memory.  Blocks of at most 128 threads.
#ifndef MCPT_SHARED_STACK
#define MCPT_SHARED_STACK 0
#endif
#ifndef MCPT_TOP_CACHE
#define MCPT_TOP_CACHE 0
#endif
#if MCPT_SHARED_STACK
#define MCPT_STACK_DECL __shared__ int sstack_[MCPT_SHARED_STACK * 128]; int* const sst_ = sstack_ + threadIdx.x;
#define MCPT_PUSH(v) do { if (sp < MCPT_SHARED_STACK) sst_[sp * 128] = (v); else w.stack[sp] = (v); ++sp; } while (0)
#define MCPT_POP() (--sp, sp < MCPT_SHARED_STACK ? sst_[sp * 128] : w.stack[sp])
#else
#define MCPT_STACK_DECL
#if MCPT_CHECKED
#define MCPT_STORE_(v) do { if (sp >= STACK_SIZE) { check_failed(CHK_STACK, sp, STACK_SIZE); sp = STACK_SIZE - 1; } w.stack[sp++] = (v); } while (0)
#else
#define MCPT_STORE_(v) (w.stack[sp++] = (v))
#endif
#if MCPT_TOP_CACHE
// the newest entry stays in a register (0 = none: the root is never pushed and leaf codes carry LEAF_FLAG): a push that
// finds the register free and the pop that follows it touch no memory at all - the pattern of every "near child is a
// leaf, far child waits" visit
#define MCPT_PUSH(v) do { if (top != 0) MCPT_STORE_(top); top = (v); } while (0)
#define MCPT_POP() (top != 0 ? take_top_(top) : w.stack[--sp])
#define MCPT_NONEMPTY ((top | sp) != 0)
#else
#define MCPT_PUSH(v) MCPT_STORE_(v)
#define MCPT_POP() (w.stack[--sp])
#endif
#endif
#ifndef MCPT_NONEMPTY
#define MCPT_NONEMPTY (sp > 0)
#endif
__device__ __forceinline__ int take_top_(int& top) { const int t = top; top = 0; return t; }

template <bool COUNT>
__device__ __forceinline__ void walk2_round(Walk2& w, const SceneDev& sc, const RayRegs& ray, float tmin) {
	MCPT_STACK_DECL
	const float rx = w.rx, ry = w.ry, rz = w.rz;
	int cur = w.cur, sp = w.sp;
#if MCPT_TOP_CACHE
	int top = w.top;
#endif
	float best = w.best;
	const float tminlo = tmin * (1.0f - 0x1p-19f);
	// per-ray constants are copied out of `w` (which lives in local memory because of its stack array): read
	// through `w` inside the loop they would be reloaded after every stack store
	const float c1x = w.c1x, c1y = w.c1y, c1z = w.c1z, c2x = w.c2x, c2y = w.c2y, c2z = w.c2z;
	// ---- inner nodes: conservative tests of both children, near child first
	while (cur >= 0 && cur != WALK_DONE) {
		MCPT_CHECK_INDEX(CHK_NODE, cur, sc.ntris - 1);
		const float4* p = reinterpret_cast<const float4*>(sc.fast_nodes + cur);
		float4 a, b, c, lk;
		ldg256_node(p, a, b);
		ldg256_node(p + 2, c, lk);
		const int2 link = make_int2(__float_as_int(lk.x), __float_as_int(lk.y));
		if (COUNT) ++w.ln;
		// plane distance as ONE fma(plane, 1/d, -o/d -+ S).  Against the reference's fl(fl(plane - o) / d) = T:
		// |t - T| <= 4.1 * 2^-24 |T| + 2.1 * 2^-24 |o/d|; the second (absolute) part is already inside c1/c2, the
		// relative part is the 2^-20 max(|tn|, |tf|) below.  The tmin and pruning tests take their relative slack
		// from constants (tminlo = tmin (1 - 2^-19); the 2^-7 pruning margin dwarfs 2^-20).
		const float tprune = __fmaf_rn(best, PRUNE_REL, PRUNE_ABS);
		float o1, o2, n0, n1, n2, f0, f1, f2;
		o1 = __fmaf_rn(a.x, rx, c1x); o2 = __fmaf_rn(a.w, rx, c2x); n0 = fminf(o1, o2); f0 = fmaxf(o1, o2);
		o1 = __fmaf_rn(a.y, ry, c1y); o2 = __fmaf_rn(b.x, ry, c2y); n1 = fminf(o1, o2); f1 = fmaxf(o1, o2);
		o1 = __fmaf_rn(a.z, rz, c1z); o2 = __fmaf_rn(b.y, rz, c2z); n2 = fminf(o1, o2); f2 = fmaxf(o1, o2);
		const float tnL = fmaxf(fmaxf(n0, n1), n2), tfL = fminf(fminf(f0, f1), f2);
		o1 = __fmaf_rn(b.z, rx, c1x); o2 = __fmaf_rn(c.y, rx, c2x); n0 = fminf(o1, o2); f0 = fmaxf(o1, o2);
		o1 = __fmaf_rn(b.w, ry, c1y); o2 = __fmaf_rn(c.z, ry, c2y); n1 = fminf(o1, o2); f1 = fmaxf(o1, o2);
		o1 = __fmaf_rn(c.x, rz, c1z); o2 = __fmaf_rn(c.w, rz, c2z); n2 = fminf(o1, o2); f2 = fmaxf(o1, o2);
		const float tnR = fmaxf(fmaxf(n0, n1), n2), tfR = fminf(fminf(f0, f1), f2);
		const float loL = __fmaf_rn(fmaxf(fabsf(tnL), fabsf(tfL)), -SLAB_SLACK, tnL);
		const float loR = __fmaf_rn(fmaxf(fabsf(tnR), fabsf(tfR)), -SLAB_SLACK, tnR);
		const bool hitL = !((tfL < loL) || (tfL < tminlo) || (tnL > tprune));
		const bool hitR = !((tfR < loR) || (tfR < tminlo) || (tnR > tprune));
		const int cl_ = link.x, cr_ = link.y;
		// one straight-line selection instead of four divergent arms (ncu r1n: the arms ran at 11 of 32 lanes)
		{
			const bool swap = loR < loL, both = hitL && hitR;
			const int nearc = (hitL && !(hitR && swap)) ? cl_ : cr_;
			if (both) MCPT_PUSH(swap ? cl_ : cr_);
			int nxt = nearc;
			if (!(hitL || hitR)) {
				nxt = WALK_DONE;
				if (MCPT_NONEMPTY) nxt = MCPT_POP();
			}
			cur = nxt;
		}
	}
	// ---- leaves: certified reference box test, then the certified triangle test
	while (cur < 0) {
		const int leaf = cur;
		cur = MCPT_NONEMPTY ? MCPT_POP() : WALK_DONE;
		int tri = leaf & 0x7fffffff;
		MCPT_CHECK_INDEX(CHK_LEAF, tri, sc.ntris);
		const float4* p = reinterpret_cast<const float4*>(sc.leaves + tri);
		float4 q0, q1;
		ldg256_leaf(p, q0, q1);
		if (COUNT) ++w.lr;
		if (leaf_box_certified(q0, q1, ray, rx, ry, rz, w.fastok, tmin, best * PRUNE_REL + PRUNE_ABS)) {
			if (COUNT) ++w.lt;
			int orig;
			TriTest r = leaf_triangle_certified(p, ray, w.fastok, tmin, best + 2.0f * WINDOW, orig);
			if (r.valid) {
				if (r.t < best) best = r.t;
				if (!(r.t - best > WINDOW)) cand_insert(w.cl, best, r.t, orig, __float_as_int(q0.w));
			}
		}
	}
	w.cur = cur; w.sp = sp; w.best = best;
#if MCPT_TOP_CACHE
	w.top = top;
#endif
}

// The reference's (t, triangle) from the finished walk; rays the fast walk cannot decide are re-traced.
template <bool COUNT>
__device__ __forceinline__ void walk2_resolve(Walk2& w, const SceneDev& sc, const RayRegs& ray, float tmin, HitRegs& hit,
                                              unsigned long long* cnt_nodes, unsigned long long* cnt_tris) {
	if (COUNT) { atomicAdd(cnt_nodes, (unsigned long long)w.ln); atomicAdd(cnt_tris, (unsigned long long)w.lt); atomicAdd(cnt_nodes + 5, (unsigned long long)w.lr); }
	hit = HitRegs{FLT_MAX, -1, 0, false};
	if (w.slow) {
		if (COUNT) atomicAdd(cnt_nodes + 6, 1ull);
		trace_reference<COUNT>(sc, ray, tmin, hit, cnt_nodes, cnt_tris);
		return;
	}
	if (w.cl.overflow || !cand_resolve(w.cl, w.best, hit)) {
		if (COUNT) atomicAdd(cnt_nodes + 7, 1ull);
		hit = HitRegs{FLT_MAX, -1, 0, false};
		trace_canonical<COUNT>(sc, ray, tmin, hit, cnt_nodes, cnt_tris);
	}
	hit.any = hit.closest >= 0;
#if MCPT_CHECKED
	if (hit.closest < -1 || hit.closest >= sc.ntris) { check_failed(CHK_HIT, hit.closest, sc.ntris); hit.closest = -1; hit.any = false; }
#endif
}

template <bool COUNT>
__device__ __forceinline__ void trace_ordered2(const SceneDev& sc, const RayRegs& ray, float tmin, HitRegs& hit,
                                               unsigned long long* cnt_nodes, unsigned long long* cnt_tris) {
	Walk2 w;
	walk2_init(w, sc, ray);
	while (!walk2_finished(w)) walk2_round<COUNT>(w, sc, ray, tmin);
	walk2_resolve<COUNT>(w, sc, ray, tmin, hit, cnt_nodes, cnt_tris);
}

}  // namespace mcpt
