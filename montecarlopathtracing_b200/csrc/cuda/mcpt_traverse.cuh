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

struct SceneDev {
	const mcpt_bvh_node* nodes_ref;
	const mcpt_triangle* tris_ref;
	const DNode* nodes;          // reference topology, two-box layout (canonical walk)
	const DNode* fast_nodes;     // SAH tree over the same leaves (fast walk)
	const DLeaf* leaves;
	const mcpt_material* materials;
	const float2* cs_table;
	int ntris, nnodes;
	int root_is_leaf;
};

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

template <bool COUNT>
__device__ __forceinline__ void trace_fast(const SceneDev& sc, const RayRegs& ray, float tmin, HitRegs& hit,
                                           unsigned long long* cnt_nodes, unsigned long long* cnt_tris) {
	if (ray.direction.x == 0.0f || ray.direction.y == 0.0f || ray.direction.z == 0.0f || sc.root_is_leaf) {
		trace_reference<COUNT>(sc, ray, tmin, hit, cnt_nodes, cnt_tris);
		return;
	}
	const float rx = 1.0f / ray.direction.x, ry = 1.0f / ray.direction.y, rz = 1.0f / ray.direction.z;
	int stack[STACK_SIZE];     // >= 0 inner node; LEAF_FLAG | triangle for a leaf
	int sp = 0;
	int cur = 0;
	float best = FLT_MAX;
	CandList cl;
	cl.n = 0; cl.overflow = false;
	cl.t0 = cl.t1 = cl.t2 = 0.f; cl.tri0 = cl.tri1 = cl.tri2 = 0; cl.rank0 = cl.rank1 = cl.rank2 = 0;
	bool any = false;
	unsigned ln = 0, lt = 0;
	const int DONE = 0x7fffffff;
	while (cur != DONE) {
		// ---- inner nodes: conservative tests of both children, near child first
		while (cur >= 0 && cur != DONE) {
			const float4* p = reinterpret_cast<const float4*>(sc.fast_nodes + cur);
			float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
			int4 link = __ldg(reinterpret_cast<const int4*>(p + 3));
			if (COUNT) ++ln;
			const float tprune = best * PRUNE_REL + PRUNE_ABS;
			// (bb - o) * (1/d): |computed - reference| <= 2^-22 |t|, covered 4x by the 2^-20 relative slack.  (A single
			// fused bb*(1/d) - o*(1/d) is cheaper but its error scales with |o/d|: rays nearly parallel to an axis
			// would then accept every box - measured 3x slower on random rays, so it is not used.)
			float o1, o2, n0, n1, n2, f0, f1, f2;
			o1 = (a.x - ray.origin.x) * rx; o2 = (a.w - ray.origin.x) * rx; n0 = fminf(o1, o2); f0 = fmaxf(o1, o2);
			o1 = (a.y - ray.origin.y) * ry; o2 = (b.x - ray.origin.y) * ry; n1 = fminf(o1, o2); f1 = fmaxf(o1, o2);
			o1 = (a.z - ray.origin.z) * rz; o2 = (b.y - ray.origin.z) * rz; n2 = fminf(o1, o2); f2 = fmaxf(o1, o2);
			float tnL = fmaxf(fmaxf(n0, n1), n2), tfL = fminf(fminf(f0, f1), f2);
			o1 = (b.z - ray.origin.x) * rx; o2 = (c.y - ray.origin.x) * rx; n0 = fminf(o1, o2); f0 = fmaxf(o1, o2);
			o1 = (b.w - ray.origin.y) * ry; o2 = (c.z - ray.origin.y) * ry; n1 = fminf(o1, o2); f1 = fmaxf(o1, o2);
			o1 = (c.x - ray.origin.z) * rz; o2 = (c.w - ray.origin.z) * rz; n2 = fminf(o1, o2); f2 = fmaxf(o1, o2);
			float tnR = fmaxf(fmaxf(n0, n1), n2), tfR = fminf(fminf(f0, f1), f2);
			const float sL = SLAB_SLACK * fmaxf(fmaxf(fabsf(tnL), fabsf(tfL)), tmin);
			const float sR = SLAB_SLACK * fmaxf(fmaxf(fabsf(tnR), fabsf(tfR)), tmin);
			const float loL = tnL - sL, loR = tnR - sR;
			// comparisons are false on NaN => not rejected (conservative)
			const bool hitL = !((tfL < loL) || (tfL < tmin - sL) || (loL > tprune));
			const bool hitR = !((tfR < loR) || (tfR < tmin - sR) || (loR > tprune));
			int cl_ = link.x < 0 ? (LEAF_FLAG | ~link.x) : link.x;
			int cr_ = link.y < 0 ? (LEAF_FLAG | ~link.y) : link.y;
			if (hitL && hitR) {
				const bool swap = loR < loL;
				const int nearc = swap ? cr_ : cl_, farc = swap ? cl_ : cr_;
				if (sp < STACK_SIZE) stack[sp++] = farc; else cl.overflow = true;
				cur = nearc;
			} else if (hitL) cur = cl_;
			else if (hitR) cur = cr_;
			else cur = sp > 0 ? stack[--sp] : DONE;
		}
		// ---- leaf: exact reference box test, then the exact triangle test
		if (cur != DONE) {
			const int tri = cur & 0x7fffffff;
			const float4* p = reinterpret_cast<const float4*>(sc.leaves + tri);
			float4 q0 = __ldg(p), q1 = __ldg(p + 1);
			if (COUNT) ++ln;
			float tn;
			if (intersect_aabb_exact(q0.x, q0.y, q0.z, q1.x, q1.y, q1.z, ray, tmin, tn) && !(tn > best * PRUNE_REL + PRUNE_ABS)) {
				if (COUNT) ++lt;
				TriTest r = leaf_triangle(sc.leaves, tri, ray, tmin);
				if (r.valid) {
					any = true;
					if (r.t < best) best = r.t;
					if (!(r.t - best > WINDOW)) cand_insert(cl, best, r.t, tri, __float_as_int(q0.w));
				}
			}
			cur = sp > 0 ? stack[--sp] : DONE;
		}
	}
	if (COUNT) { atomicAdd(cnt_nodes, (unsigned long long)ln); atomicAdd(cnt_tris, (unsigned long long)lt); }
	hit.any = any;
	if (cl.overflow || !cand_resolve(cl, best, hit)) {
		hit = HitRegs{FLT_MAX, -1, 0, false};
		trace_canonical<COUNT>(sc, ray, tmin, hit, cnt_nodes, cnt_tris);
	}
}

}  // namespace mcpt
