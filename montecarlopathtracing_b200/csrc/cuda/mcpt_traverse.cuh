// BVH traversal on the device.
//
//  trace_reference : intersectObjects (objdef.h:240-275) as written — 64-byte BVHNode records, one
//                    box test per visited node, left-first DFS with a 64-entry stack, no pruning.
//  trace_fast      : the same canonical left-first LEAF ORDER on a derived layout.  Why this is
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

#include "mcpt_device.cuh"

namespace mcpt {

// Derived node: both children's boxes + child references.  ref >= 0: inner node index; ref < 0: ~triangle.
struct __align__(16) DNode {
	float lminx, lminy, lminz, lmaxx;
	float lmaxy, lmaxz, rminx, rminy;
	float rminz, rmaxx, rmaxy, rmaxz;
	int32_t left, right, pad0, pad1;
};
static_assert(sizeof(DNode) == 64, "DNode");

// Derived triangle: everything the exact test needs, 4 x 16 B.
struct __align__(16) DTri {
	float v0x, v0y, v0z; int32_t material;
	float e1x, e1y, e1z, nx;       // e1 = -(v1 - v0)
	float e2x, e2y, e2z, ny;       // e2 = -(v2 - v0)
	float cr0, cr1, cr2, nz;       // ray-independent 2x2 determinants of (e1, e2)
};
static_assert(sizeof(DTri) == 64, "DTri");

struct SceneDev {
	const mcpt_bvh_node* nodes_ref;
	const mcpt_triangle* tris_ref;
	const DNode* nodes;
	const DTri* tris;
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

__device__ __forceinline__ void test_dtri(const DTri* __restrict__ tris, int tri, const RayRegs& ray, float tmin, HitRegs& hit) {
	const float4* p = reinterpret_cast<const float4*>(tris + tri);
	float4 q0 = __ldg(p), q1 = __ldg(p + 1), q2 = __ldg(p + 2), q3 = __ldg(p + 3);
	TriTest r = intersect_triangle(ray, q0.x, q0.y, q0.z, q1.x, q1.y, q1.z, q2.x, q2.y, q2.z, q3.x, q3.y, q3.z, q1.w,
	                               q2.w, q3.w, tmin);
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
				test_dtri(sc.tris, link.y, ray, tmin, hit);
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
__device__ __forceinline__ void trace_fast(const SceneDev& sc, const RayRegs& ray, float tmin, HitRegs& hit,
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
				test_dtri(sc.tris, ~left, ray, tmin, hit);
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
					test_dtri(sc.tris, ~right, ray, tmin, hit);
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
			test_dtri(sc.tris, ~e, ray, tmin, hit);
			next = -1;
		}
		if (next == -2) break;
		ni = next;
	}
	if (COUNT) { atomicAdd(cnt_nodes, (unsigned long long)ln); atomicAdd(cnt_tris, (unsigned long long)lt); }
}

}  // namespace mcpt
