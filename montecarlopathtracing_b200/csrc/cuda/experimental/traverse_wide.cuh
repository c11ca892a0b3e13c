// EXPERIMENTAL, measured slower than the product walk on B200 and kept only so that the numbers can be reproduced
// (DESIGN.md §5): MCPT_TRAVERSAL_WIDE8, the product walk's candidate / replay logic on 8-wide nodes with 8-bit quantised
// child boxes (after Ylitie, Karras, Laine 2017).  Bit-exact like the other walks (tests/test_gpu_parity.py), 25 % slower
// than MCPT_TRAVERSAL_FAST: the decode lands on the ALU pipe and the leaf phase runs at 2.5 of 32 lanes.
// Included by mcpt_kernels.cuh after mcpt_traverse.cuh; nothing in the product path (mcpt_render, FAST) reaches it.
// Other rejected experiments (32-byte grid-quantised nodes, postponed leaves, the non-fused slab test, four-armed child
// selection) were removed from the sources after measurement; their results are in DESIGN.md §5 and the git history.
#pragma once
#include "../mcpt_traverse.cuh"
#include "accel_wide.hpp"

namespace mcpt {

// ------------------------------------------------------------------ 8-wide compressed traversal (experimental)
// Same candidate / replay logic as trace_ordered2; inner tests run on 8-bit quantised child boxes that
// contain the true child boxes (outward rounding, verified in exact arithmetic at build time).  Plane
// distances are one FFMA each: t = q * (scale/d) + (p - o)/d.  With A = (p-o)*r, B = scale*r the computed
// value is within 2^-22 (|A| + 255 |B|) of the exact distance to the quantised plane; every near plane is
// moved down and every far plane up by s = 2^-20 (|A| + 256 |B|) PER AXIS (a per-axis slack cannot blow
// up the way a per-ray |o/d| term would: an axis the ray is parallel to has huge planes AND a huge slack,
// but they only matter when the origin is within 1e-6 of that slab's boundary).
constexpr int WSTACK = 32;

template <bool COUNT>
__device__ __forceinline__ void trace_wide(const SceneDev& sc, const RayRegs& ray, float tmin, HitRegs& hit,
                                           unsigned long long* cnt_nodes, unsigned long long* cnt_tris) {
	if (ray.direction.x == 0.0f || ray.direction.y == 0.0f || ray.direction.z == 0.0f || sc.root_is_leaf) {
		trace_reference<COUNT>(sc, ray, tmin, hit, cnt_nodes, cnt_tris);
		return;
	}
	// 2^23 * scale / d must stay finite for the byte -> float trick below: steeper rays use the binary walk
	if (!(fabsf(ray.direction.x) > 1e-20f && fabsf(ray.direction.y) > 1e-20f && fabsf(ray.direction.z) > 1e-20f)) {
		trace_ordered2<COUNT>(sc, ray, tmin, hit, cnt_nodes, cnt_tris);
		return;
	}
	const float rx = 1.0f / ray.direction.x, ry = 1.0f / ray.direction.y, rz = 1.0f / ray.direction.z;
	const bool fastok = fabsf(ray.direction.x) > 1e-30f && fabsf(ray.direction.y) > 1e-30f && fabsf(ray.direction.z) > 1e-30f;
	const bool posx = ray.direction.x >= 0.0f, posy = ray.direction.y >= 0.0f, posz = ray.direction.z >= 0.0f;
	const unsigned octinv = (posx ? 1u : 0u) | (posy ? 2u : 0u) | (posz ? 4u : 0u);
	uint2 stack[WSTACK];
	int sp = 0;
	uint2 ng = make_uint2(0u, 0x80000000u);       // virtual group holding the root (imask 0 => relative index 0)
	float best = FLT_MAX;
	CandList cl;
	cl.n = 0; cl.overflow = false;
	cl.t0 = cl.t1 = cl.t2 = 0.f; cl.tri0 = cl.tri1 = cl.tri2 = 0; cl.rank0 = cl.rank1 = cl.rank2 = 0;
	bool any = false;
	unsigned ln = 0, lt = 0, lr = 0;
	for (;;) {
		uint2 tg = make_uint2(0u, 0u);
		if (ng.y > 0x00FFFFFFu) {
			// ---- next inner child of the current group (highest bit first = nearest octant first)
			const unsigned bit = 31u - unsigned(__clz(int(ng.y)));
			const unsigned gmask = ng.y & 0xFFu;
			ng.y &= ~(1u << bit);
			if (ng.y > 0x00FFFFFFu) { if (sp < WSTACK) stack[sp++] = ng; else cl.overflow = true; }
			const unsigned slot = (bit - 24u) ^ octinv;
			const unsigned rel = unsigned(__popc(gmask & ((1u << slot) - 1u)));
			const float4* p = reinterpret_cast<const float4*>(sc.wnodes + (ng.x + rel));
			const float4 n0 = __ldg(p), n1 = __ldg(p + 1), n2 = __ldg(p + 2), n3 = __ldg(p + 3), n4 = __ldg(p + 4);
			if (COUNT) ++ln;
			const unsigned ew = __float_as_uint(n0.w);
			const float sx = __uint_as_float((ew & 0xFFu) << 23), sy = __uint_as_float(((ew >> 8) & 0xFFu) << 23),
			            sz = __uint_as_float(((ew >> 16) & 0xFFu) << 23);
			const float Bx = sx * rx, By = sy * ry, Bz = sz * rz;
			const float Ax = (n0.x - ray.origin.x) * rx, Ay = (n0.y - ray.origin.y) * ry, Az = (n0.z - ray.origin.z) * rz;
			// A byte q becomes the float 2^23 + q with ONE permute (0x4B0000qq); the 2^23 * B is folded into the
			// constant term, whose rounding costs at most half a quantisation step |B|/2 -> one extra |B| of slack.
			const float slx = SLAB_SLACK * (fabsf(Ax) + 256.0f * fabsf(Bx)) + fabsf(Bx), sly = SLAB_SLACK * (fabsf(Ay) + 256.0f * fabsf(By)) + fabsf(By),
			            slz = SLAB_SLACK * (fabsf(Az) + 256.0f * fabsf(Bz)) + fabsf(Bz);
			const float Mx = Ax - 8388608.0f * Bx, My = Ay - 8388608.0f * By, Mz = Az - 8388608.0f * Bz;
			const float Anx = Mx - slx, Afx = Mx + slx, Any = My - sly, Afy = My + sly, Anz = Mz - slz, Afz = Mz + slz;
			// near / far quantised planes per axis, chosen by the ray octant
			const unsigned qlx0 = __float_as_uint(n2.x), qlx1 = __float_as_uint(n2.y), qly0 = __float_as_uint(n2.z), qly1 = __float_as_uint(n2.w);
			const unsigned qlz0 = __float_as_uint(n3.x), qlz1 = __float_as_uint(n3.y), qhx0 = __float_as_uint(n3.z), qhx1 = __float_as_uint(n3.w);
			const unsigned qhy0 = __float_as_uint(n4.x), qhy1 = __float_as_uint(n4.y), qhz0 = __float_as_uint(n4.z), qhz1 = __float_as_uint(n4.w);
			const unsigned nx[2] = {posx ? qlx0 : qhx0, posx ? qlx1 : qhx1}, fx[2] = {posx ? qhx0 : qlx0, posx ? qhx1 : qlx1};
			const unsigned ny[2] = {posy ? qly0 : qhy0, posy ? qly1 : qhy1}, fy[2] = {posy ? qhy0 : qly0, posy ? qhy1 : qly1};
			const unsigned nz[2] = {posz ? qlz0 : qhz0, posz ? qlz1 : qhz1}, fz[2] = {posz ? qhz0 : qlz0, posz ? qhz1 : qlz1};
			const unsigned meta[2] = {__float_as_uint(n1.z), __float_as_uint(n1.w)};
			const float tprune = best * PRUNE_REL + PRUNE_ABS;
			unsigned hitmask = 0u;
#pragma unroll
			for (int c = 0; c < 8; ++c) {
				const int w = c >> 2, sh = (c & 3) * 8;
				const float tnx = __fmaf_rn(__uint_as_float(__byte_perm(nx[w], 0x4B000000u, 0x7440u | unsigned(c & 3))), Bx, Anx);
				const float tny = __fmaf_rn(__uint_as_float(__byte_perm(ny[w], 0x4B000000u, 0x7440u | unsigned(c & 3))), By, Any);
				const float tnz = __fmaf_rn(__uint_as_float(__byte_perm(nz[w], 0x4B000000u, 0x7440u | unsigned(c & 3))), Bz, Anz);
				const float tfx = __fmaf_rn(__uint_as_float(__byte_perm(fx[w], 0x4B000000u, 0x7440u | unsigned(c & 3))), Bx, Afx);
				const float tfy = __fmaf_rn(__uint_as_float(__byte_perm(fy[w], 0x4B000000u, 0x7440u | unsigned(c & 3))), By, Afy);
				const float tfz = __fmaf_rn(__uint_as_float(__byte_perm(fz[w], 0x4B000000u, 0x7440u | unsigned(c & 3))), Bz, Afz);
				const float tn = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, tmin));
				const float tf = fminf(fminf(tfx, tfy), fminf(tfz, tprune));
				const unsigned m = (meta[w] >> sh) & 0xFFu;
				if (!(tn > tf)) {                           // false on NaN => kept (conservative)
					unsigned idx = m & 31u;
					if (idx >= 24u) idx ^= octinv;
					hitmask |= (m >> 5) << idx;
				}
			}
			ng = make_uint2(__float_as_uint(n1.x), (hitmask & 0xFF000000u) | (ew >> 24));
			tg = make_uint2(__float_as_uint(n1.y), hitmask & 0x00FFFFFFu);
		}
		// ---- leaf records of this node: exact reference box test, then the exact triangle test
		while (tg.y != 0u) {
			const unsigned bit = unsigned(__ffs(int(tg.y))) - 1u;
			tg.y &= tg.y - 1u;
			const float4* p = reinterpret_cast<const float4*>(sc.wleaves + (tg.x + bit));
			const float4 q0 = __ldg(p), q1 = __ldg(p + 1);
			if (COUNT) ++lr;
			if (leaf_box_certified(q0, q1, ray, rx, ry, rz, fastok, tmin, best * PRUNE_REL + PRUNE_ABS)) {
				if (COUNT) ++lt;
				int orig;
				TriTest r = leaf_triangle_certified(p, ray, fastok, tmin, best + 2.0f * WINDOW, orig);
				if (r.valid) {
					any = true;
					if (r.t < best) best = r.t;
					if (!(r.t - best > WINDOW)) cand_insert(cl, best, r.t, orig, __float_as_int(q0.w));
				}
			}
		}
		if (ng.y <= 0x00FFFFFFu) {
			if (sp == 0) break;
			ng = stack[--sp];
		}
	}
	if (COUNT) { atomicAdd(cnt_nodes, (unsigned long long)ln); atomicAdd(cnt_tris, (unsigned long long)lt); atomicAdd(cnt_nodes + 5, (unsigned long long)lr); }
	hit.any = any;
	if (cl.overflow || !cand_resolve(cl, best, hit)) {
		hit = HitRegs{FLT_MAX, -1, 0, false};
		trace_canonical<COUNT>(sc, ray, tmin, hit, cnt_nodes, cnt_tris);
	}
}


}  // namespace mcpt
