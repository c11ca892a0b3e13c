// NEE + MIS sampling mode - NOT a reference feature: the reference
// samples the BSDF only (shade.cl:114-154) and averages the non-zero samples (history.cl:10-12).  This mode
// estimates the SAME image with less variance:
//
//   reference pixel = E[X] / P(X != 0),  X = value of one BSDF-sampled path (0 unless it ends on an emitter
//   within MAX_DEPTH).
//
// * The BSDF chain is the reference's, unchanged and with the same LCG draws, so the indicator "X != 0" - the
//   denominator - is exactly the reference's.
// * Numerator: at every vertex where the reference drew its direction from randomDirection about the shading
//   normal (diffuse, and the diffuse lobe of glossy), the emitted radiance arriving DIRECTLY is estimated by
//   two techniques combined with the balance heuristic: (a) the chain's own next hit (density p_ref, see
//   random_direction_pdf), (b) a point sampled uniformly by area on the emitter triangles (solid-angle density
//   q = d^2 / (|cos_l| A)).  The reference's per-bounce factor is w_ref = Kd cos / (2 pi) (shade.cl:120) applied
//   to directions of density p_ref, i.e. it integrates g = p_ref * w_ref; technique (b) therefore contributes
//   T * g(w) Le / (p_ref + q) and technique (a) T * w_ref Le * p_ref / (p_ref + q).
// * The glossy lobe (shade.cl:125-147) is treated the same way: its directions have density
//   p_lobe = p_ref(cos(d, r)) / A(angle(r, n)) above the surface (r = mirror direction, A = acceptance probability of
//   the reference's rejection loop, tabulated: phong_accept_probability) and carry the weight
//   w = ks pow(cos(d, r), Ns) cos(d, n) / (2 pi) (shade.cl:141-143).  Glass is a delta BSDF: no NEE through it.
// * A shadow ray counts only if its CLOSEST hit (same tie semantics as everywhere: the product walk) is the
//   sampled triangle itself, which is what "first emitter seen in that direction" means to the chain.
// * Emitters radiate from both sides and without a cosine at the light, like shade.cl:155-158.
// Output: per pixel sum of the numerator over ALL samples and the count of samples with X != 0 - the same
// seven planes as the reference-faithful mode (mcpt_read_sums); image = sum / count.  hist/count (the running
// mean of history.cl) cannot express a ratio of sums and are left untouched.
// * MCPT_SAMPLING_NEE_MIS_RATIO estimates the DENOMINATOR the same way.  The raw count is a sum of Bernoulli(P) draws
//   (relative noise sqrt((1 - P) / (N P)), P = 2 % on cbox) and dominates the image noise once the numerator is
//   well estimated.  P(X != 0) = sum over vertices of P(the next hit is an emitter and the sample value is non-zero),
//   an integral of the same form with Le replaced by the indicator [|X(w)| != 0] (history.cl:10, evaluated on the
//   value the chain would have produced in that direction): technique (a) adds the MIS weight whenever the chain
//   reaches an emitter with X != 0, technique (b) adds indicator * p / (p + q) for unoccluded emitter samples.  Plane
//   6 then holds the sum of these estimates (not an integer); the chain - and therefore what (a) sees - is unchanged.
#pragma once
#include "mcpt_kernels.cuh"

namespace mcpt {

struct MisDev {
	float4* rad;            // per path: numerator so far (rgb), w = 1 once the chain has reached an emitter with X != 0
	float* pdf[2];          // per queue item: p_ref of the direction that produced it when NEE was done at the previous vertex, else 0
	float4* sod;            // shadow rays of the current bounce: origin
	float4* sdp;            //   direction + path id
	float4* scon;           //   contribution if unoccluded (rgb) + sampled emitter triangle
	uint32_t* keys;         // coherence keys of the NEXT bounce's queue items (null: queues stay unsorted), as in WaveDev
	uint32_t* skeys;        //   ... and of this bounce's shadow rays
	float* den;             // ratio mode, per path: estimate of P(X != 0) so far (null in MCPT_SAMPLING_NEE_MIS)
	float* sden;            // ratio mode, per shadow ray: its share of that estimate if unoccluded
	float2* shit;           //   hit record written by the walk
	uint32_t* scount;       // scount[d] = shadow rays emitted at bounce d
	const int32_t* emit_tri;
	const float* phong_accept;   // acceptance probability of the glossy lobe's rejection loop (mcpt_device.cuh)
	const float* emit_cdf;  // inclusive, normalised
	int nemit;
	float emit_area;
};

struct QueueTraceJob {
	const float4* od;
	const float4* dp;
	float2* hitrec;
	const uint32_t* perm;   // walk order (coherence-sorted queue) or null
	uint32_t perm_n;        // sorted entries (WaveDev::perm_n)
	__device__ __forceinline__ bool load(unsigned& item, RayRegs& ray) {
		if (perm && item < perm_n) item = __ldcs(perm + item);
		const unsigned i = item;
		const float4 o = __ldcs(od + i), d = __ldcs(dp + i);
		ray.origin = ld4(o);
		ray.direction = f4{d.x, d.y, d.z, 0.0f};
		return true;
	}
	__device__ __forceinline__ void finish(unsigned i, const RayRegs&, const HitRegs& h) {
		__stcs(hitrec + i, make_float2(h.t, __int_as_float(h.closest)));
	}
};

__global__ void __launch_bounds__(128, MCPT_MIN_BLOCKS) k_trace_queue_persistent(SceneDev sc, const float4* od, const float4* dp, float2* hitrec,
                                                                const uint32_t* count, float tmin, unsigned* head, const uint32_t* perm, uint32_t perm_n) {
	QueueTraceJob job{od, dp, hitrec, perm, perm_n};
	persistent_walk(sc, tmin, *count, head, job);
}

__device__ __forceinline__ float u01(uint32_t& h) {      // 24-bit uniform in [0,1) from a hash chain (NOT the path's LCG)
	h = mix32(h + 0x9E3779B9u);
	return float(h >> 8) * (1.0f / 16777216.0f);
}

__global__ void __launch_bounds__(256) k_wave_shade_mis(SceneDev sc, WaveDev w, MisDev m, int depth, int max_depth) {
	const int src = depth & 1, dst = src ^ 1;
	const uint32_t total = w.qcount[depth];
	const unsigned lane = threadIdx.x & 31u;
	const uint32_t stride = gridDim.x * blockDim.x;
	for (uint32_t base = (blockIdx.x * blockDim.x + threadIdx.x) & ~31u; base < total; base += stride) {
		const uint32_t i = base + lane;
		bool alive = false, shadow = false;
		RayRegs ray;
		f4 color{0, 0, 0, 0};
		uint32_t seed = 0;
		int path = 0;
		float pdf_next = 0.0f;
		float4 s_o, s_d, s_c;
		float s_den = 0.0f;
		if (i < total) {
			const float4 o = __ldcs(w.od[src] + i), d = __ldcs(w.dp[src] + i), c = __ldcs(w.cs[src] + i);
			const float2 hr = __ldcs(w.hitrec + i);
			const float pdf_prev = __ldcs(m.pdf[src] + i);
			const float t = hr.x;
			const int closest = __float_as_int(hr.y);
			path = __float_as_int(d.w);
			ray.origin = ld4(o);
			ray.direction = f4{d.x, d.y, d.z, __int_as_float(path % w.npix)};
			color = f4{c.x, c.y, c.z, 0.0f};
			seed = uint32_t(__float_as_int(c.w));
			const f4 in_dir = ray.direction;
			const f4 T_before = color;
			f4 nn{0, 0, 0, 0}, ip{0, 0, 0, 0}, ng{0, 0, 0, 0};
			uint32_t mat = 0;
			int mtype = 0;
			if (closest >= 0) {
				const float4* tp = reinterpret_cast<const float4*>(sc.leaves + closest);
				float4 q1 = __ldg(tp + 1), q2 = __ldg(tp + 2), q3 = __ldg(tp + 3), q4 = __ldg(tp + 4);
				ng = f4{q2.w, q3.w, q4.w, 0.0f};
				nn = ng;
				mat = uint32_t(__float_as_int(q1.w));
				mtype = sc.materials[mat].type;
				ip = ray.origin + t * ray.direction;
				if (dot3(ray.direction, nn) > 0) nn = -nn;
			}
			uint32_t nee_rng = seed ^ 0xA511E9B3u;            // before the chain advances the LCG
			bool diffuse_lobe = false;
			f4 phong_axis;
			alive = shade_one(sc.materials, sc.cs_table, ray, t, nn, ip, mat, color, seed, max_depth, &diffuse_lobe, &phong_axis);
			const bool phong_lobe = phong_axis.w != 0.0f;
			const float accept = phong_lobe ? phong_accept_probability(m.phong_accept, dot3(phong_axis, nn)) : 1.0f;
			float4 rad = m.rad[path];
			if (closest >= 0 && mtype == MCPT_LIGHT) {
				// technique (a): the chain reached an emitter; `color` is the reference's sample value X
				float wgt = 1.0f;
				if (pdf_prev > 0.0f) {
					const float cl = fabsf(dot3(in_dir, ng));
					const float q = cl > 0.0f ? t * t / (cl * m.emit_area) : FLT_MAX;
					wgt = pdf_prev / (pdf_prev + q);
				}
				rad.x += color.x * wgt; rad.y += color.y * wgt; rad.z += color.z * wgt;
				if (sqrtf(dot4(color, color)) != 0.0f) {                   // history.cl:10 - the reference's own test (squares may underflow)
					rad.w = 1.0f;
					if (m.den) m.den[path] += wgt;
				}
				m.rad[path] = rad;
			}
			if (!alive) {
				if (m.den) rad.w = m.den[path];                            // ratio mode: plane 6 accumulates the estimate of P(X != 0)
				__stcs(w.sample_out + path, rad);
			} else if (diffuse_lobe || phong_lobe) {
				const f4 axis = phong_lobe ? f4{phong_axis.x, phong_axis.y, phong_axis.z, 0.0f} : nn;
				pdf_next = random_direction_pdf(dot3(ray.direction, axis)) / accept;
				if (m.nemit > 0) {
					// technique (b): one emitter point, uniform by area
					const float r0 = u01(nee_rng), r1 = u01(nee_rng), r2 = u01(nee_rng);
					int lo = 0, hi = m.nemit - 1;
					while (lo < hi) { const int mid = (lo + hi) >> 1; if (__ldg(m.emit_cdf + mid) > r0) hi = mid; else lo = mid + 1; }
					const int etri = __ldg(m.emit_tri + lo);
					const float4* ep = reinterpret_cast<const float4*>(sc.leaves + etri);
					const float4 e1q = __ldg(ep + 1), e2q = __ldg(ep + 2), e3q = __ldg(ep + 3), e4q = __ldg(ep + 4);
					const float su = sqrtf(r1), b1 = su * (1.0f - r2), b2 = su * r2;      // v0 + b1 (v1 - v0) + b2 (v2 - v0); record holds -(v1-v0), -(v2-v0)
					const f4 y{e2q.x - b1 * e3q.x - b2 * e4q.x, e2q.y - b1 * e3q.y - b2 * e4q.y, e2q.z - b1 * e3q.z - b2 * e4q.z, 0.0f};
					f4 wv = y - ip;
					wv.w = 0.0f;
					const float d2 = dot3(wv, wv);
					if (d2 > 0.0f) {
						const float dist = sqrtf(d2);
						const f4 wdir{wv.x / dist, wv.y / dist, wv.z / dist, 0.0f};
						const float cs_ = dot3(wdir, nn);
						const float cl = fabsf(wdir.x * e2q.w + wdir.y * e3q.w + wdir.z * e4q.w);
						const float ca = dot3(wdir, axis);                 // the lobe's own polar cosine (= cs_ on the diffuse lobe)
						if (cs_ > 0.0f && cl > 0.0f && ca > 0.0f) {
							const float p = random_direction_pdf(ca) / accept;
							const float q = d2 / (cl * m.emit_area);
							const MaterialDev sm = load_material(sc.materials, mat);
							const MaterialDev em = load_material(sc.materials, uint32_t(__float_as_int(e1q.w)));
							float k = cs_ / TWO_PI_F;
							f4 refl = sm.kd;
							if (phong_lobe) {                                // shade.cl:141-143 (ks shares the union slot with ka)
								refl = sm.ka_ks;
								k *= powf(ca, sm.Ns);
							}
							// X = the value the chain's sample would have had in this direction, then the balance weight
							const f4 X{T_before.x * refl.x * k * em.ka_ks.x, T_before.y * refl.y * k * em.ka_ks.y, T_before.z * refl.z * k * em.ka_ks.z, 0.0f};
							const float wb = p / (p + q);
							s_c = make_float4(X.x * wb, X.y * wb, X.z * wb, __int_as_float(etri));
							if (m.den && sqrtf(dot4(X, X)) != 0.0f) s_den = wb;
							if (s_c.x != 0.0f || s_c.y != 0.0f || s_c.z != 0.0f || s_den != 0.0f) {
								const f4 so = ip + KERNEL_EPSILON * wdir;      // the offset the reference gives its bounce rays (shade.cl:116)
								s_o = make_float4(so.x, so.y, so.z, 0.0f);
								s_d = make_float4(wdir.x, wdir.y, wdir.z, __int_as_float(path));
								shadow = true;
							}
						}
					}
				}
			}
		}
		unsigned mask = __ballot_sync(0xffffffffu, alive);
		if (mask) {
			uint32_t qb = 0;
			if (lane == 0) qb = atomicAdd(&w.qcount[depth + 1], uint32_t(__popc(mask)));
			qb = __shfl_sync(0xffffffffu, qb, 0);
			if (alive) {
				const uint32_t pos = qb + uint32_t(__popc(mask & ((1u << lane) - 1u)));
				__stcs(w.od[dst] + pos, st4(ray.origin));
				__stcs(w.dp[dst] + pos, make_float4(ray.direction.x, ray.direction.y, ray.direction.z, __int_as_float(path)));
				__stcs(w.cs[dst] + pos, make_float4(color.x, color.y, color.z, __int_as_float(int(seed))));
				__stcs(m.pdf[dst] + pos, pdf_next);
				if (m.keys) m.keys[pos] = ray_sort_key(sc, ray);
			}
		}
		mask = __ballot_sync(0xffffffffu, shadow);
		if (mask) {
			uint32_t qb = 0;
			if (lane == 0) qb = atomicAdd(&m.scount[depth], uint32_t(__popc(mask)));
			qb = __shfl_sync(0xffffffffu, qb, 0);
			if (shadow) {
				const uint32_t pos = qb + uint32_t(__popc(mask & ((1u << lane) - 1u)));
				__stcs(m.sod + pos, s_o);
				__stcs(m.sdp + pos, s_d);
				__stcs(m.scon + pos, s_c);
				if (m.sden) __stcs(m.sden + pos, s_den);
				if (m.skeys) {
					RayRegs sr;
					sr.origin = f4{s_o.x, s_o.y, s_o.z, 0.0f};
					sr.direction = f4{s_d.x, s_d.y, s_d.z, 0.0f};
					m.skeys[pos] = ray_sort_key(sc, sr);
				}
			}
		}
	}
}

// Shadow rays whose closest hit is the sampled emitter triangle add their contribution to the path.
__global__ void __launch_bounds__(256) k_wave_nee_add(MisDev m, int depth) {
	const uint32_t total = m.scount[depth];
	for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
		const float2 hr = __ldcs(m.shit + i);
		const float4 c = __ldcs(m.scon + i);
		if (__float_as_int(hr.y) != __float_as_int(c.w)) continue;
		const int path = __float_as_int(__ldcs(m.sdp + i).w);
		float4 r = m.rad[path];
		r.x += c.x; r.y += c.y; r.z += c.z;
		m.rad[path] = r;
		if (m.den) m.den[path] += __ldcs(m.sden + i);
	}
}

// Paths still alive after the last bounce cannot exist (the depth cut terminates them), so every sample has been
// written to sample_out.  Sums: numerator over ALL samples, count = samples whose chain reached an emitter.
__global__ void __launch_bounds__(256) k_wave_resolve_mis(WaveDev w, MisDev m, int nsamples, int max_depth) {
	int p = blockIdx.x * blockDim.x + threadIdx.x;
	if (p == 0) {
		unsigned long long traced = 0;
		for (int d = 0; d < max_depth; ++d) traced += (unsigned long long)w.qcount[d] + m.scount[d];
		*w.rays_traced += traced;
		*w.paths_done += (unsigned long long)nsamples * (unsigned long long)w.npix;
		*w.first_sample += uint32_t(nsamples);
	}
	if (p >= w.npix) return;
	// continue from the stored sums one sample at a time: independent of the wave split (see k_wave_resolve)
	const size_t n = size_t(w.npix);
	double s0 = w.sums[p], s1 = w.sums[n + p], s2 = w.sums[2 * n + p];
	double q0 = w.sums[3 * n + p], q1 = w.sums[4 * n + p], q2 = w.sums[5 * n + p], k = w.sums[6 * n + p];
	for (int s = 0; s < nsamples; ++s) {
		const float4 c = w.sample_out[size_t(s) * w.npix + p];
		s0 += c.x; s1 += c.y; s2 += c.z;
		q0 += double(c.x) * c.x; q1 += double(c.y) * c.y; q2 += double(c.z) * c.z;
		k += c.w;
	}
	w.sums[p] = s0; w.sums[n + p] = s1; w.sums[2 * n + p] = s2;
	w.sums[3 * n + p] = q0; w.sums[4 * n + p] = q1; w.sums[5 * n + p] = q2;
	w.sums[6 * n + p] = k;
}

}  // namespace mcpt
