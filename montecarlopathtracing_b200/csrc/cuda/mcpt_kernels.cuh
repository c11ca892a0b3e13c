// Kernels of the path-tracing hot path.
//
// Stage kernels (k_generate_rays, k_intersect, k_shade, k_history, k_display) are one-to-one with the
// reference's five OpenCL kernels and work on the reference's buffer layouts (Ray 48 B, Hit 48 B,
// colour float4, seed u32): they are the drop-in surface behind RayGeneration/SceneBase/ColorOut.
//
// Wavefront kernels (k_wave_generate, k_wave_bounce, k_wave_resolve) are the product path used by
// mcpt_render: path state lives in three float4 streams that are compacted at every bounce
// (warp ballot + prefix popcount + one atomic per warp), traversal and shading are fused so no Hit
// record is ever materialised, and finished samples are folded into the accumulators in sample order,
// which makes the running mean bit-identical to history.cl's frame-by-frame update.
#pragma once

#include "mcpt.h"
#include "mcpt_traverse.cuh"
#include "experimental/traverse_wide.cuh"   // MCPT_TRAVERSAL_WIDE8 (measured slower; not on the product path)

namespace mcpt {

struct FrameDev {               // reference-layout per-pixel buffers (stage path)
	mcpt_ray* rays;
	mcpt_hit* hits;
	int32_t* closest;
	float4* color;
	uint32_t* seeds;
	float4* hist;
	int32_t* counts;
	double* sums;               // 7 planes of n doubles
	int n;
};

template <int TRAVERSAL, bool COUNT>
__device__ __forceinline__ void trace_dispatch(const SceneDev& sc, const RayRegs& ray, float tmin, HitRegs& h, unsigned long long* counters) {
	if (TRAVERSAL == MCPT_TRAVERSAL_REFERENCE) trace_reference<COUNT>(sc, ray, tmin, h, counters, counters + 1);
	else if (TRAVERSAL == MCPT_TRAVERSAL_CANONICAL) trace_canonical<COUNT>(sc, ray, tmin, h, counters, counters + 1);
	else if (TRAVERSAL == MCPT_TRAVERSAL_WIDE8) trace_wide<COUNT>(sc, ray, tmin, h, counters, counters + 1);
	else trace_ordered2<COUNT>(sc, ray, tmin, h, counters, counters + 1);
}

__device__ __forceinline__ RayRegs load_ray(const mcpt_ray* rays, long i) {
	const float4* p = reinterpret_cast<const float4*>(rays + i);
	float4 o = p[0], d = p[1];
	return {ld4(o), ld4(d)};
}
__device__ __forceinline__ void store_ray(mcpt_ray* rays, long i, const RayRegs& r) {
	float4* p = reinterpret_cast<float4*>(rays + i);
	p[0] = st4(r.origin);
	p[1] = st4(r.direction);
	// Ray.ratio is never written by the reference kernels
}

// ------------------------------------------------------------------ K1
__global__ void k_generate_rays(CameraDev cam, mcpt_ray* rays) {
	uint32_t x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
	if (x >= cam.width || y >= cam.height) return;
	RayRegs r = generate_ray(cam, x, y);
	store_ray(rays, long(y) * cam.width + x, r);
}

__global__ void k_fill_color(float4* color, int n, float v) {
	int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) color[i] = make_float4(v, v, v, v);
}

// ------------------------------------------------------------------ K2  (intersect.cl:1-28)
// Writes the 48-byte Hit of the reference; `closest` (optional) gets the triangle whose data is in it.
template <int TRAVERSAL, bool COUNT>
__global__ void __launch_bounds__(256) k_intersect(SceneDev sc, const mcpt_ray* __restrict__ rays, long n, float tmin,
                                                    mcpt_hit* __restrict__ hits, int32_t* __restrict__ closest,
                                                    unsigned long long* counters) {
	long i = long(blockIdx.x) * blockDim.x + threadIdx.x;
	bool live = false;
	if (i < n) {
		RayRegs ray = load_ray(rays, i);
		uint32_t state = uint32_t(__float_as_int(ray.origin.w));
		if (!(state & MCPT_RAY_TERM_MASK)) {
			live = true;
			HitRegs h{FLT_MAX, -1, 0, false};
			trace_dispatch<TRAVERSAL, COUNT>(sc, ray, tmin, h, counters);
			float4 nrm = make_float4(0, 0, 0, 0), pt = make_float4(0, 0, 0, 0);
			uint32_t mat = 0;
			if (h.closest >= 0) {
				const float4* tp = reinterpret_cast<const float4*>(sc.leaves + h.closest);
				float4 q1 = __ldg(tp + 1), q2 = __ldg(tp + 2), q3 = __ldg(tp + 3), q4 = __ldg(tp + 4);
				f4 nn{q2.w, q3.w, q4.w, 0.0f};
				mat = uint32_t(__float_as_int(q1.w));
				pt = st4(ray.origin + h.t * ray.direction);            // 4 lanes, like the float4 expression at objdef.h:218
				if (h.any && dot3(ray.direction, nn) > 0) nn = -nn;    // intersect.cl:23-25
				nrm = st4(nn);
			}
			float4* hp = reinterpret_cast<float4*>(hits + i);
			hp[0] = nrm;
			hp[1] = pt;
			uint32_t tid = TRAVERSAL == MCPT_TRAVERSAL_REFERENCE ? uint32_t(h.last_touched)
			                                                     : uint32_t(h.closest >= 0 ? h.closest : 0);
			hp[2] = make_float4(h.t, __int_as_float(int(tid)), __int_as_float(int(mat)), __int_as_float(0));
			if (closest) closest[i] = h.closest;
		} else if (closest) {
			closest[i] = -1;
		}
	}
	if (COUNT) {      // live-ray counter in counters[2]
		unsigned m = __ballot_sync(0xffffffffu, live);
		if ((threadIdx.x & 31) == 0 && m) atomicAdd(counters + 2, (unsigned long long)__popc(m));
	}
}

// ------------------------------------------------------------------ K3  (shade.cl:75-206)
__global__ void __launch_bounds__(128) k_shade(SceneDev sc, mcpt_ray* rays, const mcpt_hit* __restrict__ hits,
                                                float4* color, uint32_t* seeds, int n, int max_depth) {
	int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	RayRegs ray = load_ray(rays, i);
	if (uint32_t(__float_as_int(ray.origin.w)) & MCPT_RAY_TERM_MASK) return;     // :89
	const float4* hp = reinterpret_cast<const float4*>(hits + i);
	float4 nrm = hp[0], pt = hp[1], misc = hp[2];
	f4 c = ld4(color[i]);
	uint32_t seed = seeds[i];
	shade_one(sc.materials, sc.cs_table, ray, misc.x, ld4(nrm), ld4(pt), uint32_t(__float_as_int(misc.z)), c, seed, max_depth);
	color[i] = st4(c);
	seeds[i] = seed;
	store_ray(rays, i, ray);
}

// ------------------------------------------------------------------ K4  (history.cl:3-28) + sums
__global__ void k_history(float4* color, float4* hist, int32_t* counts, double* sums, int n, int max_attempt) {
	int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	f4 c = ld4(color[i]), h = ld4(hist[i]);
	int32_t cnt = counts[i];
	f4 raw = c;
	if (history_one(c, h, cnt, max_attempt)) {
		sums[i] += raw.x; sums[size_t(n) + i] += raw.y; sums[2 * size_t(n) + i] += raw.z;
		sums[3 * size_t(n) + i] += double(raw.x) * raw.x;
		sums[4 * size_t(n) + i] += double(raw.y) * raw.y;
		sums[5 * size_t(n) + i] += double(raw.z) * raw.z;
		sums[6 * size_t(n) + i] += 1.0;
	}
	color[i] = st4(c);
	hist[i] = st4(h);
	counts[i] = cnt;
}

// ------------------------------------------------------------------ K5  (testkernel.cl:9-14)
__global__ void k_display(const float4* __restrict__ src, float4* dst, int n) {
	int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	float4 c = src[i];
	const float g = 1 / 2.2f;
	dst[i] = make_float4(cr_pow(c.x, g), cr_pow(c.y, g), cr_pow(c.z, g), 0.0f);
}

// ================================================================== wavefront path
struct WaveDev {
	float4* od[2];          // origin.xyz + state word          (double buffered)
	float4* dp[2];          // direction.xyz + path id
	float4* cs[2];          // throughput rgb + LCG state
	float4* sample_out;     // finished samples, [s * npix + pixel] = (r, g, b, 0)
	float2* hitrec;         // split pipeline: (t, closest triangle) of queue item i, written by the trace kernel
	uint32_t* qcount;       // qcount[d] = live paths entering bounce d
	uint32_t* first_sample; // device-resident global sample index of the wave's first sample
	unsigned long long* rays_traced;
	unsigned long long* paths_done;
	uint32_t* seeds;        // per-pixel persistent LCG state (seed_mode 0)
	uint32_t* sort_keys;    // coherence key of queue item i of the NEXT bounce, written by the shade kernel (null: queues stay unsorted)
	const uint32_t* perm;   // walk order of THIS bounce's queue: the walk kernel's i-th ray is item perm[i] (null: identity)
	uint32_t perm_n;        // entries of perm that were sorted (an estimate of the live count made on the host): they are a
	                        // permutation of [0, perm_n) whenever the queue is longer, so items from perm_n on walk in queue order
	uint32_t capacity;      // items every queue buffer holds (checked build: bound of queue positions and path ids)
	float4* hist;
	int32_t* counts;
	double* sums;
	int npix;
};

__global__ void __launch_bounds__(256) k_wave_generate(CameraDev cam, WaveDev w, int nsamples, int seed_mode, uint32_t run_seed) {
	int i = blockIdx.x * blockDim.x + threadIdx.x;
	int total = nsamples * w.npix;
	if (i == 0) w.qcount[0] = uint32_t(total);
	if (i >= total) return;
	int s = i / w.npix, p = i - s * w.npix;
	RayRegs r = generate_ray(cam, uint32_t(p) % cam.width, uint32_t(p) / cam.width);
	uint32_t seed = seed_mode == 0 ? w.seeds[p] : sample_seed(uint32_t(p), *w.first_sample + uint32_t(s), run_seed);
	w.od[0][i] = st4(r.origin);
	w.dp[0][i] = make_float4(r.direction.x, r.direction.y, r.direction.z, __int_as_float(i));
	w.cs[0][i] = make_float4(1.0f, 1.0f, 1.0f, __int_as_float(int(seed)));
}

// One bounce for every live path: closest hit + shade + (terminate | compact into the next queue).
template <int TRAVERSAL, bool COUNT>
__global__ void __launch_bounds__(256) k_wave_bounce(SceneDev sc, WaveDev w, int depth, int max_depth, float tmin, int seed_mode,
                                                      unsigned long long* counters) {
	const int src = depth & 1, dst = src ^ 1;
	const uint32_t total = w.qcount[depth];
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (blockIdx.x * blockDim.x >= total) return;          // whole block idle (uniform exit)
	bool alive = false;
	RayRegs ray;
	f4 color;
	uint32_t seed = 0;
	int path = 0;
	if (i < total) {
		float4 o = w.od[src][i], d = w.dp[src][i], c = w.cs[src][i];
		path = __float_as_int(d.w);
		const int pixel = path % w.npix;
		ray.origin = ld4(o);
		ray.direction = f4{d.x, d.y, d.z, __int_as_float(pixel)};
		color = f4{c.x, c.y, c.z, 0.0f};
		seed = uint32_t(__float_as_int(c.w));
		HitRegs h{FLT_MAX, -1, 0, false};
		trace_dispatch<TRAVERSAL, COUNT>(sc, ray, tmin, h, counters);
		f4 nn{0, 0, 0, 0}, ip{0, 0, 0, 0};
		uint32_t mat = 0;
		if (h.closest >= 0) {
			const float4* tp = reinterpret_cast<const float4*>(sc.leaves + h.closest);
			float4 q1 = __ldg(tp + 1), q2 = __ldg(tp + 2), q3 = __ldg(tp + 3), q4 = __ldg(tp + 4);
			nn = f4{q2.w, q3.w, q4.w, 0.0f};
			mat = uint32_t(__float_as_int(q1.w));
			ip = ray.origin + h.t * ray.direction;
			if (dot3(ray.direction, nn) > 0) nn = -nn;
		}
		alive = shade_one(sc.materials, sc.cs_table, ray, h.t, nn, ip, mat, color, seed, max_depth);
		if (!alive) {
			w.sample_out[path] = make_float4(color.x, color.y, color.z, 0.0f);
			if (seed_mode == 0) w.seeds[pixel] = seed;
		}
	}
	// queue compaction: ballot + prefix popcount, one atomic per warp
	const unsigned mask = __ballot_sync(0xffffffffu, alive);
	if (mask) {
		const int lane = threadIdx.x & 31;
		uint32_t base = 0;
		if (lane == 0) base = atomicAdd(&w.qcount[depth + 1], uint32_t(__popc(mask)));
		base = __shfl_sync(0xffffffffu, base, 0);
		if (alive) {
			const uint32_t pos = base + uint32_t(__popc(mask & ((1u << lane) - 1u)));
			w.od[dst][pos] = st4(ray.origin);
			w.dp[dst][pos] = make_float4(ray.direction.x, ray.direction.y, ray.direction.z, __int_as_float(path));
			w.cs[dst][pos] = make_float4(color.x, color.y, color.z, __int_as_float(int(seed)));
		}
	}
}

// Fold the wave's finished samples into the accumulators, in sample order (history.cl semantics).
__global__ void __launch_bounds__(256) k_wave_resolve(WaveDev w, int nsamples, int max_depth, int max_attempt) {
	int p = blockIdx.x * blockDim.x + threadIdx.x;
	if (p == 0) {
		unsigned long long traced = 0;
		for (int d = 0; d < max_depth; ++d) traced += w.qcount[d];
		*w.rays_traced += traced;
		*w.paths_done += (unsigned long long)nsamples * (unsigned long long)w.npix;
		*w.first_sample += uint32_t(nsamples);
	}
	if (p >= w.npix) return;
	f4 h = ld4(w.hist[p]);
	int32_t cnt = w.counts[p];
	// The running sums continue from their stored values, one sample at a time: ((prev + x1) + x2) + ..., so the
	// doubles do not depend on how a render is cut into waves (and equal a frame-by-frame accumulation).
	const size_t n = size_t(w.npix);
	double s0 = 0, s1 = 0, s2 = 0, q0 = 0, q1 = 0, q2 = 0, k = 0;
	bool touched = false;
	for (int s = 0; s < nsamples; ++s) {
		f4 c = ld4(w.sample_out[size_t(s) * w.npix + p]);
		f4 raw = c;
		if (history_one(c, h, cnt, max_attempt)) {
			if (!touched) {
				s0 = w.sums[p]; s1 = w.sums[n + p]; s2 = w.sums[2 * n + p];
				q0 = w.sums[3 * n + p]; q1 = w.sums[4 * n + p]; q2 = w.sums[5 * n + p];
				k = w.sums[6 * n + p];
				touched = true;
			}
			s0 += raw.x; s1 += raw.y; s2 += raw.z;
			q0 += double(raw.x) * raw.x; q1 += double(raw.y) * raw.y; q2 += double(raw.z) * raw.z;
			k += 1.0;
		}
	}
	w.hist[p] = st4(h);
	w.counts[p] = cnt;
	if (touched) {
		w.sums[p] = s0; w.sums[n + p] = s1; w.sums[2 * n + p] = s2;
		w.sums[3 * n + p] = q0; w.sums[4 * n + p] = q1; w.sums[5 * n + p] = q2;
		w.sums[6 * n + p] = k;
	}
}

__global__ void k_init_seeds(uint32_t* seeds, int n, uint32_t run_seed) {
	int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) seeds[i] = pixel_seed(uint32_t(i), run_seed);
}

}  // namespace mcpt

// ================================================================== coherence order of the bounce queues
// The walk is bound by L1 data-pipe wavefronts (ncu r2a: 90 % of peak): a divergent warp pays one wavefront per lane and
// load instruction, lanes that fetch the same node share one.  Bounce rays leave the shader in an arbitrary order, so
// the queue is walked in the order of a 23-bit key - direction on a 16 x 16 octahedral grid (high bits), origin cell on
// a 32^3 Morton grid (low bits) - found by tools/sort_probe.py to be the best of 40 candidates (x1.56 on the walk of
// diningroom bounce rays, x1.6 on veach_mis).  Only the ORDER of the walk changes: the shader reads and writes the
// queue by position as before, so results are bit-identical by construction.
namespace mcpt {
constexpr int SORT_KEY_BITS = 24;        // three 8-bit radix passes
constexpr uint32_t SORT_KEY_INVALID = 0xFFFFFFFFu;
__device__ __forceinline__ uint32_t spread5(uint32_t v) {       // 5 bits -> every third bit
	v = (v | (v << 8)) & 0x0000100Fu;
	v = (v | (v << 4)) & 0x000010C3u;
	v = (v | (v << 2)) & 0x00001249u;
	return v;
}
__device__ __forceinline__ uint32_t ray_sort_key(const SceneDev& sc, const RayRegs& ray) {
	const float cx = (ray.origin.x - sc.key_lo[0]) * sc.key_scale[0], cy = (ray.origin.y - sc.key_lo[1]) * sc.key_scale[1],
	            cz = (ray.origin.z - sc.key_lo[2]) * sc.key_scale[2];
	const uint32_t qx = uint32_t(fminf(fmaxf(cx, 0.0f), 31.0f)), qy = uint32_t(fminf(fmaxf(cy, 0.0f), 31.0f)),
	               qz = uint32_t(fminf(fmaxf(cz, 0.0f), 31.0f));
	const uint32_t cell = spread5(qx) | (spread5(qy) << 1) | (spread5(qz) << 2);
	// octahedral map of the direction: (u, v) in [-1, 1]^2
	const float dx = ray.direction.x, dy = ray.direction.y, dz = ray.direction.z;
	const float inv = rcp_approx(fabsf(dx) + fabsf(dy) + fabsf(dz) + 1e-30f);
	float u = dx * inv, v = dy * inv;
	if (dz < 0.0f) {
		const float uu = copysignf(1.0f - fabsf(v), u), vv = copysignf(1.0f - fabsf(u), v);
		u = uu; v = vv;
	}
	const uint32_t qu = uint32_t(fminf(fmaxf(u * 8.0f + 8.0f, 0.0f), 15.0f)), qv = uint32_t(fminf(fmaxf(v * 8.0f + 8.0f, 0.0f), 15.0f));
	return (((qu << 4) | qv) << 15) | cell;
}

// keys[count_next, count_now) <- invalid: those positions held live rays one bounce ago.  Everything from count_now on
// is invalid already (induction over the bounces; bounce 0 covers the whole wave), so a sort of any prefix of the key
// array puts exactly the live items first.
__global__ void __launch_bounds__(256) k_invalidate_keys(uint32_t* __restrict__ keys, const uint32_t* __restrict__ count_next,
                                                         const uint32_t* __restrict__ count_now) {
	const uint32_t lo = *count_next, hi = *count_now;
	for (uint32_t i = lo + blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += gridDim.x * blockDim.x) keys[i] = SORT_KEY_INVALID;
}

__global__ void k_iota(uint32_t* p, uint32_t n) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) p[i] = i;
}
}  // namespace mcpt

// ================================================================== persistent kernels (MCPT_TRAVERSAL_FAST)
// Rays of very different traversal length share a warp; with one thread per ray the warp runs until its
// longest ray is done (measured: 13 of 32 lanes active in the inner loop on incoherent rays).  Here each
// warp keeps pulling work: whenever fewer than REFILL_THRESHOLD lanes are still walking, the idle lanes
// first FINISH their completed rays together (hit record / shading / queue compaction - also the divergent
// part, so batching it matters) and then fetch new work items with one atomic per warp.
namespace mcpt {

#ifndef MCPT_MIN_BLOCKS
#define MCPT_MIN_BLOCKS 8        // resident 128-thread blocks per SM the persistent kernels are compiled for (8 -> 64 registers)
#endif
#ifndef MCPT_REFILL_THRESHOLD
#define MCPT_REFILL_THRESHOLD 16       // lanes still walking below which a warp refills; with coherence-sorted queues 16 measured +0.6 % over 22 (26: -0.4 %, 30: -0.9 %; profiles/r2l_ab_walk_variants.log)
#endif

// Job: bool load(unsigned& item, RayRegs&) - false when the item needs no trace (the job has dealt with it); may
//          replace the work index by the queue position it stands for (sorted queues);
//      void finish(unsigned item, const RayRegs&, const HitRegs&)
template <class Job>
__device__ __forceinline__ void persistent_walk(const SceneDev& sc, float tmin, unsigned total, unsigned* head, Job& job) {
	const unsigned FULL = 0xffffffffu;
	const unsigned lane = threadIdx.x & 31u;
	const unsigned lt_mask = (1u << lane) - 1u;
	Walk2 w;
	RayRegs ray;
	unsigned item = 0;
	bool working = false, have_result = false, exhausted = false;
	for (;;) {
		unsigned wm = __ballot_sync(FULL, working);
		if (__popc(wm) < MCPT_REFILL_THRESHOLD) {
			if (have_result) {
				HitRegs h;
				walk2_resolve<false>(w, sc, ray, tmin, h, nullptr, nullptr);
				job.finish(item, ray, h);
				have_result = false;
			}
			if (!exhausted) {
				bool want = !working;
				unsigned m = __ballot_sync(FULL, want);
				while (m) {                                    // warp-uniform loop
					const unsigned n = unsigned(__popc(m));
					unsigned base = 0;
					if (lane == unsigned(__ffs(int(m))) - 1u) base = atomicAdd(head, n);
					base = __shfl_sync(FULL, base, __ffs(int(m)) - 1);
					if (want) {
						const unsigned i = base + unsigned(__popc(m & lt_mask));
						if (i < total) {
							item = i;
							if (job.load(item, ray)) {
								walk2_init(w, sc, ray);
								working = true;
								want = false;
							}
						} else {
							want = false;
						}
					}
					if (base + n >= total) exhausted = true;   // base, n are warp-uniform
					m = exhausted ? 0u : __ballot_sync(FULL, want);
				}
			}
			wm = __ballot_sync(FULL, working);
			if (wm == 0u) break;
		}
		if (working) {
			walk2_round<false>(w, sc, ray, tmin);
			if (walk2_finished(w)) { working = false; have_result = true; }
		}
	}
}

struct IntersectJob {
	const SceneDev& sc;
	const mcpt_ray* rays;
	mcpt_hit* hits;
	int32_t* closest;
	__device__ __forceinline__ bool load(unsigned& item, RayRegs& ray) {
		const unsigned i = item;
		ray = load_ray(rays, long(i));
		if (uint32_t(__float_as_int(ray.origin.w)) & MCPT_RAY_TERM_MASK) {      // intersect.cl:16
			if (closest) closest[i] = -1;
			return false;
		}
		return true;
	}
	__device__ __forceinline__ void finish(unsigned i, const RayRegs& ray, const HitRegs& h) {
		float4 nrm = make_float4(0, 0, 0, 0), pt = make_float4(0, 0, 0, 0);
		uint32_t mat = 0;
		if (h.closest >= 0) {
			const float4* tp = reinterpret_cast<const float4*>(sc.leaves + h.closest);
			float4 q1 = __ldg(tp + 1), q2 = __ldg(tp + 2), q3 = __ldg(tp + 3), q4 = __ldg(tp + 4);
			f4 nn{q2.w, q3.w, q4.w, 0.0f};
			mat = uint32_t(__float_as_int(q1.w));
			pt = st4(ray.origin + h.t * ray.direction);
			if (h.any && dot3(ray.direction, nn) > 0) nn = -nn;
			nrm = st4(nn);
		}
		float4* hp = reinterpret_cast<float4*>(hits + i);
		hp[0] = nrm;
		hp[1] = pt;
		hp[2] = make_float4(h.t, __int_as_float(h.closest >= 0 ? h.closest : 0), __int_as_float(int(mat)), __int_as_float(0));
		if (closest) closest[i] = h.closest;
	}
};

__global__ void __launch_bounds__(128, MCPT_MIN_BLOCKS) k_intersect_persistent(SceneDev sc, const mcpt_ray* __restrict__ rays, unsigned n, float tmin,
                                                              mcpt_hit* __restrict__ hits, int32_t* __restrict__ closest, unsigned* head) {
	IntersectJob job{sc, rays, hits, closest};
	persistent_walk(sc, tmin, n, head, job);
}

struct BounceJob {
	const SceneDev& sc;
	const WaveDev& w;
	int depth, max_depth, seed_mode, src, dst;
	// per-lane path state carried from load to finish
	f4 color;
	uint32_t seed;
	int path;
	__device__ __forceinline__ bool load(unsigned& item, RayRegs& ray) {
		const unsigned i = item;
		float4 o = w.od[src][i], d = w.dp[src][i], c = w.cs[src][i];
		path = __float_as_int(d.w);
		ray.origin = ld4(o);
		ray.direction = f4{d.x, d.y, d.z, __int_as_float(path % w.npix)};
		color = f4{c.x, c.y, c.z, 0.0f};
		seed = uint32_t(__float_as_int(c.w));
		return true;
	}
	__device__ __forceinline__ void finish(unsigned, RayRegs ray, const HitRegs& h) {
		f4 nn{0, 0, 0, 0}, ip{0, 0, 0, 0};
		uint32_t mat = 0;
		if (h.closest >= 0) {
			const float4* tp = reinterpret_cast<const float4*>(sc.leaves + h.closest);
			float4 q1 = __ldg(tp + 1), q2 = __ldg(tp + 2), q3 = __ldg(tp + 3), q4 = __ldg(tp + 4);
			nn = f4{q2.w, q3.w, q4.w, 0.0f};
			mat = uint32_t(__float_as_int(q1.w));
			ip = ray.origin + h.t * ray.direction;
			if (dot3(ray.direction, nn) > 0) nn = -nn;
		}
		const bool alive = shade_one(sc.materials, sc.cs_table, ray, h.t, nn, ip, mat, color, seed, max_depth);
		if (!alive) {
			w.sample_out[path] = make_float4(color.x, color.y, color.z, 0.0f);
			if (seed_mode == 0) w.seeds[path % w.npix] = seed;
		} else {
			// queue compaction among the lanes that reach this point together: one atomic per group
			const unsigned m = __activemask();
			const unsigned lane = threadIdx.x & 31u;
			const int leader = __ffs(int(m)) - 1;
			uint32_t base = 0;
			if (int(lane) == leader) base = atomicAdd(&w.qcount[depth + 1], uint32_t(__popc(m)));
			base = __shfl_sync(m, base, leader);
			const uint32_t pos = base + uint32_t(__popc(m & ((1u << lane) - 1u)));
			w.od[dst][pos] = st4(ray.origin);
			w.dp[dst][pos] = make_float4(ray.direction.x, ray.direction.y, ray.direction.z, __int_as_float(path));
			w.cs[dst][pos] = make_float4(color.x, color.y, color.z, __int_as_float(int(seed)));
		}
	}
};

// ---- split pipeline: the persistent kernel only walks (so its warps never run the divergent shader with a
// handful of lanes: measured 8 of 32 lanes active there when fused), a second, fully populated kernel shades
// and compacts.  Costs 8 B written + 40 B re-read per ray of extra HBM traffic, on a kernel that uses 3% of it.
struct TraceJob {
	const WaveDev& w;
	int src;
	unsigned total;         // items in this bounce's queue (bound of the permutation's entries in the checked build)
	__device__ __forceinline__ bool load(unsigned& item, RayRegs& ray) {
		if (w.perm && item < w.perm_n) item = __ldcs(w.perm + item);   // walk order -> queue position; finish() writes the record there
		MCPT_CHECK_INDEX(CHK_PERM, item, total);
		const unsigned i = item;
		const float4 o = __ldcs(w.od[src] + i), d = __ldcs(w.dp[src] + i);
		ray.origin = ld4(o);
		ray.direction = f4{d.x, d.y, d.z, 0.0f};
		return true;
	}
	__device__ __forceinline__ void finish(unsigned i, const RayRegs&, const HitRegs& h) {
		__stcs(w.hitrec + i, make_float2(h.t, __int_as_float(h.closest)));
	}
};

__global__ void __launch_bounds__(128, MCPT_MIN_BLOCKS) k_wave_trace_persistent(SceneDev sc, WaveDev w, int depth, float tmin, unsigned* head) {
	const unsigned total = w.qcount[depth];
	TraceJob job{w, depth & 1, total};
	persistent_walk(sc, tmin, total, head, job);
}

// (62 registers: 4 blocks per SM.  Launch bounds for 5 / 6 blocks - 48 / 40 registers, 16 / 64 B of spills - measured
// -1 % / -3 %, profiles/r3v_shade_variants.log.)
__global__ void __launch_bounds__(256) k_wave_shade(SceneDev sc, WaveDev w, int depth, int max_depth, int seed_mode) {
	const int src = depth & 1, dst = src ^ 1;
	const uint32_t total = w.qcount[depth];
	const unsigned lane = threadIdx.x & 31u;
	const uint32_t stride = gridDim.x * blockDim.x;
	for (uint32_t bbase = blockIdx.x * blockDim.x; bbase < total; bbase += stride) {   // block-uniform trip count
		const uint32_t i = bbase + threadIdx.x;
		bool alive = false, finished = false;
		RayRegs ray;
		f4 color{0, 0, 0, 0};
		uint32_t seed = 0;
		int path = 0;
		if (i < total) {
			const float4 o = __ldcs(w.od[src] + i), d = __ldcs(w.dp[src] + i), c = __ldcs(w.cs[src] + i);
			const float2 hr = __ldcs(w.hitrec + i);
			const float t = hr.x;
			const int closest = __float_as_int(hr.y);
			path = __float_as_int(d.w);
			MCPT_CHECK_INDEX(CHK_PATH, path, w.capacity);
			const int pixel = path % w.npix;
			ray.origin = ld4(o);
			ray.direction = f4{d.x, d.y, d.z, __int_as_float(pixel)};
			color = f4{c.x, c.y, c.z, 0.0f};
			seed = uint32_t(__float_as_int(c.w));
			f4 nn{0, 0, 0, 0}, ip{0, 0, 0, 0};
			uint32_t mat = 0;
			if (closest >= 0) {
				const float4* tp = reinterpret_cast<const float4*>(sc.leaves + closest);
				float4 q1 = __ldg(tp + 1), q2 = __ldg(tp + 2), q3 = __ldg(tp + 3), q4 = __ldg(tp + 4);
				nn = f4{q2.w, q3.w, q4.w, 0.0f};
				mat = uint32_t(__float_as_int(q1.w));
				ip = ray.origin + t * ray.direction;
				if (dot3(ray.direction, nn) > 0) nn = -nn;
			}
			alive = shade_one(sc.materials, sc.cs_table, ray, t, nn, ip, mat, color, seed, max_depth);
			finished = !alive;
		}
		if (finished) {
			__stcs(w.sample_out + path, make_float4(color.x, color.y, color.z, 0.0f));
			if (seed_mode == 0) w.seeds[path % w.npix] = seed;
		}
		const unsigned mask = __ballot_sync(0xffffffffu, alive);
		if (mask) {
			uint32_t qb = 0;
			if (lane == 0) qb = atomicAdd(&w.qcount[depth + 1], uint32_t(__popc(mask)));
			qb = __shfl_sync(0xffffffffu, qb, 0);
			if (alive) {
				uint32_t pos = qb + uint32_t(__popc(mask & ((1u << lane) - 1u)));
				MCPT_CHECK_INDEX(CHK_QUEUE, pos, w.capacity);
				__stcs(w.od[dst] + pos, st4(ray.origin));
				__stcs(w.dp[dst] + pos, make_float4(ray.direction.x, ray.direction.y, ray.direction.z, __int_as_float(path)));
				__stcs(w.cs[dst] + pos, make_float4(color.x, color.y, color.z, __int_as_float(int(seed))));
				if (w.sort_keys) w.sort_keys[pos] = ray_sort_key(sc, ray);
			}
		}
	}
}

__global__ void __launch_bounds__(128, MCPT_MIN_BLOCKS) k_wave_bounce_persistent(SceneDev sc, WaveDev w, int depth, int max_depth, float tmin, int seed_mode,
                                                                unsigned* head) {
	const unsigned total = w.qcount[depth];
	BounceJob job{sc, w, depth, max_depth, seed_mode, depth & 1, (depth & 1) ^ 1, f4{0, 0, 0, 0}, 0u, 0};
	persistent_walk(sc, tmin, total, head, job);
}

// ---- path regeneration for seed_mode 0 (one LCG stream per pixel, continued from sample to sample, shade.cl:101).
// A pixel's samples are sequential by construction, so a frame-at-a-time wave has one sample of every pixel in flight
// and its late bounces run almost empty (cbox 1024^2: 1 M rays at bounce 0, 0.4 M at bounce 7).  Here a pixel whose
// path ends folds the sample into hist / counts / sums at once (history.cl's update, the same per-pixel order) and
// starts its next sample in the same launch, so every round traces one ray per unfinished pixel, whatever the
// depths of the individual paths.  Round r reads queue r & 1 and writes queue (r + 1) & 1; qcount is a ring of
// per-round counters prepared by the host.  Results are those of the frame loop bit for bit.
__global__ void __launch_bounds__(256) k_wave_shade_regen(SceneDev sc, WaveDev w, CameraDev cam, int round, int max_depth, int max_attempt,
                                                          int32_t* __restrict__ frames_left) {
	const int src = round & 1, dst = src ^ 1;
	const uint32_t total = w.qcount[round];
	const unsigned lane = threadIdx.x & 31u;
	const uint32_t stride = gridDim.x * blockDim.x;
	if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(w.rays_traced, (unsigned long long)total);   // pixel groups run concurrently
	const size_t n = size_t(w.npix);
	for (uint32_t bbase = blockIdx.x * blockDim.x; bbase < total; bbase += stride) {   // block-uniform trip count
		const uint32_t i = bbase + threadIdx.x;
		bool alive = false;
		RayRegs ray;
		f4 color{0, 0, 0, 0};
		uint32_t seed = 0;
		int pixel = 0;
		if (i < total) {
			const float4 o = __ldcs(w.od[src] + i), d = __ldcs(w.dp[src] + i), c = __ldcs(w.cs[src] + i);
			const float2 hr = __ldcs(w.hitrec + i);
			const float t = hr.x;
			const int closest = __float_as_int(hr.y);
			pixel = __float_as_int(d.w);
			ray.origin = ld4(o);
			ray.direction = f4{d.x, d.y, d.z, __int_as_float(pixel)};
			color = f4{c.x, c.y, c.z, 0.0f};
			seed = uint32_t(__float_as_int(c.w));
			f4 nn{0, 0, 0, 0}, ip{0, 0, 0, 0};
			uint32_t mat = 0;
			if (closest >= 0) {
				const float4* tp = reinterpret_cast<const float4*>(sc.leaves + closest);
				float4 q1 = __ldg(tp + 1), q2 = __ldg(tp + 2), q3 = __ldg(tp + 3), q4 = __ldg(tp + 4);
				nn = f4{q2.w, q3.w, q4.w, 0.0f};
				mat = uint32_t(__float_as_int(q1.w));
				ip = ray.origin + t * ray.direction;
				if (dot3(ray.direction, nn) > 0) nn = -nn;
			}
			alive = shade_one(sc.materials, sc.cs_table, ray, t, nn, ip, mat, color, seed, max_depth);
			if (!alive) {
				// the sample is complete: history.cl for this pixel, then its next sample (if any)
				f4 h = ld4(w.hist[pixel]);
				int32_t cnt = w.counts[pixel];
				f4 cfold = color;
				if (history_one(cfold, h, cnt, max_attempt)) {
					w.sums[pixel] += color.x; w.sums[n + pixel] += color.y; w.sums[2 * n + pixel] += color.z;
					w.sums[3 * n + pixel] += double(color.x) * color.x; w.sums[4 * n + pixel] += double(color.y) * color.y;
					w.sums[5 * n + pixel] += double(color.z) * color.z;
					w.sums[6 * n + pixel] += 1.0;
				}
				w.hist[pixel] = st4(h);
				w.counts[pixel] = cnt;
				w.seeds[pixel] = seed;
				const int32_t left = frames_left[pixel] - 1;
				frames_left[pixel] = left;
				if (left > 0) {
					ray = generate_ray(cam, uint32_t(pixel) % cam.width, uint32_t(pixel) / cam.width);
					color = f4{1.0f, 1.0f, 1.0f, 0.0f};
					alive = true;
				}
			}
		}
		const unsigned mask = __ballot_sync(0xffffffffu, alive);
		if (mask) {
			uint32_t qb = 0;
			if (lane == 0) qb = atomicAdd(&w.qcount[round + 1], uint32_t(__popc(mask)));
			qb = __shfl_sync(0xffffffffu, qb, 0);
			if (alive) {
				const uint32_t pos = qb + uint32_t(__popc(mask & ((1u << lane) - 1u)));
				__stcs(w.od[dst] + pos, st4(ray.origin));
				__stcs(w.dp[dst] + pos, make_float4(ray.direction.x, ray.direction.y, ray.direction.z, __int_as_float(pixel)));
				__stcs(w.cs[dst] + pos, make_float4(color.x, color.y, color.z, __int_as_float(int(seed))));
			}
		}
	}
}

// first sample of pixels [p0, p0 + count): queue position j holds pixel p0 + j
__global__ void __launch_bounds__(256) k_regen_start(CameraDev cam, WaveDev w, int p0, int count) {
	const int j = blockIdx.x * blockDim.x + threadIdx.x;
	if (j == 0) w.qcount[0] = uint32_t(count);
	if (j >= count) return;
	const int p = p0 + j;
	RayRegs r = generate_ray(cam, uint32_t(p) % cam.width, uint32_t(p) / cam.width);
	w.od[0][j] = st4(r.origin);
	w.dp[0][j] = make_float4(r.direction.x, r.direction.y, r.direction.z, __int_as_float(p));
	w.cs[0][j] = make_float4(1.0f, 1.0f, 1.0f, __int_as_float(int(w.seeds[p])));
}

__global__ void k_add_stats(unsigned long long* paths_done, unsigned long long paths, uint32_t* first_sample, uint32_t frames) {
	*paths_done += paths;
	*first_sample += frames;
}

__global__ void k_fill_i32(int32_t* p, int n, int32_t v) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) p[i] = v;
}

// ---- seed_mode 0 without launch boundaries: one work item = one PIXEL, all `frames` samples of it.  A lane keeps its
// pixel from ray to ray - walk, shade, then either the bounce ray or, when the path ended, the fold into hist /
// counts / sums (history.cl, the pixel's own sample order) and the next sample's primary ray - and only fetches
// another pixel when this one has no samples left.  No queues, no per-bounce launches, so none of the ~100 us a launch
// waits for its last warp batch; the price is the shader running inside the walk kernel's refill phase
// at partial lane occupancy.  Bit-identical to the frame loop: every pixel sees the same sequence of operations.
template <class Job>
__device__ __forceinline__ void persistent_walk_chain(const SceneDev& sc, float tmin, unsigned total, unsigned* head, Job& job) {
	const unsigned FULL = 0xffffffffu;
	const unsigned lane = threadIdx.x & 31u;
	const unsigned lt_mask = (1u << lane) - 1u;
	Walk2 w;
	RayRegs ray;
	bool working = false, have_result = false, exhausted = false;
	for (;;) {
		unsigned wm = __ballot_sync(FULL, working);
		if (__popc(wm) < MCPT_REFILL_THRESHOLD) {
			if (have_result) {
				HitRegs h;
				walk2_resolve<false>(w, sc, ray, tmin, h, nullptr, nullptr);
				have_result = false;
				if (job.next(ray, h)) {                        // the pixel goes on with `ray`
					walk2_init(w, sc, ray);
					working = true;
				}
			}
			if (!exhausted) {
				bool want = !working;
				unsigned m = __ballot_sync(FULL, want);
				while (m) {                                    // warp-uniform loop
					const unsigned n = unsigned(__popc(m));
					unsigned base = 0;
					if (lane == unsigned(__ffs(int(m))) - 1u) base = atomicAdd(head, n);
					base = __shfl_sync(FULL, base, __ffs(int(m)) - 1);
					if (want) {
						const unsigned i = base + unsigned(__popc(m & lt_mask));
						if (i < total) {
							job.start(i, ray);
							walk2_init(w, sc, ray);
							working = true;
						}
						want = false;
					}
					if (base + n >= total) exhausted = true;   // base, n are warp-uniform
					m = 0u;
				}
			}
			wm = __ballot_sync(FULL, working);
			if (wm == 0u) break;
		}
		if (working) {
			walk2_round<false>(w, sc, ray, tmin);
			if (walk2_finished(w)) { working = false; have_result = true; }
		}
	}
}

struct PixelChainJob {
	const SceneDev& sc;
	const WaveDev& w;
	const CameraDev& cam;
	int frames, max_depth, max_attempt;
	f4 color;
	uint32_t seed;
	int pixel, left;
	unsigned long long traced;
	__device__ __forceinline__ void start(unsigned i, RayRegs& ray) {
		pixel = int(i);
		left = frames;
		traced = 0;
		seed = w.seeds[pixel];
		color = f4{1.0f, 1.0f, 1.0f, 0.0f};
		ray = generate_ray(cam, uint32_t(pixel) % cam.width, uint32_t(pixel) / cam.width);
	}
	__device__ __forceinline__ bool next(RayRegs& ray, const HitRegs& h) {
		++traced;
		f4 nn{0, 0, 0, 0}, ip{0, 0, 0, 0};
		uint32_t mat = 0;
		if (h.closest >= 0) {
			const float4* tp = reinterpret_cast<const float4*>(sc.leaves + h.closest);
			float4 q1 = __ldg(tp + 1), q2 = __ldg(tp + 2), q3 = __ldg(tp + 3), q4 = __ldg(tp + 4);
			nn = f4{q2.w, q3.w, q4.w, 0.0f};
			mat = uint32_t(__float_as_int(q1.w));
			ip = ray.origin + h.t * ray.direction;
			if (dot3(ray.direction, nn) > 0) nn = -nn;
		}
		if (shade_one(sc.materials, sc.cs_table, ray, h.t, nn, ip, mat, color, seed, max_depth)) return true;
		// the sample is complete: history.cl for this pixel
		const size_t n = size_t(w.npix);
		f4 hs = ld4(w.hist[pixel]);
		int32_t cnt = w.counts[pixel];
		f4 cfold = color;
		if (history_one(cfold, hs, cnt, max_attempt)) {
			w.sums[pixel] += color.x; w.sums[n + pixel] += color.y; w.sums[2 * n + pixel] += color.z;
			w.sums[3 * n + pixel] += double(color.x) * color.x; w.sums[4 * n + pixel] += double(color.y) * color.y;
			w.sums[5 * n + pixel] += double(color.z) * color.z;
			w.sums[6 * n + pixel] += 1.0;
		}
		w.hist[pixel] = st4(hs);
		w.counts[pixel] = cnt;
		if (--left == 0) {
			w.seeds[pixel] = seed;
			atomicAdd(w.rays_traced, traced);
			return false;
		}
		ray = generate_ray(cam, uint32_t(pixel) % cam.width, uint32_t(pixel) / cam.width);
		color = f4{1.0f, 1.0f, 1.0f, 0.0f};
		return true;
	}
};

__global__ void __launch_bounds__(128, MCPT_MIN_BLOCKS) k_pixel_chain(SceneDev sc, WaveDev w, CameraDev cam, int frames, int max_depth,
                                                                        int max_attempt, float tmin, unsigned* head) {
	PixelChainJob job{sc, w, cam, frames, max_depth, max_attempt, f4{0, 0, 0, 0}, 0u, 0, 0, 0ull};
	persistent_walk_chain(sc, tmin, unsigned(w.npix), head, job);
}

// ---- multi-GPU merge: acc[i] += peer_0[i] + peer_1[i] + ... with the peers' accumulators read in place over
// NVLink peer mappings (mcpt_merge_sums).  Streaming, 8 B per element and peer; peers are added in index order.
constexpr int MERGE_MAX_PEERS = 15;
struct MergePeers { const double* p[MERGE_MAX_PEERS]; int n; };
__global__ void __launch_bounds__(256) k_merge_sums(double* __restrict__ acc, MergePeers peers, size_t count) {
	const size_t stride = size_t(gridDim.x) * blockDim.x;
	for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < count; i += stride) {
		double v = acc[i];
#pragma unroll 1
		for (int k = 0; k < peers.n; ++k) v += __ldcs(peers.p[k] + i);
		acc[i] = v;
	}
}

// leaf links of the fast tree as the builders write them (~triangle) -> LEAF_FLAG | triangle, the form the walk tests
__global__ void k_encode_leaf_links(DNode* nodes, int n) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const int l = nodes[i].left, r = nodes[i].right;
	if (l < 0) nodes[i].left = int32_t(0x80000000u | uint32_t(~l));
	if (r < 0) nodes[i].right = int32_t(0x80000000u | uint32_t(~r));
}

}  // namespace mcpt
