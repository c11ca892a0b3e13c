// libmcpt_cuda.so — context management and the C ABI of include/mcpt.h (sm_100a only, no CPU fallback).
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include <cub/device/device_radix_sort.cuh>

#include "mcpt.h"
#include "../common/hash128.hpp"
#include "accel_build.hpp"
#include "experimental/accel_wide.hpp"
#include "mcpt_kernels.cuh"
#include "mcpt_mis.cuh"

using namespace mcpt;

namespace {

thread_local std::string g_err;

int fail(int code, const std::string& msg) {
	g_err = msg;
	return code;
}
}  // namespace
namespace mcpt {
void set_last_error(const std::string& s) { g_err = s; }     // for the other translation units of the library
}
namespace {

#define CU(call)                                                                                       \
	do {                                                                                               \
		cudaError_t e_ = (call);                                                                       \
		if (e_ != cudaSuccess)                                                                         \
			return fail(MCPT_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));            \
	} while (0)

template <class T>
struct DevBuf {
	T* p = nullptr;
	size_t n = 0;
	cudaError_t alloc(size_t count) {
		if (p && n == count) return cudaSuccess;     // same size: keep the allocation
		release();
		n = count;
		return count ? cudaMalloc(reinterpret_cast<void**>(&p), count * sizeof(T)) : cudaSuccess;
	}
	void release() {
		if (p) cudaFree(p);
		p = nullptr;
		n = 0;
	}
	DevBuf() = default;
	DevBuf(const DevBuf&) = delete;
	DevBuf& operator=(const DevBuf&) = delete;
	~DevBuf() { release(); }
};

}  // namespace

// Acceptance probability of the glossy lobe's rejection loop (shade.cl:130-132) against theta = angle(mirror direction,
// normal), PHONG_ACCEPT_N + 1 points uniform in [0, pi/2]: the share of randomDirection's density about r that lies
// above the surface.  randomDirection draws u = k / 32768 and a uniform azimuth; a direction at polar angle alpha(u)
// about r is above the horizon for the azimuth fraction 1 (cot alpha cot theta >= 1) or acos(-cot alpha cot theta) / pi.
// Summed over the reference's own 32 768 values of u, once per process.
const std::vector<float>& phong_accept_table() {
	static std::vector<float> tab;
	static std::once_flag once;
	std::call_once(once, [] {
		constexpr int NU = 32768;
		std::vector<double> cot_alpha(NU);
		for (int k = 0; k < NU; ++k) {
			const double u = double(k) / NU, D = 1.0 - u + u * u;
			const double ca = (1.0 - u) / std::sqrt(D), sa = std::sqrt(std::max(0.0, 1.0 - ca * ca));
			cot_alpha[size_t(k)] = sa > 0.0 ? ca / sa : 1e300;
		}
		tab.assign(size_t(PHONG_ACCEPT_N) + 1, 1.0f);
		const unsigned nt = std::max(1u, std::min(8u, std::thread::hardware_concurrency()));
		std::vector<std::thread> th;
		for (unsigned t = 0; t < nt; ++t)
			th.emplace_back([&, t] {
				for (int j = int(t); j <= PHONG_ACCEPT_N; j += int(nt)) {
					const double theta = 1.57079632679489661923 * double(j) / PHONG_ACCEPT_N;
					const double ct = std::cos(theta), st = std::sin(theta);
					const double cot_theta = st > 0.0 ? ct / st : 1e300;
					double acc = 0.0;
					for (int k = 0; k < NU; ++k) {
						const double x = cot_alpha[size_t(k)] * cot_theta;
						acc += x >= 1.0 ? 1.0 : std::acos(-x) / 3.14159265358979323846;
					}
					tab[size_t(j)] = float(acc / NU);
				}
			});
		for (auto& x : th) x.join();
	});
	return tab;
}

// fn(lo, hi) over [0, n) on up to 8 host threads (the per-triangle / per-node loops of derive_scene)
template <class F>
static void parallel_ranges(int n, F&& fn) {
	const unsigned nt = unsigned(std::min<long>(std::min(8u, std::max(1u, std::thread::hardware_concurrency())), (long(n) + 16383) / 16384));
	if (nt <= 1) { fn(0, n); return; }
	std::vector<std::thread> th;
	for (unsigned t = 0; t < nt; ++t) th.emplace_back([&, t] { fn(int(long(n) * t / nt), int(long(n) * (t + 1) / nt)); });
	for (auto& x : th) x.join();
}

// Everything mcpt_upload_scene derives from the geometry alone (leaf records, two-box nodes of the reference
// topology, the binned-SAH tree of the fast walk).  Kept per process for the last scene, so that the N contexts of
// a multi-GPU run derive it once (the SAH build is the expensive part of an upload).
struct DerivedScene {
	std::vector<mcpt_triangle> tris;      // the inputs it was derived from (compared byte for byte)
	std::vector<mcpt_bvh_node> nodes;
	int nmats = 0;
	std::vector<DLeaf> dt;
	std::vector<DNode> dn, fast;
	bool contained = true;
	int fast_depth = 0;
};

struct mcpt_ctx {
	mcpt_params prm{};
	cudaStream_t stream = nullptr;
	cudaStream_t stream_g[4] = {nullptr, nullptr, nullptr, nullptr};   // streams of the regeneration loop's other pixel groups
	cudaEvent_t ev_fork = nullptr;
	int npix = 0;
	bool have_scene = false, have_camera = false;
	bool force_reference = false;     // node array violates box containment: monotonicity argument void
	CameraDev cam{};
	SceneDev scene{};
	// scene
	DevBuf<mcpt_bvh_node> nodes_ref;
	DevBuf<mcpt_triangle> tris_ref;
	DevBuf<DNode> nodes;
	DevBuf<DLeaf> wleaves;
	DevBuf<unsigned char> hot;        // fast-walk nodes followed by the leaf records: ONE range, so that one L2 access-policy window covers it
	DNode* fast_nodes = nullptr;
	DLeaf* leaves = nullptr;
	DevBuf<WNode> wnodes;
	DevBuf<mcpt_material> mats;
	DevBuf<float2> cs_table;
	// reference-layout frame buffers (stage path)
	DevBuf<mcpt_ray> rays;
	DevBuf<mcpt_hit> hits;
	DevBuf<int32_t> closest;
	DevBuf<float4> color, hist, display;
	DevBuf<uint32_t> seeds;
	DevBuf<int32_t> counts;
	DevBuf<double> sums_own;
	double* sums = nullptr;
	// wavefront path
	int wave_samples = 0;             // allocated samples per wave
	DevBuf<float4> od[2], dp[2], cs[2], sample_out;
	DevBuf<float2> hitrec;
	// NEE + MIS mode
	DevBuf<float4> mis_rad, mis_sod, mis_sdp, mis_scon;
	DevBuf<float> mis_pdf[2], emit_cdf, phong_accept, mis_den, mis_sden;
	DevBuf<float2> mis_shit;
	DevBuf<uint32_t> mis_scount, mis_shead;
	DevBuf<int32_t> emit_tri;
	int nemit = 0;
	float emit_area = 0.0f;
	MisDev mis() {
		MisDev m{};
		m.rad = mis_rad.p; m.pdf[0] = mis_pdf[0].p; m.pdf[1] = mis_pdf[1].p;
		m.sod = mis_sod.p; m.sdp = mis_sdp.p; m.scon = mis_scon.p; m.shit = mis_shit.p;
		m.scount = mis_scount.p;
		m.emit_tri = emit_tri.p; m.emit_cdf = emit_cdf.p; m.nemit = nemit; m.emit_area = emit_area;
		m.phong_accept = phong_accept.p;
		m.den = mis_den.p; m.sden = mis_sden.p;
		return m;
	}
	// coherence order of the bounce queues (mcpt_kernels.cuh): keys written by the shade kernel, sorted by CUB, the
	// resulting permutation read by the next walk
	bool sort_queues = true;               // env MCPT_SORT=0 switches it off (A/B)
	int shade_grid = 32;                   // blocks per SM of the shade kernels' grid-stride launches: 8 -> 32 measured +2.6 % on diningroom, +1.1 % cbox,
	                                       // +1.4 % veach_mis (4 blocks are resident; finer blocks even out the tail; profiles/r3w_shade_grid.log; env MCPT_SHADE_GRID for A/B)
	int mis_sort_mask = 3;                 // NEE + MIS: bit 0 main queue, bit 1 shadow queue (env MCPT_SORT=1/2/3)
	// NEE + MIS: whether the two sorts pay depends on the scene (+13 % cbox, +14 % diningroom, -9 % veach_mis), and the
	// image does not depend on the queue order, so one wave is timed each way and the faster arrangement kept.
	struct SortTrial { cudaEvent_t t0 = nullptr, t1 = nullptr; bool pending = false, have = false; double paths = 0.0, ms_per_path = 0.0; } mis_trial[2];
	int mis_sort_choice = -1;              // -1 undecided; else the mask in force until the scene or the camera changes
	bool mis_sort_tune = true;             // off when MCPT_SORT pins the mask
	bool host_tree = false;                // env MCPT_HOST_TREE=1: the fast walk's tree from the host builder instead of sah_build.cu (same bytes)
	DevBuf<uint32_t> sort_keys, sort_keys_out, sort_iota, sort_perm;
	DevBuf<uint32_t> mis_skeys, mis_sperm;    // NEE + MIS modes: keys / walk order of the shadow-ray queue
	DevBuf<unsigned char> sort_temp;
	size_t sort_temp_bytes = 0;
	// How many items to sort is only known on the device (qcount[d]); CUB wants it on the host.  Any prefix works (see
	// k_invalidate_keys), so the host uses the live fraction of the same bounce in an EARLIER wave - read back without
	// ever waiting for the wave being enqueued - plus a margin; until one is known it sorts the whole wave.
	uint32_t* h_counts = nullptr;          // pinned, 2 slots of (max_depth + 2)
	cudaEvent_t ev_counts[2] = {nullptr, nullptr};
	bool slot_pending[2] = {false, false};
	long slot_total[2] = {0, 0};
	int next_slot = 0;
	std::vector<double> live_frac;         // per bounce: live paths / wave paths (1.0 = unknown)
	DevBuf<uint32_t> rq_count, rq_head;    // per-round counter rings of the regeneration loop (seed_mode 0)
	DevBuf<int32_t> frames_left;
	DevBuf<uint32_t> qcount, qhead, first_sample;   // qhead[d]: work-fetch cursor of the persistent bounce d (+1 slot for mcpt_trace)
	int nsm = 148;
	DevBuf<unsigned long long> counters;   // 0 nodes, 1 tris, 2 live rays (stage path), 3 rays (wave), 4 paths (wave)
	// trace scratch
	DevBuf<mcpt_ray> t_rays;
	DevBuf<mcpt_hit> t_hits;
	DevBuf<int32_t> t_closest;
	int opt_flags = 0, block = 128;          // development switches: env MCPT_OPT, MCPT_BLOCK
	std::shared_ptr<const DerivedScene> derived;   // what the resident scene was derived from / into (shared by the contexts of a
	                                               // process): the key of the unchanged-geometry shortcut, the source of the 8-wide layout
	uint64_t geometry_tag = 0;               // caller's generation counter of the last upload (mcpt_upload_scene_tagged)
	uint64_t launches = 0;
	uint64_t stage_paths = 0;
	// measurement aids
	bool count_fetches = false, time_bounces = false;
	std::vector<cudaEvent_t> ev_pool;     // pairs (start, stop), ev_used of them recorded
	size_t ev_used = 0;
	double bounce_ms = 0.0;
	uint64_t bounce_launches = 0;

	~mcpt_ctx() {
		for (cudaEvent_t e : ev_counts) if (e) cudaEventDestroy(e);
		for (SortTrial& t : mis_trial) { if (t.t0) cudaEventDestroy(t.t0); if (t.t1) cudaEventDestroy(t.t1); }
		if (h_counts) cudaFreeHost(h_counts);
		for (cudaEvent_t e : ev_pool) cudaEventDestroy(e);
		if (ev_fork) cudaEventDestroy(ev_fork);
		for (cudaStream_t g : stream_g) if (g) cudaStreamDestroy(g);
		if (stream) cudaStreamDestroy(stream);
	}
	// fold recorded event pairs into bounce_ms (requires the stream to be idle)
	cudaError_t drain_events() {
		for (size_t k = 0; k + 1 < ev_used; k += 2) {
			float ms = 0.f;
			cudaError_t e = cudaEventElapsedTime(&ms, ev_pool[k], ev_pool[k + 1]);
			if (e != cudaSuccess) return e;
			bounce_ms += ms;
			++bounce_launches;
		}
		ev_used = 0;
		return cudaSuccess;
	}
	FrameDev frame() {
		return FrameDev{rays.p, hits.p, closest.p, color.p, seeds.p, hist.p, counts.p, sums, npix};
	}
	WaveDev wave() {
		WaveDev w{};
		for (int k = 0; k < 2; ++k) { w.od[k] = od[k].p; w.dp[k] = dp[k].p; w.cs[k] = cs[k].p; }
		w.sample_out = sample_out.p;
		w.hitrec = hitrec.p;
		w.qcount = qcount.p;
		w.first_sample = first_sample.p;
		w.rays_traced = counters.p + 3;
		w.paths_done = counters.p + 4;
		w.seeds = seeds.p;
		w.hist = hist.p;
		w.counts = counts.p;
		w.sums = sums;
		w.npix = npix;
		w.capacity = uint32_t(std::min<size_t>(od[0].n, 0xffffffffu));
		return w;
	}
	int traversal(int requested) const { return force_reference ? MCPT_TRAVERSAL_REFERENCE : requested; }
	// host copies kept for the lazily built 8-wide layout (MCPT_TRAVERSAL_WIDE8, experimental)
	bool have_wide = false;
	int persistent_grid(int bs) const { return nsm * (MCPT_MIN_BLOCKS * 128 / bs); }
};

// expands `stmt` with TRAV bound to the compile-time traversal id
#define MCPT_DISPATCH_TRAVERSAL(trav, stmt)                                                     \
	do {                                                                                        \
		if ((trav) == MCPT_TRAVERSAL_REFERENCE) { constexpr int TRAV = MCPT_TRAVERSAL_REFERENCE; stmt; }      \
		else if ((trav) == MCPT_TRAVERSAL_CANONICAL) { constexpr int TRAV = MCPT_TRAVERSAL_CANONICAL; stmt; } \
		else if ((trav) == MCPT_TRAVERSAL_WIDE8) { constexpr int TRAV = MCPT_TRAVERSAL_WIDE8; stmt; }   \
		else { constexpr int TRAV = MCPT_TRAVERSAL_FAST; stmt; }                                \
	} while (0)

namespace {

inline int blocks(long n, int bs) { return int((n + bs - 1) / bs); }

int check_ctx(mcpt_ctx* c, bool need_scene, bool need_camera) {
	if (!c) return fail(MCPT_ERR_ARG, "null context");
	if (need_scene && !c->have_scene) return fail(MCPT_ERR_STATE, "scene not uploaded (mcpt_upload_scene)");
	if (need_camera && !c->have_camera) return fail(MCPT_ERR_STATE, "camera not set (mcpt_set_camera)");
	cudaError_t e = cudaSetDevice(c->prm.device);
	if (e != cudaSuccess) return fail(MCPT_ERR_CUDA, std::string("cudaSetDevice: ") + cudaGetErrorString(e));
	return MCPT_OK;
}

int ensure_wave(mcpt_ctx* c, int samples) {
	if (samples <= c->wave_samples) return MCPT_OK;
	size_t n = size_t(samples) * size_t(c->npix);
	for (int k = 0; k < 2; ++k) {
		CU(c->od[k].alloc(n));
		CU(c->dp[k].alloc(n));
		CU(c->cs[k].alloc(n));
	}
	CU(c->sample_out.alloc(n));
	CU(c->hitrec.alloc(n));
	if (c->sort_queues && n < (size_t(1) << 31)) {
		if (c->prm.sampling != MCPT_SAMPLING_REFERENCE) { CU(c->mis_skeys.alloc(n)); CU(c->mis_sperm.alloc(n)); }
		CU(c->sort_keys.alloc(n));
		CU(c->sort_keys_out.alloc(n));
		CU(c->sort_iota.alloc(n));
		CU(c->sort_perm.alloc(n));
		k_iota<<<blocks(long(n), 256), 256, 0, c->stream>>>(c->sort_iota.p, uint32_t(n));
		CU(cudaMemsetAsync(c->sort_keys.p, 0xFF, sizeof(uint32_t) * n, c->stream));
		size_t bytes = 0;
		CU(cub::DeviceRadixSort::SortPairs(nullptr, bytes, c->sort_keys.p, c->sort_keys_out.p, c->sort_iota.p, c->sort_perm.p,
		                                   int(n), 0, SORT_KEY_BITS, c->stream));
		CU(c->sort_temp.alloc(bytes));
		c->sort_temp_bytes = bytes;
		if (!c->h_counts) {
			CU(cudaMallocHost(reinterpret_cast<void**>(&c->h_counts), sizeof(uint32_t) * 2 * (size_t(c->prm.max_depth) + 2)));
			for (cudaEvent_t& e : c->ev_counts) CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
			c->live_frac.assign(size_t(c->prm.max_depth) + 2, 1.0);
		}
	}
	if (c->prm.sampling != MCPT_SAMPLING_REFERENCE) {
		if (c->prm.sampling == MCPT_SAMPLING_NEE_MIS_RATIO) { CU(c->mis_den.alloc(n)); CU(c->mis_sden.alloc(n)); }
		CU(c->mis_rad.alloc(n)); CU(c->mis_sod.alloc(n)); CU(c->mis_sdp.alloc(n)); CU(c->mis_scon.alloc(n));
		CU(c->mis_pdf[0].alloc(n)); CU(c->mis_pdf[1].alloc(n)); CU(c->mis_shit.alloc(n));
		if (!c->phong_accept.p) {
			const std::vector<float>& tab = phong_accept_table();
			CU(c->phong_accept.alloc(tab.size()));
			CU(cudaMemcpyAsync(c->phong_accept.p, tab.data(), sizeof(float) * tab.size(), cudaMemcpyHostToDevice, c->stream));
		}
	}
	c->wave_samples = samples;
	return MCPT_OK;
}

int reset_accumulation(mcpt_ctx* c) {
	const int n = c->npix;
	CU(cudaMemsetAsync(c->hist.p, 0, sizeof(float4) * n, c->stream));
	CU(cudaMemsetAsync(c->counts.p, 0, sizeof(int32_t) * n, c->stream));
	CU(cudaMemsetAsync(c->sums, 0, sizeof(double) * 7 * size_t(n), c->stream));
	CU(cudaMemsetAsync(c->first_sample.p, 0, sizeof(uint32_t), c->stream));
	k_init_seeds<<<blocks(n, 256), 256, 0, c->stream>>>(c->seeds.p, n, c->prm.run_seed);
	++c->launches;
	CU(cudaGetLastError());
	return MCPT_OK;
}

// One wave of the fused pipeline: nsamples consecutive samples of every pixel.
int ensure_wide(mcpt_ctx* c);

// New scene or camera: the sort trial of the NEE + MIS mode starts over.  A trial still in flight is simply dropped
// (its events are re-recorded by the next one).
void reset_sort_trials(mcpt_ctx* c) {
	for (auto& t : c->mis_trial) { t.pending = false; t.have = false; }
	c->mis_sort_choice = -1;
}

// Folds the live counts of finished earlier waves into the per-bounce estimates (never waits).
void poll_live_counts(mcpt_ctx* c) {
	const int D = c->prm.max_depth;
	for (int s = 0; s < 2; ++s) {
		if (!c->slot_pending[s] || cudaEventQuery(c->ev_counts[s]) != cudaSuccess) continue;
		c->slot_pending[s] = false;
		const uint32_t* h = c->h_counts + size_t(s) * (size_t(D) + 2);
		for (int d = 0; d <= D; ++d) c->live_frac[size_t(d)] = double(h[d]) / double(c->slot_total[s]);
	}
	(void)cudaGetLastError();       // cudaErrorNotReady from the query is not an error
}

// This wave's live counts, for the sort sizes of later waves.  A slot is reused two waves later; by then its copy has
// long completed unless the host runs more than a wave ahead, in which case it waits here (never for the wave it has
// just enqueued).
int record_live_counts(mcpt_ctx* c, long total) {
	const int D = c->prm.max_depth;
	const int s = c->next_slot;
	if (c->slot_pending[s]) {
		CU(cudaEventSynchronize(c->ev_counts[s]));
		poll_live_counts(c);
	}
	CU(cudaMemcpyAsync(c->h_counts + size_t(s) * (size_t(D) + 2), c->qcount.p, sizeof(uint32_t) * (size_t(D) + 1),
	                   cudaMemcpyDeviceToHost, c->stream));
	CU(cudaEventRecord(c->ev_counts[s], c->stream));
	c->slot_pending[s] = true;
	c->slot_total[s] = total;
	c->next_slot = s ^ 1;
	return MCPT_OK;
}

int enqueue_wave(mcpt_ctx* c, int nsamples) {
	const int D = c->prm.max_depth;
	const long total = long(nsamples) * c->npix;
	WaveDev w = c->wave();
	// small waves stay unsorted: six more launches per bounce cost more than a few thousand rays gain
	const bool sorted = c->sort_queues && c->sort_keys.p && c->traversal(c->prm.traversal) == MCPT_TRAVERSAL_FAST &&
	                    !(c->opt_flags & (2 | 8)) && !c->count_fetches && size_t(total) <= c->sort_keys.n && total >= (1L << 18);
	if (sorted) poll_live_counts(c);
	CU(cudaMemsetAsync(c->qcount.p, 0, sizeof(uint32_t) * (D + 2), c->stream));
	CU(cudaMemsetAsync(c->qhead.p, 0, sizeof(uint32_t) * (D + 2), c->stream));
	k_wave_generate<<<blocks(total, 256), 256, 0, c->stream>>>(c->cam, w, nsamples, c->prm.seed_mode, c->prm.run_seed);
	const int trav = c->traversal(c->prm.traversal);
	if (trav == MCPT_TRAVERSAL_WIDE8) if (int r = ensure_wide(c)) return r;
	const int BS = c->block;
	const int g = blocks(total, BS);
	unsigned long long* cnt = c->counters.p;
	uint32_t perm_n = 0;             // how many entries the sort behind the current permutation covered
	for (int d = 0; d < D; ++d) {
		bool timed = false;
		if (c->time_bounces) {
			if (c->ev_used + 2 > c->ev_pool.size()) {
				if (c->ev_pool.size() >= 8192) {       // bounded pool: drain (one sync) and reuse
					CU(cudaStreamSynchronize(c->stream));
					CU(c->drain_events());
				} else {
					cudaEvent_t a, b;
					CU(cudaEventCreate(&a));
					CU(cudaEventCreate(&b));
					c->ev_pool.push_back(a);
					c->ev_pool.push_back(b);
				}
			}
			CU(cudaEventRecord(c->ev_pool[c->ev_used], c->stream));
		}
		if (c->count_fetches)
			MCPT_DISPATCH_TRAVERSAL(trav, (k_wave_bounce<TRAV, true><<<g, BS, 0, c->stream>>>(c->scene, w, d, D, c->prm.tmin, c->prm.seed_mode, cnt)));
		else if (trav == MCPT_TRAVERSAL_FAST && !(c->opt_flags & 2)) {
			if (c->opt_flags & 8) {        // fused walk + shade in one persistent kernel (the split pipeline measured faster)
				k_wave_bounce_persistent<<<std::min(g, c->persistent_grid(BS)), BS, 0, c->stream>>>(c->scene, w, d, D, c->prm.tmin, c->prm.seed_mode, c->qhead.p + d);
			} else {
				w.perm = (sorted && d > 0) ? c->sort_perm.p : nullptr;       // primary rays are coherent as generated
				w.perm_n = perm_n;
				w.sort_keys = (sorted && d + 1 < D) ? c->sort_keys.p : nullptr;
				k_wave_trace_persistent<<<std::min(g, c->persistent_grid(BS)), BS, 0, c->stream>>>(c->scene, w, d, c->prm.tmin, c->qhead.p + d);
				if (c->time_bounces) {      // the timed kernel is the walk; the shader runs after the stop event
					CU(cudaEventRecord(c->ev_pool[c->ev_used + 1], c->stream));
					c->ev_used += 2;
					timed = true;
				}
				k_wave_shade<<<std::min(blocks(total, 256), c->nsm * c->shade_grid), 256, 0, c->stream>>>(c->scene, w, d, D, c->prm.seed_mode);
				++c->launches;
				if (w.sort_keys) {
					// the next bounce's walk order: live items first (their keys are fresh), sorted by coherence key
					k_invalidate_keys<<<c->nsm * 2, 256, 0, c->stream>>>(c->sort_keys.p, c->qcount.p + d + 1, c->qcount.p + d);
					long n_est = long(c->live_frac[size_t(d) + 1] * 1.02 * double(total)) + 4096;
					if (n_est > total) n_est = total;
					CU(cub::DeviceRadixSort::SortPairs(c->sort_temp.p, c->sort_temp_bytes, c->sort_keys.p, c->sort_keys_out.p,
					                                   c->sort_iota.p, c->sort_perm.p, int(n_est), 0, SORT_KEY_BITS, c->stream));
					perm_n = uint32_t(n_est);
					c->launches += 6;       // invalidate + CUB's histogram, exclusive sum and 3 onesweep passes
				}
			}
		}
		else
			MCPT_DISPATCH_TRAVERSAL(trav, (k_wave_bounce<TRAV, false><<<g, BS, 0, c->stream>>>(c->scene, w, d, D, c->prm.tmin, c->prm.seed_mode, cnt)));
		if (c->time_bounces && !timed) {
			CU(cudaEventRecord(c->ev_pool[c->ev_used + 1], c->stream));
			c->ev_used += 2;
		}
	}
	k_wave_resolve<<<blocks(c->npix, 256), 256, 0, c->stream>>>(w, nsamples, D, c->prm.max_attempt);
	c->launches += uint64_t(D) + 2;
	if (sorted)
		if (int r = record_live_counts(c, total)) return r;
	CU(cudaGetLastError());
	return MCPT_OK;
}

// seed_mode 0, `frames` more samples of every pixel with path regeneration (k_wave_shade_regen): rounds of
// (walk, shade + fold + regenerate) until no pixel has samples left.  The host cannot know the number of rounds (it
// is the longest chain of path lengths of any pixel), so rounds are enqueued in chunks and the live count is read
// back after each chunk; rounds past the end find an empty queue and return at once.
#ifndef MCPT_REGEN_DEFAULT
#define MCPT_REGEN_DEFAULT 1
#endif
constexpr int REGEN_CHUNK = 1024;
constexpr int REGEN_MAX_GROUPS = 4;
constexpr int REGEN_CHAIN_FRAMES = 1 << 30;   // samples per pixel and launch of k_pixel_chain: measured best = all of them (4: -18%, 16: -9%, 64: -3%)
#ifndef MCPT_REGEN_GROUPS
#define MCPT_REGEN_GROUPS 2
#endif
constexpr int REGEN_GROUPS = MCPT_REGEN_GROUPS;      // even: the queue parity carries over from chunk to chunk

int enqueue_regen(mcpt_ctx* c, int frames) {
	const int D = c->prm.max_depth;
	const int n = c->npix;
	if (int r = ensure_wave(c, 1)) return r;
	// Several pixel groups, each on its own stream: a round's launches are small (one ray per unfinished pixel), and a persistent
	// kernel's tail - its last warps finishing alone - is a third of such a launch; the other group's kernels fill it.
	int G = (n >= (1 << 18) && !(c->opt_flags & 32)) ? REGEN_GROUPS : 1;
	if (const char* e = std::getenv("MCPT_REGEN_GROUPS")) G = std::max(1, std::min(REGEN_MAX_GROUPS, std::atoi(e)));   // development
	if (G > n) G = 1;
	for (int k = 1; k < G; ++k)
		if (!c->stream_g[k]) CU(cudaStreamCreateWithFlags(&c->stream_g[k], cudaStreamNonBlocking));
	if (G > 1 && !c->ev_fork) CU(cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
	const int RING = REGEN_CHUNK + 2;
	CU(c->rq_count.alloc(size_t(REGEN_MAX_GROUPS) * RING));
	CU(c->rq_head.alloc(size_t(REGEN_MAX_GROUPS) * RING));
	CU(c->frames_left.alloc(size_t(n)));
	const int BS = c->block;
	k_fill_i32<<<blocks(n, 256), 256, 0, c->stream>>>(c->frames_left.p, n, frames);
	CU(cudaMemsetAsync(c->rq_count.p, 0, sizeof(uint32_t) * REGEN_MAX_GROUPS * RING, c->stream));
	CU(cudaMemsetAsync(c->rq_head.p, 0, sizeof(uint32_t) * REGEN_MAX_GROUPS * RING, c->stream));
	++c->launches;
	struct Group { WaveDev w; uint32_t* head; cudaStream_t st; int count; uint32_t live; } grp[REGEN_MAX_GROUPS];
	if (G > 1) CU(cudaEventRecord(c->ev_fork, c->stream));
	for (int k = 0; k < G; ++k) {
		const int p0 = int(long(n) * k / G);
		const int cnt = int(long(n) * (k + 1) / G) - p0;
		WaveDev w = c->wave();
		for (int b2 = 0; b2 < 2; ++b2) { w.od[b2] += p0; w.dp[b2] += p0; w.cs[b2] += p0; }
		w.hitrec += p0;
		w.qcount = c->rq_count.p + size_t(k) * RING;
		grp[k] = Group{w, c->rq_head.p + size_t(k) * RING, k == 0 ? c->stream : c->stream_g[k], cnt, uint32_t(cnt)};
		if (k > 0) CU(cudaStreamWaitEvent(grp[k].st, c->ev_fork, 0));      // the other streams start after the set-up above
		k_regen_start<<<blocks(cnt, 256), 256, 0, grp[k].st>>>(c->cam, w, p0, cnt);
		++c->launches;
	}
	const long max_rounds = long(frames) * D;        // every path ends within D rounds of its start
	long done_rounds = 0;
	bool any = true;
	while (any && done_rounds < max_rounds) {
		const long want = std::min<long>(REGEN_CHUNK, max_rounds - done_rounds);
		const int rounds = int((want + 1) & ~1L);        // even: the queue parity carries over to the next chunk
		for (int r = 0; r < rounds; ++r) {
			for (int k = 0; k < G; ++k) {
				if (!grp[k].live) continue;
				k_wave_trace_persistent<<<std::min(blocks(grp[k].count, BS), c->persistent_grid(BS)), BS, 0, grp[k].st>>>(
					c->scene, grp[k].w, r, c->prm.tmin, grp[k].head + r);
			}
			for (int k = 0; k < G; ++k) {
				if (!grp[k].live) continue;
				k_wave_shade_regen<<<std::min(blocks(grp[k].count, 256), c->nsm * 8), 256, 0, grp[k].st>>>(
					c->scene, grp[k].w, c->cam, r, D, c->prm.max_attempt, c->frames_left.p);
				c->launches += 2;
			}
		}
		for (int k = 0; k < G; ++k) {
			if (!grp[k].live) continue;
			uint32_t* q = grp[k].w.qcount;
			CU(cudaMemcpyAsync(&grp[k].live, q + rounds, sizeof(uint32_t), cudaMemcpyDeviceToHost, grp[k].st));
			// carry the queue over: round `rounds` of this chunk is round 0 of the next
			CU(cudaMemcpyAsync(q, q + rounds, sizeof(uint32_t), cudaMemcpyDeviceToDevice, grp[k].st));
			CU(cudaMemsetAsync(q + 1, 0, sizeof(uint32_t) * (RING - 1), grp[k].st));
			CU(cudaMemsetAsync(grp[k].head, 0, sizeof(uint32_t) * RING, grp[k].st));
		}
		for (int k = 0; k < G; ++k) CU(cudaStreamSynchronize(grp[k].st));
		done_rounds += rounds;
		any = false;
		for (int k = 0; k < G; ++k) any = any || grp[k].live != 0;
	}
	if (any) return fail(MCPT_ERR_STATE, "mcpt_render: regeneration loop did not drain (internal error)");
	// bookkeeping the frame loop does in k_wave_resolve
	const unsigned long long paths = (unsigned long long)frames * (unsigned long long)n;
	k_add_stats<<<1, 1, 0, c->stream>>>(grp[0].w.paths_done, paths, grp[0].w.first_sample, uint32_t(frames));
	++c->launches;
	CU(cudaGetLastError());
	return MCPT_OK;
}

// seed_mode 0, `frames` more samples of every pixel in ONE launch (k_pixel_chain): no queues, nothing to wait for.
int enqueue_pixel_chain(mcpt_ctx* c, int frames) {
	if (int r = ensure_wave(c, 1)) return r;          // only for a valid WaveDev; the queues stay unused
	CU(c->rq_head.alloc(size_t(REGEN_MAX_GROUPS) * (REGEN_CHUNK + 2)));
	WaveDev w = c->wave();
	const int BS = 128;
	// A pixel's samples are one sequential chain, so a launch ends with its slowest pixel; splitting the run into
	// shorter launches to bound that tail measured slower than the tail itself, so one launch takes all samples.
	int per_launch = REGEN_CHAIN_FRAMES;
	if (const char* e = std::getenv("MCPT_CHAIN_FRAMES")) per_launch = std::max(1, std::atoi(e));     // development
	CU(cudaMemsetAsync(c->rq_head.p, 0, sizeof(uint32_t) * (REGEN_CHUNK + 2), c->stream));
	int slot = 0;
	for (int done = 0; done < frames;) {
		const int f = std::min(per_launch, frames - done);
		if (slot == REGEN_CHUNK + 2) {                      // recycle the work-fetch cursors
			CU(cudaMemsetAsync(c->rq_head.p, 0, sizeof(uint32_t) * (REGEN_CHUNK + 2), c->stream));
			slot = 0;
		}
		k_pixel_chain<<<std::min(blocks(c->npix, BS), c->persistent_grid(BS)), BS, 0, c->stream>>>(
			c->scene, w, c->cam, f, c->prm.max_depth, c->prm.max_attempt, c->prm.tmin, c->rq_head.p + slot);
		++slot;
		++c->launches;
		done += f;
	}
	const unsigned long long paths = (unsigned long long)frames * (unsigned long long)c->npix;
	k_add_stats<<<1, 1, 0, c->stream>>>(w.paths_done, paths, w.first_sample, uint32_t(frames));
	c->launches += 2;
	CU(cudaGetLastError());
	return MCPT_OK;
}

int ensure_wide(mcpt_ctx* c) {
	if (c->have_wide) return MCPT_OK;
	std::vector<WNode> wide;
	std::vector<DLeaf> wide_leaves;
	build_wide_bvh8(c->derived->fast, c->derived->dt.data(), int(c->derived->dt.size()), wide, wide_leaves);
	CU(cudaStreamSynchronize(c->stream));
	CU(c->wnodes.alloc(std::max<size_t>(1, wide.size())));
	CU(c->wleaves.alloc(std::max<size_t>(1, wide_leaves.size())));
	if (!wide.empty()) CU(cudaMemcpyAsync(c->wnodes.p, wide.data(), sizeof(WNode) * wide.size(), cudaMemcpyHostToDevice, c->stream));
	if (!wide_leaves.empty())
		CU(cudaMemcpyAsync(c->wleaves.p, wide_leaves.data(), sizeof(DLeaf) * wide_leaves.size(), cudaMemcpyHostToDevice, c->stream));
	CU(cudaStreamSynchronize(c->stream));
	c->scene.wnodes = c->wnodes.p;
	c->scene.wleaves = c->wleaves.p;
	c->have_wide = true;
	return MCPT_OK;
}

// One wave of the NEE + MIS mode: per bounce walk -> shade (chain + emitter sample) -> walk the shadow rays -> add.
int enqueue_wave_mis(mcpt_ctx* c, int nsamples) {
	const int D = c->prm.max_depth;
	const long total = long(nsamples) * c->npix;
	WaveDev w = c->wave();
	MisDev m = c->mis();
	CU(cudaMemsetAsync(c->qcount.p, 0, sizeof(uint32_t) * (D + 2), c->stream));
	CU(cudaMemsetAsync(c->qhead.p, 0, sizeof(uint32_t) * (D + 2), c->stream));
	CU(cudaMemsetAsync(c->mis_scount.p, 0, sizeof(uint32_t) * (D + 2), c->stream));
	CU(cudaMemsetAsync(c->mis_shead.p, 0, sizeof(uint32_t) * (D + 2), c->stream));
	CU(cudaMemsetAsync(c->mis_rad.p, 0, sizeof(float4) * size_t(total), c->stream));
	CU(cudaMemsetAsync(c->mis_pdf[0].p, 0, sizeof(float) * size_t(total), c->stream));
	if (c->mis_den.p) CU(cudaMemsetAsync(c->mis_den.p, 0, sizeof(float) * size_t(total), c->stream));
	k_wave_generate<<<blocks(total, 256), 256, 0, c->stream>>>(c->cam, w, nsamples, c->prm.seed_mode, c->prm.run_seed);
	const int BS = 128;
	const int pg = std::min(blocks(total, BS), c->persistent_grid(BS));
	const int sg = std::min(blocks(total, 256), c->nsm * c->shade_grid);
	// coherence-sorted queues as in enqueue_wave: the main queue of the next bounce and this bounce's shadow-ray queue
	const bool sorted = c->sort_queues && c->sort_keys.p && c->mis_skeys.p && size_t(total) <= c->sort_keys.n && total >= (1L << 18);
	if (sorted) poll_live_counts(c);
	// the arrangement of this wave: the decided one; while undecided, one timed wave with both sorts, then one without
	int mask = c->mis_sort_mask, trial = -1;
	if (sorted && c->mis_sort_tune) {
		if (c->mis_sort_choice < 0) {
			for (auto& t : c->mis_trial) {
				if (!t.pending || cudaEventQuery(t.t1) != cudaSuccess) continue;
				float ms = 0.0f;
				CU(cudaEventElapsedTime(&ms, t.t0, t.t1));
				t.pending = false; t.have = true; t.ms_per_path = double(ms) / t.paths;
			}
			(void)cudaGetLastError();
			if (c->mis_trial[0].have && c->mis_trial[1].have) c->mis_sort_choice = c->mis_trial[0].ms_per_path <= c->mis_trial[1].ms_per_path ? 3 : 0;
		}
		if (c->mis_sort_choice >= 0) mask = c->mis_sort_choice;
		else if (c->live_frac[1] < 1.0) {         // live counts known: the sorts run at their steady-state sizes
			for (int k = 0; k < 2 && trial < 0; ++k)
				if (!c->mis_trial[k].have && !c->mis_trial[k].pending) trial = k;
			if (trial >= 0) {
				auto& t = c->mis_trial[trial];
				if (!t.t0) { CU(cudaEventCreate(&t.t0)); CU(cudaEventCreate(&t.t1)); }
				mask = trial == 0 ? 3 : 0;
				CU(cudaEventRecord(t.t0, c->stream));
			}
		}
	}
	auto estimate = [&](int d) {      // live paths entering bounce d, from the previous waves (+2 % and a constant; the walk copes with more)
		long n = long(c->live_frac[size_t(d)] * 1.02 * double(total)) + 4096;
		return n > total ? total : n;
	};
	uint32_t perm_n = 0;
	for (int d = 0; d < D; ++d) {
		const bool sort_main = sorted && (mask & 1), sort_shadow = sorted && (mask & 2);
		const uint32_t* perm = (sort_main && d > 0) ? c->sort_perm.p : nullptr;
		m.keys = (sort_main && d + 1 < D) ? c->sort_keys.p : nullptr;
		m.skeys = sort_shadow ? c->mis_skeys.p : nullptr;
		k_trace_queue_persistent<<<pg, BS, 0, c->stream>>>(c->scene, w.od[d & 1], w.dp[d & 1], w.hitrec, c->qcount.p + d, c->prm.tmin, c->qhead.p + d, perm, perm_n);
		const long ns = sorted ? estimate(d) : 0;          // shadow rays of bounce d: at most one per live path
		if (sort_shadow) {       // the sorted prefix of the shadow queue starts invalid; live rays overwrite theirs; the main queue's
			// invalid tail is maintained bounce by bounce like in the reference mode (k_invalidate_keys)
			CU(cudaMemsetAsync(c->mis_skeys.p, 0xFF, sizeof(uint32_t) * size_t(ns), c->stream));
		}
		k_wave_shade_mis<<<sg, 256, 0, c->stream>>>(c->scene, w, m, d, D);
		const uint32_t* sperm = nullptr;
		if (sorted) {
			if (sort_shadow) {
				CU(cub::DeviceRadixSort::SortPairs(c->sort_temp.p, c->sort_temp_bytes, c->mis_skeys.p, c->sort_keys_out.p, c->sort_iota.p,
				                                   c->mis_sperm.p, int(ns), 0, SORT_KEY_BITS, c->stream));
				sperm = c->mis_sperm.p;
				c->launches += 5;       // CUB's histogram, exclusive sum and 3 onesweep passes
			}
			if (m.keys) {
				k_invalidate_keys<<<c->nsm * 2, 256, 0, c->stream>>>(c->sort_keys.p, c->qcount.p + d + 1, c->qcount.p + d);
				CU(cub::DeviceRadixSort::SortPairs(c->sort_temp.p, c->sort_temp_bytes, c->sort_keys.p, c->sort_keys_out.p, c->sort_iota.p,
				                                   c->sort_perm.p, int(estimate(d + 1)), 0, SORT_KEY_BITS, c->stream));
				perm_n = uint32_t(estimate(d + 1));
				c->launches += 6;
			}
		}
		k_trace_queue_persistent<<<pg, BS, 0, c->stream>>>(c->scene, m.sod, m.sdp, m.shit, c->mis_scount.p + d, c->prm.tmin, c->mis_shead.p + d, sperm, uint32_t(ns));
		k_wave_nee_add<<<sg, 256, 0, c->stream>>>(m, d);
	}
	k_wave_resolve_mis<<<blocks(c->npix, 256), 256, 0, c->stream>>>(w, m, nsamples, D);
	if (trial >= 0) {
		auto& t = c->mis_trial[trial];
		CU(cudaEventRecord(t.t1, c->stream));
		t.pending = true;
		t.paths = double(total);
	}
	if (sorted)
		if (int r = record_live_counts(c, total)) return r;
	c->launches += uint64_t(4 * D) + 2;
	CU(cudaGetLastError());
	return MCPT_OK;
}

int launch_intersect(mcpt_ctx* c, const mcpt_ray* d_rays, long n, mcpt_hit* d_hits, int32_t* d_closest, int traversal, bool count) {
	const int trav = c->traversal(traversal);
	if (trav == MCPT_TRAVERSAL_WIDE8) if (int r = ensure_wide(c)) return r;
	const int BS = c->block;
	const int g = blocks(n, BS);
	unsigned long long* cnt = c->counters.p;
	if (trav < 0 || trav > 3) return fail(MCPT_ERR_ARG, "unknown traversal");
	if (count)
		MCPT_DISPATCH_TRAVERSAL(trav, (k_intersect<TRAV, true><<<g, BS, 0, c->stream>>>(c->scene, d_rays, n, c->prm.tmin, d_hits, d_closest, cnt)));
	else if (trav == MCPT_TRAVERSAL_FAST && !(c->opt_flags & 2) && n < (1L << 32)) {
		unsigned* head = c->qhead.p + c->prm.max_depth + 1;
		CU(cudaMemsetAsync(head, 0, sizeof(unsigned), c->stream));
		k_intersect_persistent<<<std::min(g, c->persistent_grid(BS)), BS, 0, c->stream>>>(c->scene, d_rays, unsigned(n), c->prm.tmin, d_hits, d_closest, head);
	} else
		MCPT_DISPATCH_TRAVERSAL(trav, (k_intersect<TRAV, false><<<g, BS, 0, c->stream>>>(c->scene, d_rays, n, c->prm.tmin, d_hits, d_closest, cnt)));
	++c->launches;
	CU(cudaGetLastError());
	return MCPT_OK;
}

}  // namespace

// ================================================================== C ABI
extern "C" {

const char* mcpt_last_error(void) { return g_err.c_str(); }
int mcpt_abi_version(void) { return 1; }

mcpt_ctx* mcpt_create(const mcpt_params* params) {
	if (!params || params->width <= 0 || params->height <= 0 || params->max_depth <= 0 || params->max_depth > 0xFFFF ||
	    params->max_attempt < 0 || long(params->width) * params->height > (1L << 28)) {
		fail(MCPT_ERR_ARG, "mcpt_create: bad parameters");
		return nullptr;
	}
	int ndev = 0;
	cudaError_t e = cudaGetDeviceCount(&ndev);
	if (e != cudaSuccess || ndev == 0) {
		fail(MCPT_ERR_CUDA, std::string("no CUDA device: ") + cudaGetErrorString(e));
		return nullptr;
	}
	if (params->device < 0 || params->device >= ndev) { fail(MCPT_ERR_ARG, "mcpt_create: device ordinal out of range"); return nullptr; }
	cudaDeviceProp prop{};
	if ((e = cudaGetDeviceProperties(&prop, params->device)) != cudaSuccess || (e = cudaSetDevice(params->device)) != cudaSuccess) {
		fail(MCPT_ERR_CUDA, cudaGetErrorString(e));
		return nullptr;
	}
	if (prop.major != 10) {
		fail(MCPT_ERR_CUDA, "this library carries sm_100a code only; device is sm_" + std::to_string(prop.major * 10 + prop.minor));
		return nullptr;
	}
	mcpt_ctx* c = new mcpt_ctx();
	c->prm = *params;
	if (c->prm.tmin == 0.0f) c->prm.tmin = 0.001f;
	if (c->prm.seed_mode == 0) c->prm.samples_in_flight = 1;
	if (c->prm.sampling != MCPT_SAMPLING_REFERENCE && c->prm.sampling != MCPT_SAMPLING_NEE_MIS && c->prm.sampling != MCPT_SAMPLING_NEE_MIS_RATIO) {
		fail(MCPT_ERR_ARG, "mcpt_create: unknown sampling mode");
		delete c;
		return nullptr;
	}
	if (c->prm.sampling != MCPT_SAMPLING_REFERENCE && c->prm.seed_mode != 1) {
		fail(MCPT_ERR_ARG, "mcpt_create: MCPT_SAMPLING_NEE_MIS needs seed_mode 1 (it renders in waves of whole samples)");
		delete c;
		return nullptr;
	}
	c->npix = params->width * params->height;
	c->nsm = prop.multiProcessorCount;
	if (const char* e = std::getenv("MCPT_OPT")) c->opt_flags = std::atoi(e);
	if (const char* e = std::getenv("MCPT_BLOCK")) { int b = std::atoi(e); if (b == 64 || b == 128) c->block = b; }
	if (const char* e = std::getenv("MCPT_SORT")) { c->sort_queues = std::atoi(e) != 0; if (c->sort_queues) c->mis_sort_mask = std::atoi(e) & 3; c->mis_sort_tune = false; }      // development A/B
	if (const char* e = std::getenv("MCPT_SHADE_GRID")) c->shade_grid = std::max(1, std::min(1 << 16, std::atoi(e)));
	if (const char* e = std::getenv("MCPT_HOST_TREE")) c->host_tree = std::atoi(e) != 0;
	const size_t n = size_t(c->npix);
	bool ok = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) == cudaSuccess;
	ok = ok && c->rays.alloc(n) == cudaSuccess && c->hits.alloc(n) == cudaSuccess && c->closest.alloc(n) == cudaSuccess &&
	     c->color.alloc(n) == cudaSuccess && c->hist.alloc(n) == cudaSuccess && c->display.alloc(n) == cudaSuccess &&
	     c->seeds.alloc(n) == cudaSuccess && c->counts.alloc(n) == cudaSuccess && c->sums_own.alloc(7 * n) == cudaSuccess &&
	     c->qcount.alloc(size_t(params->max_depth) + 2) == cudaSuccess && c->qhead.alloc(size_t(params->max_depth) + 2) == cudaSuccess && c->first_sample.alloc(1) == cudaSuccess &&
	     c->counters.alloc(8) == cudaSuccess && c->mis_scount.alloc(size_t(params->max_depth) + 2) == cudaSuccess &&
	     c->mis_shead.alloc(size_t(params->max_depth) + 2) == cudaSuccess;
	if (ok) {
		c->sums = c->sums_own.p;
		ok = cudaMemsetAsync(c->rays.p, 0, sizeof(mcpt_ray) * n, c->stream) == cudaSuccess &&
		     cudaMemsetAsync(c->hits.p, 0, sizeof(mcpt_hit) * n, c->stream) == cudaSuccess &&
		     cudaMemsetAsync(c->counters.p, 0, sizeof(unsigned long long) * 8, c->stream) == cudaSuccess &&
		     reset_accumulation(c) == MCPT_OK && cudaStreamSynchronize(c->stream) == cudaSuccess;
	}
	if (!ok) {
		if (g_err.empty()) fail(MCPT_ERR_CUDA, std::string("mcpt_create: ") + cudaGetErrorString(cudaGetLastError()));
		delete c;
		return nullptr;
	}
	return c;
}

void mcpt_destroy(mcpt_ctx* c) {
	if (!c) return;
	cudaSetDevice(c->prm.device);
	cudaStreamSynchronize(c->stream);
	delete c;
}

void* mcpt_stream(mcpt_ctx* c) { return c ? (void*)c->stream : nullptr; }

int mcpt_synchronize(mcpt_ctx* c) {
	if (int r = check_ctx(c, false, false)) return r;
	CU(cudaStreamSynchronize(c->stream));
	return MCPT_OK;
}

// Emitter triangles (material type LIGHT) with an area CDF, for the NEE + MIS mode.
static int build_emitters(mcpt_ctx* c, const mcpt_triangle* tris, int ntris, const mcpt_material* mats) {
	std::vector<int32_t> list;
	std::vector<double> area;
	double total = 0.0;
	for (int i = 0; i < ntris; ++i) {
		if (mats[tris[i].material_id].type != MCPT_LIGHT) continue;
		const mcpt_float4 &a = tris[i].v[0], &b = tris[i].v[1], &d = tris[i].v[2];
		const double ux = double(b.x) - a.x, uy = double(b.y) - a.y, uz = double(b.z) - a.z;
		const double vx = double(d.x) - a.x, vy = double(d.y) - a.y, vz = double(d.z) - a.z;
		const double cx = uy * vz - uz * vy, cy = uz * vx - ux * vz, cz = ux * vy - uy * vx;
		const double ar = 0.5 * std::sqrt(cx * cx + cy * cy + cz * cz);
		if (!(ar > 0.0) || !std::isfinite(ar)) continue;
		list.push_back(i);
		total += ar;
		area.push_back(total);
	}
	std::vector<float> cdf(list.size());
	for (size_t k = 0; k < list.size(); ++k) cdf[k] = float(area[k] / total);
	if (!cdf.empty()) cdf.back() = 1.0f;
	c->nemit = int(list.size());
	c->emit_area = float(total);
	CU(c->emit_tri.alloc(std::max<size_t>(1, list.size())));
	CU(c->emit_cdf.alloc(std::max<size_t>(1, list.size())));
	if (!list.empty()) {
		CU(cudaMemcpyAsync(c->emit_tri.p, list.data(), sizeof(int32_t) * list.size(), cudaMemcpyHostToDevice, c->stream));
		CU(cudaMemcpyAsync(c->emit_cdf.p, cdf.data(), sizeof(float) * list.size(), cudaMemcpyHostToDevice, c->stream));
		CU(cudaStreamSynchronize(c->stream));     // the host vectors go out of scope
	}
	return MCPT_OK;
}

// memcmp(a, b, n) == 0 on up to 8 threads: the cache checks below compare ~80 MB per upload, which one core does
// in about 10 ms - as long as the copies they save.
static bool bytes_equal(const void* a, const void* b, size_t n) {
	const size_t chunk = size_t(4) << 20;
	if (n < 2 * chunk) return std::memcmp(a, b, n) == 0;
	const unsigned nt = unsigned(std::min<size_t>(8, (n + chunk - 1) / chunk));
	std::atomic<bool> same{true};
	std::vector<std::thread> th;
	for (unsigned t = 0; t < nt; ++t) {
		const size_t lo = n * t / nt, hi = n * (t + 1) / nt;
		th.emplace_back([&, lo, hi] {
			if (std::memcmp(static_cast<const char*>(a) + lo, static_cast<const char*>(b) + lo, hi - lo) != 0) same = false;
		});
	}
	for (auto& x : th) x.join();
	return same;
}

static std::mutex g_derived_mutex;
static std::shared_ptr<const DerivedScene> g_derived;

// tree_device >= 0: the binned-SAH tree is built on that device (sah_build.cu; same bytes as the host builder)
static int derive_scene(const mcpt_triangle* tris, int ntris, const mcpt_material* mats, int nmats, const mcpt_bvh_node* nodes,
                 int nnodes, DerivedScene& out, int tree_device = -1, float* tree_ms = nullptr) {
	// ---- validate + derive the traversal layout (host side, once per scene)
	std::vector<DLeaf>& dt = out.dt;
	dt.resize(size_t(ntris));
	std::atomic<bool> bad_material{false};
	parallel_ranges(ntris, [&](int lo, int hi) {
	std::memset(dt.data() + lo, 0, size_t(hi - lo) * sizeof(DLeaf));
	for (int i = lo; i < hi; ++i) {
		const mcpt_triangle& t = tris[i];
		if (t.material_id < 0 || t.material_id >= nmats) { bad_material = true; return; }
		DLeaf& d = dt[i];
		d.v0x = t.v[0].x; d.v0y = t.v[0].y; d.v0z = t.v[0].z; d.material = t.material_id;
		// e1 = -(v1 - v0), e2 = -(v2 - v0): the matrix rows of objdef.h:194-199
		d.e1x = -(t.v[1].x - t.v[0].x); d.e1y = -(t.v[1].y - t.v[0].y); d.e1z = -(t.v[1].z - t.v[0].z);
		d.e2x = -(t.v[2].x - t.v[0].x); d.e2y = -(t.v[2].y - t.v[0].y); d.e2z = -(t.v[2].z - t.v[0].z);
		d.cr0 = d.e1y * d.e2z - d.e2y * d.e1z;
		d.cr1 = d.e1x * d.e2z - d.e2x * d.e1z;
		d.cr2 = d.e1x * d.e2y - d.e2x * d.e1y;
		d.nx = t.nx; d.ny = t.ny; d.nz = t.nz;
		std::memcpy(&d.pad, &i, sizeof(int));     // own index: leaf records are self-describing in any order
	}
	});
	if (bad_material) return fail(MCPT_ERR_ARG, "mcpt_upload_scene: triangle material id out of range");
	for (int m = 0; m < nmats; ++m)
		if (mats[m].kd.w != 0.0f || mats[m].ka_ks.w != 0.0f)
			return fail(MCPT_ERR_ARG, "mcpt_upload_scene: material w lanes must be 0 (as loadObject produces them)");
	auto is_leaf = [&](int i) { return nodes[i].left == nodes[i].right; };
	// canonical rank of every leaf = its position in the reference's left-first DFS (objdef.h:252-273);
	// the leaf's box is copied from the reference leaf record (exact floats of the reference box test)
	{
		std::vector<int> st;
		std::vector<char> seen(size_t(ntris), 0);
		st.push_back(0);
		int rank = 0;
		long guard = 0;
		while (!st.empty()) {
			int i = st.back();
			st.pop_back();
			if (i < 0 || i >= nnodes || ++guard > 4L * nnodes) return fail(MCPT_ERR_ARG, "mcpt_upload_scene: malformed node links");
			if (is_leaf(i)) {
				int t = nodes[i].left;
				if (t < 0 || t >= ntris || seen[t]) return fail(MCPT_ERR_ARG, "mcpt_upload_scene: leaf triangle index out of range or repeated");
				seen[t] = 1;
				DLeaf& d = dt[t];
				d.rank = rank++;
				d.bminx = nodes[i].bbmin.x; d.bminy = nodes[i].bbmin.y; d.bminz = nodes[i].bbmin.z;
				d.bmaxx = nodes[i].bbmax.x; d.bmaxy = nodes[i].bbmax.y; d.bmaxz = nodes[i].bbmax.z;
			} else {
				st.push_back(nodes[i].right);
				st.push_back(nodes[i].left);
			}
		}
		if (rank != ntris) return fail(MCPT_ERR_ARG, "mcpt_upload_scene: the node array does not reach every triangle exactly once");
	}
	bool& contained = out.contained;
	contained = true;
	const bool root_leaf = is_leaf(0);
	std::vector<DNode>& dn = out.dn;                          // indexed by reference node id (leaf slots unused)
	dn.assign(root_leaf ? 1 : size_t(nnodes), DNode{});
	std::atomic<int> node_error{0};
	std::atomic<bool> nested{true};
	parallel_ranges(nnodes, [&](int lo, int hi) {
	if (!root_leaf) std::memset(dn.data() + lo, 0, size_t(hi - lo) * sizeof(DNode));
	for (int i = lo; i < hi; ++i) {
		const mcpt_bvh_node& nd = nodes[i];
		if (nd.left < 0 || nd.right < 0) { node_error = 1; return; }
		if (is_leaf(i)) {
			if (nd.left >= ntris) { node_error = 2; return; }
			continue;
		}
		if (nd.left >= nnodes || nd.right >= nnodes) { node_error = 3; return; }
		const mcpt_bvh_node& L = nodes[nd.left];
		const mcpt_bvh_node& R = nodes[nd.right];
		for (const mcpt_bvh_node* ch : {&L, &R}) {
			if (!(ch->bbmin.x >= nd.bbmin.x && ch->bbmin.y >= nd.bbmin.y && ch->bbmin.z >= nd.bbmin.z &&
			      ch->bbmax.x <= nd.bbmax.x && ch->bbmax.y <= nd.bbmax.y && ch->bbmax.z <= nd.bbmax.z))
				nested = false;
		}
		if (root_leaf) continue;
		DNode& d = dn[i];
		d.lminx = L.bbmin.x; d.lminy = L.bbmin.y; d.lminz = L.bbmin.z; d.lmaxx = L.bbmax.x; d.lmaxy = L.bbmax.y; d.lmaxz = L.bbmax.z;
		d.rminx = R.bbmin.x; d.rminy = R.bbmin.y; d.rminz = R.bbmin.z; d.rmaxx = R.bbmax.x; d.rmaxy = R.bbmax.y; d.rmaxz = R.bbmax.z;
		d.left = is_leaf(nd.left) ? ~L.left : nd.left;
		d.right = is_leaf(nd.right) ? ~R.left : nd.right;
	}
	});
	if (node_error == 1) return fail(MCPT_ERR_ARG, "mcpt_upload_scene: negative child index");
	if (node_error == 2) return fail(MCPT_ERR_ARG, "mcpt_upload_scene: leaf triangle index out of range");
	if (node_error == 3) return fail(MCPT_ERR_ARG, "mcpt_upload_scene: child index out of range");
	contained = nested;
	std::vector<DNode>& fast = out.fast;
	fast.clear();
	int device_depth = 0;                 // the device builder reports the depth of its tree (inner levels, root = 1)
	if (root_leaf) fast.assign(1, DNode{});
	else if (tree_device >= 0) {
		DevBuf<DLeaf> d_leaves;
		CU(cudaSetDevice(tree_device));
		CU(d_leaves.alloc(size_t(ntris)));
		CU(cudaMemcpy(d_leaves.p, dt.data(), sizeof(DLeaf) * size_t(ntris), cudaMemcpyHostToDevice));
		fast.assign(size_t(ntris) - 1, DNode{});
		if (int r = build_sah_bvh2_device(tree_device, d_leaves.p, dt.data(), ntris, fast.data(), tree_ms, &device_depth)) return r;
	} else build_sah_bvh2(dt.data(), ntris, fast);
	// the fast walk pushes at most one entry per level and does not check its stack: verify the depth here
	int& fast_depth = out.fast_depth;
	fast_depth = device_depth;
	if (!root_leaf && device_depth == 0) {
		std::vector<std::pair<int, int>> todo{{0, 1}};
		while (!todo.empty()) {
			auto [node, d] = todo.back();
			todo.pop_back();
			fast_depth = std::max(fast_depth, d);
			if (fast[size_t(node)].left >= 0) todo.push_back({fast[size_t(node)].left, d + 1});
			if (fast[size_t(node)].right >= 0) todo.push_back({fast[size_t(node)].right, d + 1});
		}
	}
	return MCPT_OK;
}

// ---- the derived layout as a relocatable blob (scene cache): header | DLeaf[ntris] | DNode[ndn] | DNode[nfast]
struct DerivedHeader {
	char magic[8];
	uint32_t version, sizeof_dnode, sizeof_dleaf, builder_id;
	int32_t ntris, nnodes, nmats, contained, fast_depth, pad;
	uint64_t ndn, nfast;
	uint64_t input_hash[2];      // of the Triangle[] and BVHNode[] bytes the layout was derived from
};
static_assert(sizeof(DerivedHeader) == 80, "DerivedHeader");
constexpr char DERIVED_MAGIC[8] = {'M', 'C', 'P', 'T', 'D', 'R', 'V', '1'};
constexpr uint32_t DERIVED_VERSION = 1;

static Hash128 geometry_hash(const mcpt_triangle* tris, int ntris, const mcpt_bvh_node* nodes, int nnodes) {
	Hash128 h;
	hash_update(h, tris, sizeof(mcpt_triangle) * size_t(ntris));
	hash_update(h, nodes, sizeof(mcpt_bvh_node) * size_t(nnodes));
	return h;
}

static size_t derived_blob_bytes(const DerivedScene& d) {
	return sizeof(DerivedHeader) + d.dt.size() * sizeof(DLeaf) + (d.dn.size() + d.fast.size()) * sizeof(DNode);
}

static void write_derived_blob(const DerivedScene& d, int nnodes, const Hash128& geom, unsigned char* out) {
	DerivedHeader h{};
	std::memcpy(h.magic, DERIVED_MAGIC, 8);
	h.version = DERIVED_VERSION;
	h.sizeof_dnode = sizeof(DNode);
	h.sizeof_dleaf = sizeof(DLeaf);
	h.builder_id = FAST_TREE_BUILDER_ID;
	h.ntris = int32_t(d.dt.size());
	h.nnodes = nnodes;
	h.nmats = d.nmats;
	h.contained = d.contained ? 1 : 0;
	h.fast_depth = d.fast_depth;
	h.ndn = d.dn.size();
	h.nfast = d.fast.size();
	h.input_hash[0] = geom.a; h.input_hash[1] = geom.b;
	std::memcpy(out, &h, sizeof(h));
	out += sizeof(h);
	std::memcpy(out, d.dt.data(), d.dt.size() * sizeof(DLeaf));
	out += d.dt.size() * sizeof(DLeaf);
	std::memcpy(out, d.dn.data(), d.dn.size() * sizeof(DNode));
	out += d.dn.size() * sizeof(DNode);
	std::memcpy(out, d.fast.data(), d.fast.size() * sizeof(DNode));
}

// Accepts a blob only if it was derived from exactly these triangles and nodes (content hash) by this library version,
// and if every index in it stays inside its array: the walk kernels follow child links without bounds checks.
static int read_derived_blob(const void* blob, size_t bytes, const mcpt_triangle* tris, int ntris, int nmats,
                             const mcpt_bvh_node* nodes, int nnodes, DerivedScene& out) {
	if (!blob || bytes < sizeof(DerivedHeader)) return fail(MCPT_ERR_ARG, "derived layout: blob too small");
	DerivedHeader h;
	std::memcpy(&h, blob, sizeof(h));
	if (std::memcmp(h.magic, DERIVED_MAGIC, 8) != 0 || h.version != DERIVED_VERSION || h.sizeof_dnode != sizeof(DNode) ||
	    h.sizeof_dleaf != sizeof(DLeaf) || h.builder_id != FAST_TREE_BUILDER_ID)
		return fail(MCPT_ERR_ARG, "derived layout: written by another library version");
	if (h.ntris != ntris || h.nnodes != nnodes || h.nmats != nmats) return fail(MCPT_ERR_ARG, "derived layout: belongs to another scene (counts differ)");
	const bool root_leaf = nodes[0].left == nodes[0].right;
	const uint64_t want_dn = root_leaf ? 1 : uint64_t(nnodes), want_fast = uint64_t(std::max(1, ntris - 1));
	if (h.ndn != want_dn || h.nfast != want_fast ||
	    bytes != sizeof(DerivedHeader) + uint64_t(ntris) * sizeof(DLeaf) + (h.ndn + h.nfast) * sizeof(DNode))
		return fail(MCPT_ERR_ARG, "derived layout: section sizes do not match the scene");
	const Hash128 geom = geometry_hash(tris, ntris, nodes, nnodes);
	if (geom.a != h.input_hash[0] || geom.b != h.input_hash[1])
		return fail(MCPT_ERR_ARG, "derived layout: stale (triangles or nodes changed since it was derived)");
	const unsigned char* p = static_cast<const unsigned char*>(blob) + sizeof(DerivedHeader);
	out.dt.resize(size_t(ntris));
	std::memcpy(out.dt.data(), p, out.dt.size() * sizeof(DLeaf));
	p += out.dt.size() * sizeof(DLeaf);
	out.dn.resize(size_t(h.ndn));
	std::memcpy(out.dn.data(), p, out.dn.size() * sizeof(DNode));
	p += out.dn.size() * sizeof(DNode);
	out.fast.resize(size_t(h.nfast));
	std::memcpy(out.fast.data(), p, out.fast.size() * sizeof(DNode));
	out.contained = h.contained != 0;
	out.fast_depth = h.fast_depth;
	out.nmats = nmats;
	for (int i = 0; i < ntris; ++i)
		if (out.dt[size_t(i)].rank < 0 || out.dt[size_t(i)].rank >= ntris || out.dt[size_t(i)].material < 0 || out.dt[size_t(i)].material >= nmats)
			return fail(MCPT_ERR_ARG, "derived layout: leaf record out of range");
	if (!root_leaf) {
		// reference-topology nodes: inner slots only (leaf slots stay zero); fast tree: preorder, so a child's index
		// exceeds its parent's (no cycles) and every triangle hangs below exactly one link
		auto ref_ok = [&](int32_t r, int64_t lo, int64_t hi) { return r < 0 ? ~r < ntris : (r > lo && r < hi); };
		for (int i = 0; i < nnodes; ++i) {
			if (nodes[i].left == nodes[i].right) continue;
			const DNode& d = out.dn[size_t(i)];
			if (!ref_ok(d.left, -1, nnodes) || !ref_ok(d.right, -1, nnodes)) return fail(MCPT_ERR_ARG, "derived layout: node link out of range");
		}
		std::vector<char> seen(size_t(ntris), 0);
		long leaves = 0;
		for (size_t i = 0; i < out.fast.size(); ++i) {
			for (int32_t r : {out.fast[i].left, out.fast[i].right}) {
				if (!ref_ok(r, int64_t(i), int64_t(out.fast.size()))) return fail(MCPT_ERR_ARG, "derived layout: fast-tree link out of range");
				if (r < 0) {
					if (seen[size_t(~r)]) return fail(MCPT_ERR_ARG, "derived layout: triangle linked twice");
					seen[size_t(~r)] = 1;
					++leaves;
				}
			}
		}
		if (leaves != ntris) return fail(MCPT_ERR_ARG, "derived layout: fast tree does not reach every triangle");
		if (h.fast_depth < 1 || h.fast_depth > ntris) return fail(MCPT_ERR_ARG, "derived layout: implausible depth");
	}
	return MCPT_OK;
}

static int upload_scene(mcpt_ctx* c, const mcpt_triangle* tris, int ntris, const mcpt_material* mats, int nmats,
                        const mcpt_bvh_node* nodes, int nnodes, uint64_t tag, const void* blob = nullptr, size_t blob_bytes = 0);

static int check_scene_args(const mcpt_triangle* tris, int ntris, const mcpt_material* mats, int nmats, const mcpt_bvh_node* nodes, int nnodes) {
	if (!tris || !mats || !nodes || ntris <= 0 || nmats <= 0 || nnodes != 2 * ntris - 1)
		return fail(MCPT_ERR_ARG, "mcpt_upload_scene: expected ntris > 0, nmats > 0 and nnodes == 2*ntris-1");
	return MCPT_OK;
}

int mcpt_derive_scene(const mcpt_triangle* tris, int ntris, const mcpt_material* mats, int nmats, const mcpt_bvh_node* nodes,
                      int nnodes, void** blob, size_t* bytes) {
	if (!blob || !bytes) return fail(MCPT_ERR_ARG, "mcpt_derive_scene: null output");
	if (int r = check_scene_args(tris, ntris, mats, nmats, nodes, nnodes)) return r;
	DerivedScene d;
	if (int r = derive_scene(tris, ntris, mats, nmats, nodes, nnodes, d)) return r;
	d.nmats = nmats;
	const size_t n = derived_blob_bytes(d);
	unsigned char* out = static_cast<unsigned char*>(std::malloc(n));
	if (!out) return fail(MCPT_ERR_ARG, "mcpt_derive_scene: out of host memory");
	write_derived_blob(d, nnodes, geometry_hash(tris, ntris, nodes, nnodes), out);
	*blob = out;
	*bytes = n;
	return MCPT_OK;
}

int mcpt_derive_scene_device(int device, const mcpt_triangle* tris, int ntris, const mcpt_material* mats, int nmats,
                             const mcpt_bvh_node* nodes, int nnodes, void** blob, size_t* bytes, float* tree_ms) {
	if (!blob || !bytes) return fail(MCPT_ERR_ARG, "mcpt_derive_scene_device: null output");
	if (int r = check_scene_args(tris, ntris, mats, nmats, nodes, nnodes)) return r;
	DerivedScene d;
	if (int r = derive_scene(tris, ntris, mats, nmats, nodes, nnodes, d, device, tree_ms)) return r;
	d.nmats = nmats;
	const size_t n = derived_blob_bytes(d);
	unsigned char* out = static_cast<unsigned char*>(std::malloc(n));
	if (!out) return fail(MCPT_ERR_ARG, "mcpt_derive_scene_device: out of host memory");
	write_derived_blob(d, nnodes, geometry_hash(tris, ntris, nodes, nnodes), out);
	*blob = out;
	*bytes = n;
	return MCPT_OK;
}

void mcpt_free_blob(void* blob) { std::free(blob); }

int mcpt_upload_scene_derived(mcpt_ctx* c, const mcpt_triangle* tris, int ntris, const mcpt_material* mats, int nmats,
                              const mcpt_bvh_node* nodes, int nnodes, const void* blob, size_t bytes, uint64_t geometry_tag) {
	if (!blob || bytes == 0) return fail(MCPT_ERR_ARG, "mcpt_upload_scene_derived: null blob");
	return upload_scene(c, tris, ntris, mats, nmats, nodes, nnodes, geometry_tag, blob, bytes);
}

int mcpt_upload_scene(mcpt_ctx* c, const mcpt_triangle* tris, int ntris, const mcpt_material* mats, int nmats,
                      const mcpt_bvh_node* nodes, int nnodes) {
	return upload_scene(c, tris, ntris, mats, nmats, nodes, nnodes, 0);
}

int mcpt_upload_scene_tagged(mcpt_ctx* c, const mcpt_triangle* tris, int ntris, const mcpt_material* mats, int nmats,
                             const mcpt_bvh_node* nodes, int nnodes, uint64_t geometry_tag) {
	return upload_scene(c, tris, ntris, mats, nmats, nodes, nnodes, geometry_tag);
}

static int upload_scene(mcpt_ctx* c, const mcpt_triangle* tris, int ntris, const mcpt_material* mats, int nmats,
                        const mcpt_bvh_node* nodes, int nnodes, uint64_t tag, const void* blob, size_t blob_bytes) {
	if (int r = check_ctx(c, false, false)) return r;
	if (int r = check_scene_args(tris, ntris, mats, nmats, nodes, nnodes)) return r;
	std::fill(c->live_frac.begin(), c->live_frac.end(), 1.0);     // sort sizes of the bounce queues: unknown again for new inputs
	reset_sort_trials(c);
	// ---- a layout derived earlier (scene cache): validated against these very arrays before anything else, never trusted
	std::shared_ptr<DerivedScene> from_blob;
	if (blob) {
		from_blob = std::make_shared<DerivedScene>();
		if (int r = read_derived_blob(blob, blob_bytes, tris, ntris, nmats, nodes, nnodes, *from_blob)) return r;
		for (int i = 0; i < ntris; ++i)
			if (tris[i].material_id < 0 || tris[i].material_id >= nmats) return fail(MCPT_ERR_ARG, "mcpt_upload_scene: triangle material id out of range");
	}
	// ---- unchanged geometry (same bytes as the previous upload): the derived layout on the device is still
	// valid; only the caller's arrays are copied again (what OpenCL's CL_MEM_COPY_HOST_PTR upload would do)
	// A non-zero geometry_tag equal to the previous upload's is the caller's word that triangles and nodes are the same
	// bytes (a generation counter); without one the two arrays are compared (80 MB for the diningroom stand-in).
	const bool same_tag = tag != 0 && tag == c->geometry_tag;
	if (c->have_scene && c->derived && c->derived->tris.size() == size_t(ntris) && c->derived->nodes.size() == size_t(nnodes) &&
	    (same_tag || (bytes_equal(c->derived->tris.data(), tris, sizeof(mcpt_triangle) * ntris) &&
	                  bytes_equal(c->derived->nodes.data(), nodes, sizeof(mcpt_bvh_node) * nnodes)))) {
		c->geometry_tag = tag;
		for (int m = 0; m < nmats; ++m)
			if (mats[m].kd.w != 0.0f || mats[m].ka_ks.w != 0.0f)
				return fail(MCPT_ERR_ARG, "mcpt_upload_scene: material w lanes must be 0 (as loadObject produces them)");
		for (int i = 0; i < ntris; ++i)
			if (tris[i].material_id < 0 || tris[i].material_id >= nmats)
				return fail(MCPT_ERR_ARG, "mcpt_upload_scene: triangle material id out of range");
		CU(cudaStreamSynchronize(c->stream));
		CU(c->mats.alloc(nmats));
		CU(cudaMemcpyAsync(c->nodes_ref.p, nodes, sizeof(mcpt_bvh_node) * nnodes, cudaMemcpyHostToDevice, c->stream));
		CU(cudaMemcpyAsync(c->tris_ref.p, tris, sizeof(mcpt_triangle) * ntris, cudaMemcpyHostToDevice, c->stream));
		CU(cudaMemcpyAsync(c->mats.p, mats, sizeof(mcpt_material) * nmats, cudaMemcpyHostToDevice, c->stream));
		CU(cudaStreamSynchronize(c->stream));
		c->scene.materials = c->mats.p;
		return build_emitters(c, tris, ntris, mats);
	}
	// ---- validate + derive the traversal layout (host side, once per scene and process)
	std::shared_ptr<const DerivedScene> derived;
	{
		std::lock_guard<std::mutex> lock(g_derived_mutex);
		if (g_derived && g_derived->nmats == nmats && g_derived->tris.size() == size_t(ntris) && g_derived->nodes.size() == size_t(nnodes) &&
		    bytes_equal(g_derived->tris.data(), tris, sizeof(mcpt_triangle) * ntris) &&
		    bytes_equal(g_derived->nodes.data(), nodes, sizeof(mcpt_bvh_node) * nnodes)) {
			derived = g_derived;
		} else {
			auto fresh = from_blob ? from_blob : std::make_shared<DerivedScene>();
			if (!from_blob)
				if (int r = derive_scene(tris, ntris, mats, nmats, nodes, nnodes, *fresh, c->host_tree ? -1 : c->prm.device)) return r;
			fresh->tris.assign(tris, tris + ntris);
			fresh->nodes.assign(nodes, nodes + nnodes);
			fresh->nmats = nmats;
			derived = g_derived = fresh;
		}
	}
	for (int m = 0; m < nmats; ++m)
		if (mats[m].kd.w != 0.0f || mats[m].ka_ks.w != 0.0f)
			return fail(MCPT_ERR_ARG, "mcpt_upload_scene: material w lanes must be 0 (as loadObject produces them)");
	const std::vector<DLeaf>& dt = derived->dt;
	const std::vector<DNode>& dn = derived->dn;
	const std::vector<DNode>& fast = derived->fast; // leaf links as ~triangle; re-encoded on the device after the copy
	const bool contained = derived->contained;
	const bool root_leaf = nodes[0].left == nodes[0].right;
	const int fast_depth = derived->fast_depth;
	// azimuth table of randomDirection (shade.cl:43): angle evaluated in double, narrowed, then cos/sin
	std::vector<float2> cs(32768);
	for (int r = 0; r < 32768; ++r) {
		float deg = float(2 * 3.14159265358979323846 / 32768 * double(uint32_t(r)));
		cs[r] = make_float2(float(std::cos(double(deg))), float(std::sin(double(deg))));
	}
	const bool fast_too_deep = fast_depth > STACK_SIZE;
	// the experimental 8-wide layout is derived on first use (ensure_wide), not on every upload
	c->have_wide = false;
	CU(cudaStreamSynchronize(c->stream));
	const size_t hot_nodes_bytes = (sizeof(DNode) * fast.size() + 255) / 256 * 256;
	const size_t hot_leaf_bytes = (sizeof(DLeaf) * size_t(ntris) + 255) / 256 * 256;
	const size_t hot_bytes = hot_nodes_bytes + hot_leaf_bytes;
	CU(c->hot.alloc(hot_bytes));
	c->fast_nodes = reinterpret_cast<DNode*>(c->hot.p);
	c->leaves = reinterpret_cast<DLeaf*>(c->hot.p + hot_nodes_bytes);
	CU(cudaMemcpyAsync(c->fast_nodes, fast.data(), sizeof(DNode) * fast.size(), cudaMemcpyHostToDevice, c->stream));
	// leaf links as LEAF_FLAG | triangle: no decode in the walk's loop
	k_encode_leaf_links<<<blocks(long(fast.size()), 256), 256, 0, c->stream>>>(c->fast_nodes, int(fast.size()));
	CU(c->nodes_ref.alloc(nnodes));
	CU(c->tris_ref.alloc(ntris));
	CU(c->nodes.alloc(dn.size()));
	CU(c->mats.alloc(nmats));
	CU(c->cs_table.alloc(32768));
	CU(cudaMemcpyAsync(c->nodes_ref.p, nodes, sizeof(mcpt_bvh_node) * nnodes, cudaMemcpyHostToDevice, c->stream));
	CU(cudaMemcpyAsync(c->tris_ref.p, tris, sizeof(mcpt_triangle) * ntris, cudaMemcpyHostToDevice, c->stream));
	CU(cudaMemcpyAsync(c->nodes.p, dn.data(), sizeof(DNode) * dn.size(), cudaMemcpyHostToDevice, c->stream));
	CU(cudaMemcpyAsync(c->leaves, dt.data(), sizeof(DLeaf) * ntris, cudaMemcpyHostToDevice, c->stream));
	CU(cudaMemcpyAsync(c->mats.p, mats, sizeof(mcpt_material) * nmats, cudaMemcpyHostToDevice, c->stream));
	CU(cudaMemcpyAsync(c->cs_table.p, cs.data(), sizeof(float2) * 32768, cudaMemcpyHostToDevice, c->stream));
	CU(cudaStreamSynchronize(c->stream));
	c->force_reference = !contained;
	float max_abs = 0.0f;
	for (float v : {nodes[0].bbmin.x, nodes[0].bbmin.y, nodes[0].bbmin.z, nodes[0].bbmax.x, nodes[0].bbmax.y, nodes[0].bbmax.z})
		max_abs = std::max(max_abs, std::fabs(v));
	if (!(max_abs < 1e30f) || fast_too_deep) c->force_reference = true;
	// Keep the walk's working set resident in L2: the per-bounce ray / path-state streams (hundreds of MB)
	// otherwise evict it, and every evicted node costs a whole warp a DRAM round trip (ncu: 20% L2 misses,
	// 0.6 DRAM node fetches per ray).  Measured on B200: no gain (-1%), so it is opt-in: MCPT_OPT bit 2.
	if (c->opt_flags & 4) {
		cudaDeviceProp prop;
		CU(cudaGetDeviceProperties(&prop, c->prm.device));
		const size_t carve = std::min<size_t>(size_t(prop.persistingL2CacheMaxSize), hot_bytes);
		if (carve > 0 && prop.accessPolicyMaxWindowSize > 0) {
			CU(cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, carve));
			cudaStreamAttrValue attr{};
			attr.accessPolicyWindow.base_ptr = c->hot.p;
			attr.accessPolicyWindow.num_bytes = std::min<size_t>(hot_bytes, size_t(prop.accessPolicyMaxWindowSize));
			attr.accessPolicyWindow.hitRatio = float(std::min(1.0, double(carve) / double(attr.accessPolicyWindow.num_bytes)));
			attr.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
			attr.accessPolicyWindow.missProp = cudaAccessPropertyNormal;
			CU(cudaStreamSetAttribute(c->stream, cudaStreamAttributeAccessPolicyWindow, &attr));
		}
	}
	c->derived = derived;
	c->geometry_tag = tag;
	c->scene = SceneDev{c->nodes_ref.p, c->tris_ref.p, c->nodes.p, c->fast_nodes, c->wnodes.p, c->wleaves.p, c->leaves, c->mats.p, c->cs_table.p, ntris, nnodes, root_leaf ? 1 : 0, c->opt_flags, max_abs};
	for (int a = 0; a < 3; ++a) {
		const float lo = a == 0 ? nodes[0].bbmin.x : a == 1 ? nodes[0].bbmin.y : nodes[0].bbmin.z;
		const float hi = a == 0 ? nodes[0].bbmax.x : a == 1 ? nodes[0].bbmax.y : nodes[0].bbmax.z;
		c->scene.key_lo[a] = lo;
		c->scene.key_scale[a] = (hi > lo && std::isfinite(32.0f / (hi - lo))) ? 32.0f / (hi - lo) : 0.0f;
	}
	c->have_scene = true;
	return build_emitters(c, tris, ntris, mats);
}

int mcpt_set_camera(mcpt_ctx* c, const mcpt_camera* cam) {
	if (int r = check_ctx(c, false, false)) return r;
	if (!cam) return fail(MCPT_ERR_ARG, "mcpt_set_camera: null camera");
	std::fill(c->live_frac.begin(), c->live_frac.end(), 1.0);
	reset_sort_trials(c);
	CameraDev d{};
	d.center = f4{cam->center.x, cam->center.y, cam->center.z, cam->center.w};
	d.direction = f4{cam->direction.x, cam->direction.y, cam->direction.z, cam->direction.w};
	d.up = f4{cam->up.x, cam->up.y, cam->up.z, cam->up.w};
	d.horizontal = f4{cam->horizontal.x, cam->horizontal.y, cam->horizontal.z, cam->horizontal.w};
	d.arg = cam->arg;
	d.distance = 0.5f / float(std::tan(double(cam->arg / 2)));     // rayGenerator.cl:16, uniform over the image
	d.camera_type = cam->camera_type;
	d.width = uint32_t(c->prm.width);
	d.height = uint32_t(c->prm.height);
	c->cam = d;
	c->have_camera = true;
	return MCPT_OK;
}

// ---------------------------------------------------------------- stage level
int mcpt_generate_rays(mcpt_ctx* c) {
	if (int r = check_ctx(c, false, true)) return r;
	dim3 bs(32, 8), gs((c->prm.width + 31) / 32, (c->prm.height + 7) / 8);
	k_generate_rays<<<gs, bs, 0, c->stream>>>(c->cam, c->rays.p);
	++c->launches;
	CU(cudaGetLastError());
	return MCPT_OK;
}

int mcpt_reset_color(mcpt_ctx* c) {
	if (int r = check_ctx(c, false, false)) return r;
	k_fill_color<<<blocks(c->npix, 256), 256, 0, c->stream>>>(c->color.p, c->npix, 1.0f);
	++c->launches;
	CU(cudaGetLastError());
	return MCPT_OK;
}

int mcpt_intersect(mcpt_ctx* c, int traversal) {
	if (int r = check_ctx(c, true, false)) return r;
	// the live-ray counter (counters[2]) feeds stats.rays_traced for the stage path
	const int trav = c->traversal(traversal);
	const int g = blocks(c->npix, 128);
	if (trav < 0 || trav > 3) return fail(MCPT_ERR_ARG, "unknown traversal");
	if (trav == MCPT_TRAVERSAL_WIDE8) if (int r = ensure_wide(c)) return r;
	MCPT_DISPATCH_TRAVERSAL(trav, (k_intersect<TRAV, true><<<g, 128, 0, c->stream>>>(c->scene, c->rays.p, c->npix, c->prm.tmin, c->hits.p, c->closest.p, c->counters.p)));
	++c->launches;
	CU(cudaGetLastError());
	return MCPT_OK;
}

int mcpt_shade(mcpt_ctx* c) {
	if (int r = check_ctx(c, true, false)) return r;
	k_shade<<<blocks(c->npix, 128), 128, 0, c->stream>>>(c->scene, c->rays.p, c->hits.p, c->color.p, c->seeds.p, c->npix, c->prm.max_depth);
	++c->launches;
	CU(cudaGetLastError());
	return MCPT_OK;
}

int mcpt_accumulate(mcpt_ctx* c) {
	if (int r = check_ctx(c, false, false)) return r;
	k_history<<<blocks(c->npix, 256), 256, 0, c->stream>>>(c->color.p, c->hist.p, c->counts.p, c->sums, c->npix, c->prm.max_attempt);
	++c->launches;
	CU(cudaGetLastError());
	return MCPT_OK;
}

int mcpt_display(mcpt_ctx* c, int from_history, float* host_rgba) {
	if (int r = check_ctx(c, false, false)) return r;
	if (!host_rgba) return fail(MCPT_ERR_ARG, "mcpt_display: null destination");
	k_display<<<blocks(c->npix, 256), 256, 0, c->stream>>>(from_history ? c->hist.p : c->color.p, c->display.p, c->npix);
	++c->launches;
	CU(cudaGetLastError());
	CU(cudaMemcpyAsync(host_rgba, c->display.p, sizeof(float4) * c->npix, cudaMemcpyDeviceToHost, c->stream));
	CU(cudaStreamSynchronize(c->stream));
	return MCPT_OK;
}

int mcpt_frame(mcpt_ctx* c) {
	if (int r = check_ctx(c, true, true)) return r;
	int r;
	if ((r = mcpt_reset_color(c)) || (r = mcpt_generate_rays(c))) return r;
	for (int d = 0; d < c->prm.max_depth; ++d)
		if ((r = mcpt_intersect(c, c->prm.traversal)) || (r = mcpt_shade(c))) return r;
	if ((r = mcpt_accumulate(c))) return r;
	c->stage_paths += uint64_t(c->npix);
	CU(cudaStreamSynchronize(c->stream));
	return MCPT_OK;
}

// ---------------------------------------------------------------- fused wavefront path
int mcpt_render_async(mcpt_ctx* c, int spp) {
	if (int r = check_ctx(c, true, true)) return r;
	if (spp < 0) return fail(MCPT_ERR_ARG, "mcpt_render: negative spp");
	if (c->prm.sampling != MCPT_SAMPLING_REFERENCE && (c->force_reference || c->scene.root_is_leaf))
		return fail(MCPT_ERR_STATE, "mcpt_render: the NEE + MIS mode needs the fast walk, which this scene cannot use (single triangle or boxes not nested)");
	int S = c->prm.seed_mode == 0 ? 1 : c->prm.samples_in_flight;
	if (S <= 0) {
		// automatic: about 128 Mi paths in flight (120 B each + 16 B of sort keys / permutation), never more than a
		// quarter of the free memory.  Large waves matter twice: every bounce launch ends with a tail in which the
		// persistent blocks run dry one by one, and the coherence sort finds closer neighbours in a denser population
		// (measured at 1080p, depth 16, sorted queues: 8 samples in flight 3 829 Mrays/s, 16: 4 070, 32: 4 202,
		// 64: 4 283; profiles/r2q_wave_size.log)
		size_t free_b = 0, total_b = 0;
		CU(cudaMemGetInfo(&free_b, &total_b));
		long paths = 128L << 20;
		if (c->wave_samples == 0 && size_t(paths) * 136 > free_b / 4) paths = long(free_b / 4 / 136);
		S = int(paths / c->npix);
		if (S < 1) S = 1;
		if (S > 64) S = 64;
	}
	// seed_mode 0 with the product walk: path regeneration instead of one wave per frame (MCPT_REGEN_DEFAULT; MCPT_OPT
	// bit 16 selects the other one)
	if (c->prm.seed_mode == 0 && spp >= 2 && c->prm.sampling == MCPT_SAMPLING_REFERENCE && !(c->opt_flags & (2 | 8)) &&
	    (((c->opt_flags & 16) != 0) != (MCPT_REGEN_DEFAULT != 0)) &&
	    c->traversal(c->prm.traversal) == MCPT_TRAVERSAL_FAST && !c->count_fetches && !c->time_bounces)
		return (c->opt_flags & 64) ? enqueue_regen(c, spp) : enqueue_pixel_chain(c, spp);      // bit 64: the queue-based loop
	if (S > spp && spp > 0) S = spp;
	if (int r = ensure_wave(c, S)) return r;
	for (int done = 0; done < spp;) {
		int s = spp - done < S ? spp - done : S;
		if (int r = (c->prm.sampling != MCPT_SAMPLING_REFERENCE ? enqueue_wave_mis(c, s) : enqueue_wave(c, s))) return r;
		done += s;
	}
	return MCPT_OK;
}

int mcpt_render(mcpt_ctx* c, int spp) {
	if (int r = mcpt_render_async(c, spp)) return r;
	CU(cudaStreamSynchronize(c->stream));
	return MCPT_OK;
}

int mcpt_set_first_sample(mcpt_ctx* c, uint32_t first) {
	if (int r = check_ctx(c, false, false)) return r;
	// seed_mode 0 never reads the sample index (one LCG stream per pixel): ranks that "shard" it would all draw the
	// same samples, silently.  Only the bookkeeping value 0 (a reset) is accepted there.
	if (c->prm.seed_mode == 0 && first != 0)
		return fail(MCPT_ERR_STATE, "mcpt_set_first_sample: seed_mode 0 has one stream per pixel and cannot be sharded by sample index (use seed_mode 1)");
	CU(cudaMemcpyAsync(c->first_sample.p, &first, sizeof(first), cudaMemcpyHostToDevice, c->stream));
	CU(cudaStreamSynchronize(c->stream));
	return MCPT_OK;
}

int mcpt_reset_accumulation(mcpt_ctx* c) {
	if (int r = check_ctx(c, false, false)) return r;
	if (int r = reset_accumulation(c)) return r;
	CU(cudaStreamSynchronize(c->stream));
	return MCPT_OK;
}

// ---------------------------------------------------------------- buffer I/O
#define MCPT_IO(name, type, member, dir_read)                                                                     \
	int name(mcpt_ctx* c, type* buf) {                                                                            \
		if (int r = check_ctx(c, false, false)) return r;                                                         \
		if (!buf) return fail(MCPT_ERR_ARG, #name ": null buffer");                                               \
		if (dir_read) CU(cudaMemcpyAsync((void*)buf, c->member.p, sizeof(*c->member.p) * c->npix, cudaMemcpyDeviceToHost, c->stream)); \
		else CU(cudaMemcpyAsync(c->member.p, (const void*)buf, sizeof(*c->member.p) * c->npix, cudaMemcpyHostToDevice, c->stream));    \
		CU(cudaStreamSynchronize(c->stream));                                                                     \
		return MCPT_OK;                                                                                           \
	}
MCPT_IO(mcpt_read_rays, mcpt_ray, rays, true)
MCPT_IO(mcpt_write_rays, const mcpt_ray, rays, false)
MCPT_IO(mcpt_write_hits, const mcpt_hit, hits, false)
MCPT_IO(mcpt_read_color, float, color, true)
MCPT_IO(mcpt_write_color, const float, color, false)
MCPT_IO(mcpt_read_seeds, uint32_t, seeds, true)
MCPT_IO(mcpt_write_seeds, const uint32_t, seeds, false)
#undef MCPT_IO

int mcpt_read_hits(mcpt_ctx* c, mcpt_hit* dst, int32_t* closest) {
	if (int r = check_ctx(c, false, false)) return r;
	if (!dst) return fail(MCPT_ERR_ARG, "mcpt_read_hits: null buffer");
	CU(cudaMemcpyAsync(dst, c->hits.p, sizeof(mcpt_hit) * c->npix, cudaMemcpyDeviceToHost, c->stream));
	if (closest) CU(cudaMemcpyAsync(closest, c->closest.p, sizeof(int32_t) * c->npix, cudaMemcpyDeviceToHost, c->stream));
	CU(cudaStreamSynchronize(c->stream));
	return MCPT_OK;
}

int mcpt_read_history(mcpt_ctx* c, float* hist_rgba, int32_t* counts) {
	if (int r = check_ctx(c, false, false)) return r;
	if (!hist_rgba) return fail(MCPT_ERR_ARG, "mcpt_read_history: null buffer");
	CU(cudaMemcpyAsync(hist_rgba, c->hist.p, sizeof(float4) * c->npix, cudaMemcpyDeviceToHost, c->stream));
	if (counts) CU(cudaMemcpyAsync(counts, c->counts.p, sizeof(int32_t) * c->npix, cudaMemcpyDeviceToHost, c->stream));
	CU(cudaStreamSynchronize(c->stream));
	return MCPT_OK;
}

int mcpt_read_sums(mcpt_ctx* c, double* dst) {
	if (int r = check_ctx(c, false, false)) return r;
	if (!dst) return fail(MCPT_ERR_ARG, "mcpt_read_sums: null buffer");
	CU(cudaMemcpyAsync(dst, c->sums, sizeof(double) * 7 * size_t(c->npix), cudaMemcpyDeviceToHost, c->stream));
	CU(cudaStreamSynchronize(c->stream));
	return MCPT_OK;
}

// ---- checkpoint / resume of the progressive state (the reference keeps hist / counts / seeds device-resident and
// dumps once, colorout.cpp:23-24,49-50,63-68): together with mcpt_write_seeds (seed_mode 0) or
// mcpt_set_first_sample (seed_mode 1) these restore a context to where mcpt_read_history / mcpt_read_sums left it.
int mcpt_write_sums(mcpt_ctx* c, const double* src) {
	if (int r = check_ctx(c, false, false)) return r;
	if (!src) return fail(MCPT_ERR_ARG, "mcpt_write_sums: null buffer");
	CU(cudaMemcpyAsync(c->sums, src, sizeof(double) * 7 * size_t(c->npix), cudaMemcpyHostToDevice, c->stream));
	CU(cudaStreamSynchronize(c->stream));
	return MCPT_OK;
}

int mcpt_write_history(mcpt_ctx* c, const float* hist_rgba, const int32_t* counts) {
	if (int r = check_ctx(c, false, false)) return r;
	if (!hist_rgba || !counts) return fail(MCPT_ERR_ARG, "mcpt_write_history: null buffer");
	CU(cudaMemcpyAsync(c->hist.p, hist_rgba, sizeof(float4) * c->npix, cudaMemcpyHostToDevice, c->stream));
	CU(cudaMemcpyAsync(c->counts.p, counts, sizeof(int32_t) * c->npix, cudaMemcpyHostToDevice, c->stream));
	CU(cudaStreamSynchronize(c->stream));
	return MCPT_OK;
}

int mcpt_get_first_sample(mcpt_ctx* c, uint32_t* first) {
	if (int r = check_ctx(c, false, false)) return r;
	if (!first) return fail(MCPT_ERR_ARG, "mcpt_get_first_sample: null output");
	CU(cudaMemcpyAsync(first, c->first_sample.p, sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
	CU(cudaStreamSynchronize(c->stream));
	return MCPT_OK;
}

int mcpt_bind_sums(mcpt_ctx* c, double* device_ptr) {
	if (int r = check_ctx(c, false, false)) return r;
	CU(cudaStreamSynchronize(c->stream));
	c->sums = device_ptr ? device_ptr : c->sums_own.p;
	return MCPT_OK;
}

// ---------------------------------------------------------------- multi-GPU merge over peer memory
// dst.sums += src_0.sums + src_1.sums + ... (in this order, so the doubles equal a host-side merge in context order).
// One kernel on dst's device reads the peers' accumulators in place through NVLink peer mappings; peers that cannot
// be mapped are staged through a device buffer with cudaMemcpyPeer.  No host copy of the 56 B/pixel sums.
int mcpt_merge_sums(mcpt_ctx* dst, mcpt_ctx* const* srcs, int nsrc) {
	if (int r = check_ctx(dst, false, false)) return r;
	if (nsrc < 0 || nsrc > MERGE_MAX_PEERS || (nsrc > 0 && !srcs))
		return fail(MCPT_ERR_ARG, "mcpt_merge_sums: expected 0..15 source contexts");
	const size_t count = 7 * size_t(dst->npix);
	MergePeers peers{};
	std::vector<DevBuf<double>> staged(static_cast<size_t>(nsrc));
	for (int k = 0; k < nsrc; ++k) {
		mcpt_ctx* s = srcs[k];
		if (!s || s == dst || s->npix != dst->npix) return fail(MCPT_ERR_ARG, "mcpt_merge_sums: sources must be other contexts of the same resolution");
		CU(cudaSetDevice(s->prm.device));
		CU(cudaStreamSynchronize(s->stream));              // the peer's samples are complete
		CU(cudaSetDevice(dst->prm.device));
		const double* ptr = s->sums;
		if (s->prm.device != dst->prm.device) {
			int can = 0;
			CU(cudaDeviceCanAccessPeer(&can, dst->prm.device, s->prm.device));
			cudaError_t e = can ? cudaDeviceEnablePeerAccess(s->prm.device, 0) : cudaErrorPeerAccessUnsupported;
			if (e == cudaErrorPeerAccessAlreadyEnabled) { e = cudaSuccess; (void)cudaGetLastError(); }
			if (e != cudaSuccess) {                         // no peer mapping: stage the plane set on dst's device
				(void)cudaGetLastError();
				CU(staged[size_t(k)].alloc(count));
				CU(cudaMemcpyPeerAsync(staged[size_t(k)].p, dst->prm.device, s->sums, s->prm.device, sizeof(double) * count, dst->stream));
				ptr = staged[size_t(k)].p;
			}
		}
		peers.p[k] = ptr;
	}
	peers.n = nsrc;
	if (nsrc > 0) {
		k_merge_sums<<<dst->persistent_grid(256), 256, 0, dst->stream>>>(dst->sums, peers, count);
		CU(cudaGetLastError());
	}
	CU(cudaStreamSynchronize(dst->stream));
	return MCPT_OK;
}

// ---------------------------------------------------------------- microbench / ad-hoc queries
int mcpt_trace_device(mcpt_ctx* c, const mcpt_ray* d_rays, long n, mcpt_hit* d_hits, int32_t* d_closest, int traversal) {
	if (int r = check_ctx(c, true, false)) return r;
	if (!d_rays || !d_hits || n < 0) return fail(MCPT_ERR_ARG, "mcpt_trace_device: bad arguments");
	if (n == 0) return MCPT_OK;
	return launch_intersect(c, d_rays, n, d_hits, d_closest, traversal, false);
}

int mcpt_trace_count(mcpt_ctx* c, const mcpt_ray* d_rays, long n, int traversal) {
	if (int r = check_ctx(c, true, false)) return r;
	if (!d_rays || n <= 0) return fail(MCPT_ERR_ARG, "mcpt_trace_count: bad arguments");
	if (c->t_hits.n < size_t(n)) CU(c->t_hits.alloc(n));
	CU(cudaMemsetAsync(c->counters.p, 0, sizeof(unsigned long long) * 3, c->stream));
	CU(cudaMemsetAsync(c->counters.p + 5, 0, sizeof(unsigned long long) * 3, c->stream));
	if (int r = launch_intersect(c, d_rays, n, c->t_hits.p, nullptr, traversal, true)) return r;
	CU(cudaStreamSynchronize(c->stream));
	return MCPT_OK;
}

int mcpt_trace(mcpt_ctx* c, const mcpt_ray* rays, long n, mcpt_hit* hits, int32_t* closest, int traversal) {
	if (int r = check_ctx(c, true, false)) return r;
	if (n < 0 || (n > 0 && (!rays || !hits))) return fail(MCPT_ERR_ARG, "mcpt_trace: bad arguments");
	if (n == 0) return MCPT_OK;
	// chunked so that a 64 Mi-ray batch needs bounded scratch; copies and kernels are stream ordered
	const long CH = 8L << 20;
	const long cap = n < CH ? n : CH;
	if (c->t_rays.n < size_t(cap)) { CU(c->t_rays.alloc(cap)); CU(c->t_hits.alloc(cap)); CU(c->t_closest.alloc(cap)); }
	if (c->t_hits.n < size_t(cap)) CU(c->t_hits.alloc(cap));
	for (long off = 0; off < n; off += CH) {
		const long m = n - off < CH ? n - off : CH;
		CU(cudaMemcpyAsync(c->t_rays.p, rays + off, sizeof(mcpt_ray) * m, cudaMemcpyHostToDevice, c->stream));
		if (int r = launch_intersect(c, c->t_rays.p, m, c->t_hits.p, c->t_closest.p, traversal, false)) return r;
		CU(cudaMemcpyAsync(hits + off, c->t_hits.p, sizeof(mcpt_hit) * m, cudaMemcpyDeviceToHost, c->stream));
		if (closest) CU(cudaMemcpyAsync(closest + off, c->t_closest.p, sizeof(int32_t) * m, cudaMemcpyDeviceToHost, c->stream));
	}
	CU(cudaStreamSynchronize(c->stream));
	return MCPT_OK;
}

// ---------------------------------------------------------------- allocation-free variant of mcpt_trace
// No counterpart in the reference: added for callers that must not call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  mcpt_trace stages the
// host rays / hits through device scratch it allocates on first use and k_intersect_persistent needs a zeroed
// work-fetch cursor; here all of that lives in ONE file-scope device array of fixed size, held by every GPU that loads
// the library until the process exits:
//   [0, 256)                       the cursor (cleared by every call)
//   + MAX * 48 B rays, + MAX * 48 B hits (cleared by every call: terminated rays leave their record untouched),
//   + MAX * 4 B closest-triangle ids
// Every call on a device shares the array and finds the previous call's contents in it; calls that overlap on the GPU
// (two host threads, two contexts of one device) race.  Calls are synchronous, so one caller at a time is safe.
constexpr long NOALLOC_MAX_RAYS = MCPT_TRACE_NOALLOC_MAX_RAYS;
constexpr size_t NOALLOC_HEADER = 256;
constexpr size_t NOALLOC_BYTES = NOALLOC_HEADER + size_t(NOALLOC_MAX_RAYS) * (sizeof(mcpt_ray) + sizeof(mcpt_hit) + sizeof(int32_t));
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_BYTES];

// The same walk as k_intersect_persistent, with rays / hits / closest / cursor carved out of the scratch array.
__global__ void __launch_bounds__(128, MCPT_MIN_BLOCKS) k_intersect_persistent_static(SceneDev sc, unsigned char* scratch, unsigned n, float tmin) {
	const mcpt_ray* rays = reinterpret_cast<const mcpt_ray*>(scratch + NOALLOC_HEADER);
	mcpt_hit* hits = reinterpret_cast<mcpt_hit*>(scratch + NOALLOC_HEADER + size_t(NOALLOC_MAX_RAYS) * sizeof(mcpt_ray));
	int32_t* closest = reinterpret_cast<int32_t*>(scratch + NOALLOC_HEADER + size_t(NOALLOC_MAX_RAYS) * (sizeof(mcpt_ray) + sizeof(mcpt_hit)));
	IntersectJob job{sc, rays, hits, closest};
	persistent_walk(sc, tmin, n, reinterpret_cast<unsigned*>(scratch), job);
}

// Device address of the array on `device` (which must be current): in host code the array's NAME is only a stand-in, a
// kernel handed it would fault.  Looked up once per device.
static int noalloc_scratch(int device, unsigned char** out) {
	static std::mutex m;
	static unsigned char* table[64] = {};
	if (device < 0 || device >= 64) return fail(MCPT_ERR_ARG, "mcpt_trace_noalloc: device ordinal out of range");
	std::lock_guard<std::mutex> lock(m);
	if (!table[device]) {
		void* p = nullptr;
		CU(cudaGetSymbolAddress(&p, g_noalloc_scratch));
		table[device] = static_cast<unsigned char*>(p);
	}
	*out = table[device];
	return MCPT_OK;
}

int mcpt_trace_noalloc(mcpt_ctx* c, const mcpt_ray* rays, long n, mcpt_hit* hits, int32_t* closest) {
	if (int r = check_ctx(c, true, false)) return r;          // cudaSetDevice(ctx device): the lookup below needs it current
	if (n < 0 || (n > 0 && (!rays || !hits))) return fail(MCPT_ERR_ARG, "mcpt_trace_noalloc: bad arguments");
	if (n > NOALLOC_MAX_RAYS)
		return fail(MCPT_ERR_ARG, "mcpt_trace_noalloc: " + std::to_string(n) + " rays exceed the maximum of " +
		            std::to_string(NOALLOC_MAX_RAYS) + " per call (MCPT_TRACE_NOALLOC_MAX_RAYS; use mcpt_trace or split the batch)");
	if (n == 0) return MCPT_OK;
	unsigned char* scratch = nullptr;
	if (int r = noalloc_scratch(c->prm.device, &scratch)) return r;
	mcpt_ray* d_rays = reinterpret_cast<mcpt_ray*>(scratch + NOALLOC_HEADER);
	mcpt_hit* d_hits = reinterpret_cast<mcpt_hit*>(scratch + NOALLOC_HEADER + size_t(NOALLOC_MAX_RAYS) * sizeof(mcpt_ray));
	int32_t* d_closest = reinterpret_cast<int32_t*>(scratch + NOALLOC_HEADER + size_t(NOALLOC_MAX_RAYS) * (sizeof(mcpt_ray) + sizeof(mcpt_hit)));
	CU(cudaMemsetAsync(scratch, 0, NOALLOC_HEADER, c->stream));                       // the cursor starts at zero
	CU(cudaMemsetAsync(d_hits, 0, sizeof(mcpt_hit) * size_t(n), c->stream));            // no stale records of the last call
	CU(cudaMemcpyAsync(d_rays, rays, sizeof(mcpt_ray) * size_t(n), cudaMemcpyHostToDevice, c->stream));
	if (c->force_reference)      // scene the product walk cannot take (boxes not nested): the reference-order kernel, same scratch
		k_intersect<MCPT_TRAVERSAL_REFERENCE, false><<<blocks(n, 128), 128, 0, c->stream>>>(c->scene, d_rays, n, c->prm.tmin, d_hits, d_closest, nullptr);
	else
		k_intersect_persistent_static<<<std::min(blocks(n, 128), c->persistent_grid(128)), 128, 0, c->stream>>>(c->scene, scratch, unsigned(n), c->prm.tmin);
	++c->launches;
	CU(cudaGetLastError());
	CU(cudaMemcpyAsync(hits, d_hits, sizeof(mcpt_hit) * size_t(n), cudaMemcpyDeviceToHost, c->stream));
	if (closest) CU(cudaMemcpyAsync(closest, d_closest, sizeof(int32_t) * size_t(n), cudaMemcpyDeviceToHost, c->stream));
	CU(cudaStreamSynchronize(c->stream));
	return MCPT_OK;
}

int mcpt_check_failures(int device, uint64_t* out4) {
#if MCPT_CHECKED
	if (!out4) return fail(MCPT_ERR_ARG, "mcpt_check_failures: null output");
	unsigned long long h[4] = {0, 0, 0, 0};
	CU(cudaSetDevice(device));
	CU(cudaDeviceSynchronize());
	CU(cudaMemcpyFromSymbol(h, g_check, sizeof(h)));
	for (int k = 0; k < 4; ++k) out4[k] = h[k];
	return MCPT_OK;
#else
	(void)device; (void)out4;
	return fail(MCPT_ERR_STATE, "mcpt_check_failures: this is the product build (bounds checks are compiled in with -DMCPT_CHECKED=1, tools/checked_run.sh)");
#endif
}

int mcpt_phong_accept_table(float* out, int n) {
	const std::vector<float>& tab = phong_accept_table();
	if (!out || n != int(tab.size())) return fail(MCPT_ERR_ARG, "mcpt_phong_accept_table: expected a buffer of " + std::to_string(tab.size()) + " floats");
	std::memcpy(out, tab.data(), sizeof(float) * tab.size());
	return MCPT_OK;
}

int mcpt_set_profiling(mcpt_ctx* c, int count_fetches, int time_bounces) {
	if (int r = check_ctx(c, false, false)) return r;
	CU(cudaStreamSynchronize(c->stream));
	CU(c->drain_events());
	c->count_fetches = count_fetches != 0;
	c->time_bounces = time_bounces != 0;
	return MCPT_OK;
}

int mcpt_get_bounce_time(mcpt_ctx* c, double* ms, uint64_t* launches) {
	if (int r = check_ctx(c, false, false)) return r;
	CU(cudaStreamSynchronize(c->stream));
	CU(c->drain_events());
	if (ms) *ms = c->bounce_ms;
	if (launches) *launches = c->bounce_launches;
	c->bounce_ms = 0.0;
	c->bounce_launches = 0;
	return MCPT_OK;
}

int mcpt_get_stats(mcpt_ctx* c, mcpt_stats* out) {
	if (int r = check_ctx(c, false, false)) return r;
	if (!out) return fail(MCPT_ERR_ARG, "mcpt_get_stats: null output");
	unsigned long long h[8];
	CU(cudaMemcpyAsync(h, c->counters.p, sizeof(h), cudaMemcpyDeviceToHost, c->stream));
	CU(cudaStreamSynchronize(c->stream));
	out->nodes_fetched = h[0];
	out->tris_fetched = h[1];
	out->leaves_fetched = h[5];
	out->slow_rays = h[6];
	out->replayed_rays = h[7];
	out->rays_traced = h[2] + h[3];
	out->paths = h[4] + c->stage_paths;
	out->kernel_launches = c->launches;
	return MCPT_OK;
}

int mcpt_reset_stats(mcpt_ctx* c) {
	if (int r = check_ctx(c, false, false)) return r;
	CU(cudaMemsetAsync(c->counters.p, 0, sizeof(unsigned long long) * 8, c->stream));
	CU(cudaStreamSynchronize(c->stream));
	c->launches = 0;
	c->stage_paths = 0;
	return MCPT_OK;
}

int mcpt_get_queue_sorting(mcpt_ctx* c, int* out) {
	if (int r = check_ctx(c, false, false)) return r;
	if (!out) return fail(MCPT_ERR_ARG, "mcpt_get_queue_sorting: null output");
	if (!c->sort_queues) *out = 0;
	else if (c->prm.sampling == MCPT_SAMPLING_REFERENCE) *out = 1;
	else if (!c->mis_sort_tune) *out = c->mis_sort_mask;
	else *out = c->mis_sort_choice;
	return MCPT_OK;
}

}  // extern "C"
