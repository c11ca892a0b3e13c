// mcpt — headless command-line renderer: the reference's main.cpp / OpenCLApp.cpp with the OpenCL layer
// replaced by the CUDA C ABI (include/mcpt.h) and the GLUT window replaced by file output.
//
//   reference                               here
//   Config::CONFIG()        config.cpp:70   mcpt_host_load_config(./config.json or --config, --configid)
//   OpenCL::init()          OpenCLApp.cpp:36 App::init(): loadObject -> HLBVH -> mcpt_create/upload/set_camera
//   OpenCL::update()        OpenCLApp.cpp:57 App::update(): mcpt_render(batch) (one call renders many frames)
//   ColorOut (attempt+1 frames, then <objname>.hdr) colorout.cpp:55-68   the loop in main(); also <objname>.png
//
// Multi-GPU (--gpus N, one host thread per device): each device renders a contiguous block of sample indices
// (seed_mode 1) into its own accumulators; the sums are merged by one kernel over NVLink peer memory (mcpt_merge_sums).
// --procs N: the same split with one PROCESS per device (forked before the first CUDA call, each seeing only its own
// device), because a process that sees eight devices spends ~5 s in its first CUDA call; the workers write their sums
// into a shared mapping and the parent adds them in rank order - the same doubles as mcpt_merge_sums.
// The torchrun path (bench.py) merges with one NCCL reduce instead.
//
// Scene cache: what loadObject + the BVH build + the layout derivation produce is kept in
// <directory><objname>.<bvhtype>.mcptcache, keyed by the content of the OBJ / MTL files (--no-cache turns it off).
// Checkpoints (colorout.cpp:23-24,49-50 keeps the progressive state on the device for the whole run): --checkpoint FILE
// [--checkpoint-every K] writes hist / counts / sums / seeds every K frames, --resume FILE continues bit for bit.
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <algorithm>
#include <atomic>
#include <mutex>
#include <vector>

#include <sys/mman.h>
#include <sys/wait.h>
#include <unistd.h>

#include "mcpt.h"
#include "mcpt_host.h"

namespace MCPT {
namespace App {

struct State {
	mcpt_host_config cfg{};
	mcpt_triangle* tris = nullptr;
	mcpt_material* mats = nullptr;
	int* mat_ids = nullptr;
	int ntris = 0, nmats = 0;
	std::vector<mcpt_bvh_node> nodes;
	void* derived = nullptr;     // traversal layout blob (mcpt_derive_scene / scene cache); nullptr = derive at upload
	size_t derived_bytes = 0;
	bool use_cache = true;       // --no-cache
	const char* cache_state = "off";
	double load_s = 0.0;         // OBJ/MTL (or cache) -> triangles, materials, nodes, layout blob
	mutable double create_s = 0.0, upload_s = 0.0;   // slowest device: mcpt_create (first CUDA call of its thread / process), scene upload
	mutable std::mutex timing_mutex;
	std::vector<mcpt_ctx*> ctx;
	int spp = 0;             // config "attempt"
	uint32_t run_seed = 1;
	bool reference_seeding = true;
	bool mis = false;        // --mis: NEE + MIS sampling mode (same image, far less variance; needs order-independent seeds)
	bool mis_ratio = false;  // --mis-ratio: as --mis, with the denominator P(sample != 0) estimated by NEE + MIS as well
	bool host_bvh = false;   // --host-bvh: build the LBVH with the kept host code instead of the device builder
	bool overlap_init = false;   // --overlap-init (multi-GPU): every device starts rendering as soon as IT is ready
	double first_ready_s = 0.0, all_ready_s = 0.0;
};

[[noreturn]] void die(const char* what, const char* detail) {
	std::fprintf(stderr, "mcpt: %s: %s\n", what, detail ? detail : "");
	std::exit(1);
}

// One model of the BVH quality harness (BVH::TEST::test / testmodel, bvhtest.cpp:448-613): same printed lines,
// metrics from mcpt_bvh_metrics.  All three "bvhtype" branches of the reference end in the same three numbers.
void test_model(const mcpt_host_config& cfg, const char* objname, bool with_camera, bool host_bvh) {
	mcpt_triangle* tris = nullptr;
	mcpt_material* mats = nullptr;
	int* ids = nullptr;
	int ntris = 0, nmats = 0;
	if (mcpt_host_load_object_mt(cfg.directory, objname, 1, 0, &tris, &ntris, &mats, &nmats, &ids) < 0)
		die("loadObj", mcpt_host_last_error());
	std::printf("%s %d\n%s\n", objname, ntris, cfg.bvhtype);
	std::vector<mcpt_bvh_node> nodes(size_t(2 * ntris - 1));
	const std::string bvhtype = cfg.bvhtype;
	if (host_bvh) {
		if (mcpt_host_build_bvh(tris, ntris, cfg.bvhtype, nodes.data()) < 0) die("BVH", mcpt_host_last_error());
	} else {
		if (bvhtype != "hlbvh" && bvhtype != "treelet" && bvhtype != "treeletGPU") die("BVH", "BVH Not Implemented");
		if (mcpt_build_lbvh(0, tris, ntris, nodes.data(), nullptr) < 0) die("BVH (device)", mcpt_last_error());
		if (bvhtype != "hlbvh" && mcpt_optimize_treelet(0, nodes.data(), int(nodes.size()), nodes.data(), nullptr) < 0)
			die("treelet (device)", mcpt_last_error());
	}
	mcpt_bvh_quality q{};
	if (mcpt_bvh_metrics(0, nodes.data(), int(nodes.size()), tris, ntris, with_camera ? &cfg.camera : nullptr,
	                     cfg.width, cfg.height, &q, nullptr, nullptr) < 0)
		die("metrics", mcpt_last_error());
	std::printf("SAH: %g\nEPO_GPU: %g\n", q.sah, q.epo);       // bvhtest.cpp:476,487
	if (with_camera) std::printf("LCV: %g\n", q.lcv);           // :494
	mcpt_host_free(tris);
	mcpt_host_free(mats);
	mcpt_host_free(ids);
}

mcpt_ctx* make_context(const State& s, int g, int ngpus) {
	mcpt_params p{};
	p.width = s.cfg.width;
	p.height = s.cfg.height;
	p.max_depth = s.cfg.maxdepth;
	p.max_attempt = s.cfg.attempt;
	p.device = g;
	p.seed_mode = (ngpus > 1 || !s.reference_seeding || s.mis) ? 1 : 0;
	p.sampling = s.mis ? (s.mis_ratio ? MCPT_SAMPLING_NEE_MIS_RATIO : MCPT_SAMPLING_NEE_MIS) : MCPT_SAMPLING_REFERENCE;
	p.run_seed = s.run_seed;
	p.traversal = MCPT_TRAVERSAL_FAST;
	auto t0 = std::chrono::steady_clock::now();
	mcpt_ctx* c = mcpt_create(&p);
	if (!c) die("mcpt_create", mcpt_last_error());
	auto t1 = std::chrono::steady_clock::now();
	int rc = s.derived ? mcpt_upload_scene_derived(c, s.tris, s.ntris, s.mats, s.nmats, s.nodes.data(), int(s.nodes.size()),
	                                                s.derived, s.derived_bytes, 0)
	                   : mcpt_upload_scene(c, s.tris, s.ntris, s.mats, s.nmats, s.nodes.data(), int(s.nodes.size()));
	if (rc && s.derived) {
		// a cached layout this library does not accept (written by another version of the builder): derive it afresh
		std::fprintf(stderr, "mcpt: cached traversal layout not used (%s)\n", mcpt_last_error());
		rc = mcpt_upload_scene(c, s.tris, s.ntris, s.mats, s.nmats, s.nodes.data(), int(s.nodes.size()));
	}
	if (rc || mcpt_set_camera(c, &s.cfg.camera)) die("scene upload", mcpt_last_error());
	auto t2 = std::chrono::steady_clock::now();
	{
		std::lock_guard<std::mutex> lock(s.timing_mutex);
		s.create_s = std::max(s.create_s, std::chrono::duration<double>(t1 - t0).count());
		s.upload_s = std::max(s.upload_s, std::chrono::duration<double>(t2 - t1).count());
	}
	return c;
}

// config -> triangles, materials, BVH nodes and the traversal layout blob: from the scene cache when its key (content
// hash of OBJ + MTL + bvhtype) matches, else loadObject (parallel) + the BVH build + mcpt_derive_scene, then stored.
void load_scene(State& s) {
	auto t0 = std::chrono::steady_clock::now();
	if (!s.cfg.opencl) die("config", "Not Implemented");            // raygeneration.cpp:36-38 throws the same
	const std::string bvhtype = s.cfg.bvhtype;
	mcpt_scene_key key{};
	char path[1024] = "";
	if (s.use_cache) {
		if (mcpt_host_scene_key(s.cfg.directory, s.cfg.objname, s.cfg.bvhtype, 1, &key) < 0) die("loadObject", mcpt_host_last_error());
		if (mcpt_host_cache_path(s.cfg.directory, s.cfg.objname, s.cfg.bvhtype, path, sizeof(path)) < 0) die("scene cache", mcpt_host_last_error());
		mcpt_bvh_node* nodes = nullptr;
		int nnodes = 0;
		const int rc = mcpt_host_cache_load(path, &key, &s.tris, &s.ntris, &s.mats, &s.nmats, &s.mat_ids, &nodes, &nnodes,
		                                    &s.derived, &s.derived_bytes);
		if (rc < 0) die("scene cache", mcpt_host_last_error());
		if (rc == 0) {
			s.nodes.assign(nodes, nodes + nnodes);
			mcpt_host_free(nodes);
			if (s.derived_bytes == 0) { mcpt_host_free(s.derived); s.derived = nullptr; }
			s.cache_state = "hit";
			s.load_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
			return;
		}
		s.cache_state = "miss";
	}
	if (mcpt_host_load_object_mt(s.cfg.directory, s.cfg.objname, 1, 0, &s.tris, &s.ntris, &s.mats, &s.nmats, &s.mat_ids) < 0)
		die("loadObject", mcpt_host_last_error());
	s.nodes.resize(size_t(2 * s.ntris - 1));
	// built on GPU 0 (same bytes as the host build, ~100x faster); --host-bvh keeps the host code
	if (!s.host_bvh && (bvhtype == "hlbvh" || bvhtype == "treelet" || bvhtype == "treeletGPU")) {
		if (mcpt_build_lbvh(0, s.tris, s.ntris, s.nodes.data(), nullptr) < 0) die("BVH (device)", mcpt_last_error());
		// both treelet spellings: the device optimiser, whose output equals TreeletBVH<CPU>'s byte for byte
		if (bvhtype != "hlbvh" && mcpt_optimize_treelet(0, s.nodes.data(), int(s.nodes.size()), s.nodes.data(), nullptr) < 0)
			die("treelet (device)", mcpt_last_error());
	} else if (mcpt_host_build_bvh(s.tris, s.ntris, s.cfg.bvhtype, s.nodes.data()) < 0) die("BVH", mcpt_host_last_error());
	if (s.use_cache) {
		if (mcpt_derive_scene(s.tris, s.ntris, s.mats, s.nmats, s.nodes.data(), int(s.nodes.size()), &s.derived, &s.derived_bytes))
			die("scene layout", mcpt_last_error());
		// a read-only scene directory is not an error: the next start simply derives again
		if (mcpt_host_cache_store(path, &key, s.tris, s.ntris, s.mats, s.nmats, s.mat_ids, s.nodes.data(), int(s.nodes.size()),
		                          s.derived, s.derived_bytes) < 0)
			std::fprintf(stderr, "mcpt: scene cache not written: %s\n", mcpt_host_last_error());
	}
	s.load_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

void init(State& s, int ngpus) {
	if (!s.tris) load_scene(s);
	if (!s.cfg.has_camera) die("config", "no camera");
	if (s.cfg.res_x != s.cfg.width || s.cfg.res_y != s.cfg.height)
		die("config", "camera.resolution must equal width/height (SURVEY A.5)");
	// one host thread per device: context creation, buffers and uploads overlap; the geometry-derived layout
	// (SAH tree of the fast walk) is built by the first thread and shared through the library's per-process cache
	s.ctx.assign(size_t(ngpus), nullptr);
	if (s.overlap_init) return;                                     // run_overlapped() creates them
	std::vector<std::thread> th;
	for (int g = 0; g < ngpus; ++g) th.emplace_back([&s, g, ngpus] { s.ctx[size_t(g)] = make_context(s, g, ngpus); });
	for (auto& t : th) t.join();
}

// ---- checkpoint file: header | hist float4[W*H] | counts int32[W*H] | sums double[7*W*H] | seeds uint32[W*H]
struct CheckpointHeader {
	char magic[8];               // "MCPTCKP1"
	int32_t width, height, maxdepth, seed_mode;
	uint32_t run_seed, frames_done, first_sample, sampling;
};

void write_checkpoint(const State& s, const std::string& file, int frames_done) {
	mcpt_ctx* c = s.ctx[0];
	const size_t npix = size_t(s.cfg.width) * size_t(s.cfg.height);
	std::vector<float> hist(4 * npix);
	std::vector<int32_t> counts(npix);
	std::vector<double> sums(7 * npix);
	std::vector<uint32_t> seeds(npix);
	CheckpointHeader h{};
	std::memcpy(h.magic, "MCPTCKP1", 8);
	h.width = s.cfg.width; h.height = s.cfg.height; h.maxdepth = s.cfg.maxdepth;
	h.seed_mode = (!s.reference_seeding || s.mis) ? 1 : 0;
	h.run_seed = s.run_seed; h.frames_done = uint32_t(frames_done); h.sampling = s.mis ? (s.mis_ratio ? 2u : 1u) : 0u;
	if (mcpt_read_history(c, hist.data(), counts.data()) || mcpt_read_sums(c, sums.data()) || mcpt_read_seeds(c, seeds.data()) ||
	    mcpt_get_first_sample(c, &h.first_sample))
		die("checkpoint", mcpt_last_error());
	const std::string tmp = file + ".tmp";
	std::FILE* f = std::fopen(tmp.c_str(), "wb");
	if (!f) die("checkpoint: cannot write", tmp.c_str());
	bool ok = std::fwrite(&h, sizeof(h), 1, f) == 1 && std::fwrite(hist.data(), sizeof(float), hist.size(), f) == hist.size() &&
	          std::fwrite(counts.data(), sizeof(int32_t), npix, f) == npix && std::fwrite(sums.data(), sizeof(double), sums.size(), f) == sums.size() &&
	          std::fwrite(seeds.data(), sizeof(uint32_t), npix, f) == npix;
	ok = (std::fclose(f) == 0) && ok;
	if (!ok || std::rename(tmp.c_str(), file.c_str()) != 0) die("checkpoint: write failed", file.c_str());
}

// returns the number of frames the checkpointed run had finished
int read_checkpoint(State& s, const std::string& file) {
	mcpt_ctx* c = s.ctx[0];
	const size_t npix = size_t(s.cfg.width) * size_t(s.cfg.height);
	std::FILE* f = std::fopen(file.c_str(), "rb");
	if (!f) die("resume: cannot open", file.c_str());
	CheckpointHeader h{};
	std::vector<float> hist(4 * npix);
	std::vector<int32_t> counts(npix);
	std::vector<double> sums(7 * npix);
	std::vector<uint32_t> seeds(npix);
	if (std::fread(&h, sizeof(h), 1, f) != 1 || std::memcmp(h.magic, "MCPTCKP1", 8) != 0) die("resume", "not a checkpoint file");
	const int seed_mode = (!s.reference_seeding || s.mis) ? 1 : 0;
	if (h.width != s.cfg.width || h.height != s.cfg.height || h.maxdepth != s.cfg.maxdepth || h.seed_mode != seed_mode ||
	    h.run_seed != s.run_seed || h.sampling != (s.mis ? (s.mis_ratio ? 2u : 1u) : 0u))
		die("resume", "checkpoint was written with another size, depth, seed or seeding / sampling mode");
	const bool ok = std::fread(hist.data(), sizeof(float), hist.size(), f) == hist.size() && std::fread(counts.data(), sizeof(int32_t), npix, f) == npix &&
	                std::fread(sums.data(), sizeof(double), sums.size(), f) == sums.size() && std::fread(seeds.data(), sizeof(uint32_t), npix, f) == npix;
	std::fclose(f);
	if (!ok) die("resume", "truncated checkpoint");
	if (mcpt_write_history(c, hist.data(), counts.data()) || mcpt_write_sums(c, sums.data())) die("resume", mcpt_last_error());
	if (seed_mode == 0) { if (mcpt_write_seeds(c, seeds.data())) die("resume", mcpt_last_error()); }
	else if (mcpt_set_first_sample(c, h.first_sample)) die("resume", mcpt_last_error());
	return int(h.frames_done);
}

// Renders `frames` more frames on every device (device g takes its share of the sample indices).
void update(State& s, int first_frame, int frames) {
	const int G = int(s.ctx.size());
	std::vector<std::thread> th;
	for (int g = 0; g < G; ++g) {
		int base = frames / G, rem = frames % G;
		int count = base + (g < rem ? 1 : 0);
		int first = first_frame + g * base + (g < rem ? g : rem);
		th.emplace_back([&, g, first, count] {
			if (G > 1 && mcpt_set_first_sample(s.ctx[g], uint32_t(first))) die("mcpt_set_first_sample", mcpt_last_error());
			if (mcpt_render(s.ctx[g], count)) die("mcpt_render", mcpt_last_error());
		});
	}
	for (auto& t : th) t.join();
}

// --overlap-init: CUDA brings the devices of a process up one after the other (about a second each on an 8-GPU
// box), so with a static split everybody waits for the last one.  Here each device's thread creates its context and
// then pulls blocks of sample indices from a shared counter until none are left: the device that is ready first
// renders most.  Same samples, same sums (seed_mode 1 makes a sample a function of its index alone).
void run_overlapped(State& s, int frames, int batch, std::chrono::steady_clock::time_point t0) {
	const int G = int(s.ctx.size());
	std::atomic<int> next{0};
	std::vector<double> ready(size_t(G), 0.0);
	std::vector<std::thread> th;
	for (int g = 0; g < G; ++g) {
		th.emplace_back([&, g] {
			mcpt_ctx* c = make_context(s, g, G);
			s.ctx[size_t(g)] = c;
			ready[size_t(g)] = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
			for (;;) {
				const int first = next.fetch_add(batch);
				if (first >= frames) break;
				const int count = std::min(batch, frames - first);
				if (mcpt_set_first_sample(c, uint32_t(first))) die("mcpt_set_first_sample", mcpt_last_error());
				if (mcpt_render(c, count)) die("mcpt_render", mcpt_last_error());
			}
		});
	}
	for (auto& t : th) t.join();
	s.first_ready_s = *std::min_element(ready.begin(), ready.end());
	s.all_ready_s = *std::max_element(ready.begin(), ready.end());
}

// ---- --procs N: one process per device
std::vector<pid_t> g_unreaped;       // workers whose results are in but whose processes are still exiting
void reap_workers() {
	for (pid_t p : g_unreaped) { int st = 0; waitpid(p, &st, 0); }
	g_unreaped.clear();
}
struct WorkerShared { int turn; int next_block; };   // --overlap-init: whose turn it is to initialise CUDA; next unrendered frame
struct WorkerResult {
	int ok;
	uint64_t rays, paths;
	double ready_s, done_s;      // since t0 (CLOCK_MONOTONIC is shared by the processes of a box)
	double create_s, upload_s;   // mcpt_create (the process's first CUDA call), scene upload
	char error[200];
};

// the g-th device this process may use: the g-th entry of an inherited CUDA_VISIBLE_DEVICES list, else ordinal g
std::string visible_device(int g) {
	const char* v = std::getenv("CUDA_VISIBLE_DEVICES");
	if (!v || !*v) return std::to_string(g);
	std::string list = v;
	size_t pos = 0;
	for (int k = 0; k < g; ++k) {
		pos = list.find(',', pos);
		if (pos == std::string::npos) die("--procs", "fewer devices in CUDA_VISIBLE_DEVICES than processes");
		++pos;
	}
	size_t end = list.find(',', pos);
	return list.substr(pos, end == std::string::npos ? std::string::npos : end - pos);
}

// Forks `procs` workers (no CUDA call has happened in this process: the scene came from the cache, or from a helper
// child that built it).  Worker g renders its block of sample indices on its own device and writes the seven sum planes
// into its slice of a shared mapping; the parent adds the slices in rank order ((s0 + s1) + s2 ... per element, the order
// of mcpt_merge_sums) on host threads.
void run_processes(State& s, int procs, int frames, std::chrono::steady_clock::time_point t0, std::vector<double>& acc,
                   mcpt_stats& tot, double& all_ready_s, double& first_ready_s, double& render_done_s) {
	const size_t npix = size_t(s.cfg.width) * size_t(s.cfg.height);
	const size_t plane_bytes = 7 * npix * sizeof(double);
	const size_t bytes = size_t(procs) * plane_bytes + size_t(procs) * sizeof(WorkerResult) + sizeof(WorkerShared);
	void* shm = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS, -1, 0);
	if (shm == MAP_FAILED) die("--procs", "mmap of the shared accumulators failed");
	double* sums = static_cast<double*>(shm);
	WorkerResult* res = reinterpret_cast<WorkerResult*>(static_cast<char*>(shm) + size_t(procs) * plane_bytes);
	WorkerShared* shared = reinterpret_cast<WorkerShared*>(res + procs);        // zero-filled by mmap
	const bool staggered = s.overlap_init;
	const int block = 32;                    // frames per pull in the staggered mode
	std::vector<pid_t> pids;
	for (int g = 0; g < procs; ++g) {
		const int base = frames / procs, rem = frames % procs;
		const int count = base + (g < rem ? 1 : 0);
		const int first = g * base + (g < rem ? g : rem);
		const std::string dev = visible_device(g);
		pid_t pid = fork();
		if (pid < 0) die("--procs", "fork failed");
		if (pid == 0) {
			WorkerResult& r = res[g];
			bool holds_turn = false;
			auto bail = [&](const char* what, const char* detail) {
				std::snprintf(r.error, sizeof(r.error), "%s: %s", what, detail ? detail : "");
				if (holds_turn) __sync_fetch_and_add(&shared->turn, 1);      // never leave the others waiting
				_exit(2);
			};
			setenv("CUDA_VISIBLE_DEVICES", dev.c_str(), 1);
			if (staggered) {
				// The driver initialises one process at a time whatever we do (eight starting at once each take 9.4 s on an
				// 8 x B200 box, one alone 0.7 s: profiles/r2s_startup_probe_8gpu.log), so the workers queue up explicitly:
				// worker g makes its first CUDA call when worker g - 1 has its context, and starts rendering at once.
				while (*const_cast<volatile int*>(&shared->turn) != g) usleep(200);
				holds_turn = true;
			}
			mcpt_params p{};
			p.width = s.cfg.width; p.height = s.cfg.height; p.max_depth = s.cfg.maxdepth; p.max_attempt = s.cfg.attempt;
			p.device = 0; p.seed_mode = 1; p.sampling = s.mis ? (s.mis_ratio ? MCPT_SAMPLING_NEE_MIS_RATIO : MCPT_SAMPLING_NEE_MIS) : MCPT_SAMPLING_REFERENCE;
			p.run_seed = s.run_seed; p.traversal = MCPT_TRAVERSAL_FAST;
			auto w0 = std::chrono::steady_clock::now();
			mcpt_ctx* c = mcpt_create(&p);
			if (!c) bail("mcpt_create", mcpt_last_error());
			r.create_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - w0).count();
			if (staggered) { __sync_fetch_and_add(&shared->turn, 1); holds_turn = false; }
			int rc = s.derived ? mcpt_upload_scene_derived(c, s.tris, s.ntris, s.mats, s.nmats, s.nodes.data(), int(s.nodes.size()), s.derived, s.derived_bytes, 0)
			                   : mcpt_upload_scene(c, s.tris, s.ntris, s.mats, s.nmats, s.nodes.data(), int(s.nodes.size()));
			if (rc && s.derived) rc = mcpt_upload_scene(c, s.tris, s.ntris, s.mats, s.nmats, s.nodes.data(), int(s.nodes.size()));
			if (rc || mcpt_set_camera(c, &s.cfg.camera)) bail("scene upload", mcpt_last_error());
			r.ready_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
			r.upload_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - w0).count() - r.create_s;
			if (staggered) {
				// blocks of sample indices from a shared counter: whoever is ready first renders most (seed_mode 1 makes a
				// sample a function of its index alone, so the set of samples - not their owner - defines the image)
				for (;;) {
					const int b0 = __sync_fetch_and_add(&shared->next_block, block);
					if (b0 >= frames) break;
					if (mcpt_set_first_sample(c, uint32_t(b0)) || mcpt_render(c, std::min(block, frames - b0))) bail("mcpt_render", mcpt_last_error());
				}
			} else if (count > 0 && (mcpt_set_first_sample(c, uint32_t(first)) || mcpt_render(c, count))) bail("mcpt_render", mcpt_last_error());
			if (mcpt_read_sums(c, sums + size_t(g) * 7 * npix)) bail("readback", mcpt_last_error());
			mcpt_stats st{};
			mcpt_get_stats(c, &st);
			r.rays = st.rays_traced; r.paths = st.paths;
			r.done_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
			__sync_synchronize();          // the sums and the record above are visible before the flag
			*const_cast<volatile int*>(&r.ok) = 1;
			_exit(0);      // no destructors: the driver tears the context down with the process
		}
		pids.push_back(pid);
	}
	// A worker's sums are complete when it sets `ok` (its last store before _exit); its process then spends another
	// second or two tearing its CUDA context down, which the image does not have to wait for: poll the flags, notice
	// early deaths through waitpid(WNOHANG), reap everybody at the very end (reap_workers).
	bool failed = false;
	std::vector<char> finished(size_t(procs), 0), reaped(size_t(procs), 0);
	for (int left = procs; left > 0 && !failed;) {
		for (int g = 0; g < procs; ++g) {
			if (finished[size_t(g)]) continue;
			if (*const_cast<volatile int*>(&res[g].ok)) { finished[size_t(g)] = 1; --left; continue; }
			int status = 0;
			if (!reaped[size_t(g)] && waitpid(pids[size_t(g)], &status, WNOHANG) == pids[size_t(g)]) {
				reaped[size_t(g)] = 1;
				if (!*const_cast<volatile int*>(&res[g].ok)) {
					std::fprintf(stderr, "mcpt: worker %d failed: %s\n", g, res[g].error[0] ? res[g].error : "(killed)");
					failed = true;
				}
			}
		}
		if (left > 0) usleep(500);
	}
	if (failed) std::exit(1);
	__sync_synchronize();
	for (int g = 0; g < procs; ++g) if (!reaped[size_t(g)]) g_unreaped.push_back(pids[size_t(g)]);
	all_ready_s = first_ready_s = res[0].ready_s;
	render_done_s = res[0].done_s;
	for (int g = 0; g < procs; ++g) {
		all_ready_s = std::max(all_ready_s, res[g].ready_s);
		first_ready_s = std::min(first_ready_s, res[g].ready_s);
		render_done_s = std::max(render_done_s, res[g].done_s);
		tot.rays_traced += res[g].rays;
		tot.paths += res[g].paths;
		s.create_s = std::max(s.create_s, res[g].create_s);
		s.upload_s = std::max(s.upload_s, res[g].upload_s);
	}
	acc.resize(7 * npix);
	const unsigned nt = std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
	std::vector<std::thread> th;
	for (unsigned t = 0; t < nt; ++t)
		th.emplace_back([&, t] {
			const size_t lo = 7 * npix * t / nt, hi = 7 * npix * (t + 1) / nt;
			for (size_t i = lo; i < hi; ++i) {
				double v = sums[i];
				for (int g = 1; g < procs; ++g) v += sums[size_t(g) * 7 * npix + i];
				acc[i] = v;
			}
		});
	for (auto& x : th) x.join();
	munmap(shm, bytes);
}

}  // namespace App
}  // namespace MCPT

int main(int argc, char** argv) {
	using namespace MCPT::App;
	std::string config = "./config.json", out_prefix, checkpoint_file, resume_file;
	int configid = -1, gpus = 1, spp_override = -1, batch = 64, procs = 0, checkpoint_every = 0, stop_after = -1;
	State s;
	for (int i = 1; i < argc; ++i) {
		std::string a = argv[i];
		auto next = [&]() -> const char* { if (i + 1 >= argc) die("missing value for", a.c_str()); return argv[++i]; };
		if (a == "--config") config = next();
		else if (a == "--configid") configid = std::atoi(next());
		else if (a == "--gpus") gpus = std::atoi(next());
		else if (a == "--procs") procs = std::atoi(next());
		else if (a == "--spp") spp_override = std::atoi(next());
		else if (a == "--out") out_prefix = next();
		else if (a == "--seed") s.run_seed = uint32_t(std::strtoul(next(), nullptr, 10));
		else if (a == "--batch") batch = std::atoi(next());
		else if (a == "--order-independent-seeds") s.reference_seeding = false;
		else if (a == "--host-bvh") s.host_bvh = true;
		else if (a == "--mis") s.mis = true;
		else if (a == "--mis-ratio") s.mis = s.mis_ratio = true;
		else if (a == "--overlap-init") s.overlap_init = true;
		else if (a == "--no-cache") s.use_cache = false;
		else if (a == "--checkpoint") checkpoint_file = next();
		else if (a == "--checkpoint-every") checkpoint_every = std::atoi(next());
		else if (a == "--resume") resume_file = next();
		else if (a == "--stop-after") stop_after = std::atoi(next());      // leave after this many frames (an interrupted run)
		else die("unknown option (use --config --configid --gpus --procs --spp --out --seed --batch --order-independent-seeds --host-bvh --mis --mis-ratio --overlap-init --no-cache --checkpoint --checkpoint-every --resume --stop-after)", a.c_str());
	}
	if (mcpt_host_load_config(config.c_str(), configid, &s.cfg) < 0) die("config", mcpt_host_last_error());
	if (s.cfg.testall) {                                     // main.cpp:12-15 -> BVH::TEST::testall(), bvhtest.cpp:615-650
		for (int i = 0; i < s.cfg.nobjs; ++i) {
			test_model(s.cfg, s.cfg.objs[i], false, s.host_bvh);
			std::printf("\n");
		}
		return 0;
	}
	if (s.cfg.testbvh) {                                     // main.cpp:16-19 -> BVH::TEST::test(), bvhtest.cpp:448-530
		test_model(s.cfg, s.cfg.objname, s.cfg.has_camera != 0, s.host_bvh);
		return 0;
	}
	if (spp_override >= 0) s.cfg.attempt = spp_override;
	if (procs > 1) gpus = procs;
	if (gpus < 1) gpus = 1;
	if (gpus == 1) s.overlap_init = false;
	if ((!checkpoint_file.empty() || !resume_file.empty()) && gpus != 1)
		die("checkpoint", "--checkpoint / --resume work on one device (multi-GPU runs keep per-device sums)");
	if (!checkpoint_file.empty() && checkpoint_every <= 0) checkpoint_every = 64;
	auto t0 = std::chrono::steady_clock::now();
	// the reference accumulates while attemptCount <= attempt, i.e. attempt + 1 frames (colorout.cpp:55-62)
	const int frames = stop_after >= 0 ? std::min(stop_after, s.cfg.attempt + 1) : s.cfg.attempt + 1;
	const size_t npix = size_t(s.cfg.width) * size_t(s.cfg.height);
	std::vector<float> img(4 * npix, 0.0f);
	std::vector<double> acc;
	mcpt_stats tot{};
	std::chrono::steady_clock::time_point t1, t2, t3;
	double first_ready_s = 0.0, all_ready_s = 0.0;
	if (procs > 1) {
		if (!s.cfg.has_camera) die("config", "no camera");
		if (s.cfg.res_x != s.cfg.width || s.cfg.res_y != s.cfg.height) die("config", "camera.resolution must equal width/height (SURVEY A.5)");
		// The parent must not touch CUDA before it forks.  A cache hit needs none; on a miss a helper child loads,
		// builds (device LBVH) and stores, and the parent then reads the file it wrote.
		char tmpdir[] = "/tmp/mcpt_cache_XXXXXX";
		const bool private_cache = !s.use_cache;
		if (private_cache) {
			if (!mkdtemp(tmpdir)) die("--procs", "mkdtemp failed");
			setenv("MCPT_CACHE_DIR", tmpdir, 1);
			s.use_cache = true;
		}
		mcpt_scene_key key{};
		char path[1024] = "";
		if (mcpt_host_scene_key(s.cfg.directory, s.cfg.objname, s.cfg.bvhtype, 1, &key) < 0) die("loadObject", mcpt_host_last_error());
		if (mcpt_host_cache_path(s.cfg.directory, s.cfg.objname, s.cfg.bvhtype, path, sizeof(path)) < 0) die("scene cache", mcpt_host_last_error());
		const char* state = "hit";
		for (int attempt = 0;; ++attempt) {
			mcpt_bvh_node* nodes = nullptr;
			int nnodes = 0;
			const int rc = mcpt_host_cache_load(path, &key, &s.tris, &s.ntris, &s.mats, &s.nmats, &s.mat_ids, &nodes, &nnodes, &s.derived, &s.derived_bytes);
			if (rc < 0) die("scene cache", mcpt_host_last_error());
			if (rc == 0) {
				s.nodes.assign(nodes, nodes + nnodes);
				mcpt_host_free(nodes);
				break;
			}
			if (attempt > 0) die("--procs: scene cache unusable after a fresh build", mcpt_host_last_error());
			state = "miss";
			pid_t pid = fork();
			if (pid < 0) die("--procs", "fork failed");
			if (pid == 0) {
				setenv("CUDA_VISIBLE_DEVICES", visible_device(0).c_str(), 1);     // the device builders run on the first device
				load_scene(s);               // this is the child's own copy of the state
				_exit(std::strcmp(s.cache_state, "miss") == 0 ? 0 : 3);
			}
			int status = 0;
			waitpid(pid, &status, 0);
			if (!WIFEXITED(status) || WEXITSTATUS(status) != 0) die("--procs", "the scene preparation process failed");
		}
		s.cache_state = private_cache ? "off" : state;
		if (private_cache) {
			std::remove(path);
			rmdir(tmpdir);
		}
		s.load_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
		double render_done_s = 0.0;
		run_processes(s, procs, frames, t0, acc, tot, all_ready_s, first_ready_s, render_done_s);
		// render_s counts from the moment the first worker could render (staggered start) / all could (static split)
		t1 = t0 + std::chrono::duration_cast<std::chrono::steady_clock::duration>(std::chrono::duration<double>(s.overlap_init ? first_ready_s : all_ready_s));
		t2 = t0 + std::chrono::duration_cast<std::chrono::steady_clock::duration>(std::chrono::duration<double>(render_done_s));
	} else {
		init(s, gpus);
		t1 = std::chrono::steady_clock::now();
		first_ready_s = all_ready_s = std::chrono::duration<double>(t1 - t0).count();
		int done = resume_file.empty() ? 0 : read_checkpoint(s, resume_file);
		if (done > frames) die("resume", "the checkpoint holds more frames than this run asks for");
		if (s.overlap_init) {
			run_overlapped(s, frames, std::min(batch, 32), t0);
			first_ready_s = s.first_ready_s;
			all_ready_s = s.all_ready_s;
		} else {
			// one device: one call (seed_mode 0 regenerates paths inside the library and wants the whole run at once),
			// cut only where a checkpoint is due
			if (gpus == 1) batch = checkpoint_every > 0 ? checkpoint_every : frames;
			while (done < frames) {
				int n = std::min(batch * gpus, frames - done);
				update(s, done, n);
				done += n;
				if (checkpoint_every > 0) write_checkpoint(s, checkpoint_file, done);
			}
		}
		t2 = std::chrono::steady_clock::now();
	}
	if (procs <= 1 && gpus == 1 && !s.mis) {
		if (mcpt_read_history(s.ctx[0], img.data(), nullptr)) die("readback", mcpt_last_error());
		mcpt_get_stats(s.ctx[0], &tot);
	} else {
		if (procs <= 1) {
			// one merge on GPU 0 over NVLink peer memory, one readback (the sums never pass through host memory per GPU)
			acc.assign(7 * npix, 0.0);
			if (s.ctx.size() > 1 && mcpt_merge_sums(s.ctx[0], s.ctx.data() + 1, int(s.ctx.size()) - 1)) die("merge", mcpt_last_error());
			if (mcpt_read_sums(s.ctx[0], acc.data())) die("readback", mcpt_last_error());
			for (mcpt_ctx* c : s.ctx) {
				mcpt_stats st{};
				mcpt_get_stats(c, &st);
				tot.rays_traced += st.rays_traced;
				tot.paths += st.paths;
			}
		}
		for (size_t p = 0; p < npix; ++p) {
			double n = acc[6 * npix + p] > 0 ? acc[6 * npix + p] : 1.0;
			for (int k = 0; k < 3; ++k) img[4 * p + k] = float(acc[k * npix + p] / n);
		}
	}
	t3 = std::chrono::steady_clock::now();
	std::string base = out_prefix.empty() ? std::string(s.cfg.objname) : out_prefix;
	if (mcpt_host_write_hdr((base + ".hdr").c_str(), s.cfg.width, s.cfg.height, img.data()) < 0 ||
	    mcpt_host_write_png((base + ".png").c_str(), s.cfg.width, s.cfg.height, img.data()) < 0)
		die("outputPicture", mcpt_host_last_error());
	auto t4 = std::chrono::steady_clock::now();
	double sec = std::chrono::duration<double>(t2 - t1).count();
	std::printf("Finished Attempting\n");                                    // colorout.cpp:65
	std::printf("{\"scene\": \"%s%s\", \"width\": %d, \"height\": %d, \"frames\": %d, \"gpus\": %d, \"procs\": %d, \"scene_cache\": \"%s\", \"load_s\": %.3f, \"create_s\": %.3f, \"upload_s\": %.3f, \"init_s\": %.3f, "
	            "\"render_s\": %.3f, \"merge_read_s\": %.3f, \"total_s\": %.3f, \"total_with_files_s\": %.3f, \"first_gpu_ready_s\": %.3f, \"all_gpus_ready_s\": %.3f, \"mrays_per_s\": %.1f, \"mpaths_per_s\": %.1f, \"out\": \"%s.hdr\"}\n",
	            s.cfg.directory, s.cfg.objname, s.cfg.width, s.cfg.height, frames, gpus, procs > 1 ? procs : 1, s.cache_state, s.load_s, s.create_s, s.upload_s,
	            std::chrono::duration<double>(t1 - t0).count(), sec, std::chrono::duration<double>(t3 - t2).count(), std::chrono::duration<double>(t3 - t0).count(),
	            std::chrono::duration<double>(t4 - t0).count(), first_ready_s, all_ready_s, tot.rays_traced / sec / 1e6, tot.paths / sec / 1e6, base.c_str());
	for (mcpt_ctx* c : s.ctx) mcpt_destroy(c);
	reap_workers();
	mcpt_host_free(s.tris);
	mcpt_host_free(s.mats);
	mcpt_host_free(s.mat_ids);
	if (s.derived) mcpt_free_blob(s.derived);
	return 0;
}
