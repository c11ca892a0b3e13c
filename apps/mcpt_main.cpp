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
// (seed_mode 1) into its own accumulators; the sums are merged on the host at the end.  The torchrun path
// (bench.py) merges with one NCCL reduce instead.
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <algorithm>
#include <atomic>
#include <vector>

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
	std::vector<mcpt_ctx*> ctx;
	int spp = 0;             // config "attempt"
	uint32_t run_seed = 1;
	bool reference_seeding = true;
	bool mis = false;        // --mis: NEE + MIS sampling mode (same image, far less variance; needs order-independent seeds)
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
	p.sampling = s.mis ? MCPT_SAMPLING_NEE_MIS : MCPT_SAMPLING_REFERENCE;
	p.run_seed = s.run_seed;
	p.traversal = MCPT_TRAVERSAL_FAST;
	mcpt_ctx* c = mcpt_create(&p);
	if (!c) die("mcpt_create", mcpt_last_error());
	if (mcpt_upload_scene(c, s.tris, s.ntris, s.mats, s.nmats, s.nodes.data(), int(s.nodes.size())) ||
	    mcpt_set_camera(c, &s.cfg.camera))
		die("scene upload", mcpt_last_error());
	return c;
}

void init(State& s, int ngpus) {
	if (!s.cfg.opencl) die("config", "Not Implemented");            // raygeneration.cpp:36-38 throws the same
	if (mcpt_host_load_object_mt(s.cfg.directory, s.cfg.objname, 1, 0, &s.tris, &s.ntris, &s.mats, &s.nmats, &s.mat_ids) < 0)
		die("loadObject", mcpt_host_last_error());
	s.nodes.resize(size_t(2 * s.ntris - 1));
	// built on GPU 0 (same bytes as the host build, ~100x faster); --host-bvh keeps the host code
	const std::string bvhtype = s.cfg.bvhtype;
	if (!s.host_bvh && (bvhtype == "hlbvh" || bvhtype == "treelet" || bvhtype == "treeletGPU")) {
		if (mcpt_build_lbvh(0, s.tris, s.ntris, s.nodes.data(), nullptr) < 0) die("BVH (device)", mcpt_last_error());
		// both treelet spellings: the device optimiser, whose output equals TreeletBVH<CPU>'s byte for byte
		if (bvhtype != "hlbvh" && mcpt_optimize_treelet(0, s.nodes.data(), int(s.nodes.size()), s.nodes.data(), nullptr) < 0)
			die("treelet (device)", mcpt_last_error());
	} else if (mcpt_host_build_bvh(s.tris, s.ntris, s.cfg.bvhtype, s.nodes.data()) < 0) die("BVH", mcpt_host_last_error());
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

}  // namespace App
}  // namespace MCPT

int main(int argc, char** argv) {
	using namespace MCPT::App;
	std::string config = "./config.json", out_prefix;
	int configid = -1, gpus = 1, spp_override = -1, batch = 64;
	State s;
	for (int i = 1; i < argc; ++i) {
		std::string a = argv[i];
		auto next = [&]() -> const char* { if (i + 1 >= argc) die("missing value for", a.c_str()); return argv[++i]; };
		if (a == "--config") config = next();
		else if (a == "--configid") configid = std::atoi(next());
		else if (a == "--gpus") gpus = std::atoi(next());
		else if (a == "--spp") spp_override = std::atoi(next());
		else if (a == "--out") out_prefix = next();
		else if (a == "--seed") s.run_seed = uint32_t(std::strtoul(next(), nullptr, 10));
		else if (a == "--batch") batch = std::atoi(next());
		else if (a == "--order-independent-seeds") s.reference_seeding = false;
		else if (a == "--host-bvh") s.host_bvh = true;
		else if (a == "--mis") s.mis = true;
		else if (a == "--overlap-init") s.overlap_init = true;
		else die("unknown option (use --config --configid --gpus --spp --out --seed --batch --order-independent-seeds --host-bvh --mis --overlap-init)", a.c_str());
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
	if (gpus < 1) gpus = 1;
	if (gpus == 1) s.overlap_init = false;
	auto t0 = std::chrono::steady_clock::now();
	init(s, gpus);
	auto t1 = std::chrono::steady_clock::now();
	// the reference accumulates while attemptCount <= attempt, i.e. attempt + 1 frames (colorout.cpp:55-62)
	const int frames = s.cfg.attempt + 1;
	if (s.overlap_init) {
		run_overlapped(s, frames, std::min(batch, 32), t0);
	} else {
		// one device: one call (seed_mode 0 regenerates paths inside the library and wants the whole run at once)
		if (gpus == 1) batch = frames;
		for (int done = 0; done < frames;) {
			int n = std::min(batch * gpus, frames - done);
			update(s, done, n);
			done += n;
		}
	}
	auto t2 = std::chrono::steady_clock::now();
	const size_t npix = size_t(s.cfg.width) * size_t(s.cfg.height);
	std::vector<float> img(4 * npix, 0.0f);
	mcpt_stats tot{};
	if (gpus == 1 && !s.mis) {
		if (mcpt_read_history(s.ctx[0], img.data(), nullptr)) die("readback", mcpt_last_error());
		mcpt_get_stats(s.ctx[0], &tot);
	} else {
		// one merge on GPU 0 over NVLink peer memory, one readback (the sums never pass through host memory per GPU)
		std::vector<double> acc(7 * npix, 0.0);
		if (s.ctx.size() > 1 && mcpt_merge_sums(s.ctx[0], s.ctx.data() + 1, int(s.ctx.size()) - 1)) die("merge", mcpt_last_error());
		if (mcpt_read_sums(s.ctx[0], acc.data())) die("readback", mcpt_last_error());
		for (mcpt_ctx* c : s.ctx) {
			mcpt_stats st{};
			mcpt_get_stats(c, &st);
			tot.rays_traced += st.rays_traced;
			tot.paths += st.paths;
		}
		for (size_t p = 0; p < npix; ++p) {
			double n = acc[6 * npix + p] > 0 ? acc[6 * npix + p] : 1.0;
			for (int k = 0; k < 3; ++k) img[4 * p + k] = float(acc[k * npix + p] / n);
		}
	}
	auto t3 = std::chrono::steady_clock::now();
	std::string base = out_prefix.empty() ? std::string(s.cfg.objname) : out_prefix;
	if (mcpt_host_write_hdr((base + ".hdr").c_str(), s.cfg.width, s.cfg.height, img.data()) < 0 ||
	    mcpt_host_write_png((base + ".png").c_str(), s.cfg.width, s.cfg.height, img.data()) < 0)
		die("outputPicture", mcpt_host_last_error());
	double sec = std::chrono::duration<double>(t2 - t1).count();
	std::printf("Finished Attempting\n");                                    // colorout.cpp:65
	std::printf("{\"scene\": \"%s%s\", \"width\": %d, \"height\": %d, \"frames\": %d, \"gpus\": %d, \"init_s\": %.3f, "
	            "\"render_s\": %.3f, \"merge_read_s\": %.3f, \"total_s\": %.3f, \"first_gpu_ready_s\": %.3f, \"all_gpus_ready_s\": %.3f, \"mrays_per_s\": %.1f, \"mpaths_per_s\": %.1f, \"out\": \"%s.hdr\"}\n",
	            s.cfg.directory, s.cfg.objname, s.cfg.width, s.cfg.height, frames, gpus,
	            std::chrono::duration<double>(t1 - t0).count(), sec, std::chrono::duration<double>(t3 - t2).count(), std::chrono::duration<double>(t3 - t0).count(),
	            s.overlap_init ? s.first_ready_s : std::chrono::duration<double>(t1 - t0).count(),
	            s.overlap_init ? s.all_ready_s : std::chrono::duration<double>(t1 - t0).count(), tot.rays_traced / sec / 1e6, tot.paths / sec / 1e6, base.c_str());
	for (mcpt_ctx* c : s.ctx) mcpt_destroy(c);
	mcpt_host_free(s.tris);
	mcpt_host_free(s.mats);
	mcpt_host_free(s.mat_ids);
	return 0;
}
