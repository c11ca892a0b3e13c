// Replacement of the reference's MonteCarloPathTracing/OpenCLApp.cpp: the same two entry points
// (MCPT::OpenCL::init / update, OpenCLApp.h:4-5) over libmcpt_cuda.so instead of oclbasic.cpp + kernels/*.cl.
//
// This file is compiled AGAINST THE REFERENCE'S OWN SOURCES (config.cpp, thirdpartywrapper.cpp, auxiliary.cpp,
// BVH/hlbvh.cpp, BVH/treeletBVH.cpp, tiny_obj_loader, stb_image_write - where they lie under /root/reference) by
// oracle/build_ref.sh into oracle/_ref/mcpt_dropin, and tests/test_gpu_dropin.py holds its output file equal to
// build/mcpt's: it is the proof that include/mcpt.h is a drop-in for the reference's device layer, and the code
// INTEGRATION.md §1 shows to a maintainer.  Reference lines are cited per statement.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <ctime>
#include <iostream>
#include <memory>
#include <string>
#include <vector>
#undef M_PI                       // glibc's macro collides with `constexpr double M_PI` in oclbasic.h:192

#include "OpenCLApp.h"
#include "config.h"
#include "thirdpartywrapper.h"    // loadObject; pulls oclbasic.h for the host float4 math (normalize, cross)
#include "auxiliary.h"            // parseCamera
#include "BVH/hlbvh.h"
#include "BVH/treeletBVH.h"
#include "stb_image_write.h"

#include "mcpt.h"

using namespace MCPT;

static_assert(sizeof(Triangle) == sizeof(mcpt_triangle) && sizeof(Material) == sizeof(mcpt_material) &&
              sizeof(BVHNode) == sizeof(mcpt_bvh_node) && sizeof(Camera) == sizeof(mcpt_camera),
              "include/mcpt_types.h mirrors objdef.h:21-99 byte for byte");

namespace {
mcpt_ctx* ctx = nullptr;
int frames_done = 0;          // colorout.cpp:29 attemptCount
int frames_pending = 0;       // update() calls not yet handed to the device
int batch = 1;                // frames per mcpt_render_async call (MCPT_DROPIN_BATCH; 1 = one launch group per update())
bool written = false;

void flush() {
	if (frames_pending == 0) return;
	if (mcpt_render_async(ctx, frames_pending)) throw mcpt_last_error();   // enqueue only: the stream keeps the order
	frames_pending = 0;
}
}  // namespace

namespace MCPT::OpenCL {

void init() {                                                               // OpenCLApp.cpp:36-55
	mcpt_params p{};
	p.width = Config::WIDTH();                                              // :38
	p.height = Config::HEIGHT();                                            // :39
	p.max_depth = Config::MAXDEPTH();                                       // -D MAX_DEPTH,   scenebuild.cpp:163
	p.max_attempt = Config::MAXATTEPMT();                                   // -D MAX_ATTEMPT, colorout.cpp:35
	p.seed_mode = 0;                                                        // one LCG stream per pixel, shade.cl:101
	const char* seed = std::getenv("MCPT_SEED");                            // srand(time(NULL)), scenebuild.cpp:116
	p.run_seed = seed ? uint32_t(std::strtoul(seed, nullptr, 10)) : uint32_t(time(nullptr));
	p.traversal = MCPT_TRAVERSAL_FAST;
	if (const char* b = std::getenv("MCPT_DROPIN_BATCH")) batch = std::max(1, std::atoi(b));
	ctx = mcpt_create(&p);
	if (!ctx) throw mcpt_last_error();                                      // the reference throws C strings too

	auto [tr, mat, ids] = ThirdPartyWrapper::loadObject(Config::GETDIRECTORY(), Config::GETOBJNAME());   // :47
	for (size_t i = 0; i < ids.size(); ++i) {                               // SceneCL ctor, scenebuild.cpp:58-62
		auto& t = tr[i];
		t.normal = normalize(cross(t.v[1] - t.v[0], t.v[2] - t.v[0]));
		t.materialID.w = ids[i];
	}
	// scenebuild.cpp:66-79: the reference's own host builders, unchanged
	const std::string bvhtype = Config::BVHTYPE();
	std::vector<BVHNode> nodes;
	if (bvhtype == "hlbvh") nodes = BVH::HLBVH<BVH::CPU>(tr).releaseBVH();
	else if (bvhtype == "treelet" || bvhtype == "treeletGPU")               // treeletGPU: kernels/treeletBVH.cl is racy; its
		nodes = BVH::TreeletBVH<BVH::CPU>(BVH::HLBVH<BVH::CPU>(tr).releaseBVH()).releaseBVH();   // CPU twin defines the result
	else throw "BVH Not Implemented";                                       // scenebuild.cpp:78
	if (mcpt_upload_scene(ctx, reinterpret_cast<const mcpt_triangle*>(tr.data()), int(tr.size()),
	                      reinterpret_cast<const mcpt_material*>(mat.data()), int(mat.size()),
	                      reinterpret_cast<const mcpt_bvh_node*>(nodes.data()), int(nodes.size())))
		throw mcpt_last_error();
	Camera cam = Auxiliary::parseCamera(Config::GETCAMERA());               // raygeneration.cpp:44-45
	if (mcpt_set_camera(ctx, reinterpret_cast<const mcpt_camera*>(&cam))) throw mcpt_last_error();
}

void update() {                                                             // OpenCLApp.cpp:57-82: one frame = one sample per pixel
	if (Config::TESTBVH()) return;                                          // :58
	if (frames_done <= Config::MAXATTEPMT()) {                              // colorout.cpp:55: attempt+1 frames accumulate
		++frames_done;
		if (++frames_pending >= batch) flush();
		return;
	}
	if (!written) {                                                         // colorout.cpp:63-68, thirdpartywrapper.cpp:14-23
		flush();
		std::vector<float> img(size_t(4) * Config::WIDTH() * Config::HEIGHT());
		if (mcpt_read_history(ctx, img.data(), nullptr)) throw mcpt_last_error();   // blocking, like readBuffer(CL_TRUE)
		std::cout << "Finished Attempting" << std::endl;
		stbi_flip_vertically_on_write(true);
		stbi_write_hdr((Config::GETOBJNAME() + ".hdr").c_str(), Config::WIDTH(), Config::HEIGHT(), 4, img.data());
		written = true;
	}
	// on-screen image (testkernel.cl): mcpt_display(ctx, /*from_history=*/1, host_rgba) -> glTexSubImage2D; headless here
}

}  // namespace MCPT::OpenCL

// Used by the headless main (the reference's idle loop runs until the window closes).
bool mcpt_dropin_finished() { return written; }
void mcpt_dropin_shutdown() { mcpt_destroy(ctx); ctx = nullptr; }
