// TEST INFRASTRUCTURE — not part of the product.  Only tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline / --impl reference legs may load the library this file builds.
//
// Unity translation unit that compiles the REFERENCE's own host sources where they lie under
// /root/reference (nothing is copied into this repo) and exports a small C ABI so the tests can
// compare our host code against the reference's:
//   * ThirdPartyWrapper::loadObject      (thirdpartywrapper.cpp:25-99)  + tinyobj
//   * SceneCL ctor normals / material id (scenebuild.cpp:58-62)        [restated inline: that file
//                                          cannot be compiled, it needs a live OpenCL context]
//   * BVH::HLBVH<CPU>                    (BVH/hlbvh.cpp:92-200)
//   * BVH::TreeletBVH<CPU>               (BVH/treeletBVH.cpp:30-365)
//   * Auxiliary::parseCamera             (auxiliary.cpp:20-71)
//   * stbi_write_hdr with vertical flip  (thirdpartywrapper.cpp:14-23)
//   * BVH::TEST::SAH / EPO / LCV          (bvhtest.cpp:97-115, :120-281, :324-444)  [the CPU metrics; EPO_GPU
//                                          and test()/testall() compile but need a live OpenCL context]
// Build recipe: oracle/build_ref.sh (g++ -std=gnu++17 -ffp-contract=off; see that file for the
// one-line patch of the MSVC-only constructor spelling in hlbvh.cpp).
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <fstream>
#include <functional>
#include <iostream>
#include <memory>
#include <queue>
#include <sstream>
#include <string>
#include <tuple>
#include <unordered_set>
#include <vector>
// glibc's M_PI macro collides with `constexpr double M_PI` in the reference's oclbasic.h:192
#undef M_PI

#include "oclbasic.h"
#include "config.h"

// ---- stubs for the OpenCL runtime statics the compiled files reference but never reach ----
namespace MCPT {
std::vector<cl::Platform> OpenCLBasic::platforms;
std::vector<cl::Device> OpenCLBasic::devices;
cl::Context OpenCLBasic::context;
cl::CommandQueue OpenCLBasic::queue;
bool OpenCLBasic::has_init = false;
cl::Program OpenCLBasic::createProgramFromFileWithHeader(const std::string&, const std::string&, std::string) {
	throw "ref_shim: no OpenCL device";
}
cl::Kernel OpenCLBasic::createKernel(cl::Program&, const std::string&) { throw "ref_shim: no OpenCL device"; }
void OpenCLBasic::enqueueReadBuffer(const cl::Buffer&, size_t, size_t, void*) { throw "ref_shim: no OpenCL device"; }
void OpenCLBasic::enqueue1DKernelWithGroupCount(cl::Kernel&, size_t, size_t, const std::vector<cl::Event>*, cl::Event*) {
	throw "ref_shim: no OpenCL device";
}
float OpenCLBasic::timeCost(const cl::Event&, int) { return 0.f; }
// reached only from bvhtest.cpp's EPO_GPU / test() / testall(), which the shim never calls
void OpenCLBasic::init() { throw "ref_shim: no OpenCL device"; }
cl::Context& OpenCLBasic::getContext() { return context; }
cl::CommandQueue& OpenCLBasic::getQueue() { return queue; }
void OpenCLBasic::enqueueNDRange(cl::Kernel&, cl::NDRange, cl::NDRange, const std::vector<cl::Event>*, cl::Event*) {
	throw "ref_shim: no OpenCL device";
}
void OpenCLBasic::readBuffer(const cl::Buffer&, void*) { throw "ref_shim: no OpenCL device"; }
namespace Config {
static int g_w = 0, g_h = 0;
int WIDTH() { return g_w; }
int HEIGHT() { return g_h; }
nlohmann::json GETCAMERA() { return nlohmann::json(); }
std::string GETDIRECTORY() { return ""; }
std::string GETOBJNAME() { return ""; }
nlohmann::json GETOBJS() { return nlohmann::json::array(); }
std::string BVHTYPE() { return "hlbvh"; }
}
}

#include "BVH/bvh.cpp"
#include "hlbvh_patched.cpp"      // oracle/_ref/src/, written by build_ref.sh from BVH/hlbvh.cpp
#include "BVH/treeletBVH.cpp"
#include "auxiliary.cpp"
#include "thirdpartywrapper.cpp"
#include "bvhtest.cpp"
#include "stb_image_write.h"

using namespace MCPT;

extern "C" {

// loadObject + the per-triangle normal/material-id step of the SceneCL constructor.
// Returns malloc'ed arrays the caller frees with ref_free().
int ref_load_object(const char* directory, const char* objname, int with_normals,
                    void** tris, int* ntris, void** mats, int* nmats, int** matids) {
	try {
		auto [tr, mat, ids] = ThirdPartyWrapper::loadObject(directory, objname);
		if (with_normals) {
			for (size_t i = 0; i < ids.size(); ++i) {   // scenebuild.cpp:58-62
				auto& t = tr[i];
				t.normal = normalize(cross(t.v[1] - t.v[0], t.v[2] - t.v[0]));
				t.materialID.w = ids[i];
			}
		}
		*ntris = (int)tr.size();
		*nmats = (int)mat.size();
		*tris = malloc(tr.size() * sizeof(Triangle) + 1);
		*mats = malloc(mat.size() * sizeof(Material) + 1);
		*matids = (int*)malloc(ids.size() * sizeof(int) + 1);
		memcpy(*tris, tr.data(), tr.size() * sizeof(Triangle));
		memcpy(*mats, mat.data(), mat.size() * sizeof(Material));
		memcpy(*matids, ids.data(), ids.size() * sizeof(int));
		return 0;
	} catch (...) { return -1; }
}
void ref_free(void* p) { free(p); }

// HLBVH<CPU> over n triangles -> 2n-1 nodes written to out_nodes (caller-allocated).
int ref_hlbvh(const void* tris, int n, void* out_nodes) {
	try {
		std::vector<Triangle> t((const Triangle*)tris, (const Triangle*)tris + n);
		BVH::HLBVH<BVH::CPU> b(t);
		auto& nodes = b.getBVH();
		memcpy(out_nodes, nodes.data(), nodes.size() * sizeof(BVHNode));
		return (int)nodes.size();
	} catch (...) { return -1; }
}

// TreeletBVH<CPU> over an existing node array (same slot count out).
int ref_treelet(const void* nodes_in, int nn, void* out_nodes) {
	try {
		std::vector<BVHNode> v((const BVHNode*)nodes_in, (const BVHNode*)nodes_in + nn);
		BVH::TreeletBVH<BVH::CPU> b(v);
		auto& nodes = b.getBVH();
		memcpy(out_nodes, nodes.data(), nodes.size() * sizeof(BVHNode));
		return (int)nodes.size();
	} catch (...) { return -1; }
}

// parseCamera on a JSON text of the "camera" object.
int ref_parse_camera(const char* json_text, void* out_camera) {
	try {
		auto j = nlohmann::json::parse(std::string(json_text));
		Camera c = Auxiliary::parseCamera(j);
		memcpy(out_camera, &c, sizeof(Camera));
		return 0;
	} catch (...) { return -1; }
}

// outputPicture's file-writing half (thirdpartywrapper.cpp:21-22) on host float4 data.
int ref_write_hdr(const char* path, int w, int h, const float* rgba) {
	stbi_flip_vertically_on_write(true);
	return stbi_write_hdr(path, w, h, 4, rgba) ? 0 : -1;
}

// BVH quality metrics of the research harness, CPU versions (bvhtest.cpp).  LCV casts Config::WIDTH() x HEIGHT()
// rays, hence the two extra arguments.
double ref_sah(const void* nodes, int nn) {
	std::vector<BVHNode> v((const BVHNode*)nodes, (const BVHNode*)nodes + nn);
	return BVH::TEST::SAH(v);
}
double ref_epo(const void* nodes, int nn, const void* tris, int nt) {
	try {
		std::vector<BVHNode> v((const BVHNode*)nodes, (const BVHNode*)nodes + nn);
		std::vector<Triangle> t((const Triangle*)tris, (const Triangle*)tris + nt);
		std::streambuf* old = std::cout.rdbuf(nullptr);      // the reference prints progress lines
		float r = BVH::TEST::EPO(v, t);
		std::cout.rdbuf(old);
		return r;
	} catch (...) { return -1.0; }
}
double ref_lcv(const void* nodes, int nn, const void* camera, int width, int height) {
	std::vector<BVHNode> v((const BVHNode*)nodes, (const BVHNode*)nodes + nn);
	Camera c;
	memcpy(&c, camera, sizeof(Camera));
	Config::g_w = width;
	Config::g_h = height;
	return BVH::TEST::LCV(v, c);
}

int ref_sizeof(int which) {
	switch (which) {
	case 0: return sizeof(Camera);
	case 1: return sizeof(Ray);
	case 2: return sizeof(Hit);
	case 3: return sizeof(Triangle);
	case 4: return sizeof(Material);
	case 5: return sizeof(BVHNode);
	}
	return -1;
}
}
