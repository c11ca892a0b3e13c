// TEST INFRASTRUCTURE — not part of the product.  Compiles the REFERENCE'S OWN OpenCL kernel sources as C++ and
// exposes them through a small C ABI, so that tests can hold the oracle's restatement (oracle/mcpt_oracle.cpp) and,
// through it, the CUDA kernels against the reference's code:
//   kernels/rayGenerator.cl  generateRay      kernels/intersect.cl  intersectRays (+ objdef.h:101-276)
//   kernels/shade.cl         shade            kernels/history.cl    func          kernels/testkernel.cl  func
//   kernels/EPO.cl           calculateEPO (BVH quality harness, SURVEY f-4)
// The sources are read where they lie under /root/reference; oracle/build_ref.sh writes syntactically adapted copies
// to oracle/_ref/src/ (git-ignored) - the adaptations are listed there - and oracle/clc_shim.hpp supplies the OpenCL C
// vocabulary, including the conventions chosen for OpenCL's implementation-defined built-ins.
// One work item = one call with clc::g_item set; loops over work items run under OpenMP (work items are independent).
#include <omp.h>

#include "clc_shim.hpp"

#define __kernel
#define __global
#define __write_only

namespace refcl {
using namespace clc;

static int g_max_depth = 0, g_max_attempt = 0;
#define MAX_DEPTH (refcl::g_max_depth)
#define MAX_ATTEMPT (refcl::g_max_attempt)

#include "objdef_cl.h"            // oracle/_ref/src: objdef.h with the OpenCL branch selected

namespace raygen {
#include "rayGenerator_cl.inc"
}
namespace isect {
#include "intersect_cl.inc"
}
namespace shading {
#include "shade_cl.inc"
}
namespace hist {
#include "history_cl.inc"
}
namespace disp {
#include "testkernel_cl.inc"
}
namespace epo {
#include "EPO_cl.inc"
}

static_assert(sizeof(Camera) == 80 && sizeof(Ray) == 48 && sizeof(Hit) == 48 && sizeof(Triangle) == 64 &&
              sizeof(Material) == 48 && sizeof(BVHNode) == 64, "record layouts of objdef.h");

inline void item1(size_t i, size_t n) { clc::g_item = clc::WorkItem{{i, 0, 0}, {n, 1, 1}}; }
inline void item2(size_t x, size_t y, size_t w, size_t h) { clc::g_item = clc::WorkItem{{x, y, 0}, {w, h, 1}}; }

}  // namespace refcl

using namespace refcl;

extern "C" {

int refcl_abi_version(void) { return 1; }

// RayGeneration::generateRay (raygeneration.cpp:29-35): NDRange (width, height)
void refcl_generate_rays(void* camera, int width, int height, void* rays) {
#pragma omp parallel for schedule(static)
	for (int y = 0; y < height; ++y)
		for (int x = 0; x < width; ++x) {
			item2(size_t(x), size_t(y), size_t(width), size_t(height));
			raygen::generateRay(static_cast<Camera*>(camera), static_cast<Ray*>(rays));
		}
}

// SceneCL::intersect (scenebuild.cpp:103-133): NDRange (n)
void refcl_intersect(void* rays, void* nodes, void* tris, void* hits, long n, float tmin) {
#pragma omp parallel for schedule(dynamic, 256)
	for (long i = 0; i < n; ++i) {
		item1(size_t(i), size_t(n));
		isect::intersectRays(static_cast<Ray*>(rays), static_cast<BVHNode*>(nodes), static_cast<Triangle*>(tris), static_cast<Hit*>(hits), tmin);
	}
}

// SceneCL::shade (scenebuild.cpp:135-160), program built with -D MAX_DEPTH=<maxdepth> (:163)
void refcl_shade(void* materials, void* rays, void* hits, void* color, void* seeds, long n, int max_depth) {
	g_max_depth = max_depth;
#pragma omp parallel for schedule(dynamic, 256)
	for (long i = 0; i < n; ++i) {
		item1(size_t(i), size_t(n));
		shading::shade(static_cast<Material*>(materials), static_cast<Ray*>(rays), static_cast<Hit*>(hits), static_cast<float4*>(color), static_cast<uint*>(seeds));
	}
}

// ColorOut history kernel (colorout.cpp:40-62), program built with -D MAX_ATTEMPT=<attempt> (:35)
void refcl_history(void* new_colors, void* histories, void* counts, int width, int height, int max_attempt) {
	g_max_attempt = max_attempt;
#pragma omp parallel for schedule(static)
	for (int y = 0; y < height; ++y)
		for (int x = 0; x < width; ++x) {
			item2(size_t(x), size_t(y), size_t(width), size_t(height));
			hist::func(static_cast<float4*>(new_colors), static_cast<float4*>(histories), static_cast<int*>(counts));
		}
}

// display kernel (testkernel.cl): the GL image is a float4 raster here
void refcl_display(void* color, void* out, int width, int height) {
	clc::Image2D img{static_cast<float4*>(out), size_t(width)};
#pragma omp parallel for schedule(static)
	for (int y = 0; y < height; ++y)
		for (int x = 0; x < width; ++x) {
			item2(size_t(x), size_t(y), size_t(width), size_t(height));
			disp::func(&img, static_cast<float4*>(color));
		}
}

// OpenCL::update (OpenCLApp.cpp:57-82) `frames` times: colour reset to (1,1,1,1) (:63), generateRay, MAXDEPTH x
// (intersect, shade) (:69-72), history (colorout.cpp:55-62).  seeds / hist / counts persist across calls like the
// reference's buffers; *rays_traced counts the live rays entering intersect (the bench metric).
void refcl_update(void* camera, int width, int height, void* nodes, void* tris, void* materials, int max_depth,
                  int max_attempt, int frames, float tmin, void* seeds, void* hist, void* counts,
                  unsigned long long* rays_traced) {
	const long n = long(width) * height;
	Ray* rays = new Ray[size_t(n)];
	Hit* hits = new Hit[size_t(n)];
	float4* color = new float4[size_t(n)];
	std::memset(static_cast<void*>(hits), 0, sizeof(Hit) * size_t(n));
	unsigned long long traced = 0;
	for (int f = 0; f < frames; ++f) {
#pragma omp parallel for schedule(static)
		for (long i = 0; i < n; ++i) color[i] = make_float4(1.0f, 1.0f, 1.0f, 1.0f);
		refcl_generate_rays(camera, width, height, rays);
		for (int d = 0; d < max_depth; ++d) {
			unsigned long long live = 0;
#pragma omp parallel for schedule(static) reduction(+ : live)
			for (long i = 0; i < n; ++i) live += (rays[i].term_depth.w & 0xFF000000) ? 0u : 1u;
			traced += live;
			refcl_intersect(rays, nodes, tris, hits, n, tmin);
			refcl_shade(materials, rays, hits, color, seeds, n, max_depth);
		}
		refcl_history(color, hist, counts, width, height, max_attempt);
	}
	if (rays_traced) *rays_traced += traced;
	delete[] rays;
	delete[] hits;
	delete[] color;
}

// EPO_GPU's kernel launch (bvhtest.cpp:296-300): NDRange (triCount).  The kernel's point buffers hold 8 vertices;
// callers must not pass scenes where a clipped polygon reaches 9 (oracle_epo_max_polygon tells).
void refcl_epo(void* nodes, void* tris, void* epo, void* area, int ntris) {
#pragma omp parallel for schedule(dynamic, 64)
	for (int i = 0; i < ntris; ++i) {
		item1(size_t(i), size_t(ntris));
		epo::calculateEPO(static_cast<BVHNode*>(nodes), static_cast<Triangle*>(tris), static_cast<float*>(epo), static_cast<float*>(area), uint(ntris));
	}
}

void refcl_set_threads(int n) { if (n > 0) omp_set_num_threads(n); }

}  // extern "C"
