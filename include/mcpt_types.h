/* mcpt_types.h — plain-old-data records shared by host code, the C ABI and the CUDA kernels.
 *
 * Every struct is BYTE-IDENTICAL to the record of the same name in the reference's
 * MonteCarloPathTracing/objdef.h:21-99 (sizes checked by static_assert below and against the
 * reference's own headers in tests/test_host.py and tests/test_abi.py), so a reference-side caller can hand its
 * std::vector<Triangle>/<Material>/<BVHNode> storage to the C ABI unchanged.
 *
 *   reference type   objdef.h    size   notes
 *   Camera           :21-27       80    arg = vertical fov in radians; cameraType 0 = perspective
 *   Ray              :29-39       48    origin.w aliases the int state word, direction.w the pixel id
 *   Hit              :41-48       48
 *   Triangle         :50-56       64    normal.w aliases the int material id
 *   Material         :69-79       48    ka and ks share storage
 *   BVHNode          :91-99       64    leaf <=> left == right (= triangle index)
 */
#ifndef MCPT_TYPES_H
#define MCPT_TYPES_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mcpt_float4 { float x, y, z, w; } mcpt_float4;

typedef struct mcpt_camera {
	mcpt_float4 center, direction, up, horizontal;
	float arg;            /* fov * pi / 180 (host pi = 3.14159265358, auxiliary.cpp:42) */
	float tmin;
	uint32_t camera_type; /* 0 perspective, 1 orthographic */
	float pad_;
} mcpt_camera;

/* Ray.state (reference: term_depth.w, objdef.h:32): bits 0-15 depth, 16-23 inside-glass flag,
 * 24-31 terminated flag. */
#define MCPT_RAY_DEPTH_MASK  0x0000FFFFu
#define MCPT_RAY_INSIDE_MASK 0x00FF0000u
#define MCPT_RAY_TERM_MASK   0xFF000000u

typedef struct mcpt_ray {
	float ox, oy, oz;
	int32_t state;        /* aliases origin.w */
	float dx, dy, dz;
	int32_t pixel;        /* aliases direction.w */
	mcpt_float4 ratio;    /* unused by the reference kernels */
} mcpt_ray;

typedef struct mcpt_hit {
	mcpt_float4 normal;
	mcpt_float4 point;
	float t;
	uint32_t triangle_id; /* see include/mcpt.h: "closest" vs "last touched" semantics */
	uint32_t material_id;
	uint32_t pad_;
} mcpt_hit;

typedef struct mcpt_triangle {
	mcpt_float4 v[3];
	float nx, ny, nz;
	int32_t material_id;  /* aliases normal.w */
} mcpt_triangle;

enum { MCPT_DIFFUSE = 1, MCPT_GLOSSY = 2, MCPT_TRANSPARENT = 3, MCPT_LIGHT = 4 };

typedef struct mcpt_material {
	int32_t type;
	float Ni;
	float Ns;
	float pad_;
	mcpt_float4 kd;
	mcpt_float4 ka_ks;    /* emission for LIGHT, specular for GLOSSY */
} mcpt_material;

typedef struct mcpt_bvh_node {
	mcpt_float4 bbmin;
	mcpt_float4 bbmax;
	mcpt_float4 pad_;
	int32_t parent;
	int32_t left;
	int32_t right;
	int32_t pad2_;
} mcpt_bvh_node;

#ifdef __cplusplus
}
static_assert(sizeof(mcpt_camera) == 80, "Camera layout");
static_assert(sizeof(mcpt_ray) == 48, "Ray layout");
static_assert(sizeof(mcpt_hit) == 48, "Hit layout");
static_assert(sizeof(mcpt_triangle) == 64, "Triangle layout");
static_assert(sizeof(mcpt_material) == 48, "Material layout");
static_assert(sizeof(mcpt_bvh_node) == 64, "BVHNode layout");
#endif

#endif /* MCPT_TYPES_H */
