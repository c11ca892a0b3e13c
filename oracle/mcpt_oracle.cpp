// TEST INFRASTRUCTURE — CPU oracle for the path-tracing hot path.  NOT part of the product:
// only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load the library built from this file (oracle/Makefile -> oracle/liboracle.so).
//
// It restates, function for function, the reference's OpenCL kernels, which cannot be compiled
// here (OpenCL C, no OpenCL platform, no clang):
//   K1 generateRay        kernels/rayGenerator.cl:1-31
//   K2 intersectRays      kernels/intersect.cl:1-28  + objdef.h:102-275 (det2x2..intersectObjects)
//   K3 shade              kernels/shade.cl:1-206
//   K4 history            kernels/history.cl:3-28
//   K5 display            kernels/testkernel.cl:9-14
//   frame loop            OpenCLApp.cpp:57-82, colorout.cpp:40-73
//
// PARITY PINNING: the reference has no tests, golden vectors or fixtures for this path and no
// OpenCL platform exists here, so the pin is the reference's OWN kernel source: oracle/build_ref.sh compiles
// kernels/*.cl and the device half of objdef.h as C++ (oracle/ref_kernels.cpp, oracle/clc_shim.hpp; three syntactic
// adaptations, listed in build_ref.sh) into oracle/_ref/libmcpt_refcl.so, and tests/test_reference_kernels.py holds
// every buffer this restatement produces - rays, hits incl. the raw triangleID field, colour, seeds, history,
// counts, display - bit-equal to it on the three scenes, all four material types, random and adversarial rays.
// What that pin cannot cover is OpenCL's implementation-defined part (next paragraph): both sides use the same
// stated conventions for the built-ins.  The host side is pinned by the reference's own host code (oracle/_ref/
// libmcpt_ref.so): struct layouts, scene loading, material classification, camera basis, LBVH and treelet node
// arrays, .hdr bytes, SAH / EPO / LCV.  tests/golden/ holds outputs of this oracle and of those libraries.
//
// Implementation-defined OpenCL behaviour is fixed here once (and mirrored by the CUDA kernels):
//   * no FMA contraction (build with -ffp-contract=off); + - * / sqrt are IEEE round-to-nearest
//   * dot(a,b) accumulates left to right, 3 lanes for .s012 operands, 4 lanes for float4 operands
//   * normalize(v) = v / sqrt(dot(v,v)); length(v) = sqrt(dot(v,v)); cross() has w = 0
//   * tan, cos, sin, pow = the correctly rounded float result (evaluated in double, narrowed)
//   * fmin/fmax return the non-NaN operand (std::fmin/std::fmax)
//   * M_PI inside kernels is the double 3.14159265358979323846 (the host-side 3.14159265358 of
//     oclbasic.h:192 is only visible to host code)
//   * RNG seeding: the reference uses srand(time(NULL)); rand() per pixel (scenebuild.cpp:116-119),
//     which is irreproducible; here seed = hash(pixel, run_seed) & 0x7FFF (MSVC rand() range).
//     A second mode re-seeds every (pixel, frame) so samples are order-independent (used by the
//     multi-GPU sample sharding); the reference-faithful mode keeps one LCG stream per pixel.
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>
#include <omp.h>

#include "../include/mcpt_types.h"

namespace {

constexpr float KERNEL_EPSILON = 1e-5f;            // objdef.h:16
constexpr double KERNEL_PI = 3.14159265358979323846;

struct f4 { float x, y, z, w; };
inline f4 operator+(f4 a, f4 b) { return {a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w}; }
inline f4 operator-(f4 a, f4 b) { return {a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w}; }
inline f4 operator*(f4 a, f4 b) { return {a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w}; }
inline f4 operator*(float s, f4 a) { return {s * a.x, s * a.y, s * a.z, s * a.w}; }
inline f4 operator*(f4 a, float s) { return {a.x * s, a.y * s, a.z * s, a.w * s}; }
inline f4 operator/(f4 a, float s) { return {a.x / s, a.y / s, a.z / s, a.w / s}; }
inline f4 operator-(f4 a) { return {-a.x, -a.y, -a.z, -a.w}; }
inline float dot3(f4 a, f4 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline float dot4(f4 a, f4 b) { return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w; }
inline f4 normalize4(f4 a) { return a / std::sqrt(dot4(a, a)); }
inline f4 cross4(f4 a, f4 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x, 0.0f}; }
inline f4 ld(const mcpt_float4& v) { return {v.x, v.y, v.z, v.w}; }
inline mcpt_float4 st(f4 v) { return {v.x, v.y, v.z, v.w}; }
inline float as_float(int32_t i) { float f; std::memcpy(&f, &i, 4); return f; }

inline float cr_tan(float x) { return float(std::tan(double(x))); }
inline float cr_cos(float x) { return float(std::cos(double(x))); }
inline float cr_sin(float x) { return float(std::sin(double(x))); }
inline float cr_pow(float x, float y) { return float(std::pow(double(x), double(y))); }

// ------------------------------------------------------------------ K1  rayGenerator.cl:1-31
void generate_ray(const mcpt_camera& cam, uint32_t idx, uint32_t idy, uint32_t width, uint32_t height, mcpt_ray& out) {
	uint32_t id = idy * width + idx;
	float px = float(idx) / width;                // size_t -> float conversion of width
	float py = float(idy) / height;
	float ratio = width * 1.0f / height;
	f4 o, d;
	if (cam.camera_type == 0) {
		o = ld(cam.center);
		float temp1 = px - 0.5f;
		float temp2 = py - 0.5f;
		float distance = 0.5f / cr_tan(cam.arg / 2);
		f4 dd = ld(cam.direction) * distance + temp1 * ld(cam.horizontal) * ratio + temp2 * ld(cam.up);
		d = normalize4(dd);
	} else {
		o = ld(cam.center) + (px - 0.5f) * (cam.arg) * ld(cam.horizontal) * ratio + (py - 0.5f) * (cam.arg) * ld(cam.up);
		d = normalize4(ld(cam.direction));
	}
	out.ox = o.x; out.oy = o.y; out.oz = o.z;
	out.dx = d.x; out.dy = d.y; out.dz = d.z;
	out.state = 0;
	out.pixel = int32_t(id);
	out.ratio = mcpt_float4{0, 0, 0, 0};          // never written by the reference (uninitialised buffer)
}

// ------------------------------------------------------------------ objdef.h:102-177 (literal)
inline float det2x2(float a, float b, float c, float d) { return a * d - b * c; }
inline float det3x3(float a1, float a2, float a3, float b1, float b2, float b3, float c1, float c2, float c3) {
	return a1 * det2x2(b2, b3, c2, c3) - b1 * det2x2(a2, a3, c2, c3) + c1 * det2x2(a2, a3, b2, b3);
}
inline float det4x4(float a1, float a2, float a3, float a4, float b1, float b2, float b3, float b4,
                    float c1, float c2, float c3, float c4, float d1, float d2, float d3, float d4) {
	return a1 * det3x3(b2, b3, b4, c2, c3, c4, d2, d3, d4) - b1 * det3x3(a2, a3, a4, c2, c3, c4, d2, d3, d4) +
	       c1 * det3x3(a2, a3, a4, b2, b3, b4, d2, d3, d4) - d1 * det3x3(a2, a3, a4, b2, b3, b4, c2, c3, c4);
}
// Only the nine entries of the inverse that intersectTriangle reads (s0,s1,s2,s4,s5,s6,s8,s9,sa)
// are evaluated; each is the literal expression of objdef.h:157-170.
bool inverse_used_entries(const float m[16], float s[11]) {
	float a1 = m[0], a2 = m[1], a3 = m[2], a4 = m[3], b1 = m[4], b2 = m[5], b3 = m[6], b4 = m[7];
	float c1 = m[8], c2 = m[9], c3 = m[10], c4 = m[11], d1 = m[12], d2 = m[13], d3 = m[14], d4 = m[15];
	float det = det4x4(a1, a2, a3, a4, b1, b2, b3, b4, c1, c2, c3, c4, d1, d2, d3, d4);
	if (std::fabs(det) < KERNEL_EPSILON) return false;
	s[0] = det3x3(b2, b3, b4, c2, c3, c4, d2, d3, d4) / det;
	s[1] = -det3x3(a2, a3, a4, c2, c3, c4, d2, d3, d4) / det;
	s[2] = det3x3(a2, a3, a4, b2, b3, b4, d2, d3, d4) / det;
	s[4] = -det3x3(b1, b3, b4, c1, c3, c4, d1, d3, d4) / det;
	s[5] = det3x3(a1, a3, a4, c1, c3, c4, d1, d3, d4) / det;
	s[6] = -det3x3(a1, a3, a4, b1, b3, b4, d1, d3, d4) / det;
	s[8] = det3x3(b1, b2, b4, c1, c2, c4, d1, d2, d4) / det;
	s[9] = -det3x3(a1, a2, a4, c1, c2, c4, d1, d2, d4) / det;
	s[10] = det3x3(a1, a2, a4, b1, b2, b4, d1, d2, d4) / det;
	return true;
}

// The same nine entries in reduced form: row/column 4 of the matrix is (0,0,0,1), so every cofactor
// collapses to one 2x2 determinant of the 3x3 block; multiplications by 0/1 and additions of 0 are
// exact, so the surviving operations are the same and in the same order.  This is the form the CUDA
// kernels evaluate; tests/test_oracle.py checks literal == reduced bit for bit on random input.
bool inverse_reduced(const float m[16], float s[11]) {
	float a1 = m[0], a2 = m[1], a3 = m[2], b1 = m[4], b2 = m[5], b3 = m[6], c1 = m[8], c2 = m[9], c3 = m[10];
	float T1 = b2 * c3 - c2 * b3;
	float T2 = a2 * c3 - c2 * a3;
	float T3 = a2 * b3 - b2 * a3;
	float det = a1 * T1 - b1 * T2 + c1 * T3;
	if (std::fabs(det) < KERNEL_EPSILON) return false;
	s[0] = T1 / det;
	s[1] = -T2 / det;
	s[2] = T3 / det;
	s[4] = -(b1 * c3 - c1 * b3) / det;
	s[5] = (a1 * c3 - c1 * a3) / det;
	s[6] = -(a1 * b3 - b1 * a3) / det;
	s[8] = (b1 * c2 - c1 * b2) / det;
	s[9] = -(a1 * c2 - c1 * a2) / det;
	s[10] = (a1 * b2 - b1 * a2) / det;
	return true;
}

struct RayRegs {       // the Ray as the kernel holds it: float4 origin/direction with aliased w words
	f4 origin, direction;
};
inline RayRegs load_ray(const mcpt_ray& r) {
	return {{r.ox, r.oy, r.oz, as_float(r.state)}, {r.dx, r.dy, r.dz, as_float(r.pixel)}};
}

// objdef.h:178-221
template <bool LITERAL>
bool intersect_triangle(const RayRegs& ray, const mcpt_triangle& tr, float tmin, mcpt_hit& hit) {
	f4 realN{tr.nx, tr.ny, tr.nz, 0};
	f4 Rd = ray.direction;
	if (std::fabs(dot3(realN, Rd)) < KERNEL_EPSILON) return false;
	f4 v0 = ld(tr.v[0]);
	f4 AB = ld(tr.v[1]) - v0;
	f4 AC = ld(tr.v[2]) - v0;
	f4 A_Ro = v0 - ray.origin;
	float m[16] = {Rd.x, Rd.y, Rd.z, 0, -AB.x, -AB.y, -AB.z, 0, -AC.x, -AC.y, -AC.z, 0, 0, 0, 0, 1};
	float s[11];
	if (!(LITERAL ? inverse_used_entries(m, s) : inverse_reduced(m, s))) return false;
	float t = dot3(A_Ro, f4{s[0], s[4], s[8], 0});
	float b = dot3(A_Ro, f4{s[1], s[5], s[9], 0});
	float c = dot3(A_Ro, f4{s[2], s[6], s[10], 0});
	if (b < 0 || c < 0 || b + c > 1 || t <= tmin) return false;
	if (hit.t - t >= KERNEL_EPSILON) {
		hit.t = t;
		hit.normal = mcpt_float4{tr.nx, tr.ny, tr.nz, 0.0f};
		hit.material_id = uint32_t(tr.material_id);
		hit.point = st(ray.origin + t * ray.direction);
		hit.pad_ = 1;   // oracle-private marker "this call replaced the hit"; cleared by the caller
	}
	return true;
}

// objdef.h:223-237
inline bool intersect_aabb(const mcpt_float4& bbmin, const mcpt_float4& bbmax, const RayRegs& r, float tmin) {
	f4 ori = r.origin, dir = r.direction;
	float o1x = (bbmin.x - ori.x) / dir.x, o1y = (bbmin.y - ori.y) / dir.y, o1z = (bbmin.z - ori.z) / dir.z;
	float o2x = (bbmax.x - ori.x) / dir.x, o2y = (bbmax.y - ori.y) / dir.y, o2z = (bbmax.z - ori.z) / dir.z;
	float tnear = std::fmax(std::fmax(std::fmin(o1x, o2x), std::fmin(o1y, o2y)), std::fmin(o1z, o2z));
	float tfar = std::fmin(std::fmin(std::fmax(o1x, o2x), std::fmax(o1y, o2y)), std::fmax(o1z, o2z));
	if (tfar < tnear || tfar < tmin) return false;
	return true;
}

struct TraceStats { uint64_t nodes = 0, tris = 0; };

// objdef.h:240-275 — left-first DFS, 64-entry stack, no pruning by hit.t.
// closest = index of the triangle whose data was copied into the Hit (SURVEY §8c definition); the
// raw hit.triangle_id keeps the reference's "last triangle intersected anywhere" behaviour (F7).
template <bool LITERAL>
bool intersect_objects(const mcpt_bvh_node* nodes, const mcpt_triangle* tris, const RayRegs& ray, mcpt_hit& hit,
                       float tmin, int32_t& closest, TraceStats* stats) {
	int stack[64] = {0};
	int sp = 1;
	bool is_intersect = false;
	while (sp > 0) {
		const mcpt_bvh_node* node = &nodes[stack[--sp]];
		for (;;) {
			if (stats) ++stats->nodes;
			if (!intersect_aabb(node->bbmin, node->bbmax, ray, tmin)) break;
			int l = node->left, r = node->right;
			if (l == r) {
				if (stats) ++stats->tris;
				hit.pad_ = 0;
				if (intersect_triangle<LITERAL>(ray, tris[l], tmin, hit)) {
					is_intersect = true;
					hit.triangle_id = uint32_t(l);
					if (hit.pad_) closest = l;
				}
				hit.pad_ = 0;
				break;
			}
			stack[sp++] = r;
			node = &nodes[l];
		}
	}
	return is_intersect;
}

// intersect.cl:1-28.  Returns false when the ray is terminated (hits[id] is then left untouched).
template <bool LITERAL>
bool intersect_ray(const mcpt_ray& r, const mcpt_bvh_node* nodes, const mcpt_triangle* tris, float tmin,
                   mcpt_hit& out, int32_t& closest, TraceStats* stats) {
	if (uint32_t(r.state) & MCPT_RAY_TERM_MASK) return false;
	mcpt_hit hit;
	std::memset(&hit, 0, sizeof(hit));
	hit.t = FLT_MAX;
	closest = -1;
	RayRegs ray = load_ray(r);
	bool is_intersect = intersect_objects<LITERAL>(nodes, tris, ray, hit, tmin, closest, stats);
	if (is_intersect && dot3(ray.direction, ld(hit.normal)) > 0) hit.normal = st(-ld(hit.normal));
	out = hit;
	return true;
}

// ------------------------------------------------------------------ K3  shade.cl
inline uint32_t lcg_random(uint32_t* x) {          // shade.cl:1-6
	uint32_t temp = *x;
	temp = temp * 1103515245u + 12345u;
	*x = temp;
	return (temp >> 16) & 0x00007FFFu;
}

f4 mirror_direction(f4 normal, f4 incoming) {      // shade.cl:19-25
	normal.w = 0;
	incoming.w = 0;
	f4 ans = incoming - 2 * dot4(normal, incoming) * normal;
	ans.w = 0;
	return normalize4(ans);
}

bool transmitted_direction(f4 normal, f4 incoming, float index_i, float index_t, f4* transmitted) {   // :27-38
	normal.w = 0;
	incoming.w = 0;
	float ita_r = index_i / index_t;
	float n_dot_i = -dot4(normal, incoming);
	float temp = 1.0f - ita_r * ita_r * (1 - n_dot_i * n_dot_i);
	if (temp < 0.0f) return false;
	*transmitted = normalize4((ita_r * n_dot_i - std::sqrt(temp)) * normal + ita_r * incoming);
	return true;
}

f4 random_direction(f4 normal, uint32_t* seed) {   // shade.cl:40-59
	normal.w = 0;
	float deg = float(2 * KERNEL_PI / 32768 * lcg_random(seed));
	float deg2 = lcg_random(seed) * 1.0f / 32768;
	float s = std::sqrt(deg2);
	f4 axis1 = (normal.z == 0) ? f4{0, 0, 1.0f, 0} : f4{1, 0, 0, 0};
	f4 axis2 = normalize4(cross4(axis1, normal));
	axis1 = normalize4(cross4(axis2, normal));
	return normalize4(cr_cos(deg) * s * axis1 + cr_sin(deg) * s * axis2 + (1 - deg2) * normal);
}

float calc_fresnel(f4 normal, f4 incoming, float ior) {   // shade.cl:69-73
	float k = cr_pow((ior - 1) / (ior + 1), 2.0f);
	return k + (1 - k) * cr_pow(1 - std::fabs(dot3(normal, incoming)), 5.0f);
}

// shade.cl:75-206 for one work-item.  `color` and `seed` are this pixel's slots.
void shade_one(const mcpt_material* materials, mcpt_ray& ray_slot, const mcpt_hit& my_hit, f4& color,
               uint32_t& seed_slot, int max_depth) {
	if (uint32_t(ray_slot.state) & MCPT_RAY_TERM_MASK) return;
	if (my_hit.t >= FLT_MAX) {
		color = f4{0, 0, 0, 0};
		ray_slot.state = int32_t(uint32_t(ray_slot.state) | MCPT_RAY_TERM_MASK);
		return;
	}
	const mcpt_ray my_ray = ray_slot;
	const f4 my_dir{my_ray.dx, my_ray.dy, my_ray.dz, as_float(my_ray.pixel)};
	const f4 n = ld(my_hit.normal), ip = ld(my_hit.point);
	uint32_t seed = seed_slot;
	const mcpt_material* mat = materials + my_hit.material_id;
	const float two_pi = float(2 * KERNEL_PI);
	f4 new_dir{0, 0, 0, 0}, new_org{0, 0, 0, 0};
	int32_t new_state = 0;
	switch (mat->type) {
	case MCPT_DIFFUSE: {
		new_dir = random_direction(n, &seed);
		new_org = ip + KERNEL_EPSILON * new_dir;
		new_state = my_ray.state + 1;
		color = color * ld(mat->kd) * dot3(new_dir, n) / two_pi;
		seed_slot = seed;
		break;
	}
	case MCPT_GLOSSY: {
		if (lcg_random(&seed) & 1u) {
			f4 reflect = mirror_direction(n, my_dir);
			new_dir = random_direction(reflect, &seed);
			while (dot3(new_dir, n) <= 0) new_dir = random_direction(reflect, &seed);
			new_org = ip + KERNEL_EPSILON * new_dir;
			new_state = my_ray.state + 1;
			color = color * ld(mat->ka_ks) * cr_pow(dot3(new_dir, reflect), mat->Ns) * dot3(new_dir, n) / two_pi;
			seed_slot = seed;
		} else {
			new_dir = random_direction(n, &seed);
			new_org = ip + KERNEL_EPSILON * new_dir;
			new_state = my_ray.state + 1;
			color = color * ld(mat->kd) * dot3(new_dir, n) / two_pi;
			seed_slot = seed;
		}
		break;
	}
	case MCPT_LIGHT:
		ray_slot.state = int32_t(uint32_t(ray_slot.state) | MCPT_RAY_TERM_MASK);
		color = color * ld(mat->ka_ks);
		return;
	case MCPT_TRANSPARENT: {
		bool in_obj = (uint32_t(my_ray.state) & MCPT_RAY_INSIDE_MASK) != 0;
		float i = in_obj ? mat->Ni : 1.0f;
		float t = in_obj ? 1.0f : mat->Ni;
		bool transmit = transmitted_direction(n, my_dir, i, t, &new_dir);
		if (!transmit) {
			new_org = ip;
			new_dir = mirror_direction(n, my_dir);
			new_state = my_ray.state + 1;
		} else {
			float fresnel = calc_fresnel(n, new_dir, mat->Ni);
			new_org = ip;
			new_state = my_ray.state + 1;
			if ((lcg_random(&seed) * 1.0f / 32768) >= fresnel) {
				new_state = int32_t(uint32_t(new_state) ^ MCPT_RAY_INSIDE_MASK);
			} else {
				f4 m = mirror_direction(n, my_dir);
				new_dir.x = m.x; new_dir.y = m.y; new_dir.z = m.z;
			}
			seed_slot = seed;
		}
		break;
	}
	default:
		break;   // the reference prints "Crash!!!" and stores an uninitialised ray
	}
	if ((uint32_t(new_state) & MCPT_RAY_DEPTH_MASK) >= uint32_t(max_depth)) {
		color = f4{0, 0, 0, 0};
		new_state = int32_t(uint32_t(new_state) | MCPT_RAY_TERM_MASK);
	}
	ray_slot.ox = new_org.x; ray_slot.oy = new_org.y; ray_slot.oz = new_org.z;
	ray_slot.dx = new_dir.x; ray_slot.dy = new_dir.y; ray_slot.dz = new_dir.z;
	ray_slot.state = new_state;
	ray_slot.pixel = my_ray.pixel;
	// ray_slot.ratio: the reference stores uninitialised private memory here; left unchanged.
}

// ------------------------------------------------------------------ K4  history.cl:3-28
inline void history_one(f4& new_color, f4& hist, int32_t& count, int max_attempt) {
	f4 now = new_color;
	if (std::sqrt(dot4(now, now)) == 0 || count >= max_attempt) {
		new_color = hist;
		return;
	}
	now = (now + hist * float(count)) / float(count + 1);
	hist = now;
	new_color = now;
	hist.w = 0.0f;
	++count;
}

// ------------------------------------------------------------------ K5  testkernel.cl:9-14
inline f4 display_one(f4 col) {
	col.w = 0.0f;
	col.x = cr_pow(col.x, 1 / 2.2f);
	col.y = cr_pow(col.y, 1 / 2.2f);
	col.z = cr_pow(col.z, 1 / 2.2f);
	return col;
}

// ------------------------------------------------------------------ deterministic seeding
inline uint32_t mix32(uint32_t x) {               // lowbias32 finaliser
	x ^= x >> 16; x *= 0x7feb352du;
	x ^= x >> 15; x *= 0x846ca68bu;
	x ^= x >> 16;
	return x;
}
inline uint32_t pixel_seed(uint32_t pixel, uint32_t run_seed) {
	return mix32(pixel ^ mix32(run_seed)) & 0x7FFFu;
}
inline uint32_t sample_seed(uint32_t pixel, uint32_t sample, uint32_t run_seed) {
	return mix32(pixel ^ mix32(sample ^ mix32(run_seed)));
}

}  // namespace

// ================================================================== C exports (ctypes)
extern "C" {

int oracle_abi_version(void) { return 1; }
// torchrun exports OMP_NUM_THREADS=1; the timed CPU baseline asks for all host cores explicitly.
void oracle_set_threads(int n) { if (n > 0) omp_set_num_threads(n); }
int oracle_max_threads(void) { return omp_get_max_threads(); }

void oracle_generate_rays(const mcpt_camera* cam, int width, int height, mcpt_ray* rays) {
#pragma omp parallel for schedule(static)
	for (int y = 0; y < height; ++y)
		for (int x = 0; x < width; ++x)
			generate_ray(*cam, uint32_t(x), uint32_t(y), uint32_t(width), uint32_t(height), rays[size_t(y) * size_t(width) + size_t(x)]);
}

// mode 0 = literal 4x4 cofactor inverse (objdef.h as written); 1 = reduced form (must be identical).
// hits[i] is written only for live rays (like the kernel); closest[i] = -1 on miss / dead ray.
// counters (optional, 2 x u64): nodes fetched, leaf triangles fetched, summed over rays.
void oracle_intersect(const mcpt_ray* rays, long n, const mcpt_bvh_node* nodes, const mcpt_triangle* tris, float tmin,
                      mcpt_hit* hits, int32_t* closest, int mode, uint64_t* counters) {
	uint64_t tot_nodes = 0, tot_tris = 0;
#pragma omp parallel for schedule(dynamic, 256) reduction(+ : tot_nodes, tot_tris)
	for (long i = 0; i < n; ++i) {
		TraceStats st;
		int32_t c = -1;
		mcpt_hit h;
		bool live = mode == 0 ? intersect_ray<true>(rays[i], nodes, tris, tmin, h, c, counters ? &st : nullptr)
		                      : intersect_ray<false>(rays[i], nodes, tris, tmin, h, c, counters ? &st : nullptr);
		if (live) hits[i] = h;
		if (closest) closest[i] = live ? c : -1;
		tot_nodes += st.nodes;
		tot_tris += st.tris;
	}
	if (counters) { counters[0] = tot_nodes; counters[1] = tot_tris; }
}

void oracle_shade(const mcpt_material* materials, mcpt_ray* rays, const mcpt_hit* hits, float* color_rgba,
                  uint32_t* seeds, long n, int max_depth) {
#pragma omp parallel for schedule(static)
	for (long i = 0; i < n; ++i) {
		f4 c{color_rgba[4 * i], color_rgba[4 * i + 1], color_rgba[4 * i + 2], color_rgba[4 * i + 3]};
		shade_one(materials, rays[i], hits[i], c, seeds[i], max_depth);
		color_rgba[4 * i] = c.x; color_rgba[4 * i + 1] = c.y; color_rgba[4 * i + 2] = c.z; color_rgba[4 * i + 3] = c.w;
	}
}

void oracle_history(float* new_colors, float* histories, int32_t* counts, long n, int max_attempt) {
	for (long i = 0; i < n; ++i) {
		f4 c{new_colors[4 * i], new_colors[4 * i + 1], new_colors[4 * i + 2], new_colors[4 * i + 3]};
		f4 h{histories[4 * i], histories[4 * i + 1], histories[4 * i + 2], histories[4 * i + 3]};
		history_one(c, h, counts[i], max_attempt);
		new_colors[4 * i] = c.x; new_colors[4 * i + 1] = c.y; new_colors[4 * i + 2] = c.z; new_colors[4 * i + 3] = c.w;
		histories[4 * i] = h.x; histories[4 * i + 1] = h.y; histories[4 * i + 2] = h.z; histories[4 * i + 3] = h.w;
	}
}

void oracle_display(const float* color, float* out, long n) {
	for (long i = 0; i < n; ++i) {
		f4 c = display_one(f4{color[4 * i], color[4 * i + 1], color[4 * i + 2], color[4 * i + 3]});
		out[4 * i] = c.x; out[4 * i + 1] = c.y; out[4 * i + 2] = c.z; out[4 * i + 3] = c.w;
	}
}

void oracle_init_seeds(uint32_t* seeds, long n, uint32_t run_seed) {
	for (long i = 0; i < n; ++i) seeds[i] = pixel_seed(uint32_t(i), run_seed);
}
uint32_t oracle_sample_seed(uint32_t pixel, uint32_t sample, uint32_t run_seed) { return sample_seed(pixel, sample, run_seed); }

// Triangle-test self check: literal vs reduced inverse on caller-provided matrices (n x 16 floats).
// Returns the number of inputs whose accept/reject decision or any of the nine used entries differ bitwise.
long oracle_inverse_mismatches(const float* mats, long n) {
	long bad = 0;
	for (long i = 0; i < n; ++i) {
		float a[11] = {0}, b[11] = {0};
		bool ra = inverse_used_entries(mats + 16 * i, a), rb = inverse_reduced(mats + 16 * i, b);
		if (ra != rb) { ++bad; continue; }
		if (ra && std::memcmp(a, b, sizeof(a)) != 0) {
			// signed zeros compare equal as values; count only value differences
			bool diff = false;
			for (int k = 0; k < 11; ++k) if (!(a[k] == b[k])) diff = true;
			if (diff) ++bad;
		}
	}
	return bad;
}

// Whole render = OpenCLApp.cpp:57-82 repeated `frames` times (the reference runs attempt+1 frames and
// counts at most `max_attempt` non-zero samples per pixel, colorout.cpp:55-62, history.cl:15).
// Pixels are independent, so the loop nest is pixel-outer / frame-inner; per-pixel results are
// identical to the frame-outer order of the reference.
//   seed_mode 0: one LCG stream per pixel, seeded once (reference-faithful persistence)
//   seed_mode 1: seed re-derived for every (pixel, first_sample + frame)
// Outputs (any may be NULL): hist = running mean exactly as history.cl computes it (float4);
// counts = non-zero samples; sum/sumsq = double RGB sums over the counted samples (for the 4-sigma test);
// rays_traced = live rays that entered intersect (the Mrays/s numerator).
void oracle_render(const mcpt_camera* cam, int width, int height, const mcpt_bvh_node* nodes,
                   const mcpt_triangle* tris, const mcpt_material* materials, int max_depth, int max_attempt,
                   int frames, int seed_mode, uint32_t run_seed, uint32_t first_sample, float tmin,
                   float* hist_out, int32_t* counts_out, double* sum_out, double* sumsq_out,
                   uint64_t* rays_traced) {
	const long n = long(width) * long(height);
	uint64_t traced = 0;
#pragma omp parallel for schedule(dynamic, 64) reduction(+ : traced)
	for (long p = 0; p < n; ++p) {
		uint32_t seed = pixel_seed(uint32_t(p), run_seed);
		f4 hist{0, 0, 0, 0};
		int32_t count = 0;
		double s[3] = {0, 0, 0}, s2[3] = {0, 0, 0};
		for (int f = 0; f < frames; ++f) {
			if (seed_mode == 1) seed = sample_seed(uint32_t(p), first_sample + uint32_t(f), run_seed);
			f4 color{1.0f, 1.0f, 1.0f, 1.0f};
			mcpt_ray ray;
			generate_ray(*cam, uint32_t(p % width), uint32_t(p / width), uint32_t(width), uint32_t(height), ray);
			mcpt_hit hit;
			std::memset(&hit, 0, sizeof(hit));
			for (int d = 0; d < max_depth; ++d) {
				int32_t c;
				if (intersect_ray<false>(ray, nodes, tris, tmin, hit, c, nullptr)) ++traced;
				shade_one(materials, ray, hit, color, seed, max_depth);
			}
			bool counted = !(std::sqrt(dot4(color, color)) == 0 || count >= max_attempt);
			if (counted) {
				s[0] += color.x; s[1] += color.y; s[2] += color.z;
				s2[0] += double(color.x) * color.x; s2[1] += double(color.y) * color.y; s2[2] += double(color.z) * color.z;
			}
			history_one(color, hist, count, max_attempt);
		}
		if (hist_out) { hist_out[4 * p] = hist.x; hist_out[4 * p + 1] = hist.y; hist_out[4 * p + 2] = hist.z; hist_out[4 * p + 3] = hist.w; }
		if (counts_out) counts_out[p] = count;
		if (sum_out) { sum_out[3 * p] = s[0]; sum_out[3 * p + 1] = s[1]; sum_out[3 * p + 2] = s[2]; }
		if (sumsq_out) { sumsq_out[3 * p] = s2[0]; sumsq_out[3 * p + 1] = s2[1]; sumsq_out[3 * p + 2] = s2[2]; }
	}
	if (rays_traced) *rays_traced = traced;
}


// ---------------------------------------------------------------------------------------------------------
// BVH quality metrics of the research harness: restatement of kernels/EPO.cl:29-196 (per-triangle
// EPO area, depth-first with the left child popped first) and of the host-side sums of EPO_GPU (bvhtest.cpp:283-321).
// SAH and LCV have CPU originals in the reference (bvhtest.cpp:97-115, :324-444) which oracle/_ref runs as they are.
// Deviation: the kernel's point buffers hold 8 vertices, a triangle clipped by six planes can have 9 (the kernel
// then writes out of bounds); this restatement holds 10 and agrees whenever the kernel is well defined.

struct V3 { float s[3]; };

static int epo_point_in_box(const float* p, const mcpt_float4& lo, const mcpt_float4& hi) {      // EPO.cl:29-40
	return p[0] >= lo.x && p[0] <= hi.x && p[1] >= lo.y && p[1] <= hi.y && p[2] >= lo.z && p[2] <= hi.z;
}
static float epo_half_cross_len(const float* a, const float* b, const float* o) {               // length(cross(a-o, b-o)) * 0.5f
	const float e1x = a[0] - o[0], e1y = a[1] - o[1], e1z = a[2] - o[2];
	const float e2x = b[0] - o[0], e2y = b[1] - o[1], e2z = b[2] - o[2];
	const float cx = e1y * e2z - e1z * e2y, cy = e1z * e2x - e1x * e2z, cz = e1x * e2y - e1y * e2x;
	return std::sqrt(cx * cx + cy * cy + cz * cz) * 0.5f;
}
static int g_epo_max_polygon = 0;       // largest vertex count any clip pass produced (the kernel's buffers hold 8)
static void epo_clip(V3* pts, int* size, int axis, float pos, int arg);
static void epo_clip_tracked(V3* pts, int* size, int axis, float pos, int arg) {
	epo_clip(pts, size, axis, pos, arg);
	if (*size > g_epo_max_polygon) {
#pragma omp critical
		if (*size > g_epo_max_polygon) g_epo_max_polygon = *size;
	}
}
static void epo_clip(V3* pts, int* size, int axis, float pos, int arg) {                       // roundTr, EPO.cl:43-89
	if (*size == 0) return;
	V3 buf[10];
	int inside[10];
	const int n = *size;
	for (int i = 0; i < n; ++i) buf[i] = pts[i];
	*size = 0;
	for (int i = 0; i < n; ++i) inside[i] = arg > 0 ? (buf[i].s[axis] >= pos) : (buf[i].s[axis] <= pos);
	for (int i = 0; i < n; ++i) {
		const int j = (i + 1 == n) ? 0 : i + 1;
		if (!inside[i] && !inside[j]) continue;
		if (inside[i] && inside[j]) { pts[(*size)++] = buf[i]; continue; }
		if (inside[i]) pts[(*size)++] = buf[i];
		V3 d{{buf[j].s[0] - buf[i].s[0], buf[j].s[1] - buf[i].s[1], buf[j].s[2] - buf[i].s[2]}};
		const float t = (pos - buf[i].s[axis]) / d.s[axis];
		pts[(*size)++] = V3{{buf[i].s[0] + t * d.s[0], buf[i].s[1] + t * d.s[1], buf[i].s[2] + t * d.s[2]}};
	}
}
static float epo_intersect_area(const mcpt_triangle& tr, const mcpt_float4& lo, const mcpt_float4& hi) {   // INTERSECT, EPO.cl:102-130
	const float* v0 = &tr.v[0].x; const float* v1 = &tr.v[1].x; const float* v2 = &tr.v[2].x;
	if (epo_point_in_box(v0, lo, hi) && epo_point_in_box(v1, lo, hi) && epo_point_in_box(v2, lo, hi))
		return epo_half_cross_len(v1, v2, v0);
	V3 pts[10] = {{{v0[0], v0[1], v0[2]}}, {{v1[0], v1[1], v1[2]}}, {{v2[0], v2[1], v2[2]}}};
	int n = 3;
	epo_clip_tracked(pts, &n, 0, lo.x, 1); epo_clip_tracked(pts, &n, 1, lo.y, 1); epo_clip_tracked(pts, &n, 2, lo.z, 1);
	epo_clip_tracked(pts, &n, 0, hi.x, -1); epo_clip_tracked(pts, &n, 1, hi.y, -1); epo_clip_tracked(pts, &n, 2, hi.z, -1);
	float ans = 0.0f;                                                                           // pArea, EPO.cl:92-100
	if (n < 2) return ans;
	for (int i = 1; i < n - 1; ++i) ans += epo_half_cross_len(pts[i].s, pts[i + 1].s, pts[0].s);
	return ans;
}

// calculateEPO, EPO.cl:133-196: one work item per LEAF slot gid (node numPrims-1+gid).  Returns -1 when a path or
// the to-do stack would exceed the kernel's 128 entries (the kernel only prints a warning and corrupts its stack).
int oracle_epo_per_leaf(const mcpt_bvh_node* bvh, int nnodes, const mcpt_triangle* tris, int ntris, float* epo, float* area) {
	const int prims = ntris;
	if (nnodes != 2 * prims - 1) return -2;
	int status = 0;
#pragma omp parallel for schedule(dynamic, 64)
	for (int gid = 0; gid < prims; ++gid) {
		const int me = gid + prims - 1;
		const mcpt_triangle tr = tris[bvh[me].left];
		int ancestor[128], an = 0, todo[128], ts = 0;
		bool bad = false;
		ancestor[an++] = me;
		for (int p = bvh[me].parent; p != -1; p = bvh[p].parent) {
			if (an >= 128) { bad = true; break; }
			ancestor[an++] = p;
		}
		float sum = 0.0f;
		todo[ts++] = 0;
		while (ts > 0 && !bad) {
			const int now = todo[--ts];
			bool skip = false;
			for (int i = 0; i < an; ++i) if (now == ancestor[i]) { skip = true; break; }
			bool descend = skip;
			if (!skip) {
				const float a = epo_intersect_area(tr, bvh[now].bbmin, bvh[now].bbmax);
				if (a > 0) {
					sum += a * (now >= prims - 1 ? 1.0f : 1.2f);
					descend = true;
				}
			}
			if (descend && bvh[now].left != bvh[now].right) {
				if (ts + 2 > 128) { bad = true; break; }
				todo[ts++] = bvh[now].right;
				todo[ts++] = bvh[now].left;
			}
		}
		if (bad) {
#pragma omp atomic write
			status = -1;
		}
		epo[gid] = sum;
		area[gid] = epo_half_cross_len(&tr.v[1].x, &tr.v[2].x, &tr.v[0].x);
	}
	return status;
}

// largest clipped polygon since the last reset: above 8 the reference kernel itself is undefined (buffer overflow)
int oracle_epo_max_polygon(int reset) {
	const int m = g_epo_max_polygon;
	if (reset) g_epo_max_polygon = 0;
	return m;
}

// host half of EPO_GPU (bvhtest.cpp:309-320): sequential double sums, quotient
double oracle_epo_total(const float* epo, const float* area, int n) {
	double count = 0.0, a = 0.0;
	for (int i = 0; i < n; ++i) count += epo[i];
	for (int i = 0; i < n; ++i) a += area[i];
	return count / a;
}

}  // extern "C"
