// Device-side arithmetic of the path tracer: the reference kernels' math with the implementation-
// defined parts of OpenCL C pinned down (no FMA contraction: the translation unit is compiled with
// -fmad=false; IEEE round-to-nearest + - * / sqrt; dot() accumulates left to right; fmin/fmax ignore
// NaN).  Every function cites the reference lines it computes.
#pragma once

#include <cfloat>
#include <cstdint>
#include <cuda_runtime.h>

#include "mcpt_types.h"

namespace mcpt {

constexpr float KERNEL_EPSILON = 1e-5f;   // objdef.h:16
constexpr float TWO_PI_F = 6.28318530717958647692f;   // (float)(2*M_PI), shade.cl:120

// IEEE a / b.  The compiler's divider leaves its fast path (a ~60-instruction subroutine, paid by the
// whole warp) whenever an operand is zero, and zero numerators are everywhere here (w lanes, axis-aligned
// normals and edge cross products, black colour channels).  0 / b is a signed zero for every non-zero,
// non-NaN b, so that case is answered directly and the divider sees 1 / b instead.
#ifndef MCPT_FDIV_BYPASS
#define MCPT_FDIV_BYPASS 3   // bit 0: scalar divisions (triangle tests, +3% measured); bit 1: the shader's f4 divisions (+1%); bit 2: branch on the w lane only (no gain)
#endif
__device__ __forceinline__ float fdiv(float a, float b) {
#if !(MCPT_FDIV_BYPASS & 1)
	return a / b;
#endif
	const bool z = a == 0.0f && fabsf(b) > 0.0f;
	// the select is opaque PTX: written in C++ the compiler folds `z ? 0 : (z ? 1 : a) / b` back into a / b
	float num;
	asm("{ .reg .pred p; setp.ne.s32 p, %2, 0; selp.f32 %0, 0f3F800000, %1, p; }" : "=f"(num) : "f"(a), "r"(int(z)));
	const float q = num / b;
	return z ? __int_as_float((__float_as_int(a) ^ __float_as_int(b)) & int(0x80000000u)) : q;
}

struct f4 { float x, y, z, w; };
__device__ __forceinline__ f4 operator+(f4 a, f4 b) { return {a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w}; }
__device__ __forceinline__ f4 operator-(f4 a, f4 b) { return {a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w}; }
__device__ __forceinline__ f4 operator*(f4 a, f4 b) { return {a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w}; }
__device__ __forceinline__ f4 operator*(float s, f4 a) { return {s * a.x, s * a.y, s * a.z, s * a.w}; }
__device__ __forceinline__ f4 operator*(f4 a, float s) { return {a.x * s, a.y * s, a.z * s, a.w * s}; }
// The shader's vector divisions (normalisations: three per randomDirection) through one double-precision reciprocal:
// x = a / s of two floats is either a float itself or further than 2^-49 (relative) from every float and every midpoint
// between floats (a - m s is a non-zero multiple of the last bit of the 49-bit product m s), and double(a) * RN(1 / double(s))
// is within 2^-51.9 of x, so narrowing it rounds like the IEEE quotient - for every finite a and s, denormal results
// included (their grid is coarser) - without the divider's subroutine and its slow path for small operands
// (tools/src/div_two_pi_check.c enumerates the constant-divisor case and samples this one).
#ifndef MCPT_DRCP_DIV
#define MCPT_DRCP_DIV 3          // bit 0: the normalisations, bit 1: the division by 2 pi; 0 = the hardware divider.  Mrays/s, 0 -> 3
                                 // (profiles/r3q_scene_rates.log): diningroom 4 245 -> 4 366, its seed_mode 0 pixel chains 3 158 -> 3 610,
                                 // cbox 4 651 -> 4 781, veach_mis 7 124 -> 6 852 (its shader is bound by the conversions' pipe)
#endif
__device__ __forceinline__ f4 operator/(f4 a, float s) {
#if MCPT_DRCP_DIV & 1
	const double r = __drcp_rn(double(s));
	return {float(double(a.x) * r), float(double(a.y) * r), float(double(a.z) * r), float(double(a.w) * r)};
#elif MCPT_FDIV_BYPASS & 2
	return {fdiv(a.x, s), fdiv(a.y, s), fdiv(a.z, s), fdiv(a.w, s)};
#else
	return {a.x / s, a.y / s, a.z / s, a.w / s};
#endif
}
// Division of a path's throughput by 2 pi (shade.cl:120,143).  Throughputs shrink by ~2^-7 per bounce and the Phong
// lobe's pow(cos, Ns = 4000) adds factors of 2^-100 at once: on the diningroom stand-in 5 % of the live rays carry a
// channel below 2^-120 at bounce 8 and 13 % a denormal one at bounce 14 (tools/throughput_probe.py), where the hardware's
// divider takes its slow path - a ~100-instruction subroutine the whole warp waits for while one or two lanes run it
// (19 % of k_wave_shade's warp instructions at a mid-depth bounce, ncu r3i).  The divisor is a constant, so:
// float(double(a) * RN(1 / double(2 pi))) carries the bits of a / (float)(2 pi) for EVERY float a - enumerated over all
// 2^32 of them (tools/src/div_two_pi_check.c, tests/test_div_two_pi.py).  Three instructions, no slow path.
__device__ __forceinline__ float fdiv_two_pi(float a) {
#if MCPT_DRCP_DIV & 2
	return float(double(a) * 0x1.45f306446f9b4p-3);
#else
	return fdiv(a, TWO_PI_F);
#endif
}
__device__ __forceinline__ f4 div_two_pi(f4 a) { return {fdiv_two_pi(a.x), fdiv_two_pi(a.y), fdiv_two_pi(a.z), fdiv_two_pi(a.w)}; }
__device__ __forceinline__ f4 operator-(f4 a) { return {-a.x, -a.y, -a.z, -a.w}; }
__device__ __forceinline__ float dot3(f4 a, f4 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ float dot4(f4 a, f4 b) { return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w; }
__device__ __forceinline__ f4 normalize4(f4 a) { return a / sqrtf(dot4(a, a)); }
__device__ __forceinline__ f4 cross4(f4 a, f4 b) {
	return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x, 0.0f};
}
__device__ __forceinline__ f4 ld4(const mcpt_float4& v) { return {v.x, v.y, v.z, v.w}; }
__device__ __forceinline__ f4 ld4(float4 v) { return {v.x, v.y, v.z, v.w}; }
__device__ __forceinline__ float4 st4(f4 v) { return make_float4(v.x, v.y, v.z, v.w); }

// Transcendentals: evaluated in double and narrowed, like the oracle does with libm (the stated convention for OpenCL's
// implementation-defined pow).  CUDA's double pow is accurate to 2 ulp(double), glibc's to < 1, so the narrowed floats
// agree unless the exact result lies within ~2 ulp(double) of a float rounding boundary: a 2^-27 event per call, i.e.
// "bit-identical glossy lobes" is a statement with that probability attached, not a proof (pow5 / sqr below are exact
// products in double and carry no such caveat).  tests/test_gpu_parity.py compares k_display with testkernel.cl and
// tests/test_gpu_headline.py whole glossy renders bit for bit.
__device__ __forceinline__ float cr_pow(float x, float y) { return float(pow(double(x), double(y))); }
__device__ __forceinline__ float cr_pow5(float x) {     // pow(x, 5) for calcFresnel, shade.cl:72
	double d = double(x), d2 = d * d;
	return float(d2 * d2 * d);
}
__device__ __forceinline__ float cr_sqr(float x) { double d = double(x); return float(d * d); }   // pow(x, 2)

// ------------------------------------------------------------------ camera / K1 (rayGenerator.cl:1-31)
struct CameraDev {
	f4 center, direction, up, horizontal;
	float arg;
	float distance;         // 0.5f / tan(arg / 2), uniform over the image: evaluated once on the host
	uint32_t camera_type;
	uint32_t width, height;
};

struct RayRegs {            // Ray as the kernels hold it: float4 origin/direction with the aliased w words
	f4 origin, direction;
};

__device__ __forceinline__ RayRegs generate_ray(const CameraDev& cam, uint32_t idx, uint32_t idy) {
	uint32_t id = idy * cam.width + idx;
	float px = float(idx) / float(cam.width);
	float py = float(idy) / float(cam.height);
	float ratio = float(cam.width) * 1.0f / float(cam.height);
	RayRegs r;
	if (cam.camera_type == 0) {
		float temp1 = px - 0.5f;
		float temp2 = py - 0.5f;
		f4 dd = cam.direction * cam.distance + temp1 * cam.horizontal * ratio + temp2 * cam.up;
		r.origin = cam.center;
		r.direction = normalize4(dd);
	} else {
		r.origin = cam.center + (px - 0.5f) * (cam.arg) * cam.horizontal * ratio + (py - 0.5f) * (cam.arg) * cam.up;
		r.direction = normalize4(cam.direction);
	}
	r.origin.w = __int_as_float(0);
	r.direction.w = __int_as_float(int(id));
	return r;
}

// ------------------------------------------------------------------ box test (objdef.h:223-237)
// Exact restatement: six IEEE divisions, NaN-ignoring fmin/fmax.  tnear is returned for pruning.
__device__ __forceinline__ bool intersect_aabb_exact(float bminx, float bminy, float bminz, float bmaxx, float bmaxy,
                                                     float bmaxz, const RayRegs& r, float tmin, float& tnear_out) {
	float o1x = (bminx - r.origin.x) / r.direction.x, o1y = (bminy - r.origin.y) / r.direction.y,
	      o1z = (bminz - r.origin.z) / r.direction.z;
	float o2x = (bmaxx - r.origin.x) / r.direction.x, o2y = (bmaxy - r.origin.y) / r.direction.y,
	      o2z = (bmaxz - r.origin.z) / r.direction.z;
	float tnear = fmaxf(fmaxf(fminf(o1x, o2x), fminf(o1y, o2y)), fminf(o1z, o2z));
	float tfar = fminf(fminf(fmaxf(o1x, o2x), fmaxf(o1y, o2y)), fmaxf(o1z, o2z));
	tnear_out = tnear;
	if (tfar < tnear || tfar < tmin) return false;
	return true;
}

// ------------------------------------------------------------------ triangle test (objdef.h:178-221)
// The reference solves [Rd | -AB | -AC] (t,b,c)^T = A - Ro through a generic 4x4 cofactor inverse whose
// last row/column is (0,0,0,1); every cofactor collapses exactly (x0, x1, +0 are exact) to a 2x2
// determinant of the 3x3 block, so the expressions below are the surviving operations in the same
// order (the oracle checks literal == reduced bit for bit).  e1 = -AB, e2 = -AC; cr = the three
// ray-independent 2x2 determinants (e1.y*e2.z - e2.y*e1.z, e1.x*e2.z - e2.x*e1.z, e1.x*e2.y - e2.x*e1.y),
// precomputed on the host in the same float arithmetic.
struct TriTest { float t; bool valid; };

__device__ __forceinline__ TriTest intersect_triangle(const RayRegs& ray, float v0x, float v0y, float v0z, float e1x,
                                                      float e1y, float e1z, float e2x, float e2y, float e2z, float cr0,
                                                      float cr1, float cr2, float nx, float ny, float nz, float tmin) {
	TriTest out{0.0f, false};
	const float a1 = ray.direction.x, a2 = ray.direction.y, a3 = ray.direction.z;
	if (fabsf(nx * a1 + ny * a2 + nz * a3) < KERNEL_EPSILON) return out;      // :186
	const float T2 = a2 * e2z - e2y * a3;
	const float T3 = a2 * e1z - e1y * a3;
	const float det = a1 * cr0 - e1x * T2 + e2x * T3;                          // det4x4, :150
	if (fabsf(det) < KERNEL_EPSILON) return out;                               // :152
	const float Ax = v0x - ray.origin.x, Ay = v0y - ray.origin.y, Az = v0z - ray.origin.z;
	// the comparisons of :209 are evaluated lazily; they are a pure OR, so the outcome is the same
	const float s1 = fdiv(-T2, det), s5 = fdiv(a1 * e2z - e2x * a3, det), s9 = fdiv(-(a1 * e2y - e2x * a2), det);
	const float b = Ax * s1 + Ay * s5 + Az * s9;
	if (b < 0) return out;
	const float s2 = fdiv(T3, det), s6 = fdiv(-(a1 * e1z - e1x * a3), det), s10 = fdiv(a1 * e1y - e1x * a2, det);
	const float c = Ax * s2 + Ay * s6 + Az * s10;
	if (c < 0 || b + c > 1) return out;
	const float s0 = fdiv(cr0, det), s4 = fdiv(-cr1, det), s8 = fdiv(cr2, det);
	const float t = Ax * s0 + Ay * s4 + Az * s8;
	if (t <= tmin) return out;
	out.t = t;
	out.valid = true;
	return out;
}

// ------------------------------------------------------------------ K3 helpers (shade.cl:1-73)
__device__ __forceinline__ uint32_t lcg_random(uint32_t& x) {               // :1-6
	x = x * 1103515245u + 12345u;
	return (x >> 16) & 0x00007FFFu;
}

__device__ __forceinline__ f4 mirror_direction(f4 normal, f4 incoming) {    // :19-25
	normal.w = 0;
	incoming.w = 0;
	f4 ans = incoming - 2 * dot4(normal, incoming) * normal;
	ans.w = 0;
	return normalize4(ans);
}

__device__ __forceinline__ bool transmitted_direction(f4 normal, f4 incoming, float index_i, float index_t, f4& out) {   // :27-38
	normal.w = 0;
	incoming.w = 0;
	float ita_r = index_i / index_t;
	float n_dot_i = -dot4(normal, incoming);
	float temp = 1.0f - ita_r * ita_r * (1 - n_dot_i * n_dot_i);
	if (temp < 0.0f) return false;
	out = normalize4((ita_r * n_dot_i - sqrtf(temp)) * normal + ita_r * incoming);
	return true;
}

// cos/sin of the 32768 possible azimuths 2*pi*r/32768 (shade.cl:43): tabulated on the host as the
// correctly rounded float of cos/sin of the FLOAT angle, interleaved (cos, sin).
__device__ __forceinline__ f4 random_direction(f4 normal, uint32_t& seed, const float2* __restrict__ cs_table) {   // :40-59
	normal.w = 0;
	uint32_t r1 = lcg_random(seed);
	float deg2 = float(lcg_random(seed)) * 1.0f / 32768.0f;
	float s = sqrtf(deg2);
	f4 axis1 = (normal.z == 0) ? f4{0, 0, 1.0f, 0} : f4{1, 0, 0, 0};
	f4 axis2 = normalize4(cross4(axis1, normal));
	axis1 = normalize4(cross4(axis2, normal));
	float2 cs = __ldg(cs_table + r1);
	return normalize4(cs.x * s * axis1 + cs.y * s * axis2 + (1 - deg2) * normal);
}

__device__ __forceinline__ float calc_fresnel(f4 normal, f4 incoming, float ior) {   // :69-73
	float k = cr_sqr((ior - 1) / (ior + 1));
	return k + (1 - k) * cr_pow5(1 - fabsf(dot3(normal, incoming)));
}

// shade.cl:75-206 for one work item.  ray.origin.w carries the state word.  Returns true when the
// path goes on (ray updated), false when it terminated (color final).  `color` has 4 lanes like the
// reference's float4.
struct MaterialDev {
	int32_t type;
	float Ni, Ns, pad;
	f4 kd, ka_ks;
};

__device__ __forceinline__ MaterialDev load_material(const mcpt_material* __restrict__ mats, uint32_t id) {
	const float4* p = reinterpret_cast<const float4*>(mats + id);
	float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
	MaterialDev m;
	m.type = __float_as_int(a.x);
	m.Ni = a.y; m.Ns = a.z; m.pad = a.w;
	m.kd = ld4(b);
	m.ka_ks = ld4(c);
	return m;
}

__device__ __forceinline__ bool shade_one(const mcpt_material* __restrict__ mats, const float2* __restrict__ cs_table,
                                          RayRegs& ray, float hit_t, f4 n, f4 ip, uint32_t material_id, f4& color,
                                          uint32_t& seed_slot, int max_depth, bool* diffuse_lobe = nullptr,
                                          f4* phong_axis = nullptr) {
	// phong_axis (NEE + MIS mode): receives the mirror direction with w = 1 when the glossy lobe was taken, else w = 0
	if (phong_axis) *phong_axis = f4{0, 0, 0, 0};
	if (diffuse_lobe) *diffuse_lobe = false;
	const uint32_t state = uint32_t(__float_as_int(ray.origin.w));
	if (hit_t >= FLT_MAX) {                                                   // :92-96
		color = f4{0, 0, 0, 0};
		ray.origin.w = __int_as_float(int(state | MCPT_RAY_TERM_MASK));
		return false;
	}
	const f4 my_dir = ray.direction;
	uint32_t seed = seed_slot;
	const MaterialDev mat = load_material(mats, material_id);
	f4 new_dir{0, 0, 0, 0}, new_org{0, 0, 0, 0};
	int32_t new_state = 0;
	// Diffuse (:114-123) and both glossy lobes (:124-154) share ONE call site of randomDirection - the expensive
	// part (three normalisations) - so the lanes of a warp run it together whatever their material; every lane
	// still draws its random numbers in the reference's order.
	if (mat.type == MCPT_DIFFUSE || mat.type == MCPT_GLOSSY) {
		const bool specular = mat.type == MCPT_GLOSSY && (lcg_random(seed) & 1u);
		f4 axis = n;
		if (specular) axis = mirror_direction(n, my_dir);
		do {
			new_dir = random_direction(axis, seed, cs_table);
		} while (specular && dot3(new_dir, n) <= 0);
		new_org = ip + KERNEL_EPSILON * new_dir;
		new_state = int32_t(state) + 1;
		if (specular) color = div_two_pi(color * mat.ka_ks * cr_pow(dot3(new_dir, axis), mat.Ns) * dot3(new_dir, n));
		else color = div_two_pi(color * mat.kd * dot3(new_dir, n));
		if (diffuse_lobe) *diffuse_lobe = !specular;
		if (phong_axis && specular) *phong_axis = f4{axis.x, axis.y, axis.z, 1.0f};
		seed_slot = seed;
	} else
	switch (mat.type) {
	case MCPT_LIGHT:                                                          // :155-158
		ray.origin.w = __int_as_float(int(state | MCPT_RAY_TERM_MASK));
		color = color * mat.ka_ks;
		return false;
	case MCPT_TRANSPARENT: {                                                  // :159-193
		bool in_obj = (state & MCPT_RAY_INSIDE_MASK) != 0;
		float i = in_obj ? mat.Ni : 1.0f;
		float t = in_obj ? 1.0f : mat.Ni;
		bool transmit = transmitted_direction(n, my_dir, i, t, new_dir);
		new_org = ip;
		new_state = int32_t(state) + 1;
		if (!transmit) {
			new_dir = mirror_direction(n, my_dir);
		} else {
			float fresnel = calc_fresnel(n, new_dir, mat.Ni);
			if ((float(lcg_random(seed)) * 1.0f / 32768.0f) >= fresnel) {
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
		break;   // the reference prints "Crash!!!" (shade.cl:195) and stores an uninitialised ray
	}
	bool alive = true;
	if ((uint32_t(new_state) & MCPT_RAY_DEPTH_MASK) >= uint32_t(max_depth)) {   // :199-202
		color = f4{0, 0, 0, 0};
		new_state = int32_t(uint32_t(new_state) | MCPT_RAY_TERM_MASK);
		alive = false;
	}
	const float pixel_w = ray.direction.w;
	ray.origin = f4{new_org.x, new_org.y, new_org.z, __int_as_float(new_state)};
	ray.direction = f4{new_dir.x, new_dir.y, new_dir.z, pixel_w};
	return alive;
}

// ------------------------------------------------------------------ NEE + MIS sampling mode (not in the reference)
// Solid-angle density of randomDirection (shade.cl:40-59) about `normal`, as a function of c = cos(theta):
// it draws u ~ U[0,1), phi ~ U[0, 2 pi) and returns the direction of (sqrt(u) cos phi, sqrt(u) sin phi, 1 - u), so
// c = (1-u)/sqrt(D), D = 1 - u + u^2, and p = |du/dc| / (2 pi) = D^(3/2) / (pi (1 + u)); inverting c(u) gives
// 1 - u = 2c / (c + sqrt(4 - 3 c^2)) (the cancellation-free form of SURVEY's expression).  The reference's u and
// phi live on 2^15-point grids; the continuous density is their limit (relative difference ~1e-4).
// Glossy lobe (shade.cl:125-147): randomDirection about the mirror direction r, redrawn until the direction lies above
// the surface.  Its density is random_direction_pdf(cos(d, r)) / A on the upper hemisphere of n, A(theta) = the loop's
// acceptance probability = the share of the lobe above the horizon, a function of theta = angle(r, n) alone
// (A(0) = 1, A(pi/2) = 1/2).  The host tabulates it (PHONG_ACCEPT_N + 1 points, uniform in theta, summed over the
// reference's own 2^15 values of u with the azimuth integrated in closed form); linear interpolation error < 1e-6.
constexpr int PHONG_ACCEPT_N = 1024;
__device__ __forceinline__ float phong_accept_probability(const float* __restrict__ tab, float cos_rn) {
	const float theta = acosf(fminf(fmaxf(cos_rn, 0.0f), 1.0f));
	const float x = theta * (float(PHONG_ACCEPT_N) / 1.57079632679f);
	const int i = min(int(x), PHONG_ACCEPT_N - 1);
	const float a = __ldg(tab + i), b = __ldg(tab + i + 1);
	return a + (b - a) * (x - float(i));
}

__device__ __forceinline__ float random_direction_pdf(float c) {
	if (!(c > 0.0f)) return 0.0f;
	c = fminf(c, 1.0f);
	const float v = 2.0f * c / (c + sqrtf(4.0f - 3.0f * c * c));
	const float u = 1.0f - v;
	const float D = 1.0f - u + u * u;
	return D * sqrtf(D) / (3.14159265358979f * (1.0f + u));
}

// ------------------------------------------------------------------ K4 (history.cl:3-28)
__device__ __forceinline__ bool history_one(f4& new_color, f4& hist, int32_t& count, int max_attempt) {
	f4 now = new_color;
	if (sqrtf(dot4(now, now)) == 0 || count >= max_attempt) {
		new_color = hist;
		return false;
	}
	now = (now + hist * float(count)) / float(count + 1);
	hist = now;
	new_color = now;
	hist.w = 0.0f;
	++count;
	return true;
}

// ------------------------------------------------------------------ deterministic seeding (replaces
// srand(time(NULL)); rand() per pixel, scenebuild.cpp:116-119)
__host__ __device__ __forceinline__ uint32_t mix32(uint32_t x) {
	x ^= x >> 16; x *= 0x7feb352du;
	x ^= x >> 15; x *= 0x846ca68bu;
	x ^= x >> 16;
	return x;
}
__host__ __device__ __forceinline__ uint32_t pixel_seed(uint32_t pixel, uint32_t run_seed) {
	return mix32(pixel ^ mix32(run_seed)) & 0x7FFFu;     // MSVC rand() range
}
__host__ __device__ __forceinline__ uint32_t sample_seed(uint32_t pixel, uint32_t sample, uint32_t run_seed) {
	return mix32(pixel ^ mix32(sample ^ mix32(run_seed)));
}

}  // namespace mcpt
