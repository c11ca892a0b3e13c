// TEST INFRASTRUCTURE — OpenCL C compatibility layer for compiling the REFERENCE'S OWN kernel sources
// (kernels/*.cl and the device half of objdef.h) as C++ with g++, so that the oracle's restatement can be checked
// against the reference's code instead of against a reading of it.  Only oracle/ref_kernels.cpp includes this file.
//
// What it supplies: the OpenCL vector types the kernels use (float2/3/4/16, int2/4, uint) with .x/.y/.z/.w, the
// .sN and multi-lane swizzles that appear in the sources (.s012, .s048, .s159, .s26a), component-wise operators,
// make_floatN for the `(floatN)(...)` literals (rewritten by build_ref.sh, the one construct C++ cannot parse), the
// work-item functions, vstore16, and the math built-ins.
//
// The built-ins are where OpenCL is implementation-defined; they follow the conventions written at the top of
// oracle/mcpt_oracle.cpp (which the CUDA kernels mirror): no FMA contraction (-ffp-contract=off), IEEE + - * / sqrt,
// dot() accumulating left to right over the operand's lanes, normalize(v) = v / sqrt(dot(v, v)),
// length(v) = sqrt(dot(v, v)), cross() with w = 0, fmin/fmax returning the non-NaN operand, and tan / cos / sin /
// pow / acos / asin = the correctly rounded float result (evaluated in double, narrowed).  M_PI is the double
// constant, as in an OpenCL build with cl_khr_fp64.
#pragma once
#include <cfloat>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <cstring>

namespace clc {

typedef unsigned int uint;
struct float3;

// a view of lanes A, B, C of the enclosing vector (the union member sits at offset 0 of the lane array)
template <int N, int A, int B, int C>
struct Swizzle3 {
	float lane[N];                                    // overlays the N lanes of the enclosing vector
	inline operator float3() const;
	inline Swizzle3& operator=(const float3& v);      // copy assignment stays the trivial one (see build_ref.sh)
};

struct float2 { float x, y; };
struct int2 { int x, y; };

struct float3 {
	union {
		struct { float x, y, z; };
		float s[3];
	};
};

struct float4 {
	union {
		struct { float x, y, z, w; };
		float s[4];
		Swizzle3<4, 0, 1, 2> s012;
	};
};

struct int4 {
	union {
		struct { int x, y, z, w; };
		int s[4];
	};
};

struct float16 {
	union {
		struct { float s0, s1, s2, s3, s4, s5, s6, s7, s8, s9, sa, sb, sc, sd, se, sf; };
		float s[16];
		Swizzle3<16, 0, 4, 8> s048;
		Swizzle3<16, 1, 5, 9> s159;
		Swizzle3<16, 2, 6, 10> s26a;
	};
};

template <int N, int A, int B, int C>
inline Swizzle3<N, A, B, C>::operator float3() const {
	float3 r;
	r.x = lane[A]; r.y = lane[B]; r.z = lane[C];
	return r;
}
template <int N, int A, int B, int C>
inline Swizzle3<N, A, B, C>& Swizzle3<N, A, B, C>::operator=(const float3& v) {
	lane[A] = v.x; lane[B] = v.y; lane[C] = v.z;
	return *this;
}

// ---- vector literals: (floatN)(...) takes scalars and shorter vectors in any mix
struct Lanes {
	float v[16];
	int n = 0;
	void put(float a) { v[n++] = a; }
	void put(double a) { v[n++] = float(a); }
	void put(int a) { v[n++] = float(a); }
	void put(unsigned a) { v[n++] = float(a); }
	void put(long a) { v[n++] = float(a); }
	void put(unsigned long a) { v[n++] = float(a); }
	void put(const float2& a) { put(a.x); put(a.y); }
	void put(const float3& a) { put(a.x); put(a.y); put(a.z); }
	void put(const float4& a) { put(a.x); put(a.y); put(a.z); put(a.w); }
	template <int N, int A, int B, int C>
	void put(const Swizzle3<N, A, B, C>& a) { put(float3(a)); }
};
template <class... T>
inline Lanes gather(const T&... a) {
	Lanes l;
	(l.put(a), ...);
	return l;
}
template <class... T> inline float2 make_float2(const T&... a) { Lanes l = gather(a...); float2 r; r.x = l.v[0]; r.y = l.v[1]; return r; }
template <class... T> inline float3 make_float3(const T&... a) { Lanes l = gather(a...); float3 r; r.x = l.v[0]; r.y = l.v[1]; r.z = l.v[2]; return r; }
template <class... T> inline float4 make_float4(const T&... a) { Lanes l = gather(a...); float4 r; r.x = l.v[0]; r.y = l.v[1]; r.z = l.v[2]; r.w = l.v[3]; return r; }
template <class... T> inline float16 make_float16(const T&... a) { Lanes l = gather(a...); float16 r; for (int i = 0; i < 16; ++i) r.s[i] = l.v[i]; return r; }
template <class A, class B> inline int2 make_int2(const A& a, const B& b) { int2 r; r.x = int(a); r.y = int(b); return r; }

// ---- component-wise arithmetic (scalar operands widen to every lane, integers convert to float first)
#define CLC_VEC_OPS(V, N)                                                                                             \
	inline V operator+(const V& a, const V& b) { V r; for (int i = 0; i < N; ++i) r.s[i] = a.s[i] + b.s[i]; return r; } \
	inline V operator-(const V& a, const V& b) { V r; for (int i = 0; i < N; ++i) r.s[i] = a.s[i] - b.s[i]; return r; } \
	inline V operator*(const V& a, const V& b) { V r; for (int i = 0; i < N; ++i) r.s[i] = a.s[i] * b.s[i]; return r; } \
	inline V operator/(const V& a, const V& b) { V r; for (int i = 0; i < N; ++i) r.s[i] = a.s[i] / b.s[i]; return r; } \
	inline V operator-(const V& a) { V r; for (int i = 0; i < N; ++i) r.s[i] = -a.s[i]; return r; }                     \
	inline V operator*(const V& a, float b) { V r; for (int i = 0; i < N; ++i) r.s[i] = a.s[i] * b; return r; }         \
	inline V operator*(float a, const V& b) { V r; for (int i = 0; i < N; ++i) r.s[i] = a * b.s[i]; return r; }         \
	inline V operator/(const V& a, float b) { V r; for (int i = 0; i < N; ++i) r.s[i] = a.s[i] / b; return r; }         \
	inline V operator+(const V& a, float b) { V r; for (int i = 0; i < N; ++i) r.s[i] = a.s[i] + b; return r; }
CLC_VEC_OPS(float3, 3)
CLC_VEC_OPS(float4, 4)
#undef CLC_VEC_OPS
template <int N, int A, int B, int C> inline float3 operator-(const Swizzle3<N, A, B, C>& a) { return -float3(a); }

// ---- geometric built-ins
inline float dot(const float3& a, const float3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline float dot(const float4& a, const float4& b) { return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w; }
inline float length(const float3& a) { return std::sqrt(dot(a, a)); }
inline float length(const float4& a) { return std::sqrt(dot(a, a)); }
inline float3 normalize(const float3& a) { return a / std::sqrt(dot(a, a)); }
inline float4 normalize(const float4& a) { return a / std::sqrt(dot(a, a)); }
inline float4 cross(const float4& a, const float4& b) {
	float4 r;
	r.x = a.y * b.z - a.z * b.y; r.y = a.z * b.x - a.x * b.z; r.z = a.x * b.y - a.y * b.x; r.w = 0.0f;
	return r;
}
inline float3 cross(const float3& a, const float3& b) {
	float3 r;
	r.x = a.y * b.z - a.z * b.y; r.y = a.z * b.x - a.x * b.z; r.z = a.x * b.y - a.y * b.x;
	return r;
}

// ---- math built-ins (float in, float out)
inline float fabs(float a) { return std::fabs(a); }
inline float sqrt(float a) { return std::sqrt(a); }
inline float fmin(float a, float b) { return std::fmin(a, b); }
inline float fmax(float a, float b) { return std::fmax(a, b); }
inline float4 fmin(const float4& a, const float4& b) { float4 r; for (int i = 0; i < 4; ++i) r.s[i] = std::fmin(a.s[i], b.s[i]); return r; }
inline float4 fmax(const float4& a, const float4& b) { float4 r; for (int i = 0; i < 4; ++i) r.s[i] = std::fmax(a.s[i], b.s[i]); return r; }
inline float tan(float a) { return float(std::tan(double(a))); }
inline float cos(float a) { return float(std::cos(double(a))); }
inline float sin(float a) { return float(std::sin(double(a))); }
inline float acos(float a) { return float(std::acos(double(a))); }
inline float asin(float a) { return float(std::asin(double(a))); }
inline float pow(float a, float b) { return float(std::pow(double(a), double(b))); }

inline void vstore16(const float16& m, size_t offset, float* p) { std::memcpy(p + 16 * offset, m.s, sizeof(m.s)); }

// ---- work-item functions: the harness sets these before every kernel invocation
struct WorkItem { size_t id[3], size[3]; };
inline thread_local WorkItem g_item = {{0, 0, 0}, {1, 1, 1}};
inline size_t get_global_id(int d) { return g_item.id[d]; }
inline size_t get_global_size(int d) { return g_item.size[d]; }

// ---- images: a write-only image2d_t is a float4 raster the harness owns
struct Image2D { float4* data; size_t width; };
typedef Image2D* image2d_t;
inline void write_imagef(image2d_t img, const int2& at, const float4& c) { img->data[size_t(at.y) * img->width + size_t(at.x)] = c; }

}  // namespace clc
