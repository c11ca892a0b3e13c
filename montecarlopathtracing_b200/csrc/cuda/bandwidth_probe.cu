// Measurement aid for bench.py: streaming-read bandwidth of a buffer of a given size, measured on the device
// the walk runs on.  32 MiB stays in the 126 MB L2 (after the first sweep) and gives the L2 ceiling the BVH walk's
// node and triangle fetches live under; a multi-GiB buffer gives the HBM ceiling.  Not part of the render path.
#include <cuda_runtime.h>

#include <cstdint>
#include <string>

#include "mcpt.h"

namespace mcpt {
void set_last_error(const std::string& s);     // mcpt_cuda.cu
namespace {

__global__ void __launch_bounds__(256) k_read_sweep(const uint4* __restrict__ buf, size_t n16, int iters, unsigned* sink) {
	unsigned acc = 0;
	const size_t stride = size_t(gridDim.x) * blockDim.x;
	for (int it = 0; it < iters; ++it) {
		size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
		for (; i + 3 * stride < n16; i += 4 * stride) {      // four independent 16-byte loads in flight per thread
			const uint4 a = __ldcg(buf + i), b = __ldcg(buf + i + stride), c = __ldcg(buf + i + 2 * stride), d = __ldcg(buf + i + 3 * stride);
			acc ^= a.x ^ a.y ^ a.z ^ a.w ^ b.x ^ b.y ^ b.z ^ b.w ^ c.x ^ c.y ^ c.z ^ c.w ^ d.x ^ d.y ^ d.z ^ d.w;
		}
		for (; i < n16; i += stride) {
			const uint4 a = __ldcg(buf + i);
			acc ^= a.x ^ a.y ^ a.z ^ a.w;
		}
	}
	if (acc == 0x9E3779B9u) *sink = acc;                     // keeps the loads alive
}

}  // namespace
}  // namespace mcpt

using namespace mcpt;

#define TRY_(x)                                                                                         \
	do {                                                                                                \
		cudaError_t e_ = (x);                                                                           \
		if (e_ != cudaSuccess) {                                                                        \
			set_last_error(std::string(#x) + ": " + cudaGetErrorString(e_));                            \
			if (buf) cudaFree(buf);                                                                     \
			return MCPT_ERR_CUDA;                                                                       \
		}                                                                                               \
	} while (0)

extern "C" int mcpt_measure_read_bandwidth(int device, size_t bytes, int iters, float* gb_per_s) {
	void* buf = nullptr;
	if (!gb_per_s || bytes < 4096 || iters < 1) {
		set_last_error("mcpt_measure_read_bandwidth: expected bytes >= 4096, iters >= 1 and a result pointer");
		return MCPT_ERR_ARG;
	}
	TRY_(cudaSetDevice(device));
	cudaDeviceProp prop{};
	TRY_(cudaGetDeviceProperties(&prop, device));
	const size_t n16 = bytes / 16;
	TRY_(cudaMalloc(&buf, n16 * 16 + 16));
	TRY_(cudaMemset(buf, 1, n16 * 16 + 16));
	unsigned* sink = reinterpret_cast<unsigned*>(static_cast<char*>(buf) + n16 * 16);
	const int grid = prop.multiProcessorCount * 8;
	cudaEvent_t e0, e1;
	TRY_(cudaEventCreate(&e0));
	TRY_(cudaEventCreate(&e1));
	k_read_sweep<<<grid, 256>>>(static_cast<const uint4*>(buf), n16, 2, sink);        // warm-up: fills L2 where it fits
	float best = 0.0f;
	for (int rep = 0; rep < 5; ++rep) {
		TRY_(cudaEventRecord(e0));
		k_read_sweep<<<grid, 256>>>(static_cast<const uint4*>(buf), n16, iters, sink);
		TRY_(cudaEventRecord(e1));
		TRY_(cudaEventSynchronize(e1));
		float ms = 0.0f;
		TRY_(cudaEventElapsedTime(&ms, e0, e1));
		if (ms > 0.0f) {
			const float g = float(double(n16) * 16.0 * iters / (double(ms) * 1e-3) / 1e9);
			if (g > best) best = g;
		}
	}
	TRY_(cudaGetLastError());
	cudaEventDestroy(e0);
	cudaEventDestroy(e1);
	cudaFree(buf);
	*gb_per_s = best;
	return MCPT_OK;
}
