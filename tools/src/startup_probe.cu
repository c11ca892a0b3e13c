// Development aid: where a process's CUDA start-up time goes (tools/gpu_run13_8gpu.sh).  Prints one JSON line.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
__global__ void k_touch(int* p) { if (p) *p = 1; }
int main() {
	const double t0 = now();
	int n = 0;
	cudaError_t e = cudaGetDeviceCount(&n);                 // driver initialisation (cuInit)
	const double t1 = now();
	cudaSetDevice(0);
	cudaFree(0);                                            // primary context of device 0
	const double t2 = now();
	void* p = nullptr;
	cudaMalloc(&p, size_t(1) << 30);
	const double t3 = now();
	k_touch<<<1, 1>>>(static_cast<int*>(p));                // first launch: module load
	cudaDeviceSynchronize();
	const double t4 = now();
	void* q = nullptr;
	cudaMalloc(&q, size_t(16) << 30);
	cudaMemset(q, 0, size_t(16) << 30);
	cudaDeviceSynchronize();
	const double t5 = now();
	const char* vis = std::getenv("CUDA_VISIBLE_DEVICES");
	std::printf("{\"visible\": \"%s\", \"devices\": %d, \"err\": %d, \"cuInit_s\": %.3f, \"context_s\": %.3f, \"malloc_1g_s\": %.3f, \"first_launch_s\": %.3f, \"malloc_memset_16g_s\": %.3f, \"total_s\": %.3f}\n",
	            vis ? vis : "", n, int(e), t1 - t0, t2 - t1, t3 - t2, t4 - t3, t5 - t4, t5 - t0);
	return 0;
}
