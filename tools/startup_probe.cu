// tools/startup_probe.cu -- where the ~0.7-1.0 s of CUDA start-up of `bin/matFact` go (development
// aid): wall clock of the first runtime calls of a fresh process, in the order the library makes them.
//   nvcc -gencode arch=compute_100a,code=sm_100a -o build/startup_probe tools/startup_probe.cu
#include <cstdio>
#include <ctime>
#include <cuda_runtime.h>

static double now()
{
	timespec ts;
	clock_gettime(CLOCK_MONOTONIC, &ts);
	return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

__global__ void k_nop(int *p) { if (p) *p = 1; }

#define LAP(name, call)                                              \
	do {                                                         \
		const double t0 = now();                             \
		call;                                                \
		printf("%-28s %8.1f ms\n", name, 1e3 * (now() - t0)); \
	} while (0)

int main()
{
	int n = 0;
	void *p = nullptr, *h = nullptr;
	cudaStream_t st;
	const double t0 = now();
	LAP("cudaGetDeviceCount", cudaGetDeviceCount(&n));
	LAP("cudaSetDevice(0)", cudaSetDevice(0));
	LAP("cudaFree(0)", cudaFree(0));
	LAP("cudaStreamCreate", cudaStreamCreate(&st));
	LAP("cudaMallocAsync 1 MB + sync", (cudaMallocAsync(&p, 1 << 20, st), cudaStreamSynchronize(st)));
	LAP("first kernel + sync", (k_nop<<<1, 32, 0, st>>>((int *)p), cudaStreamSynchronize(st)));
	LAP("cudaHostAlloc 16 MB", cudaHostAlloc(&h, 16 << 20, cudaHostAllocDefault));
	LAP("cudaMalloc 8 GB", cudaMalloc(&p, (size_t)8 << 30));
	LAP("cudaMallocAsync 8 GB + sync", (cudaMallocAsync(&p, (size_t)8 << 30, st), cudaStreamSynchronize(st)));
	printf("%-28s %8.1f ms  (%d devices, last error: %s)\n", "total", 1e3 * (now() - t0), n,
	       cudaGetErrorString(cudaGetLastError()));
	return 0;
}
