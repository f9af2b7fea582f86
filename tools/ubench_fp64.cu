/*
 * tools/ubench_fp64.cu -- latency probes for the instructions the accumulation chains are
 * made of (DADD, DMUL, SHFL, LDS), one warp on one SM, clock64 around unrolled sequences.
 * Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -o bin/ubench_fp64 tools/ubench_fp64.cu
 * Output: cycles per step for each pattern; profiles/r01_ubench_fp64.txt keeps one run.
 */
#include <cstdio>
#include <cuda_runtime.h>

#define FULL 0xffffffffu
constexpr int N = 16;       /* steps per unrolled block */
constexpr int REP = 256;    /* blocks */

__global__ void probe(const double *__restrict__ in, double *out, long long *cyc, int warps_active)
{
	const int lane = threadIdx.x & 31;
	const int warp = threadIdx.x >> 5;
	__shared__ double sm[32 * 64];
	for (int i = threadIdx.x; i < 32 * 64; i += blockDim.x)
		sm[i] = in[i % 1024];
	__syncthreads();
	double x[N];
#pragma unroll
	for (int r = 0; r < N; r++)
		x[r] = in[r * 32 + lane];
	double be = in[512 + lane];
	double acc = in[lane];
	long long t0, t1;
	int slot = 0;

	/* 0: dependent DADD chain */
	t0 = clock64();
	for (int it = 0; it < REP; it++) {
#pragma unroll
		for (int r = 0; r < N; r++)
			acc = __dadd_rn(acc, x[r]);
	}
	t1 = clock64();
	if (threadIdx.x == 0) cyc[slot] = t1 - t0;
	slot++;

	/* 1: dependent DMUL chain */
	t0 = clock64();
	for (int it = 0; it < REP; it++) {
#pragma unroll
		for (int r = 0; r < N; r++)
			acc = __dmul_rn(acc, x[r]);
	}
	t1 = clock64();
	if (threadIdx.x == 0) cyc[slot] = t1 - t0;
	slot++;
	acc = in[lane] + acc * 1e-300;

	/* 2: acc += c * x[r], c in a register (independent DMULs) */
	{
		double c = be;
		t0 = clock64();
		for (int it = 0; it < REP; it++) {
#pragma unroll
			for (int r = 0; r < N; r++)
				acc = __dadd_rn(acc, __dmul_rn(c, x[r]));
			c = __dadd_rn(c, 1e-9);
		}
		t1 = clock64();
		if (threadIdx.x == 0) cyc[slot] = t1 - t0;
		slot++;
	}

	/* 3: acc += shfl(be, r) * x[r]  (the chain as written in chain_rows) */
	t0 = clock64();
	for (int it = 0; it < REP; it++) {
#pragma unroll
		for (int r = 0; r < N; r++)
			acc = __dadd_rn(acc, __dmul_rn(__shfl_sync(FULL, be, r), x[r]));
		be = __dadd_rn(be, 1e-9);
	}
	t1 = clock64();
	if (threadIdx.x == 0) cyc[slot] = t1 - t0;
	slot++;

	/* 4: products first, then the additions */
	t0 = clock64();
	for (int it = 0; it < REP; it++) {
		double p[N];
#pragma unroll
		for (int r = 0; r < N; r++)
			p[r] = __dmul_rn(__shfl_sync(FULL, be, r), x[r]);
#pragma unroll
		for (int r = 0; r < N; r++)
			acc = __dadd_rn(acc, p[r]);
		be = __dadd_rn(be, 1e-9);
	}
	t1 = clock64();
	if (threadIdx.x == 0) cyc[slot] = t1 - t0;
	slot++;

	/* 5: e broadcast from shared memory (same address for all lanes), x from registers */
	t0 = clock64();
	for (int it = 0; it < REP; it++) {
		const double *e = sm + (it & 63) * 16;
#pragma unroll
		for (int r = 0; r < N; r++)
			acc = __dadd_rn(acc, __dmul_rn(e[r], x[r]));
	}
	t1 = clock64();
	if (threadIdx.x == 0) cyc[slot] = t1 - t0;
	slot++;

	/* 6: x from shared memory (lane-strided), e by shuffle: the ring variant */
	t0 = clock64();
	for (int it = 0; it < REP; it++) {
		const double *xs = sm + (it & 3) * 512 + lane;
#pragma unroll
		for (int r = 0; r < N; r++)
			acc = __dadd_rn(acc, __dmul_rn(__shfl_sync(FULL, be, r), xs[r * 32]));
		be = __dadd_rn(be, 1e-9);
	}
	t1 = clock64();
	if (threadIdx.x == 0) cyc[slot] = t1 - t0;
	slot++;

	/* 7: dependent SHFL chain (32-bit) */
	{
		int v = lane;
		t0 = clock64();
		for (int it = 0; it < REP; it++) {
#pragma unroll
			for (int r = 0; r < N; r++)
				v = __shfl_sync(FULL, v, (v + r) & 31);
		}
		t1 = clock64();
		if (threadIdx.x == 0) cyc[slot] = t1 - t0;
		slot++;
		acc += v * 1e-300;
	}

	/* 8: dependent LDS chain */
	{
		int v = lane;
		const int *smi = (const int *)sm;
		t0 = clock64();
		for (int it = 0; it < REP; it++) {
#pragma unroll
			for (int r = 0; r < N; r++)
				v = smi[(v & 1023)] & 1023;
		}
		t1 = clock64();
		if (threadIdx.x == 0) cyc[slot] = t1 - t0;
		slot++;
		acc += v * 1e-300;
	}

	/* 9: dependent global load chain (L2 hits after the first pass) */
	{
		long long v = lane;
		const long long *g = (const long long *)in;
		t0 = clock64();
		for (int it = 0; it < 16; it++) {
#pragma unroll
			for (int r = 0; r < N; r++)
				v = __ldcg(g + 2048 + (v & 1023)) & 1023;
		}
		t1 = clock64();
		if (threadIdx.x == 0) cyc[slot] = (t1 - t0) * (REP / 16);
		slot++;
		acc += v * 1e-300;
	}
	(void)warps_active;
	out[warp * 32 + lane] = acc;
}

int main()
{
	double *in, *out;
	long long *cyc;
	cudaMalloc(&in, 8192 * sizeof(double));
	cudaMalloc(&out, 1024 * sizeof(double));
	cudaMalloc(&cyc, 16 * sizeof(long long));
	double h[8192];
	for (int i = 0; i < 8192; i++)
		h[i] = 1.0 + 1e-6 * (i % 97);
	long long hl[8192];
	cudaMemcpy(in, h, sizeof(h), cudaMemcpyHostToDevice);
	/* second half of the buffer doubles as the pointer-chase table of probe 9 */
	for (int i = 0; i < 1024; i++)
		hl[i] = (i * 37 + 11) & 1023;
	cudaMemcpy((long long *)in + 2048, hl, 1024 * sizeof(long long), cudaMemcpyHostToDevice);
	static const char *name[] = {
		"DADD chain", "DMUL chain", "acc += c*x[r] (register c)", "acc += shfl(e,r)*x[r] (as written)",
		"products first, then DADD chain", "e from smem broadcast, x regs", "e by shfl, x from smem",
		"SHFL dependent", "LDS dependent", "LDG.cg dependent (L2)"};
	for (int warps = 1; warps <= 8; warps *= 2) {
		probe<<<1, 32 * warps>>>(in, out, cyc, warps);
		probe<<<1, 32 * warps>>>(in, out, cyc, warps);
		long long c[16];
		cudaMemcpy(c, cyc, sizeof(c), cudaMemcpyDeviceToHost);
		cudaError_t e = cudaDeviceSynchronize();
		if (e != cudaSuccess) {
			printf("cuda error %s\n", cudaGetErrorString(e));
			return 1;
		}
		printf("warps per CTA = %d (warp 0 timed)\n", warps);
		for (int s = 0; s < 10; s++)
			printf("  %-40s %7.2f cycles/step\n", name[s], (double)c[s] / (N * REP));
	}
	return 0;
}
