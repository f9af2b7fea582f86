/*
 * tools/ubench_l2_gather.cu -- what the B200 delivers for the two ways the loop can fetch a
 * factor row per nonzero: a gather from L2 (table resident in L2) and a gather from a panel
 * resident in shared memory.  Development aid; its numbers go into profiles/ and are the
 * measured ceilings used by bench.py's roofline.l2_gather (instead of a B300 figure).
 *
 *   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_l2_gather tools/ubench_l2_gather.cu
 *
 * Tests (all chip-wide, 148 SMs, time by CUDA events, best of 5):
 *   ldg  ROWB  rows of ROWB bytes gathered at random from a 3.2 MB table, LDG.128 to registers
 *   cpa  ROWB  the same with 16-byte cp.async into a shared-memory ring (not read back)
 *   lds  MODE  LDS.128 from a 128 KB shared-memory panel:
 *        0  a quarter-warp reads one random 128-byte row (conflict-free by construction)
 *        1  every lane reads 16 B of its own random 64-byte row   (rows 64 B apart)
 *        2  the same, rows 80 B apart (padded: 8 bank groups by row number)
 *        3  every lane reads 16 B of its own random 32-byte row
 *        4  lanes read consecutive 16-byte pieces (streaming, no conflicts)
 *        5  LDS.64: lane = feature, a warp reads one random 256-byte row
 *        6  LDS.128, all lanes the same address (broadcast)      -- counted as 16 B per warp
 *        7  LDS.64,  all lanes the same address (broadcast)      -- counted as 8 B per warp
 *        8  LDS.128, lane = row, rows 17 pieces apart (the transposed read of the dot phase)
 *        9  the same with 15 of 32 lanes active                  -- counted as 15 x 16 B per warp
 *       10  SHFL of a double (two 32-bit shuffles)               -- counted as 8 B x 32 lanes
 */
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__device__ __forceinline__ unsigned lcg(unsigned &s) { s = s * 1664525u + 1013904223u; return s >> 8; }

template <int PIECES>   /* 16-byte pieces per lane and row: row = 32*16*PIECES bytes */
__global__ void __launch_bounds__(256) k_ldg(const double2 *__restrict__ tab, int rows, int iters, double *sink)
{
	unsigned s = (blockIdx.x * 256 + threadIdx.x) / 32 * 2654435761u + 12345u;   /* per warp */
	const int lane = threadIdx.x & 31;
	double ax = 0.0;
	for (int it = 0; it < iters; it += 4) {
		double2 v[4][PIECES];
#pragma unroll
		for (int u = 0; u < 4; u++) {
			const unsigned r = lcg(s) % (unsigned)rows;
#pragma unroll
			for (int p = 0; p < PIECES; p++)
				v[u][p] = tab[(size_t)r * (32 * PIECES) + p * 32 + lane];
		}
#pragma unroll
		for (int u = 0; u < 4; u++)
#pragma unroll
			for (int p = 0; p < PIECES; p++)
				ax += v[u][p].x + v[u][p].y;
	}
	if (ax == 12345.678)
		*sink = ax;
}

/* the same rows with 8-byte loads (lane = feature): 6 loads of 256 B per 1536-byte row;
 * DEPTH rows in flight per warp, WPB warps per CTA */
template <int DEPTH>
__global__ void k_ldg64(const double *__restrict__ tab, int rows, int iters, double *sink)
{
	unsigned s = (blockIdx.x * blockDim.x + threadIdx.x) / 32 * 2654435761u + 12345u;
	const int lane = threadIdx.x & 31;
	double ax = 0.0;
	for (int it = 0; it < iters; it += DEPTH) {
		double v[DEPTH][6];
#pragma unroll
		for (int u = 0; u < DEPTH; u++) {
			const unsigned r = lcg(s) % (unsigned)rows;
#pragma unroll
			for (int p = 0; p < 6; p++)
				v[u][p] = tab[(size_t)r * 192 + p * 32 + lane];
		}
#pragma unroll
		for (int u = 0; u < DEPTH; u++)
#pragma unroll
			for (int p = 0; p < 6; p++)
				ax += v[u][p];
	}
	if (ax == 12345.678)
		*sink = ax;
}

template <int PIECES>
__global__ void __launch_bounds__(256) k_cpa(const double2 *__restrict__ tab, int rows, int iters, double *sink)
{
	extern __shared__ __align__(16) double2 ring[];      /* per warp: 4 stages x 4 rows */
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	double2 *my = ring + (size_t)warp * 16 * 32 * PIECES;
	unsigned s = (blockIdx.x * 256 + threadIdx.x) / 32 * 2654435761u + 12345u;
	for (int it = 0; it < iters; it += 4) {
		const int stage = (it >> 2) & 3;
#pragma unroll
		for (int u = 0; u < 4; u++) {
			const unsigned r = lcg(s) % (unsigned)rows;
#pragma unroll
			for (int p = 0; p < PIECES; p++) {
				const unsigned dst = (unsigned)__cvta_generic_to_shared(my + ((stage * 4 + u) * PIECES + p) * 32 + lane);
				asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(tab + (size_t)r * (32 * PIECES) + p * 32 + lane) : "memory");
			}
		}
		asm volatile("cp.async.commit_group;" ::: "memory");
		asm volatile("cp.async.wait_group 3;" ::: "memory");
	}
	asm volatile("cp.async.wait_group 0;" ::: "memory");
	if (my[lane].x == 12345.678)
		*sink = my[lane].x;
}

/* the tile gathers of the dot kernels: one cp.async instruction = 128-byte segments of 4 random
 * 1600-byte rows (8 lanes x 16 B each); STAGES x 8 instructions (32 rows) in flight per warp */
template <int STAGES>
__global__ void k_cpa128(const double2 *__restrict__ tab, int rows, int iters, double *sink)
{
	extern __shared__ __align__(16) double2 ring[];      /* per warp: STAGES x 32 rows x 8 pieces */
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	double2 *my = ring + (size_t)warp * STAGES * 256;
	unsigned s = ((blockIdx.x * blockDim.x + threadIdx.x) / 8) * 2654435761u + 12345u;   /* per 8 lanes */
	for (int it = 0; it < iters; it++) {
		const int stage = it % STAGES;
		const unsigned seg = lcg(s) % 12u;                /* which 128 bytes of the row */
#pragma unroll
		for (int u = 0; u < 8; u++) {
			const unsigned r = lcg(s) % (unsigned)rows;
			const unsigned dst = (unsigned)__cvta_generic_to_shared(my + (stage * 32 + u * 4 + (lane >> 3)) * 8 + (lane & 7));
			asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(tab + (size_t)r * 100 + seg * 8 + (lane & 7)) : "memory");
		}
		asm volatile("cp.async.commit_group;" ::: "memory");
		if (STAGES == 2) asm volatile("cp.async.wait_group 1;" ::: "memory");
		else if (STAGES == 4) asm volatile("cp.async.wait_group 3;" ::: "memory");
		else asm volatile("cp.async.wait_group 7;" ::: "memory");
	}
	asm volatile("cp.async.wait_group 0;" ::: "memory");
	if (my[lane].x == 12345.678)
		*sink = my[lane].x;
}

template <int MODE>
__global__ void __launch_bounds__(1024) k_lds(int iters, double *sink)
{
	extern __shared__ __align__(16) double2 panel[];     /* 128 KB = 8192 pieces of 16 B */
	for (int q = threadIdx.x; q < 8192; q += blockDim.x)
		panel[q] = make_double2(q, 1.0);
	__syncthreads();
	const int lane = threadIdx.x & 31;
	unsigned s = (blockIdx.x * 1024 + threadIdx.x) * 2654435761u + 777u;
	if (MODE == 0 || MODE == 5 || MODE >= 6)
		s = (blockIdx.x * 1024 + threadIdx.x) / (MODE == 0 ? 8 : 32) * 2654435761u + 777u;
	double ax = 0.0, ay = 0.0;
	for (int it = 0; it < iters; it += 8) {
		double2 v[8];
#pragma unroll
		for (int u = 0; u < 8; u++) {
			const unsigned r = lcg(s);
			unsigned piece;
			if (MODE == 0) piece = (r % 1024u) * 8 + (lane & 7);                    /* 128-byte rows   */
			else if (MODE == 1) piece = (r % 2048u) * 4 + (u & 3);                  /* 64-byte rows    */
			else if (MODE == 2) piece = (r % 1638u) * 5 + (u & 3);                  /* 80-byte stride  */
			else if (MODE == 3) piece = (r % 4096u) * 2 + (u & 1);                  /* 32-byte rows    */
			else if (MODE == 4) piece = ((r % 200u) * 32 + lane) % 8192u;           /* streaming       */
			else if (MODE == 6 || MODE == 7) piece = r % 8192u;
			else if (MODE == 8 || MODE == 9) piece = ((r % 200u) + lane * 17) % 8192u;
			else piece = 0;
			if (MODE == 6 || MODE == 7) piece = __shfl_sync(0xffffffffu, piece, 0);
			if (MODE == 10) {
				v[u] = make_double2(__shfl_sync(0xffffffffu, ax + u, (lane + r) & 31), 0.0);
			} else if (MODE == 7) {
				v[u] = make_double2(reinterpret_cast<const double *>(panel)[piece], 0.0);
			} else if (MODE == 9) {
				v[u] = make_double2(0.0, 0.0);
				if (lane < 15) v[u] = panel[piece];
			} else
			if (MODE == 5) {
				const double *p8 = reinterpret_cast<const double *>(panel) + (r % 512u) * 32 + lane;
				v[u] = make_double2(*p8, 0.0);
			} else
				v[u] = panel[piece];
		}
#pragma unroll
		for (int u = 0; u < 8; u++) { ax += v[u].x; ay += v[u].y; }
	}
	if (ax + ay == 12345.678)
		*sink = ax;
}

static float time_best(void (*launch)(void *), void *arg)
{
	cudaEvent_t a, b;
	CK(cudaEventCreate(&a));
	CK(cudaEventCreate(&b));
	float best = 1e30f;
	for (int rep = 0; rep < 6; rep++) {
		CK(cudaEventRecord(a));
		launch(arg);
		CK(cudaEventRecord(b));
		CK(cudaEventSynchronize(b));
		float ms;
		CK(cudaEventElapsedTime(&ms, a, b));
		if (rep && ms < best) best = ms;
	}
	CK(cudaGetLastError());
	return best;
}

struct Ctx { const double2 *tab; int rows, iters, blocks; double *sink; int which; };

template <int P> static void run_ldg(void *p) { Ctx *c = (Ctx *)p; k_ldg<P><<<c->blocks, 256>>>(c->tab, c->rows, c->iters, c->sink); }
template <int P> static void run_cpa(void *p) { Ctx *c = (Ctx *)p; k_cpa<P><<<c->blocks, 256, 8 * 16 * 32 * P * 16>>>(c->tab, c->rows, c->iters, c->sink); }
template <int D> static void run_ldg64(void *p) { Ctx *c = (Ctx *)p; k_ldg64<D><<<c->blocks, c->which>>>(c->tab ? (const double *)c->tab : nullptr, c->rows, c->iters, c->sink); }
template <int S> static void run_cpa128(void *p) { Ctx *c = (Ctx *)p; k_cpa128<S><<<c->blocks, c->which, (size_t)(c->which / 32) * S * 4096>>>(c->tab, c->rows, c->iters, c->sink); }
template <int M> static void run_lds(void *p) { Ctx *c = (Ctx *)p; k_lds<M><<<c->blocks, 1024, 131072>>>(c->iters, c->sink); }

int main()
{
	int dev = 0, sms = 0, khz = 0;
	CK(cudaSetDevice(dev));
	CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
	CK(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev));
	printf("SMs %d, max SM clock %.0f MHz\n", sms, khz / 1e3);
	const size_t tab_bytes = 3276800;    /* 2000 x 200 doubles: R of the inst60000 shape */
	double2 *tab;
	double *sink;
	CK(cudaMalloc(&tab, tab_bytes));
	CK(cudaMemset(tab, 0, tab_bytes));
	CK(cudaMalloc(&sink, 8));
	Ctx c;
	c.tab = tab; c.sink = sink;

	/* L2 gathers: 8 warps per CTA; CTAs per SM by the ring (cpa) or 8 (ldg) */
	const int iters = 2048;
	c.iters = iters;
#define LDG(P) do { c.rows = (int)(tab_bytes / (512 * P)); c.blocks = sms * 8; CK(cudaFuncSetAttribute(k_ldg<P>, cudaFuncAttributeMaxDynamicSharedMemorySize, 0)); \
		float ms = time_best(run_ldg<P>, &c); double bytes = (double)c.blocks * 8 * iters * 512.0 * P; \
		printf("ldg  row %5d B: %8.1f GB/s  (%.1f B/clk/SM at max clock)\n", 512 * P, bytes / ms / 1e6, bytes / ms / 1e6 / sms / (khz / 1e6)); } while (0)
	LDG(1); LDG(2); LDG(3);
#define CPA(P, PERSM) do { c.rows = (int)(tab_bytes / (512 * P)); c.blocks = sms * PERSM; CK(cudaFuncSetAttribute(k_cpa<P>, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 16 * 32 * P * 16)); \
		float ms = time_best(run_cpa<P>, &c); double bytes = (double)c.blocks * 8 * iters * 512.0 * P; \
		printf("cpa  row %5d B, %d CTAs/SM: %8.1f GB/s  (%.1f B/clk/SM)\n", 512 * P, PERSM, bytes / ms / 1e6, bytes / ms / 1e6 / sms / (khz / 1e6)); } while (0)
	CPA(1, 3); CPA(2, 1); CPA(3, 1); CPA(1, 1);
#define LDG64(D, WARPS_SM) do { c.rows = (int)(tab_bytes / 1536); c.which = 128; c.blocks = sms * (WARPS_SM) / 4; \
		float ms = time_best(run_ldg64<D>, &c); double bytes = (double)c.blocks * 4 * iters * 1536.0; \
		printf("ldg64 row 1536 B, %d rows in flight per warp, %d warps/SM: %8.1f GB/s  (%.1f B/clk/SM)\n", D, WARPS_SM, bytes / ms / 1e6, bytes / ms / 1e6 / sms / (khz / 1e6)); } while (0)
#define CPA128(S, WARPS, PERSM) do { c.rows = (int)(tab_bytes / 1600); c.which = 32 * WARPS; c.blocks = sms * PERSM; \
		CK(cudaFuncSetAttribute(k_cpa128<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, WARPS * S * 4096)); \
		float ms = time_best(run_cpa128<S>, &c); double bytes = (double)c.blocks * WARPS * iters * 4096.0; \
		printf("cpa128 (4 rows x 128 B per instruction), %d stages, %d x %d warps/SM: %8.1f GB/s  (%.1f B/clk/SM)\n", S, PERSM, WARPS, bytes / ms / 1e6, bytes / ms / 1e6 / sms / (khz / 1e6)); } while (0)
	CPA128(2, 8, 2); CPA128(2, 8, 1); CPA128(4, 8, 1); CPA128(4, 8, 2);
	LDG64(4, 64); LDG64(4, 32); LDG64(8, 16); LDG64(8, 12); LDG64(8, 24); LDG64(2, 64);
	{
		/* the same gathers from a 96 MB table (L of the inst60000 shape): L2 + HBM */
		const size_t big_bytes = 96000000;
		double2 *big;
		CK(cudaMalloc(&big, big_bytes));
		CK(cudaMemset(big, 0, big_bytes));
		c.tab = big;
		c.rows = (int)(big_bytes / 1536); c.blocks = sms * 8;
		float ms = time_best(run_ldg<3>, &c); double bytes = (double)c.blocks * 8 * iters * 1536.0;
		printf("ldg  row  1536 B, 96 MB table: %8.1f GB/s  (%.1f B/clk/SM)\n", bytes / ms / 1e6, bytes / ms / 1e6 / sms / (khz / 1e6));
		c.blocks = sms;
		ms = time_best(run_cpa<3>, &c); bytes = (double)c.blocks * 8 * iters * 1536.0;
		printf("cpa  row  1536 B, 96 MB table, 1 CTA/SM: %8.1f GB/s  (%.1f B/clk/SM)\n", bytes / ms / 1e6, bytes / ms / 1e6 / sms / (khz / 1e6));
		c.tab = tab;
		CK(cudaFree(big));
	}

	/* shared-memory gathers: one CTA of 1024 threads per SM, 128 KB panel */
	c.blocks = sms;
	c.iters = 8192;
#define LDS(M, BYTES) do { CK(cudaFuncSetAttribute(k_lds<M>, cudaFuncAttributeMaxDynamicSharedMemorySize, 131072)); \
		float ms = time_best(run_lds<M>, &c); double bytes = (double)c.blocks * 1024 * c.iters * BYTES; \
		printf("lds  mode %d: %8.1f GB/s  (%.1f B/clk/SM at max clock)\n", M, bytes / ms / 1e6, bytes / ms / 1e6 / sms / (khz / 1e6)); } while (0)
	LDS(0, 16.0); LDS(1, 16.0); LDS(2, 16.0); LDS(3, 16.0); LDS(4, 16.0); LDS(5, 8.0);
	printf("the next three count bytes per WARP instruction: B/clk/SM / bytes = warp instructions per clock\n");
	LDS(6, 16.0 / 32); LDS(7, 8.0 / 32); LDS(8, 16.0); LDS(9, 16.0 * 15 / 32); LDS(10, 8.0);
	return 0;
}
