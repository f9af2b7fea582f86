/*
 * csrc/mf_argmax.cuh -- masked arg-max over B = L * R^T without storing B (sm_100a).
 *
 * Replaces mat2d_prod (mat2d.c:100-113) followed by print_output
 * (matFact.c:10-27).  Every B[i][j] is produced exactly as the reference does:
 * dest = 0; dest = dest + L[i][k]*R[j][k] for ascending k, multiply and add
 * rounded separately -- so the values compared are bit-identical and no
 * "screen then refine" pass is needed.  print_output scans j ascending, skips
 * the items its cursor matches, takes the first remaining item unconditionally
 * and afterwards only a strictly greater value; that is: the maximum over the
 * unmasked, non-NaN values with the lowest index winning ties, except that a
 * NaN at the first unmasked item is never displaced.
 *
 * Tiling: a CTA of 256 threads owns 64 users and walks a contiguous range of
 * 64-item tiles ("split"); each thread keeps a 4x4 block of B in registers,
 * L/R panels of 16 features go through shared memory (k-major, so the 4
 * operands of a thread are one 32-byte read).  Per user the running best of
 * every split is written to part_v/part_j and k_argmax_final folds the splits
 * in ascending item order.
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "mf_types.h"

namespace mf {


__device__ __forceinline__ bool better(double v, int j, double bv, int bj)
{
	/* NaN never wins; ties go to the lower item */
	return (v == v) && (bj < 0 || v > bv || (v == bv && j < bj));
}

__global__ void __launch_bounds__(256)
k_masked_argmax(const ArgmaxArgs p, const double *__restrict__ L, const double *__restrict__ R)
{
	__shared__ __align__(16) double Ls[AM_KC][AM_LDS];
	/* item c of the tile sits at column c + 2*(c >> 4): 16 bytes of padding after every four
	 * threads' items, so that the 16-byte reads of a quarter-warp (8 threads x 32-byte stride)
	 * fall in 8 distinct bank groups (ncu on the unpadded layout: 41 % of the shared-memory
	 * wavefronts of this kernel were bank conflicts, profiles/r02_notes.md) */
	__shared__ __align__(16) double Rs[AM_KC][AM_LDS + 8];
	__shared__ unsigned long long mbits[AM_TI];
	__shared__ int mcur[AM_TI];
	__shared__ double red_v[AM_TI][16];
	__shared__ int red_j[AM_TI][16];

	const int tid = threadIdx.x;
	const int tx = tid & 15, ty = tid >> 4;          /* items / users */
	const int u0 = p.users_lo + (int)blockIdx.y * AM_TI;
	const int split = blockIdx.x;
	const int t_begin = p.tile_begin + split * p.tiles_per_split;
	const int t_end = min(p.tile_end, t_begin + p.tiles_per_split);
	const int F = p.F, ld = p.ld;

	/* per-user cursor into the ascending mask list, positioned at the split start */
	if (tid < AM_TI) {
		const int u = u0 + tid;
		int lo = 0;
		if (u < p.users_hi) {
			int a = p.mask_ptr[u], b = p.mask_ptr[u + 1];
			const int jstart = t_begin * AM_TJ;
			while (a < b) {
				const int mid = a + ((b - a) >> 1);
				if (p.mask_col[mid] < jstart) a = mid + 1; else b = mid;
			}
			lo = a;
		}
		mcur[tid] = lo;
	}

	double bv[4];
	int bj[4];
#pragma unroll
	for (int a = 0; a < 4; a++) { bv[a] = 0.0; bj[a] = -1; }

	/* loader mapping: 4 rows of 16 doubles per thread-quad pass */
	const int lrow = tid >> 2;            /* 0..63 */
	const int lk = (tid & 3) * 4;         /* 0,4,8,12 */

	/* The panel after the one being multiplied is already on its way from L2 into registers
	 * (lv, rv): a panel is 16 steps of 32 FP64 operations per thread, long enough to cover the
	 * load.  fetch() reads this thread's 4 features of one user row and one item row of panel
	 * (tile tt, features kk0..); rows / features outside the matrices read as 0.0. */
	double lv[4], rv[4];
	auto fetch = [&](int tt, int kk0) {
		const int kc = min(AM_KC, F - kk0);
		const int u = u0 + lrow, j = tt * AM_TJ + lrow;
#pragma unroll
		for (int q = 0; q < 4; q++)
			lv[q] = rv[q] = 0.0;
		if (u < p.users_hi) {
			const double *src = L + (size_t)(u - p.l_row0) * ld + kk0 + lk;
			if (lk + 4 <= kc) {               /* rows are 16-byte aligned, kk0 + lk is a multiple of 4 */
				const double2 x = *reinterpret_cast<const double2 *>(src);
				const double2 y = *reinterpret_cast<const double2 *>(src + 2);
				lv[0] = x.x; lv[1] = x.y; lv[2] = y.x; lv[3] = y.y;
			} else {
#pragma unroll
				for (int q = 0; q < 4; q++) if (lk + q < kc) lv[q] = src[q];
			}
		}
		if (j < p.items) {
			const double *src = R + (size_t)j * ld + kk0 + lk;
			if (lk + 4 <= kc) {
				const double2 x = *reinterpret_cast<const double2 *>(src);
				const double2 y = *reinterpret_cast<const double2 *>(src + 2);
				rv[0] = x.x; rv[1] = x.y; rv[2] = y.x; rv[3] = y.y;
			} else {
#pragma unroll
				for (int q = 0; q < 4; q++) if (lk + q < kc) rv[q] = src[q];
			}
		}
	};
	if (t_begin < t_end && F > 0)
		fetch(t_begin, 0);

	for (int t = t_begin; t < t_end; t++) {
		const int j0 = t * AM_TJ;
		double acc[4][4];
#pragma unroll
		for (int a = 0; a < 4; a++)
#pragma unroll
			for (int b = 0; b < 4; b++) acc[a][b] = 0.0;

		__syncthreads();   /* previous tile's mask and panels are no longer read */
		if (tid < AM_TI) {
			const int u = u0 + tid;
			unsigned long long bits = 0ull;
			if (u < p.users_hi) {
				int c = mcur[tid];
				const int e = p.mask_ptr[u + 1];
				while (c < e) {
					const int col = p.mask_col[c];
					if (col >= j0 + AM_TJ) break;
					bits |= 1ull << (col - j0);
					c++;
				}
				mcur[tid] = c;
			}
			mbits[tid] = bits;
		}

		for (int k0 = 0; k0 < F; k0 += AM_KC) {
			const int kc = min(AM_KC, F - k0);
			__syncthreads();   /* previous panel fully consumed */
#pragma unroll
			for (int q = 0; q < 4; q++) {
				Ls[lk + q][lrow] = lv[q];
				Rs[lk + q][lrow + 2 * (lrow >> 4)] = rv[q];
			}
			__syncthreads();
			if (k0 + AM_KC < F)
				fetch(t, k0 + AM_KC);
			else if (t + 1 < t_end)
				fetch(t + 1, 0);
			if (kc == AM_KC) {
#pragma unroll
				for (int kk = 0; kk < AM_KC; kk++) {
					const double2 a01 = *reinterpret_cast<const double2 *>(&Ls[kk][ty * 4]);
					const double2 a23 = *reinterpret_cast<const double2 *>(&Ls[kk][ty * 4 + 2]);
					const double2 b01 = *reinterpret_cast<const double2 *>(&Rs[kk][tx * 4 + 2 * (tx >> 2)]);
					const double2 b23 = *reinterpret_cast<const double2 *>(&Rs[kk][tx * 4 + 2 * (tx >> 2) + 2]);
					const double av[4] = {a01.x, a01.y, a23.x, a23.y};
					const double bw[4] = {b01.x, b01.y, b23.x, b23.y};
#pragma unroll
					for (int a = 0; a < 4; a++)
#pragma unroll
						for (int b = 0; b < 4; b++)
							acc[a][b] = __dadd_rn(acc[a][b], __dmul_rn(av[a], bw[b]));
				}
			} else {
				for (int kk = 0; kk < kc; kk++) {
#pragma unroll
					for (int a = 0; a < 4; a++)
#pragma unroll
						for (int b = 0; b < 4; b++)
							acc[a][b] = __dadd_rn(acc[a][b],
							                      __dmul_rn(Ls[kk][ty * 4 + a], Rs[kk][tx * 4 + 2 * (tx >> 2) + b]));
				}
			}
		}

		/* epilogue of the tile: masked running best, ascending j inside the thread */
#pragma unroll
		for (int a = 0; a < 4; a++) {
			const int ul = ty * 4 + a, u = u0 + ul;
			if (u >= p.users_hi) continue;
			const unsigned long long bits = mbits[ul];
			const int ff = p.first_free[u];
#pragma unroll
			for (int b = 0; b < 4; b++) {
				const int jl = tx * 4 + b, j = j0 + jl;
				if (j >= p.items || ((bits >> jl) & 1ull)) continue;
				const double v = acc[a][b];
				if (j == ff && v != v) p.nan_first[u - p.users_lo] = 1;
				if (better(v, j, bv[a], bj[a])) { bv[a] = v; bj[a] = j; }
			}
		}
	}

	/* fold the 16 threads that share a user */
#pragma unroll
	for (int a = 0; a < 4; a++) {
		red_v[ty * 4 + a][tx] = bv[a];
		red_j[ty * 4 + a][tx] = bj[a];
	}
	__syncthreads();
	if (tid < AM_TI) {
		const int u = u0 + tid;
		if (u < p.users_hi) {
			double v = 0.0;
			int j = -1;
			for (int x = 0; x < 16; x++)
				if (red_j[tid][x] >= 0 && better(red_v[tid][x], red_j[tid][x], v, j)) {
					v = red_v[tid][x];
					j = red_j[tid][x];
				}
			const size_t o = (size_t)(u - p.users_lo) * p.n_splits + split;
			p.part_v[o] = v;
			p.part_j[o] = j;
		}
	}
}

/* best[u] for u in [users_lo, users_hi): fold the splits in ascending item order. */
__global__ void __launch_bounds__(256)
k_argmax_final(const ArgmaxArgs p, int *__restrict__ best)
{
	const int ul = blockIdx.x * blockDim.x + threadIdx.x;
	if (ul >= p.users_hi - p.users_lo)
		return;
	double v = 0.0;
	int j = -1;
	for (int s = 0; s < p.n_splits; s++) {
		const size_t o = (size_t)ul * p.n_splits + s;
		const int cj = p.part_j[o];
		if (cj >= 0 && better(p.part_v[o], cj, v, j)) {
			v = p.part_v[o];
			j = cj;
		}
	}
	const int ff = p.first_free[p.users_lo + ul];
	if (ff >= 0 && p.nan_first[ul])
		j = ff;
	best[ul] = j;
}

/* The same fold without the last word: (value, item) of the launch's item block per user, for
 * the merge over the GPUs of an item-sharded arg-max (mfgpu_plan_recommend): the blocks are
 * folded in ascending item order with better(), and the NaN-at-the-first-unrated-item rule is
 * applied once, after the last block. */
__global__ void __launch_bounds__(256)
k_argmax_block(const ArgmaxArgs p, double *__restrict__ out_v, int *__restrict__ out_j)
{
	const int ul = blockIdx.x * blockDim.x + threadIdx.x;
	if (ul >= p.users_hi - p.users_lo)
		return;
	double v = 0.0;
	int j = -1;
	for (int s = 0; s < p.n_splits; s++) {
		const size_t o = (size_t)ul * p.n_splits + s;
		const int cj = p.part_j[o];
		if (cj >= 0 && better(p.part_v[o], cj, v, j)) {
			v = p.part_v[o];
			j = cj;
		}
	}
	out_v[ul] = v;
	out_j[ul] = j;
}

} // namespace mf
