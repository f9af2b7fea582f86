/*
 * csrc/mf_loop.cuh -- the per-iteration kernels of the factorization loop (sm_100a).
 *
 * One iteration of the reference (matFact.c:36-53) is, for every nonzero
 * n = (i, j, a) in file order, on the snapshot (L_old, R_old):
 *     e_n      = (alpha*2) * (a - dot(L_old[i], R_old[j]))         matFact.c:45, mat2d.c:126-139
 *     L[i][k] += e_n * R_old[j][k]                                 matFact.c:48-49
 *     R[j][k] += e_n * L_old[i][k]                                 matFact.c:50-51
 * with every multiply and add rounded separately (no FMA) and the additions
 * of one L row taken in file order of that row's nonzeros, those of one R row
 * in file order of that column's nonzeros.  The kernels keep exactly that
 * order, so one GPU reproduces the reference's doubles bit for bit:
 *
 *   k_update_L  by-user pass: stages the R rows of a chunk of one user's
 *               nonzeros in shared memory, computes e_n with one lane per
 *               nonzero walking k sequentially (the dot of mat2d.c:133-137),
 *               then accumulates L_new[i][k] with one lane per feature walking
 *               the nonzeros sequentially.  Writes e_n at its by-item position.
 *   k_update_R  by-item pass: one warp per (item, 32 features), walks the
 *               item's nonzeros in file order, gathers L_old rows coalesced.
 *   k_apply_R   N-GPU only: R_new = R_old + G  (G = rank-ordered sum of the
 *               partial gradients, the MPI_Iallreduce of matFact-mpi.c:208).
 *
 * The snapshot copies (mat2d_copy, matFact.c:38-39) do not exist: L and R are
 * double buffered, every touched row is rewritten each iteration and untouched
 * rows are equal in both buffers from the start.
 *
 * All arithmetic uses __dmul_rn/__dadd_rn/__dsub_rn, which nvcc never fuses.
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "mf_types.h"

namespace mf {


__device__ __forceinline__ double2 ld2(const double *p)
{
	return *reinterpret_cast<const double2 *>(p);
}

/* Stage `cnt` R rows (items given per lane in `j`) into Rs, rows S apart.
 * G lanes cooperate on one row with 16-byte loads; 32/G rows per pass. */
__device__ __forceinline__ void stage_rows(double *Rs, const double *__restrict__ Rold,
                                           int j, int cnt, int F2, int G, int ld, int S, int lane)
{
	const int sub = lane / G, l = lane % G, per = 32 / G;
	for (int n0 = 0; n0 < cnt; n0 += per) {
		const int n = n0 + sub;
		const int jn = __shfl_sync(0xffffffffu, j, n < cnt ? n : 0);
		if (n < cnt) {
			const double *src = Rold + (size_t)jn * ld;
			double *dst = Rs + (size_t)n * S;
			for (int q = l; q < F2; q += G)
				*reinterpret_cast<double2 *>(dst + 2 * q) = ld2(src + 2 * q);
		}
	}
}

/* e for the nonzero of this lane: sequential-k dot, mat2d.c:133-137. */
__device__ __forceinline__ double seq_dot_smem(const double *Ls, const double *Rr, int F)
{
	double s = 0.0;
	int k = 0;
	for (; k + 1 < F; k += 2) {
		const double2 l = ld2(Ls + k), r = ld2(Rr + k);
		s = __dadd_rn(s, __dmul_rn(l.x, r.x));
		s = __dadd_rn(s, __dmul_rn(l.y, r.y));
	}
	if (k < F)
		s = __dadd_rn(s, __dmul_rn(Ls[k], Rr[k]));
	return s;
}

template <int WARPS>
__global__ void __launch_bounds__(WARPS * 32)
k_update_L(const LoopArgs p, const double *__restrict__ Lold, double *__restrict__ Lnew,
           const double *__restrict__ Rold)
{
	extern __shared__ __align__(16) double smem[];
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const int F = p.F, ld = p.ld, S = p.S, CH = p.CH;
	const int F2 = (F + 1) >> 1;
	int G = 1;
	while (G < F2 && G < 32) G <<= 1;
	/* per-warp areas: Rs[CH*S], Ls[S], es[32] */
	const size_t per_warp = (size_t)CH * S + S + 32;
	double *Rs = smem + per_warp * warp;
	double *Ls = Rs + (size_t)CH * S;
	double *es = Ls + S;

	const bool is_long = (int)blockIdx.x < p.n_long;
	const int r = is_long ? (int)blockIdx.x : p.n_long + ((int)blockIdx.x - p.n_long) * WARPS + warp;
	if (r >= p.n_rows)
		return;                       /* whole warp leaves; no block barrier on the short path */
	const int i = p.row_id[r] - p.user_lo;
	const int base = p.row_ptr[r];
	const int m = p.row_ptr[r + 1] - base;
	const double *Lrow = Lold + (size_t)i * ld;
	double *Lout = Lnew + (size_t)i * ld;

	for (int q = lane; q < F2; q += 32)
		*reinterpret_cast<double2 *>(Ls + 2 * q) = ld2(Lrow + 2 * q);
	__syncwarp();

	const int c_first = is_long ? warp * CH : 0;
	const int c_step = is_long ? WARPS * CH : CH;
	for (int c0 = c_first; c0 < m; c0 += c_step) {
		const int cnt = min(CH, m - c0);
		const int pos = base + c0 + lane;
		int j = 0;
		double a = 0.0;
		if (lane < cnt) {
			j = p.col_idx[pos];
			a = p.val[pos];
		}
		stage_rows(Rs, Rold, j, cnt, F2, G, ld, S, lane);
		__syncwarp();
		if (lane < cnt) {
			const double s = seq_dot_smem(Ls, Rs + (size_t)lane * S, F);
			const double e = __dmul_rn(p.alpha2, __dsub_rn(a, s));
			p.e_csc[p.epos[pos]] = e;
			if (is_long)
				p.e_csr[pos] = e;
			else
				es[lane] = e;
		}
		__syncwarp();
		if (!is_long) {
			for (int k = lane; k < F; k += 32) {
				double acc = (c0 == 0) ? Ls[k] : Lout[k];
				const double *col = Rs + k;
#pragma unroll 4
				for (int n = 0; n < cnt; n++)
					acc = __dadd_rn(acc, __dmul_rn(es[n], col[(size_t)n * S]));
				Lout[k] = acc;
			}
			__syncwarp();
		}
	}
	if (!is_long)
		return;

	/* long row: all e_n are in e_csr; one thread per feature walks the row */
	__syncthreads();
	const int *cj = p.col_idx + base;
	const double *ce = p.e_csr + base;
	for (int k = threadIdx.x; k < F; k += WARPS * 32) {
		double acc = Lrow[k];
		const double *Rk = Rold + k;
		int n = 0;
		for (; n + 8 <= m; n += 8) {
			double x[8], e[8];
#pragma unroll
			for (int u = 0; u < 8; u++) {
				e[u] = ce[n + u];
				x[u] = Rk[(size_t)cj[n + u] * ld];
			}
#pragma unroll
			for (int u = 0; u < 8; u++)
				acc = __dadd_rn(acc, __dmul_rn(e[u], x[u]));
		}
		for (; n < m; n++)
			acc = __dadd_rn(acc, __dmul_rn(ce[n], Rk[(size_t)cj[n] * ld]));
		Lout[k] = acc;
	}
}

/* By-item pass.  One warp per (stored column c, group of 32 features).
 * single-GPU: Rout = R_new, starts from R_old[j][k]        (matFact.c:50-51)
 * N-GPU:      Rout = G (rows by item_slot), starts from 0  (matFact-mpi.c:187-205) */
template <int WARPS>
__global__ void __launch_bounds__(WARPS * 32)
k_update_R(const LoopArgs p, const double *__restrict__ Lold, const double *__restrict__ Rold,
           double *__restrict__ Rout)
{
	const int lane = threadIdx.x & 31;
	const int NG = (p.F + 31) >> 5;
	const long gw = (long)blockIdx.x * WARPS + (threadIdx.x >> 5);
	const int c = (int)(gw / NG), g = (int)(gw % NG);
	if (c >= p.n_cols)
		return;
	const int k = g * 32 + lane;
	if (k >= p.F)
		return;
	const int ld = p.ld;
	const int j = p.col_id[c];
	const int b = p.col_ptr[c], m = p.col_ptr[c + 1] - b;
	const int *ri = p.row_idx + b;
	const double *ce = p.e_csc + b;
	const double *Lk = Lold + k - (size_t)p.user_lo * ld;
	double acc = p.item_slot ? 0.0 : Rold[(size_t)j * ld + k];
	int n = 0;
	for (; n + 8 <= m; n += 8) {
		double x[8], e[8];
#pragma unroll
		for (int u = 0; u < 8; u++) {
			e[u] = ce[n + u];
			x[u] = Lk[(size_t)ri[n + u] * ld];
		}
#pragma unroll
		for (int u = 0; u < 8; u++)
			acc = __dadd_rn(acc, __dmul_rn(e[u], x[u]));
	}
	for (; n < m; n++)
		acc = __dadd_rn(acc, __dmul_rn(ce[n], Lk[(size_t)ri[n] * ld]));
	const size_t orow = p.item_slot ? (size_t)p.item_slot[j] : (size_t)j;
	Rout[orow * ld + k] = acc;
}

/* N-GPU: R_new[item] = R_old[item] + G[slot] for the touched items. */
__global__ void __launch_bounds__(256)
k_apply_R(int n_touched, int F, int ld, const int *__restrict__ touched_item,
          const double *__restrict__ Rold, const double *__restrict__ G, double *__restrict__ Rnew)
{
	const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
	const long total = (long)n_touched * F;
	if (t >= total)
		return;
	const int s = (int)(t / F), k = (int)(t % F);
	const size_t o = (size_t)touched_item[s] * ld + k;
	Rnew[o] = __dadd_rn(Rold[o], G[(size_t)s * ld + k]);
}

} // namespace mf
