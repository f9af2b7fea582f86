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
 *   k_update_L  by-user pass.  A warp owns a user row: the R rows of up to CH of
 *               its nonzeros are staged in shared memory with 16-byte cp.async,
 *               e_n is computed with one lane per nonzero walking k
 *               sequentially (the dot of mat2d.c:133-137), then L_new[i][k] is
 *               accumulated with lanes over features walking the nonzeros
 *               sequentially.  Rows longer than a threshold get a whole CTA:
 *               all warps compute dots, then every warp runs the sequential
 *               accumulation of 32 features with the register pipeline below.
 *               Writes e_n at its by-item position.
 *   k_update_R  by-item pass.  A warp owns (item, 32 features) and walks the
 *               item's nonzeros in file order; the gathered L values run two
 *               blocks of 32 rows ahead of the accumulation in registers.
 *   k_apply_R   N shards only: R_new = R_old + G (G = sum of the partial
 *               gradients, the MPI_Iallreduce of matFact-mpi.c:208).
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
#include "mf_async.cuh"

namespace mf {

constexpr unsigned FULL = 0xffffffffu;

__device__ __forceinline__ double2 ld2(const double *p)
{
	return *reinterpret_cast<const double2 *>(p);
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

constexpr int KT = 16;        /* features per staged tile of the dot phase          */
constexpr int TS = KT + 2;    /* tile row stride in doubles: TS/2 odd, so the 16-byte reads
                                 of 8 lanes (8 rows) fall in 8 distinct bank groups   */

/* Staging of the dot phase.  A chunk has cnt <= 32 nonzeros; a tile holds features
 * [k0, k0+KT) of their R rows, rows TS doubles apart, moved as 16-byte pieces (8 per row).
 * Lane l owns pieces l, l+32, ... (<= 8 of them): their source row and tile offset are fixed
 * for the chunk, so they are computed once (StagePlan) and a tile costs one cp.async each. */
struct StagePlan {
	const double *src[8];     /* row start + 2*piece-in-row, k0 not yet added */
	int dst[8];               /* offset in the tile, doubles                  */
	int kofs[8];              /* 2*piece-in-row, to test against ld - k0      */
	int n;                    /* pieces of this lane                          */
};

__device__ __forceinline__ void stage_plan(StagePlan &sp, const double *__restrict__ src, int j, int cnt,
                                           int ld, int lane)
{
	const int pieces = cnt * (KT / 2);
	sp.n = 0;
#pragma unroll
	for (int u = 0; u < 8; u++) {
		const int p = u * 32 + lane;
		const int row = min(p >> 3, cnt - 1), pc = p & 7;
		const int jn = __shfl_sync(FULL, j, row);
		sp.src[u] = src + (size_t)jn * ld + 2 * pc;
		sp.dst[u] = row * TS + 2 * pc;
		sp.kofs[u] = 2 * pc;
		if (p < pieces)
			sp.n = u + 1;
	}
}

__device__ __forceinline__ void stage_tile(double *dst, const StagePlan &sp, int k0, int ld)
{
#pragma unroll
	for (int u = 0; u < 8; u++)
		if (u < sp.n && k0 + sp.kofs[u] < ld)
			cp_async16(dst + sp.dst[u], sp.src[u] + k0);
}

/* s = (s + l[0]*r[0]) + l[1]*r[1] + ... over one tile (mat2d.c:133-137 continued across tiles) */
__device__ __forceinline__ double dot_tile(double s, const double *l, const double *r, int kc)
{
	if (kc == KT) {
#pragma unroll
		for (int k = 0; k < KT; k += 2) {
			const double2 a = ld2(l + k), b = ld2(r + k);
			s = __dadd_rn(s, __dmul_rn(a.x, b.x));
			s = __dadd_rn(s, __dmul_rn(a.y, b.y));
		}
	} else {
		for (int k = 0; k < kc; k++)
			s = __dadd_rn(s, __dmul_rn(l[k], r[k]));
	}
	return s;
}

/* Lout[k] = ((init[k] + es[0]*R[j_0][k]) + es[1]*R[j_1][k]) + ...  for the NG feature groups
 * k = kb + g*32 + lane of this lane (matFact.c:48-49 order within the row).  The R values of
 * nonzero n+1 are loaded (coalesced, 256 B per group) while those of n are being added. */
template <int NG>
__device__ __forceinline__ void row_update(const double *init, double *Lout,
                                           const double *__restrict__ Rold, int j, const double *es,
                                           int cnt, int ld, int kb, int F, int lane)
{
	double acc[NG], x[NG], xn[NG];
	const bool last_ok = kb + (NG - 1) * 32 + lane < F;   /* only the last group can be partial */
	size_t off[NG];
#pragma unroll
	for (int g = 0; g < NG; g++) {
		const bool ok = g < NG - 1 || last_ok;
		off[g] = ok ? (size_t)(kb + g * 32 + lane) : 0;    /* idle lanes read a valid address */
		acc[g] = ok ? init[off[g]] : 0.0;
	}
	{
		const double *row = Rold + (size_t)__shfl_sync(FULL, j, 0) * ld;
#pragma unroll
		for (int g = 0; g < NG; g++)
			x[g] = row[off[g]];
	}
	for (int n = 0; n < cnt; n++) {
		const double e = es[n];
		const double *row = Rold + (size_t)__shfl_sync(FULL, j, min(n + 1, cnt - 1)) * ld;
#pragma unroll
		for (int g = 0; g < NG; g++)
			xn[g] = row[off[g]];
#pragma unroll
		for (int g = 0; g < NG; g++) {
			acc[g] = __dadd_rn(acc[g], __dmul_rn(e, x[g]));
			x[g] = xn[g];
		}
	}
#pragma unroll
	for (int g = 0; g < NG; g++)
		if (g < NG - 1 || last_ok)
			Lout[off[g]] = acc[g];
}

/* ---- register pipeline of the sequential accumulations ----------------------------------
 *
 * acc = ((acc + ev[0]*x[0]) + ev[1]*x[1]) + ...  with x[n] = srck[idx[n]*ld], n < m, in order.
 * One warp, lane = feature.  The additions form one dependent chain per lane, so the
 * only way to go fast is to keep the operands ready: lane l holds (idx, ev) of nonzero
 * 32q + l of block q and broadcasts them by shuffle, and the gathered x of the next block
 * of 32 nonzeros are loaded into registers while the current block is being added.
 */
__device__ __forceinline__ void gather_block(double (&X)[32], int bi, const double *__restrict__ srck,
                                             int ld)
{
	/* unconditional loads (every lane's bi is a valid row, see chain_regs): a predicated
	 * load would be followed by a select that waits for it and serialise the gather */
#pragma unroll
	for (int r = 0; r < 32; r++)
		X[r] = srck[(size_t)__shfl_sync(FULL, bi, r) * ld];
}

__device__ __forceinline__ double chain_block(double acc, const double (&X)[32], double be, int rows)
{
	if (rows == 32) {
#pragma unroll
		for (int r = 0; r < 32; r++)
			acc = __dadd_rn(acc, __dmul_rn(__shfl_sync(FULL, be, r), X[r]));
	} else {
#pragma unroll
		for (int r = 0; r < 32; r++) {
			const double e = __shfl_sync(FULL, be, r);
			if (r < rows)
				acc = __dadd_rn(acc, __dmul_rn(e, X[r]));
		}
	}
	return acc;
}

__device__ __forceinline__ double chain_regs(double acc, int lane,
                                             const int *__restrict__ idx, const double *__restrict__ ev,
                                             int m, const double *__restrict__ srck, int ld)
{
	const int nblk = (m + 31) >> 5;
	int iA = 0, iB = 0, iC = 0;
	double eA = 0.0, eB = 0.0, eC = 0.0;
	auto load_blk = [&](int q, int &bi, double &be) {
		const int n = min(q * 32 + lane, m - 1);   /* past the end: repeat the last row (never added) */
		bi = idx[n];
		be = ev[n];
	};
	auto rows_of = [&](int q) { return min(32, m - q * 32); };
	double XA[32], XB[32];
	load_blk(0, iA, eA);
	load_blk(1, iB, eB);
	gather_block(XA, iA, srck, ld);
	for (int q = 0; q < nblk; q += 2) {
		if (q + 1 < nblk)
			gather_block(XB, iB, srck, ld);
		load_blk(q + 2, iC, eC);
		acc = chain_block(acc, XA, eA, rows_of(q));
		if (q + 2 < nblk)
			gather_block(XA, iC, srck, ld);
		iA = iC;
		eA = eC;
		load_blk(q + 3, iC, eC);
		if (q + 1 < nblk)
			acc = chain_block(acc, XB, eB, rows_of(q + 1));
		iB = iC;
		eB = eC;
	}
	return acc;
}

/* Same accumulation with blocks of B = 8 or 16 nonzeros and two small register buffers:
 * ~64 registers instead of ~170, so four times as many warps are resident.  This is the
 * variant for launches with many more (item, feature group) chains than warp slots, where
 * the other warps hide the gather latency; the 32-row variant above is for the few chains
 * that are so long that they alone decide how long the launch lasts. */
template <int B>
__device__ __forceinline__ void gather_small(double (&X)[B], int bi, int lane0,
                                             const double *__restrict__ srck, int ld)
{
#pragma unroll
	for (int r = 0; r < B; r++)
		X[r] = srck[(size_t)__shfl_sync(FULL, bi, lane0 + r) * ld];
}

template <int B>
__device__ __forceinline__ double chain_small(double acc, const double (&X)[B], double be, int lane0, int rows)
{
	if (rows >= B) {
#pragma unroll
		for (int r = 0; r < B; r++)
			acc = __dadd_rn(acc, __dmul_rn(__shfl_sync(FULL, be, lane0 + r), X[r]));
	} else {
#pragma unroll
		for (int r = 0; r < B; r++) {
			const double e = __shfl_sync(FULL, be, lane0 + r);
			if (r < rows)
				acc = __dadd_rn(acc, __dmul_rn(e, X[r]));
		}
	}
	return acc;
}

template <int B>
__device__ __forceinline__ double chain_regs_small(double acc, int lane, const int *__restrict__ idx,
                                                   const double *__restrict__ ev, int m,
                                                   const double *__restrict__ srck, int ld)
{
	constexpr int NSB = 32 / B;          /* sub-blocks per block of 32 (idx, e) pairs; even */
	const int nib = (m + 31) >> 5;
	int iCur, iNxt;
	double eCur, eNxt;
	auto load_blk = [&](int q, int &bi, double &be) {
		const int n = min(q * 32 + lane, m - 1);
		bi = idx[n];
		be = ev[n];
	};
	load_blk(0, iCur, eCur);
	load_blk(1, iNxt, eNxt);
	double X0[B], X1[B];
	gather_small<B>(X0, iCur, 0, srck, ld);
	for (int ib = 0; ib < nib; ib++) {
		const int rows_ib = min(32, m - ib * 32);
#pragma unroll
		for (int sb = 0; sb < NSB; sb++) {
			/* next sub-block in flight while this one is added */
			if (sb + 1 < NSB) {
				if (sb & 1) gather_small<B>(X0, iCur, (sb + 1) * B, srck, ld);
				else        gather_small<B>(X1, iCur, (sb + 1) * B, srck, ld);
			} else {
				gather_small<B>(X0, iNxt, 0, srck, ld);     /* NSB even: sub-block 0 lives in X0 */
			}
			const int rows = rows_ib - sb * B;
			if (rows > 0)
				acc = (sb & 1) ? chain_small<B>(acc, X1, eCur, sb * B, rows)
				               : chain_small<B>(acc, X0, eCur, sb * B, rows);
		}
		iCur = iNxt;
		eCur = eNxt;
		load_blk(ib + 2, iNxt, eNxt);
	}
	return acc;
}

/*
 * By-user pass.  Dynamic shared memory per warp: two dot-phase tiles [CH][TS], the user's
 * L row [ld], es[32].  LONG = true: block b takes stored row b (< n_long) with the whole
 * CTA; LONG = false: every warp takes one of the remaining rows.  The dot of a nonzero streams its R row through the two tiles in
 * pieces of KT features (next tile in flight while the current one is added); the
 * accumulation of L_new then re-reads the R rows coalesced from L2.
 */
template <int WARPS, bool LONG>
__global__ void __launch_bounds__(WARPS * 32)
k_update_L(const LoopArgs p, const double *__restrict__ Lold, double *__restrict__ Lnew,
           const double *__restrict__ Rold)
{
	extern __shared__ __align__(16) double smem[];
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const int F = p.F, ld = p.ld, CH = p.CH;
	const size_t per_warp = (size_t)2 * CH * TS + ld + 32;
	double *T0 = smem + per_warp * warp;
	double *T1 = T0 + (size_t)CH * TS;
	double *Ls = T1 + (size_t)CH * TS;
	double *es = Ls + ld;

	constexpr bool is_long = LONG;    /* two kernels: the register-hungry pipeline of the long rows
	                                     must not cap the occupancy of the short-row launch */
	const int r = is_long ? (int)blockIdx.x : p.n_long + (int)blockIdx.x * WARPS + warp;
	if (r >= p.n_rows)
		return;                       /* whole warp leaves; short-path blocks never use a block barrier */
	const int i = p.row_id[r] - p.user_lo;
	const int base = p.row_ptr[r];
	const int m = p.row_ptr[r + 1] - base;
	const double *Lrow = Lold + (size_t)i * ld;
	double *Lout = Lnew + (size_t)i * ld;
	const int ntiles = (F + KT - 1) / KT;

	const int c_first = is_long ? warp * CH : 0;
	const int c_step = is_long ? WARPS * CH : CH;
	if (c_first < m)                  /* joins the first tile's copy group */
		for (int q = lane; q < (ld >> 1); q += 32)
			cp_async16(Ls + 2 * q, Lrow + 2 * q);
	for (int c0 = c_first; c0 < m; c0 += c_step) {
		const int cnt = min(CH, m - c0);
		const int pos = base + c0 + lane;
		int j = 0;
		double a = 0.0;
		if (lane < cnt) {
			j = p.col_idx[pos];
			a = p.val[pos];
		}
		j = __shfl_sync(FULL, j, min(lane, cnt - 1));   /* idle lanes mirror a valid item */

		/* e_n = alpha2 * (a - dot): one lane per nonzero, k sequential, tiles double buffered */
		StagePlan sp;
		stage_plan(sp, Rold, j, cnt, ld, lane);
		stage_tile(T0, sp, 0, ld);
		cp_async_commit();
		double s = 0.0;
		for (int t = 0; t < ntiles; t++) {
			double *cur = (t & 1) ? T1 : T0;
			if (t + 1 < ntiles)
				stage_tile((t & 1) ? T0 : T1, sp, (t + 1) * KT, ld);
			cp_async_commit();
			cp_async_wait<1>();       /* everything but the tile just issued has landed */
			__syncwarp();
			if (lane < cnt)
				s = dot_tile(s, Ls + t * KT, cur + lane * TS, min(KT, F - t * KT));
			__syncwarp();             /* tile `cur` may be overwritten by the next issue */
		}
		if (lane < cnt) {
			const double e = __dmul_rn(p.alpha2, __dsub_rn(a, s));
			p.e_csc[p.epos[pos]] = e;
			if (is_long)
				p.e_csr[pos] = e;
			else
				es[lane] = e;
		}
		__syncwarp();
		if (!is_long) {
			/* L_new[i][k] chain over this chunk's nonzeros; the feature groups of a lane
			 * are independent chains and run interleaved */
			for (int kb = 0; kb < F; kb += 256) {
				const double *init = (c0 == 0) ? Ls : Lout;
				switch (min(8, (F - kb + 31) >> 5)) {
				case 1: row_update<1>(init, Lout, Rold, j, es, cnt, ld, kb, F, lane); break;
				case 2: row_update<2>(init, Lout, Rold, j, es, cnt, ld, kb, F, lane); break;
				case 3: row_update<3>(init, Lout, Rold, j, es, cnt, ld, kb, F, lane); break;
				case 4: row_update<4>(init, Lout, Rold, j, es, cnt, ld, kb, F, lane); break;
				case 5: row_update<5>(init, Lout, Rold, j, es, cnt, ld, kb, F, lane); break;
				case 6: row_update<6>(init, Lout, Rold, j, es, cnt, ld, kb, F, lane); break;
				case 7: row_update<7>(init, Lout, Rold, j, es, cnt, ld, kb, F, lane); break;
				default: row_update<8>(init, Lout, Rold, j, es, cnt, ld, kb, F, lane); break;
				}
			}
			__syncwarp();             /* es is rewritten by the next chunk */
		}
	}
	if (!is_long)
		return;

	/* whole-CTA row: every e_n is in e_csr (written above by this CTA); each warp now owns
	 * 32 features and walks the row's nonzeros in order with the register pipeline */
	__syncthreads();
	for (int kb = warp * 32; kb < F; kb += WARPS * 32) {
		const int k = kb + lane;
		const bool active = k < F;
		double acc = active ? Lrow[k] : 0.0;
		acc = chain_regs(acc, lane, p.col_idx + base, p.e_csr + base, m, Rold + (active ? k : 0), ld);
		if (active)
			Lout[k] = acc;
	}
}

/*
 * By-item pass over the stored columns [c_begin, c_end).  One warp per (column, group of
 * 32 features), lane = feature.  B = rows per register block of the pipeline (32 for the
 * launch of the few longest columns, 8 for the bulk).
 * single-GPU: Rout = R_new, starts from R_old[j][k]        (matFact.c:50-51)
 * N shards:   Rout = G (rows by item_slot), starts from 0  (matFact-mpi.c:187-205)
 */
template <int WARPS, int B>
__global__ void __launch_bounds__(WARPS * 32)
k_update_R(const LoopArgs p, int c_begin, int c_end, const double *__restrict__ Lold,
           const double *__restrict__ Rold, double *__restrict__ Rout)
{
	const int lane = threadIdx.x & 31;
	const int NG = (p.F + 31) >> 5;
	const long gw = (long)blockIdx.x * WARPS + (threadIdx.x >> 5);
	const int c = c_begin + (int)(gw / NG), g = (int)(gw % NG);
	if (c >= c_end)
		return;
	const int k = g * 32 + lane;
	const bool active = k < p.F;
	const int ld = p.ld;
	const int j = p.col_id[c];
	const int b = p.col_ptr[c], m = p.col_ptr[c + 1] - b;
	const double *Lk = Lold + (active ? k : 0) - (size_t)p.user_lo * ld;
	double acc = (p.item_slot || !active) ? 0.0 : Rold[(size_t)j * ld + k];
	if (B == 32)
		acc = chain_regs(acc, lane, p.row_idx + b, p.e_csc + b, m, Lk, ld);
	else
		acc = chain_regs_small<(B == 32 ? 8 : B)>(acc, lane, p.row_idx + b, p.e_csc + b, m, Lk, ld);
	if (active) {
		const size_t orow = p.item_slot ? (size_t)p.item_slot[j] : (size_t)j;
		Rout[orow * ld + k] = acc;
	}
}

/* N shards: R_new[item] = R_old[item] + G[slot] for the touched items. */
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
