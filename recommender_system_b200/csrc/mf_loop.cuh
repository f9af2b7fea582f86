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
 *   k_update_L      short user rows, one warp per row: the R rows of up to CH of its nonzeros
 *                   stream through shared-memory tiles (16-byte cp.async) while e_n is computed
 *                   with one lane per nonzero walking k sequentially (the dot of
 *                   mat2d.c:133-137); then L_new[i][k] is accumulated with lanes over features
 *                   walking the nonzeros in order.  Writes e_n at its by-item position.
 *   k_update_L_res  the same pass with the gathered rows resident in shared memory for both
 *                   phases (one L2 gather per nonzero instead of two); chosen when eight
 *                   warps' rows fit an SM.
 *   k_dot           the dot phase alone for the long rows, one warp per chunk of <= 32
 *                   nonzeros, so a row of thousands of nonzeros is spread over many warps.
 *   k_chain         the sequential accumulations that remain: L rows of the long users and all
 *                   R rows.  One warp per (row or column, 32 features) walks the nonzeros in
 *                   file order with the gathered operands running ahead in a shared-memory
 *                   ring (or in registers).
 *   k_chain_ws      the same for the few longest chains, one CTA of specialised warps per
 *                   chain: producers gather into a ring with cp.async + mbarriers, one
 *                   consumer warp does nothing but the ordered additions.
 *   k_apply_R       N shards only: R_new = R_old + G (G = sum of the partial
 *                   gradients, the MPI_Iallreduce of matFact-mpi.c:208).
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

/* first 128-byte boundary of the dynamic shared memory (the launches add 128 bytes of slack) */
__device__ __forceinline__ double *smem_lines(double *smem)
{
	return smem + ((128u - (smem_u32(smem) & 127u)) & 127u) / 8;
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
		sp.dst[u] = row * KT + 2 * (pc ^ (row & 7));
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

/* Tile rows of the dot phases are 128 bytes (16 features) on 128-byte boundaries of shared memory,
 * the 16-byte piece p of row r stored at position p ^ (r & 7).  Why not padded rows (stride 18):
 * a cp.async whose 8 lanes' destination straddles two 128-byte lines of shared memory is split in
 * two, and each half fetches its sectors of the SOURCE line on its own -- ncu counted 6 sectors
 * per gathered 128-byte segment instead of 4 (77 M sectors per launch for the 50 M needed, the
 * L2 at 73 % of its sector throughput).  With aligned rows a segment is one request of four
 * sectors, and the swizzle keeps the transposed read (lane = row, all lanes the same p) free of
 * bank conflicts: eight consecutive rows hit eight different 16-byte bank groups. */
__device__ __forceinline__ double dot_tile_swz(double s, const double *l, int lsw, const double *r, int rsw, int kc)
{
	if (kc == KT) {
#pragma unroll
		for (int p = 0; p < KT / 2; p++) {
			const double2 a = ld2(l + 2 * (p ^ lsw)), b = ld2(r + 2 * (p ^ rsw));
			s = __dadd_rn(s, __dmul_rn(a.x, b.x));
			s = __dadd_rn(s, __dmul_rn(a.y, b.y));
		}
	} else {
		for (int k = 0; k < kc; k++) {
			const int p = k >> 1, h = k & 1;
			s = __dadd_rn(s, __dmul_rn(l[2 * (p ^ lsw) + h], r[2 * (p ^ rsw) + h]));
		}
	}
	return s;
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
 * the next P nonzeros are in flight (coalesced loads, 256 B per group) while one is added: P = 1
 * when the launch has warps to spare and only throughput counts, P = 3 when a shard is so
 * small that a warp's own latency decides (rank mode with many ranks).  Loads are never
 * predicated (a predicated load is followed by a select that waits for it): rows shorter than
 * P fetch their last nonzero again instead. */
__device__ __forceinline__ void st_row(double *p, double v, bool cs)
{
	if (cs)
		__stcs(p, v);
	else
		*p = v;
}

template <int NG, int P>
__device__ __forceinline__ void row_update(const double *init, double *Lout,
                                           const double *__restrict__ Rold, int j, const double *es,
                                           int cnt, int ld, int kb, int F, int lane, bool cs = false)
{
	double acc[NG], x[P][NG], xn[NG];
	const bool last_ok = kb + (NG - 1) * 32 + lane < F;   /* only the last group can be partial */
	size_t off[NG];
#pragma unroll
	for (int g = 0; g < NG; g++) {
		const bool ok = g < NG - 1 || last_ok;
		off[g] = ok ? (size_t)(kb + g * 32 + lane) : 0;    /* idle lanes read a valid address */
		acc[g] = ok ? init[off[g]] : 0.0;
	}
#pragma unroll
	for (int u = 0; u < P; u++) {
		const double *row = Rold + (size_t)__shfl_sync(FULL, j, min(u, cnt - 1)) * ld;
#pragma unroll
		for (int g = 0; g < NG; g++)
			x[u][g] = row[off[g]];
	}
	int n = 0;
	for (; n + P < cnt; n++) {
		const double e = es[n];
		const double *row = Rold + (size_t)__shfl_sync(FULL, j, n + P) * ld;
#pragma unroll
		for (int g = 0; g < NG; g++)
			xn[g] = row[off[g]];
#pragma unroll
		for (int g = 0; g < NG; g++) {
			acc[g] = __dadd_rn(acc[g], __dmul_rn(e, x[0][g]));
#pragma unroll
			for (int u = 0; u + 1 < P; u++)
				x[u][g] = x[u + 1][g];
			x[P - 1][g] = xn[g];
		}
	}
	/* the last min(P, cnt) nonzeros: nothing left to fetch */
#pragma unroll
	for (int u = 0; u < P; u++)
		if (n + u < cnt) {
			const double e = es[n + u];
#pragma unroll
			for (int g = 0; g < NG; g++)
				acc[g] = __dadd_rn(acc[g], __dmul_rn(e, x[u][g]));
		}
#pragma unroll
	for (int g = 0; g < NG; g++)
		if (g < NG - 1 || last_ok)
			st_row(Lout + off[g], acc[g], cs);
}

/* ---- register pipelines of the sequential accumulations ------------------------------------
 *
 * acc = ((acc + ev[0]*x[0]) + ev[1]*x[1]) + ...  with x[n] = srck[idx[n]*ld], n < m, in order.
 * One warp, lane = feature.  The additions form one dependent chain per lane, so the only
 * way to go fast is to have the operands ready: lane l holds (idx, ev) of nonzero 32q + l of
 * index block q and broadcasts them by shuffle, and the gathered x run ahead of the chain in
 * registers.  Past the end of the list the last nonzero is loaded again (never added), so
 * no load is predicated: a predicated load is followed by a select that waits for it.
 *
 * chain_deep: sub-blocks of 16 rows, four register buffers -> 48 rows in flight ahead of
 * the chain and the index blocks two blocks (64 rows) ahead; ~170 registers, 8 warps/SM.
 * For the few chains so long that they alone decide how long a launch lasts.
 */
template <int B>
__device__ __forceinline__ void gather_rows(double (&X)[B], int bi, int lane0,
                                            const double *__restrict__ srck, int ld)
{
#pragma unroll
	for (int r = 0; r < B; r++)
		X[r] = srck[(size_t)__shfl_sync(FULL, bi, lane0 + r) * ld];
}

template <int B>
__device__ __forceinline__ double chain_rows(double acc, const double (&X)[B], double be, int lane0, int rows)
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

__device__ __forceinline__ double chain_deep(double acc, int lane, const int *__restrict__ idx,
                                             const double *__restrict__ ev, int m,
                                             const double *__restrict__ srck, int ld)
{
	auto load_blk = [&](int q, int &bi, double &be) {
		const int n = min(q * 32 + lane, m - 1);
		bi = idx[n];
		be = ev[n];
	};
	int i0, i1, i2, i3;
	double e0, e1, e2, e3;
	load_blk(0, i0, e0);
	load_blk(1, i1, e1);
	load_blk(2, i2, e2);
	load_blk(3, i3, e3);
	double X0[16], X1[16], X2[16], X3[16];
	gather_rows<16>(X0, i0, 0, srck, ld);
	gather_rows<16>(X1, i0, 16, srck, ld);
	gather_rows<16>(X2, i1, 0, srck, ld);
	/* per trip: index blocks q, q+1 = four sub-blocks; each sub-block is gathered three
	 * sub-blocks before it is added */
	for (int q = 0; q * 32 < m; q += 2) {
		const int left = m - q * 32;
		gather_rows<16>(X3, i1, 16, srck, ld);
		acc = chain_rows<16>(acc, X0, e0, 0, left);
		gather_rows<16>(X0, i2, 0, srck, ld);
		if (left > 16)
			acc = chain_rows<16>(acc, X1, e0, 16, left - 16);
		gather_rows<16>(X1, i2, 16, srck, ld);
		if (left > 32)
			acc = chain_rows<16>(acc, X2, e1, 0, left - 32);
		gather_rows<16>(X2, i3, 0, srck, ld);
		if (left > 48)
			acc = chain_rows<16>(acc, X3, e1, 16, left - 48);
		i0 = i2; e0 = e2;
		i1 = i3; e1 = e3;
		load_blk(q + 4, i2, e2);
		load_blk(q + 5, i3, e3);
	}
	return acc;
}

/* chain_bulk: sub-blocks of B = 8 rows and two register buffers: ~74 registers, so three
 * times as many warps are resident.  For launches with many more chains than warp slots,
 * where the other warps hide the gather latency. */
template <int B>
__device__ __forceinline__ double chain_bulk(double acc, int lane, const int *__restrict__ idx,
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
	gather_rows<B>(X0, iCur, 0, srck, ld);
	for (int ib = 0; ib < nib; ib++) {
		const int rows_ib = min(32, m - ib * 32);
#pragma unroll
		for (int sb = 0; sb < NSB; sb++) {
			/* next sub-block in flight while this one is added */
			if (sb + 1 < NSB) {
				if (sb & 1) gather_rows<B>(X0, iCur, (sb + 1) * B, srck, ld);
				else        gather_rows<B>(X1, iCur, (sb + 1) * B, srck, ld);
			} else {
				gather_rows<B>(X0, iNxt, 0, srck, ld);     /* NSB even: sub-block 0 lives in X0 */
			}
			const int rows = rows_ib - sb * B;
			if (rows > 0)
				acc = (sb & 1) ? chain_rows<B>(acc, X1, eCur, sb * B, rows)
				               : chain_rows<B>(acc, X0, eCur, sb * B, rows);
		}
		iCur = iNxt;
		eCur = eNxt;
		load_blk(ib + 2, iNxt, eNxt);
	}
	return acc;
}

/* chain_ring: the same accumulation with the gathered rows landing in a per-warp ring in
 * shared memory instead of registers: RS rows per stage, NST stages, stage s+NST-1 issued
 * (16-byte cp.async, a half-warp per row, so one instruction moves two rows) before stage
 * s is added.  cp.async groups complete in order and wait_group<N> leaves the N newest
 * pending, which register loads cannot express: the six scoreboard slots of a warp are
 * shared by all loads in flight, so waiting for an old block also waits for the newest one.
 * src_g = first of the group's 32 features in row 0 of the gathered factor; gw = number of
 * valid doubles of the group in a row (even, <= 32). */
/* doubles of one warp's ring: per stage RS gathered rows of 32 features and their RS e_n */
/* (the RS e_n of a stage take 16 doubles, so that every 256-byte row of the ring starts on a
 * 128-byte line: a cp.async whose destination straddles two lines is split, see dot_tile_swz) */
__host__ __device__ constexpr int ring_doubles(int RS, int NST) { return NST * (RS * 32 + 16); }

template <int RS, int NST>
__device__ __forceinline__ double chain_ring(double acc, int lane, const int *__restrict__ idx,
                                             const double *__restrict__ ev, int m,
                                             const double *__restrict__ src_g, int ld, int gw, double *ring)
{
	constexpr int D = NST - 1;            /* stages in flight ahead of the one being added */
	constexpr int SPB = 32 / RS;          /* stages per index block of 32 nonzeros        */
	constexpr int NB = (SPB - 1 + D) / SPB + 2;   /* index blocks held in registers       */
	constexpr int STAGE = RS * 32 + 16;
	int bi[NB];
	auto load_blk = [&](int q, int &i_) { i_ = idx[min(q * 32 + lane, m - 1)]; };
#pragma unroll
	for (int b = 0; b < NB; b++)
		load_blk(b, bi[b]);
	const int half = lane >> 4, piece = lane & 15;
	const bool piece_ok = 2 * piece < gw;
	auto issue = [&](int slot, int blk_i, int lane0, int s) {
		/* stage s = rows lane0 .. lane0+RS-1 of the index block whose indices are in blk_i.
		 * The e_n travel with the rows (8-byte copies) and are read back as 16-byte
		 * broadcasts: a shuffle per row costs the shared-memory pipe as much as a row. */
		double *st = ring + (size_t)slot * STAGE;
#pragma unroll
		for (int u = 0; u < RS / 2; u++) {
			const int i = __shfl_sync(FULL, blk_i, lane0 + 2 * u + half);
			if (piece_ok)
				cp_async16(st + (2 * u + half) * 32 + 2 * piece, src_g + (size_t)i * ld + 2 * piece);
		}
		if (lane < RS && s * RS + lane < m)
			cp_async8(st + RS * 32 + lane, ev + s * RS + lane);
	};
	/* prologue: stages 0 .. D-1 */
#pragma unroll
	for (int t = 0; t < D; t++) {
		if (t * RS < m)
			issue(t % NST, bi[t / SPB], (t % SPB) * RS, t);
		cp_async_commit();
	}
	int slot = 0;                          /* ring slot of the stage being added */
	for (int q = 0; q * 32 < m; q++) {
#pragma unroll
		for (int j = 0; j < SPB; j++) {
			const int s = q * SPB + j;
			if ((s + D) * RS < m)
				issue(slot == 0 ? NST - 1 : slot - 1, bi[(j + D) / SPB], ((j + D) % SPB) * RS, s + D);
			cp_async_commit();
			cp_async_wait<D>();            /* stage s has landed */
			__syncwarp();
			const int rows = m - s * RS;
			const double *st = ring + (size_t)slot * STAGE;
			const double *x = st + lane;
			if (rows >= RS) {
#pragma unroll
				for (int r = 0; r < RS; r += 2) {
					const double2 e2 = ld2(st + RS * 32 + r);
					acc = __dadd_rn(acc, __dmul_rn(e2.x, x[r * 32]));
					acc = __dadd_rn(acc, __dmul_rn(e2.y, x[(r + 1) * 32]));
				}
			} else if (rows > 0) {
#pragma unroll
				for (int r = 0; r < RS; r++)
					if (r < rows)
						acc = __dadd_rn(acc, __dmul_rn(st[RS * 32 + r], x[r * 32]));
			}
			__syncwarp();                  /* the slot is refilled by the next issue */
			slot = (slot + 1 == NST) ? 0 : slot + 1;
		}
#pragma unroll
		for (int b = 0; b + 1 < NB; b++)
			bi[b] = bi[b + 1];
		load_blk(q + NB, bi[NB - 1]);
	}
	cp_async_wait<0>();
	return acc;
}

/* wait until at most n of this thread's copy groups are pending (n is not a constant here) */
__device__ __forceinline__ void cp_async_wait_dyn(int n)
{
	switch (n) {
	case 0: cp_async_wait<0>(); break;
	case 1: cp_async_wait<1>(); break;
	case 2: cp_async_wait<2>(); break;
	case 3: cp_async_wait<3>(); break;
	case 4: cp_async_wait<4>(); break;
	case 5: cp_async_wait<5>(); break;
	case 6: cp_async_wait<6>(); break;
	default: cp_async_wait<7>(); break;
	}
}

/* e_n = alpha2 * (a_n - L_i . R_j) for the chunk's nonzero of this lane (lane < cnt): one
 * lane per nonzero walks k sequentially (mat2d.c:133-137) while the R rows stream through a
 * ring of NT tiles of `td` doubles each, NT - 1 tiles in flight while one is added (NT = 2 when
 * the other warps of the SM hide the latency, more when the shard is small).  Ls must be the
 * warp's copy of L_i; copies issued by the caller before this call join the first group. */
template <int NT>
__device__ __forceinline__ double chunk_error(double *tiles, int td, const double *Ls,
                                              const double *__restrict__ Rold, int j, double a, int cnt,
                                              int F, int ld, double alpha2, int lane)
{
	const int ntiles = (F + KT - 1) / KT;
	StagePlan sp;
	stage_plan(sp, Rold, j, cnt, ld, lane);
#pragma unroll
	for (int t = 0; t < NT - 1; t++) {
		if (t < ntiles)
			stage_tile(tiles + t * td, sp, t * KT, ld);
		cp_async_commit();
	}
	double s = 0.0;
	int slot = 0, fill = NT - 1;      /* ring slots of the tile being added / being filled */
	for (int t = 0; t < ntiles; t++) {
		if (t + NT - 1 < ntiles)
			stage_tile(tiles + fill * td, sp, (t + NT - 1) * KT, ld);
		cp_async_commit();
		cp_async_wait<NT - 1>();      /* everything but the NT - 1 newest groups has landed */
		__syncwarp();
		if (lane < cnt)
			s = dot_tile_swz(s, Ls + t * KT, 0, tiles + slot * td + lane * KT, lane & 7, min(KT, F - t * KT));
		__syncwarp();                 /* the slot may be overwritten by the next issue */
		slot = slot + 1 == NT ? 0 : slot + 1;
		fill = fill + 1 == NT ? 0 : fill + 1;
	}
	return __dmul_rn(alpha2, __dsub_rn(a, s));
}

/*
 * By-user pass of the short rows [n_long, n_rows): one warp per row does everything.
 * Dynamic shared memory per warp: NT dot-phase tiles [CH][TS], the user's L row [ld], es[32]
 * (NT must be LoopArgs::NT, which sized the launch).
 * P: rows in flight ahead of the one being added in the accumulation phase (row_update).
 */
template <int WARPS, int P, int NT>
__global__ void __launch_bounds__(WARPS * 32, WARPS == 8 ? (P > 1 ? 2 : 3) : 1)
k_update_L(const LoopArgs p, const double *__restrict__ Lold, double *__restrict__ Lnew,
           const double *__restrict__ Rold)
{
	extern __shared__ __align__(16) double smem[];
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const int F = p.F, ld = p.ld, CH = p.CH;
	const int td = CH * KT;
	const size_t per_warp = ((size_t)NT * td + ld + 32 + 15) & ~(size_t)15;
	double *tiles = smem_lines(smem) + per_warp * warp;
	double *Ls = tiles + (size_t)NT * td;
	double *es = Ls + ld;

	const int r = p.n_long + (int)blockIdx.x * WARPS + warp;
	if (r >= p.n_rows)
		return;
	const int i = p.row_id[r] - p.user_lo;
	const int base = p.row_ptr[r];
	const int m = p.row_ptr[r + 1] - base;
	const double *Lrow = Lold + (size_t)i * ld;
	double *Lout = Lnew + (size_t)i * ld;

	for (int q = lane; q < (ld >> 1); q += 32)      /* joins the first tile's copy group */
		cp_async16(Ls + 2 * q, Lrow + 2 * q);
	for (int c0 = 0; c0 < m; c0 += CH) {
		const int cnt = min(CH, m - c0);
		const int pos = base + c0 + lane;
		int j = 0;
		double a = 0.0;
		if (lane < cnt) {
			j = p.col_idx[pos];
			a = p.val[pos];
		}
		j = __shfl_sync(FULL, j, min(lane, cnt - 1));   /* idle lanes mirror a valid item */
		const double e = chunk_error<NT>(tiles, td, Ls, Rold, j, a, cnt, F, ld, p.alpha2, lane);
		if (lane < cnt) {
			p.e_csc[p.epos[pos]] = e;
			es[lane] = e;
		}
		__syncwarp();
		/* L_new[i][k] chain over this chunk's nonzeros; the feature groups of a lane are
		 * independent chains and run interleaved */
		const bool cs = p.cs_store && c0 + CH >= m;      /* the last chunk's store is final */
		for (int kb = 0; kb < F; kb += 256) {
			const double *init = (c0 == 0) ? Ls : Lout;
			switch (min(8, (F - kb + 31) >> 5)) {
			case 1: row_update<1, P>(init, Lout, Rold, j, es, cnt, ld, kb, F, lane, cs); break;
			case 2: row_update<2, P>(init, Lout, Rold, j, es, cnt, ld, kb, F, lane, cs); break;
			case 3: row_update<3, P>(init, Lout, Rold, j, es, cnt, ld, kb, F, lane, cs); break;
			case 4: row_update<4, P>(init, Lout, Rold, j, es, cnt, ld, kb, F, lane, cs); break;
			case 5: row_update<5, P>(init, Lout, Rold, j, es, cnt, ld, kb, F, lane, cs); break;
			case 6: row_update<6, P>(init, Lout, Rold, j, es, cnt, ld, kb, F, lane, cs); break;
			case 7: row_update<7, P>(init, Lout, Rold, j, es, cnt, ld, kb, F, lane, cs); break;
			default: row_update<8, P>(init, Lout, Rold, j, es, cnt, ld, kb, F, lane, cs); break;
			}
		}
		__syncwarp();                 /* es is rewritten by the next chunk */
	}
}

/* nonzeros per register sub-block of the accumulation kernel k_acc (below) */
#ifndef MF_ACC_SB
#define MF_ACC_SB 4
#endif
constexpr int ACC_SB = MF_ACC_SB;
#ifndef MF_ACC_CTAS
#define MF_ACC_CTAS 5
#endif
constexpr int ACC_CTAS = MF_ACC_CTAS;
#ifndef MF_ACC_PRED
#define MF_ACC_PRED 1
#endif
#ifndef MF_ACC_SWITCH
#define MF_ACC_SWITCH 0
#endif

/* a = ((a + e[0]*x[0]) + e[1]*x[1]) + ... over the first N nonzeros of a sub-block */
template <int N>
__device__ __forceinline__ double acc_chain_n(double a, const double (&x)[ACC_SB], const double (&e)[ACC_SB])
{
#pragma unroll
	for (int n = 0; n < N; n++)
		a = __dadd_rn(a, __dmul_rn(e[n], x[n]));
	return a;
}

/* m (1..ACC_SB) is the same in every lane: real branches, no selects in the chains */
__device__ __forceinline__ double acc_chain(double a, const double (&x)[ACC_SB], const double (&e)[ACC_SB], int m)
{
	static_assert(ACC_SB == 4 || ACC_SB == 8, "sub-blocks of 4 or 8 nonzeros");
#if !MF_ACC_SWITCH
	if (m == ACC_SB)
		return acc_chain_n<ACC_SB>(a, x, e);
#pragma unroll
	for (int n = 0; n < ACC_SB; n++)
		if (n < m)
			a = __dadd_rn(a, __dmul_rn(e[n], x[n]));
	return a;
#endif
	switch (m) {
	case 1: return acc_chain_n<1>(a, x, e);
	case 2: return acc_chain_n<2>(a, x, e);
	case 3: return acc_chain_n<3>(a, x, e);
	case 4: return acc_chain_n<4>(a, x, e);
	case 5: return acc_chain_n<(ACC_SB > 4 ? 5 : 4)>(a, x, e);
	case 6: return acc_chain_n<(ACC_SB > 4 ? 6 : 4)>(a, x, e);
	case 7: return acc_chain_n<(ACC_SB > 4 ? 7 : 4)>(a, x, e);
	default: return acc_chain_n<ACC_SB>(a, x, e);
	}
}

/*
 * k_dot_units: every dot of the by-user pass (mat2d.c:126-139, matFact.c:45) in split mode.
 *
 * A unit is up to 32 nonzeros that share at most two users: a chunk of a long row, one row of
 * at most 32 nonzeros, or TWO such rows whose lengths add up to at most 32 (lanes [0, la) row
 * A, lanes [la, la + lb) row B).  The dot costs the shared-memory pipe and the FP64 pipe one
 * instruction per warp whatever the number of active lanes (a 16-byte shared-memory load with
 * 15 of 32 lanes active takes as long as with 32, profiles/r02_ubench_l2_gather.txt), and the
 * inst60000 shape has 10..20 nonzeros per user: pairing fills the lanes.
 *
 * What bounds the pass is the latency of the gathered R rows (~2000 cycles under load), so a
 * warp is persistent and treats the tiles of ALL its units as one stream: the ring of NT tiles
 * (16 features of the unit's <= 32 R rows each, 16-byte cp.async) is filled NT - 1 tiles ahead
 * of the tile being added, across unit boundaries; the next unit's items, ratings and positions
 * are loaded one unit ahead and its descriptor two units ahead, so no dependent global load is
 * ever waited for.  Lane = nonzero walks k sequentially against the L row of its own user
 * (whose 16 features of the tile travel in the tile itself).
 * Writes e_n at its by-user and by-item positions; k_acc does the accumulations.
 * units: 8 ints per unit {ia, basea, la, ib, baseb, lb, -, -} (local user, first by-user
 * position, length of the two parts).  Needs NT - 1 <= ceil(F / 16) (the host sees to it).
 * Dynamic shared memory per warp: NT tiles [34][16] (the R rows and, as rows 32 and 33, the same 16
 * features of the two L rows: no L buffer of its own, 9.9 KB per warp at NT = 2), unit data
 * [2][32] x (rating, position, by-item position), [2][2] lengths.
 */
constexpr int DOT_ROWS = 34;   /* rows of a tile: the <= 32 gathered R rows, then the L rows of user A and user B */

__host__ __device__ constexpr size_t dot_units_doubles(int NT, int ld)
{
	(void)ld;
	return (size_t)NT * DOT_ROWS * KT + 144;       /* 144: the unit data, rounded to whole 128-byte lines */
}


#ifndef MF_DOT_MINB
#define MF_DOT_MINB 4
#endif
template <int WARPS, int NT>
__global__ void __launch_bounds__(WARPS * 32, WARPS == 4 ? MF_DOT_MINB : 512 / (WARPS * 32))
k_dot_units(const LoopArgs p, const int *__restrict__ units, int n_units,
            const double *__restrict__ Lold, const double *__restrict__ Rold)
{
	extern __shared__ __align__(16) double smem[];
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const int F = p.F, ld = p.ld;
	constexpr int td = DOT_ROWS * KT;
	/* tile rows on 128-byte lines of shared memory (the launch adds 128 bytes of slack) */
	double *tiles = smem_lines(smem) + dot_units_doubles(NT, ld) * warp;
	double *m_a = tiles + (size_t)NT * td;         /* [parity][32] */
	int *m_pos = (int *)(m_a + 64);                /* [parity][32] */
	int *m_ep = m_pos + 64;                        /* [parity][32] */
	int *m_len = m_ep + 64;                        /* [parity]{cnt, la} */
	const int nt = (F + KT - 1) / KT;
	const long W = (long)gridDim.x * WARPS, wid = (long)blockIdx.x * WARPS + warp;
	if (wid >= n_units)
		return;
	const int n_mine = (int)((n_units - wid + W - 1) / W);
	const long total = (long)n_mine * nt;

	struct Desc { int ia, basea, la, ib, baseb, lb; };
	struct LaneData { int pos, j, ep; double a; };
	auto load_desc = [&](int t, Desc &d) {
		const int *u = units + 8 * (size_t)(wid + t * W);
		const int4 x = *reinterpret_cast<const int4 *>(u);
		const int2 y = *reinterpret_cast<const int2 *>(u + 4);
		d.ia = x.x; d.basea = x.y; d.la = x.z; d.ib = x.w; d.baseb = y.x; d.lb = y.y;
	};
	auto load_lane = [&](const Desc &d, LaneData &x) {
		const int l = min(lane, d.la + d.lb - 1);      /* idle lanes mirror the last nonzero */
		x.pos = l < d.la ? d.basea + l : d.baseb + l - d.la;
		x.j = p.col_idx[x.pos];
		x.a = p.val[x.pos];
		x.ep = p.epos[x.pos];
	};

	Desc dI, dN;                /* unit being issued, the one after it */
	LaneData xI, xN;
	load_desc(0, dI);
	load_lane(dI, xI);
	if (n_mine > 1)
		load_desc(1, dN);
	/* Gathers of the unit being issued: lane l moves the 16-byte piece pc = l & 7 of the tile
	 * rows u*4 + (l >> 3), u < 8 -- one cp.async each per tile, destinations 4 tile rows apart. */
	const double *gsrc[8];
	const double *lsrc = Lold;  /* lanes 0..7: pieces of the L row of user A, lanes 8..15: of user B */
	bool l_ok = false;
	int gn = 0;                 /* pieces of this lane: u < gn */
	const int pc2 = 2 * (lane & 7);
	/* destination of this lane's piece in tile rows u*4 + (lane >> 3): row & 7 = (u & 1)*4 + (lane >> 3) */
	const uint32_t gdst_e = smem_u32(tiles) + (uint32_t)((lane >> 3) * 128 + (((lane & 7) ^ (lane >> 3)) << 4));
	const uint32_t gdst_o = smem_u32(tiles) + (uint32_t)((lane >> 3) * 128 + (((lane & 7) ^ (4 + (lane >> 3))) << 4));
	int ti = 0, ki = 0;         /* issue side: my ti-th unit, its tile ki */

	auto issue = [&](int g) {
		if (ki == 0) {
			/* enter unit ti: plan of its gathers, its data for the adding side, its L rows */
			const int par = ti & 1, cnt = dI.la + dI.lb;
			gn = 0;
#pragma unroll
			for (int u = 0; u < 8; u++) {
				const int row = u * 4 + (lane >> 3);
				const int jn = __shfl_sync(FULL, xI.j, min(row, cnt - 1));
				gsrc[u] = Rold + (size_t)jn * ld + pc2;
				if (row < cnt)
					gn = u + 1;
			}
			m_a[par * 32 + lane] = xI.a;
			m_pos[par * 32 + lane] = xI.pos;
			m_ep[par * 32 + lane] = xI.ep;
			if (lane == 0) {
				m_len[par * 2] = cnt;
				m_len[par * 2 + 1] = dI.la;
			}
			/* the 16 features of the two L rows travel in every tile, as its rows 32 and 33 */
			lsrc = Lold + (size_t)(lane < 8 ? dI.ia : dI.ib) * ld + pc2;
			l_ok = lane < 8 || (lane < 16 && dI.lb > 0);
			/* the loads of the units behind it start now and have a whole unit to land */
			if (ti + 1 < n_mine)
				load_lane(dN, xN);
			Desc d2 = dN;
			if (ti + 2 < n_mine)
				load_desc(ti + 2, d2);
			dI = dN;
			dN = d2;
		}
		{
			const int k0 = ki * KT;
			const bool lane_ok = k0 + pc2 < ld;
			const uint32_t tile = (uint32_t)g * (uint32_t)(td * 8);
#pragma unroll
			for (int u = 0; u < 8; u++)
				cp_async16_if(((u & 1) ? gdst_o : gdst_e) + tile + (uint32_t)(u * 4 * 128), gsrc[u] + k0, lane_ok && u < gn);
			cp_async16_if(gdst_e + tile + (uint32_t)(8 * 4 * 128), lsrc + k0, lane_ok && l_ok);
		}
		if (++ki == nt) {
			ki = 0;
			ti++;
			xI = xN;
		}
	};

	int gi = 0;                 /* ring slot of the next tile to issue */
	for (int g = 0; g < NT - 1; g++) {
		if (g < total) {
			issue(gi);
			gi = gi + 1 == NT ? 0 : gi + 1;
		}
		cp_async_commit();
	}
	int gc = 0;                 /* ring slot of the tile being added */
	int kc = 0, par = 0, cnt = 0;
	double s = 0.0, a = 0.0;
	int lrow = 32;              /* tile row of this lane's L row */
	for (long g = 0; g < total; g++) {
		if (g + NT - 1 < total) {
			issue(gi);
			gi = gi + 1 == NT ? 0 : gi + 1;
		}
		cp_async_commit();
		cp_async_wait<NT - 1>();       /* tile g (and the L rows that travel with a first tile) has landed */
		__syncwarp();
		if (kc == 0) {
			cnt = m_len[par * 2];
			const int la = m_len[par * 2 + 1];
			a = m_a[par * 32 + lane];
			lrow = lane < la ? 32 : 33;
			s = 0.0;
		}
		if (lane < cnt)
			s = dot_tile_swz(s, tiles + (size_t)gc * td + lrow * KT, lrow & 7, tiles + (size_t)gc * td + lane * KT, lane & 7,
			                 min(KT, F - kc * KT));
		gc = gc + 1 == NT ? 0 : gc + 1;
		if (++kc == nt) {
			if (lane < cnt) {
				const double e = __dmul_rn(p.alpha2, __dsub_rn(a, s));
				p.e_csc[m_ep[par * 32 + lane]] = e;
				p.e_csr[m_pos[par * 32 + lane]] = e;
			}
			kc = 0;
			par ^= 1;
		}
		__syncwarp();                  /* the slot and the unit data may be rewritten by the next issue */
	}
	cp_async_wait<0>();
}

/*
 * By-user pass of the short rows with the gathered R rows resident in shared memory: each
 * R row of a nonzero is read from L2 ONCE and serves both the dot (lane = nonzero, k
 * sequential) and the accumulation of L_new (lane = feature, nonzeros in file order); the
 * gathers are what bounds the loop (L2 -> SM bytes), so this halves the cost of the pass
 * against the tiled k_update_L, which streams the rows for the dot and reads them again for
 * the accumulation.  One warp per row, chunks of CR nonzeros; rows RST doubles apart with
 * RST/2 odd, so that the 16-byte reads of the dot (8 lanes = 8 rows per wavefront) are
 * conflict free.  The copies are issued as groups of 64 features (one 16-byte piece per
 * lane and row) and the dot starts as soon as the first group has landed.  The L_new
 * accumulators of the NG feature groups of a lane stay in registers over the whole row.
 * Dynamic shared memory per warp: R rows [CR][RST], the user's L row [ld], es[32].
 * Used for F <= 256 (NG <= 8).
 */
constexpr int RES_WARPS = 2;

/* s += x[0]*y[0], + x[1]*y[1], ... in order over 8 features held as four double2 pairs */
__device__ __forceinline__ double dot8(double s, const double2 (&x)[4], const double2 (&y)[4])
{
#pragma unroll
	for (int u = 0; u < 4; u++) {
		s = __dadd_rn(s, __dmul_rn(x[u].x, y[u].x));
		s = __dadd_rn(s, __dmul_rn(x[u].y, y[u].y));
	}
	return s;
}

__device__ __forceinline__ void load8(double2 (&x)[4], double2 (&y)[4], const double *l, const double *r)
{
#pragma unroll
	for (int u = 0; u < 4; u++) {
		x[u] = ld2(l + 2 * u);
		y[u] = ld2(r + 2 * u);
	}
}

/* s continued over features [k0, k1) of l . r (k0 even, 16-byte aligned operands): blocks of
 * 8 features alternate between two register sets, the next block being read from shared
 * memory while the current one is added */
__device__ __forceinline__ double dot_range(double s, const double *l, const double *r, int k0, int k1)
{
	const int nblk = (k1 - k0) >> 3;
	if (nblk > 0) {
		double2 xa[4], ya[4], xb[4], yb[4];
		load8(xa, ya, l + k0, r + k0);
		int b = 0;
		for (; b + 2 < nblk; b += 2) {
			load8(xb, yb, l + k0 + 8 * (b + 1), r + k0 + 8 * (b + 1));
			s = dot8(s, xa, ya);
			load8(xa, ya, l + k0 + 8 * (b + 2), r + k0 + 8 * (b + 2));
			s = dot8(s, xb, yb);
		}
		if (b + 1 < nblk) {
			load8(xb, yb, l + k0 + 8 * (b + 1), r + k0 + 8 * (b + 1));
			s = dot8(s, xa, ya);
			s = dot8(s, xb, yb);
		} else {
			s = dot8(s, xa, ya);
		}
	}
	for (int k = k0 + 8 * nblk; k < k1; k++)
		s = __dadd_rn(s, __dmul_rn(l[k], r[k]));
	return s;
}

/* x[g] = row[g*32 + lane] for the NG feature groups of a lane (klast: the last group's index) */
template <int NG>
__device__ __forceinline__ void load_groups(double (&x)[NG], const double *row, int lane, int klast)
{
	const double *rl = row + lane;
#pragma unroll
	for (int g = 0; g < NG - 1; g++)
		x[g] = rl[g * 32];
	x[NG - 1] = row[klast];
}

/* Gather of the R rows of a chunk into shared memory (k_update_L_res): row n
 * goes to rs_addr + n*rstb (rs_addr, sbase already include this lane's 16-byte piece); the
 * 16-byte pieces of a lane in a row are features 64t + 2*lane, t < NT_MAX.  The first NT_A
 * k-ranges form copy group A, the rest group B, so that the dot can start after A. */
template <int NG>
__device__ __forceinline__ void res_stage(uint32_t rs_addr, const char *sbase, unsigned ldb, unsigned rstb,
                                          int j, int cnt, const bool (&piece_ok)[(NG + 1) / 2])
{
	constexpr int NT_MAX = (NG + 1) / 2, NT_A = (NT_MAX + 1) / 2;
#pragma unroll
	for (int half = 0; half < 2; half++) {
		const int t0 = half ? NT_A : 0, t1 = half ? NT_MAX : NT_A;
		uint32_t drow = rs_addr;
		int n = 0;
		for (; n + 4 <= cnt; n += 4) {       /* four rows per trip, no guards */
			unsigned jn[4];
#pragma unroll
			for (int u = 0; u < 4; u++)
				jn[u] = (unsigned)__shfl_sync(FULL, j, n + u);
#pragma unroll
			for (int u = 0; u < 4; u++) {
				const char *srow = sbase + (unsigned long long)jn[u] * ldb;
#pragma unroll
				for (int t = 0; t < NT_MAX; t++)
					if (t >= t0 && t < t1)
						cp_async16_if(drow + u * rstb + 512 * t, srow + 512 * t, piece_ok[t]);
			}
			drow += 4 * rstb;
		}
		for (; n < cnt; n++) {
			const unsigned j1 = (unsigned)__shfl_sync(FULL, j, n);
			const char *srow = sbase + (unsigned long long)j1 * ldb;
#pragma unroll
			for (int t = 0; t < NT_MAX; t++)
				if (t >= t0 && t < t1)
					cp_async16_if(drow + 512 * t, srow + 512 * t, piece_ok[t]);
			drow += rstb;
		}
		cp_async_commit();
	}
}

template <int NG>
__global__ void __launch_bounds__(RES_WARPS * 32)
k_update_L_res(const LoopArgs p, const double *__restrict__ Lold, double *__restrict__ Lnew,
               const double *__restrict__ Rold)
{
	extern __shared__ __align__(16) double smem[];
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const int F = p.F, ld = p.ld, CR = p.CR, RST = p.RST;
	const size_t per_warp = (size_t)CR * RST + ld + 32;
	double *Rs = smem + per_warp * warp;
	double *Ls = Rs + (size_t)CR * RST;
	double *es = Ls + ld;

	/* the 16-byte pieces of this lane in a row: features 64t + 2*lane, t < NT; the first
	 * NT_A k-ranges form copy group A, the rest group B, and the dot starts after A */
	constexpr int NT_MAX = (NG + 1) / 2, NT_A = (NT_MAX + 1) / 2;
	const int kA = min(F, 64 * NT_A);              /* features covered by group A */
	const char *sbase = (const char *)(Rold + 2 * lane);
	const unsigned ldb = (unsigned)ld * 8u, rstb = (unsigned)RST * 8u;
	const uint32_t rs_addr = smem_u32(Rs + 2 * lane);
	bool piece_ok[NT_MAX];
#pragma unroll
	for (int t = 0; t < NT_MAX; t++)
		piece_ok[t] = 64 * t + 2 * lane < ld;
	/* only the last feature group can be partial; its idle lanes read a valid address */
	const bool ok_last = (NG - 1) * 32 + lane < F;
	const int klast = ok_last ? (NG - 1) * 32 + lane : 0;

	/* rows r, r + stride, ...: the row's (user, first position, length) and the (item, value,
	 * by-item position) of its first chunk are fetched one row ahead */
	const int stride = (int)gridDim.x * RES_WARPS;
	int r = p.n_long + (int)blockIdx.x * RES_WARPS + warp;
	if (r >= p.n_rows)
		return;
	int i = p.row_id[r] - p.user_lo, base = p.row_ptr[r], m = p.row_ptr[r + 1] - base;
	int j = 0, ep = 0;
	double a = 0.0;
	if (lane < min(CR, m)) {
		j = p.col_idx[base + lane];
		a = p.val[base + lane];
		ep = p.epos[base + lane];
	}
	while (r < p.n_rows) {
		const int rn = r + stride;
		int i2 = 0, base2 = 0, m2 = 0;
		if (rn < p.n_rows) {
			i2 = p.row_id[rn] - p.user_lo;
			base2 = p.row_ptr[rn];
			m2 = p.row_ptr[rn + 1] - base2;
		}
		const double *Lrow = Lold + (size_t)i * ld;
		double *Lout = Lnew + (size_t)i * ld;
		double acc[NG];
		for (int q = lane; q < (ld >> 1); q += 32)      /* joins the first copy group */
			cp_async16(Ls + 2 * q, Lrow + 2 * q);
		for (int c0 = 0; c0 < m; c0 += CR) {
			const int cnt = min(CR, m - c0);
			res_stage<NG>(rs_addr, sbase, ldb, rstb, j, cnt, piece_ok);
			/* the next chunk's nonzeros (of this row, or the first of the next row) */
			const double a_cur = a;
			const int ep_cur = ep;
			int jx = 0, epx = 0;
			double ax = 0.0;
			{
				const bool same = c0 + CR < m;
				const int posx = same ? base + c0 + CR + lane : base2 + lane;
				const int cntx = same ? min(CR, m - c0 - CR) : min(CR, m2);
				if (lane < cntx) {
					jx = p.col_idx[posx];
					ax = p.val[posx];
					epx = p.epos[posx];
				}
			}
			/* e_n (mat2d.c:133-137, matFact.c:45): idle lanes walk a valid row, unused */
			/* (lanes without a nonzero stay out: their quarter-warp costs no shared-memory
			 * wavefront, and that pipe is what this kernel runs against) */
			const double *rrow = Rs + lane * RST;
			cp_async_wait<1>();
			__syncwarp();
			double s = 0.0;
			if (lane < cnt)
				s = dot_range(s, Ls, rrow, 0, kA);
			cp_async_wait<0>();
			__syncwarp();
			if (lane < cnt)
				s = dot_range(s, Ls, rrow, kA, F);
			const double e = __dmul_rn(p.alpha2, __dsub_rn(a_cur, s));
			if (lane < cnt) {
				p.e_csc[ep_cur] = e;
				es[lane] = e;
			}
			__syncwarp();
			if (c0 == 0) {
#pragma unroll
				for (int g = 0; g < NG; g++)
					acc[g] = Ls[g < NG - 1 ? g * 32 + lane : klast];
			}
			/* matFact.c:48-49 in file order; the NG chains of a lane are independent; two
			 * register sets alternate so that the operands of n+1 are read while n is added */
			{
				double x0[NG], x1[NG], e0 = es[0], e1;
				load_groups<NG>(x0, Rs, lane, klast);
				int n = 0;
				for (; n + 1 < cnt; n += 2) {
					e1 = es[n + 1];
					load_groups<NG>(x1, Rs + (n + 1) * RST, lane, klast);
#pragma unroll
					for (int g = 0; g < NG; g++)
						acc[g] = __dadd_rn(acc[g], __dmul_rn(e0, x0[g]));
					const int n2 = min(n + 2, cnt - 1);
					e0 = es[n2];
					load_groups<NG>(x0, Rs + n2 * RST, lane, klast);
#pragma unroll
					for (int g = 0; g < NG; g++)
						acc[g] = __dadd_rn(acc[g], __dmul_rn(e1, x1[g]));
				}
				if (n < cnt) {
#pragma unroll
					for (int g = 0; g < NG; g++)
						acc[g] = __dadd_rn(acc[g], __dmul_rn(e0, x0[g]));
				}
			}
			__syncwarp();                 /* Rs, Ls and es are rewritten by the next copies */
			j = jx;
			a = ax;
			ep = epx;
		}
		if (m > 0) {
#pragma unroll
			for (int g = 0; g < NG - 1; g++)
				st_row(Lout + g * 32 + lane, acc[g], p.cs_store);
			if (ok_last)
				st_row(Lout + klast, acc[NG - 1], p.cs_store);
		}
		r = rn;
		i = i2;
		base = base2;
		m = m2;
	}
}

/*
 * Dots of the long rows [0, n_long): one warp per chunk of <= CH nonzeros of one row, so a
 * row of thousands of nonzeros spreads over as many warps as it has chunks.  Writes e_n at
 * its by-user and by-item positions; k_chain does the accumulations.
 * Dynamic shared memory per warp: as k_update_L.
 */
template <int WARPS, int NT>
__global__ void __launch_bounds__(WARPS * 32)
k_dot(const LoopArgs p, const double *__restrict__ Lold, const double *__restrict__ Rold)
{
	extern __shared__ __align__(16) double smem[];
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const int F = p.F, ld = p.ld, CH = p.CH;
	const int td = CH * KT;
	const size_t per_warp = ((size_t)NT * td + ld + 32 + 15) & ~(size_t)15;
	double *tiles = smem_lines(smem) + per_warp * warp;
	double *Ls = tiles + (size_t)NT * td;

	const int c = (int)blockIdx.x * WARPS + warp;
	if (c >= p.n_chunks)
		return;
	const int r = p.chunk_row[c];
	const int first = p.chunk_pos[c];
	const int cnt = min(CH, p.row_ptr[r + 1] - first);
	const double *Lrow = Lold + (size_t)(p.row_id[r] - p.user_lo) * ld;
	for (int q = lane; q < (ld >> 1); q += 32)
		cp_async16(Ls + 2 * q, Lrow + 2 * q);
	const int pos = first + lane;
	int j = 0;
	double a = 0.0;
	if (lane < cnt) {
		j = p.col_idx[pos];
		a = p.val[pos];
	}
	j = __shfl_sync(FULL, j, min(lane, cnt - 1));
	const double e = chunk_error<NT>(tiles, td, Ls, Rold, j, a, cnt, F, ld, p.alpha2, lane);
	if (lane < cnt) {
		p.e_csr[pos] = e;
		p.e_csc[p.epos[pos]] = e;
	}
}

/*
 * Sequential accumulations: one warp per (stored row or column r, group of 32 features),
 * lane = feature; side A takes the first warps_a warps of the launch, side B the rest.
 *   L side (matFact.c:48-49): out = L_new, init = L_old, src = R_old, idx = items of the row
 *   R side (matFact.c:50-51): out = R_new (or G from 0.0 for N shards, matFact-mpi.c:187-205),
 *                             init = R_old, src = L_old, idx = users of the column
 */
/* The finished chain value of feature k of row/column `own` goes to the own matrix, to the row
 * of G, or -- rank mode -- into the receive buffer of the rank that reduces this item, over NVLink
 * while the other chains of the launch are still running; with the low-latency exchange as a
 * (value, tag) pair that the reducer polls for (st_pair, mf_async.cuh). */
__device__ __forceinline__ void chain_store(const ChainSide &sd, int own, int ld, int k, double acc)
{
	const size_t orow = sd.out_slot ? (size_t)sd.out_slot[own] : (size_t)(own - sd.own_off);
	double *base = sd.out_owner ? sd.peer_out[sd.out_owner[own]] : sd.out;
	if (sd.ll_epoch) {
		const unsigned long long tag = *sd.ll_epoch + 1;      /* the iteration being computed */
		st_pair(base + (long)(tag & 1) * sd.ll_half + 2 * (orow * ld + k), acc, tag);
	} else
		base[orow * ld + k] = acc;
}

constexpr int CHAIN_BULK_REGS = 0, CHAIN_DEEP_REGS = 1, CHAIN_BULK_RING = 2, CHAIN_DEEP_RING = 3;
constexpr int RING_BULK_ROWS = 8, RING_BULK_STAGES = 3;     /* 6.2 KB per warp  */
constexpr int RING_DEEP_ROWS = 16, RING_DEEP_STAGES = 6;    /* 24 KB per warp */

template <int WARPS, int MODE>
__global__ void __launch_bounds__(WARPS * 32)
k_chain(const ChainSide A, const ChainSide B, long warps_a, int F, int ld)
{
	extern __shared__ __align__(16) double smem[];
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const int NG = (F + 31) >> 5;
	long w = (long)blockIdx.x * WARPS + warp;
	const bool is_a = w < warps_a;
	const ChainSide &sd = is_a ? A : B;
	if (!is_a)
		w -= warps_a;
	const int r = sd.begin + (int)(w / NG), g = (int)(w % NG);
	if (r >= sd.end)
		return;
	const int k = g * 32 + lane;
	const bool active = k < F;
	const int own = sd.id[r];
	const int b = sd.ptr[r], m = sd.ptr[r + 1] - b;
	double acc = (sd.init && active) ? sd.init[(size_t)(own - sd.own_off) * ld + k] : 0.0;
	if (MODE == CHAIN_BULK_RING || MODE == CHAIN_DEEP_RING) {
		const double *src_g = sd.src + g * 32 - (size_t)sd.src_off * ld;
		const int gw = min(32, ld - g * 32);
		if (MODE == CHAIN_DEEP_RING)
			acc = chain_ring<RING_DEEP_ROWS, RING_DEEP_STAGES>(
			    acc, lane, sd.idx + b, sd.e + b, m, src_g, ld, gw,
			    smem + (size_t)warp * ring_doubles(RING_DEEP_ROWS, RING_DEEP_STAGES));
		else
			acc = chain_ring<RING_BULK_ROWS, RING_BULK_STAGES>(
			    acc, lane, sd.idx + b, sd.e + b, m, src_g, ld, gw,
			    smem + (size_t)warp * ring_doubles(RING_BULK_ROWS, RING_BULK_STAGES));
	} else {
		const double *srck = sd.src + (active ? k : 0) - (size_t)sd.src_off * ld;   /* idle lanes: a valid column */
		if (MODE == CHAIN_DEEP_REGS)
			acc = chain_deep(acc, lane, sd.idx + b, sd.e + b, m, srck, ld);
		else
			acc = chain_bulk<8>(acc, lane, sd.idx + b, sd.e + b, m, srck, ld);
	}
	if (active)
		chain_store(sd, own, ld, k, acc);
}

/*
 * k_chain_ws: the same sequential accumulation for the longest chains, one CTA per (row or
 * column, group of 32 features) with the work split over specialised warps.  What bounds a
 * long chain is the latency of one warp: in k_chain ~13 instructions per nonzero plus
 * exposed load latencies (a warp's loads share six scoreboards, so waiting for an old gather
 * also waits for the newest), although only the addition has to be serial (8 cycles,
 * profiles/r01_ubench_fp64.txt).  Here
 *   warps 1, 2 (producers) take the stages of WS_ROWS = 32 nonzeros alternately: 16-byte
 *       cp.async of x[n] = src[idx[n]] (a half-warp per nonzero) and 8-byte cp.async of e[n]
 *       into a ring of WS_STAGES stages; the copies signal the stage's `full` mbarrier
 *       themselves when they land, so a producer never waits for data, only for a free
 *       slot.  The indices themselves run WS_IDX_AHEAD blocks ahead in a small ring of the
 *       warp's own (4-byte cp.async): cp.async groups complete in order, so
 *       wait_group<WS_IDX_AHEAD> means "my next block has landed" without waiting for the
 *       younger ones, which a register load cannot express;
 *   warp 0 (consumer) reads a landed stage into registers, adds e[n]*x[n] in file order
 *       (one LDS, one DMUL, one DADD per nonzero) and hands the slot back (`empty`).
 * WS_STAGES x WS_ROWS = 192 nonzeros are in flight ahead of the additions.
 */
constexpr int WS_PRODUCERS = 2, WS_ROWS = 32, WS_STAGES = 6;
constexpr int WS_THREADS = 32 * (1 + WS_PRODUCERS);
constexpr int WS_STAGE_DOUBLES = WS_ROWS * 32 + WS_ROWS;          /* x[32][32] | e[32] */
constexpr int WS_IDX_AHEAD = 4, WS_IDX_BLOCKS = 8;                 /* blocks of 32 indices, per producer */
constexpr int WS_SMEM = (WS_STAGES * WS_STAGE_DOUBLES + 2 * WS_STAGES) * (int)sizeof(double) +
                        WS_PRODUCERS * WS_IDX_BLOCKS * 32 * (int)sizeof(int);

__global__ void __launch_bounds__(WS_THREADS)
k_chain_ws(const ChainSide A, const ChainSide B, long units_a, int F, int ld)
{
	extern __shared__ __align__(16) double smem[];
	const uint32_t bars = smem_u32(smem + WS_STAGES * WS_STAGE_DOUBLES);
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const int NG = (F + 31) >> 5;
	long w = blockIdx.x;
	const bool is_a = w < units_a;
	const ChainSide &sd = is_a ? A : B;
	if (!is_a)
		w -= units_a;
	const int r = sd.begin + (int)(w / NG), g = (int)(w % NG);
	const int b = sd.ptr[r], m = sd.ptr[r + 1] - b;
	const int nst = (m + WS_ROWS - 1) / WS_ROWS;
	auto full = [&](int slot) { return bars + 8u * slot; };
	auto empty = [&](int slot) { return bars + 8u * (WS_STAGES + slot); };

	if (threadIdx.x == 0)
		for (int q = 0; q < WS_STAGES; q++) {
			mbar_init(full(q), 32);       /* the 32 lanes of the stage's producer */
			mbar_init(empty(q), 1);       /* lane 0 of the consumer                 */
		}
	__syncthreads();

	if (warp == 0) {
		/* ---- consumer ---- */
		const int k = g * 32 + lane;
		const bool active = k < F;
		const int own = sd.id[r];
		double acc = (sd.init && active) ? sd.init[(size_t)(own - sd.own_off) * ld + k] : 0.0;
		for (int s = 0; s < nst; s++) {
			const int slot = s % WS_STAGES;
			const double *st = smem + (size_t)slot * WS_STAGE_DOUBLES;
			double x[WS_ROWS];
			double2 e[WS_ROWS / 2];
			mbar_wait(full(slot), (s / WS_STAGES) & 1);
#pragma unroll
			for (int q = 0; q < WS_ROWS; q++)
				x[q] = st[q * 32 + lane];
#pragma unroll
			for (int q = 0; q < WS_ROWS / 2; q++)
				e[q] = ld2(st + WS_ROWS * 32 + 2 * q);
			const int rows = m - s * WS_ROWS;
			if (rows >= WS_ROWS) {
#pragma unroll
				for (int q = 0; q < WS_ROWS / 2; q++) {
					acc = __dadd_rn(acc, __dmul_rn(e[q].x, x[2 * q]));
					acc = __dadd_rn(acc, __dmul_rn(e[q].y, x[2 * q + 1]));
				}
			} else {
#pragma unroll
				for (int q = 0; q < WS_ROWS / 2; q++) {
					if (2 * q < rows)
						acc = __dadd_rn(acc, __dmul_rn(e[q].x, x[2 * q]));
					if (2 * q + 1 < rows)
						acc = __dadd_rn(acc, __dmul_rn(e[q].y, x[2 * q + 1]));
				}
			}
			__syncwarp();
			if (lane == 0)
				mbar_arrive(empty(slot));
		}
		if (active)
			chain_store(sd, own, ld, k, acc);
	} else {
		/* ---- producer p: stages p, p + 2, ... (stage s = nonzeros 32s .. 32s+31) ---- */
		const int p = warp - 1;
		const int half = lane >> 4, piece = lane & 15;
		const int gw = min(32, ld - g * 32);                 /* doubles of this group in a row (even) */
		const bool piece_ok = 2 * piece < gw;
		const double *src_g = sd.src + g * 32 + 2 * piece - (size_t)sd.src_off * ld;
		const int *idx = sd.idx + b;
		const double *ev = sd.e + b;
		int *iring = (int *)(smem + WS_STAGES * WS_STAGE_DOUBLES + 2 * WS_STAGES) + p * WS_IDX_BLOCKS * 32;
		auto request = [&](int t) {          /* indices of this producer's t-th stage */
			const int s = 2 * t + p;
			if (s < nst)
				cp_async4(iring + (t % WS_IDX_BLOCKS) * 32 + lane, idx + min(s * 32 + lane, m - 1));
			cp_async_commit();
		};
#pragma unroll
		for (int t = 0; t < WS_IDX_AHEAD; t++)
			request(t);
		for (int t = 0; 2 * t + p < nst; t++) {
			const int s = 2 * t + p, slot = s % WS_STAGES;
			request(t + WS_IDX_AHEAD);
			cp_async_wait<WS_IDX_AHEAD>();
			__syncwarp();
			const int *bi = iring + (t % WS_IDX_BLOCKS) * 32 + half;
			int iu[WS_ROWS / 2];             /* all indices first: the copies are compiler barriers */
#pragma unroll
			for (int u = 0; u < WS_ROWS / 2; u++)
				iu[u] = bi[2 * u];
			if (s >= WS_STAGES)
				mbar_wait(empty(slot), ((s / WS_STAGES) - 1) & 1);
			double *st = smem + (size_t)slot * WS_STAGE_DOUBLES;
			const uint32_t st_addr = smem_u32(st) + (uint32_t)(half * 32 + 2 * piece) * 8u;
#pragma unroll
			for (int u = 0; u < WS_ROWS / 2; u++)
				cp_async16_if(st_addr + u * 512u, src_g + (size_t)iu[u] * ld, piece_ok);
			const int n = s * WS_ROWS + lane;
			if (n < m)
				cp_async8(st + WS_ROWS * 32 + lane, ev + n);
			mbar_arrive_on_copies(full(slot));
			__syncwarp();                 /* ring entry t + WS_IDX_BLOCKS overwrites this block */
		}
		cp_async_wait<0>();
	}
}

/*
 * k_acc: the ordered accumulations (matFact.c:48-51) straight from L2 into registers.
 *
 * A unit is (stored row or column r, block of GW feature groups); one warp takes a unit and keeps
 * its GW x 32 running values in registers (lane = feature).  The nonzeros of the chain come in
 * blocks of 32 (index and e_n of nonzero 32q + l in lane l, handed to all lanes through a
 * 384-byte shared-memory slot of the warp), and are taken in sub-blocks of ACC_SB: the
 * sub-block's row pointers and e_n live in registers, and the groups are walked with two register
 * sets, the ACC_SB loads of group q+1 in flight while the ACC_SB additions of group q run
 * (16 x 256 bytes in flight per warp, ~20 warps per SM).  The loads are written as PTX with the
 * group offset as an immediate, in program order: left to the compiler, an unrolled group loop
 * hoists the loads of all groups and spills, and a rolled one pays two address instructions per
 * load.  Lanes past F in the last group read the doubles that follow the row (never used; the
 * factor buffers end with 32 spare doubles).
 * Units are handed out by an atomic counter (*ticket, zeroed before the launch), heaviest first:
 * side B (the columns, long chains), then side A (the rows); the next unit's position, length
 * and first block of indices are fetched while the current one is being added.
 */
template <int OFF>
__device__ __forceinline__ double ldg_imm(const double *p)
{
	double v;
	asm volatile("ld.global.nc.f64 %0, [%1+%2];" : "=d"(v) : "l"(p), "n"(OFF));
	return v;
}

/* the same under a predicate (the lanes past F of a partial last group load nothing) */
template <int OFF>
__device__ __forceinline__ double ldg_imm_if(const double *p, bool ok)
{
	double v = 0.0;
	asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %3, 0;\n\t@q ld.global.nc.f64 %0, [%1+%2];\n\t}"
	             : "+d"(v) : "l"(p), "n"(OFF), "r"((int)ok));
	return v;
}

template <int Q, bool LAST>
__device__ __forceinline__ void acc_load_q(double (&x)[ACC_SB], const double *const (&rp)[ACC_SB], bool ok_last)
{
#pragma unroll
	for (int n = 0; n < ACC_SB; n++)
		x[n] = (LAST && MF_ACC_PRED) ? ldg_imm_if<Q * 256>(rp[n], ok_last) : ldg_imm<Q * 256>(rp[n]);
}

/* all GW x ACC_SB loads of a sub-block first, then the GW chains: with ~170 registers a warp
 * keeps up to 56 loads (14 KB) in flight and twelve warps per SM cover the L2 latency
 * (measured: two register sets with one group fetched ahead left the additions waiting
 * on every group, 10.6 TB/s; profiles/r02_notes.md) */
template <int GW, int Q>
struct AccGroups {
	static __device__ __forceinline__ void load(double (&x)[GW][ACC_SB], const double *const (&rp)[ACC_SB], bool ok_last)
	{
		acc_load_q<Q, Q == GW - 1>(x[Q], rp, ok_last);
		AccGroups<GW, Q + 1>::load(x, rp, ok_last);
	}
	static __device__ __forceinline__ void add(double (&acc)[GW], const double (&x)[GW][ACC_SB],
	                                           const double (&e)[ACC_SB], int m)
	{
		acc[Q] = acc_chain(acc[Q], x[Q], e, m);
		AccGroups<GW, Q + 1>::add(acc, x, e, m);
	}
};
template <int GW> struct AccGroups<GW, GW> {
	static __device__ __forceinline__ void load(double (&)[GW][ACC_SB], const double *const (&)[ACC_SB], bool) {}
	static __device__ __forceinline__ void add(double (&)[GW], const double (&)[GW][ACC_SB], const double (&)[ACC_SB], int) {}
};

/* row pointers and e_n of the sub-block that starts at nonzero s0 of the warp's block of 32 */
__device__ __forceinline__ void acc_prep(const int *sidx, const double *se, int s0, const double *srcb, int ld,
                                         const double *(&rp)[ACC_SB], double (&e)[ACC_SB])
{
#pragma unroll
	for (int n = 0; n < ACC_SB; n += 4) {
		const int4 i4 = *reinterpret_cast<const int4 *>(sidx + s0 + n);
		rp[n] = srcb + (size_t)i4.x * ld;           /* past the end: a valid row again, never added */
		rp[n + 1] = srcb + (size_t)i4.y * ld;
		rp[n + 2] = srcb + (size_t)i4.z * ld;
		rp[n + 3] = srcb + (size_t)i4.w * ld;
	}
#pragma unroll
	for (int n = 0; n < ACC_SB; n += 2) {
		const double2 e2 = ld2(se + s0 + n);
		e[n] = e2.x;
		e[n + 1] = e2.y;
	}
}

constexpr int ACC_WARPS = 4;

template <int GW>
__device__ __forceinline__ void acc_body(const ChainSide &A, const ChainSide &B, long units_b, long units_total,
                                         int F, int ld, unsigned long long *ticket)
{
	/* per warp: 32 indices (as doubles' worth of row offsets) and 32 e_n */
	__shared__ __align__(16) int s_idx[ACC_WARPS][32];
	__shared__ __align__(16) double s_e[ACC_WARPS][32];
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const int ng = (F + 31) >> 5;
	const int nb = (ng + GW - 1) / GW;          /* group blocks per chain */

	auto take = [&]() {
		unsigned long long t = 0;
		if (lane == 0)
			t = atomicAdd(ticket, 1ull);
		return (long)__shfl_sync(FULL, t, 0);
	};
	/* position of unit w: side, stored row, group block */
	struct Unit { bool is_b; int r, gb, own, beg, m; int idx0; double e0; };
	auto describe = [&](long w, Unit &u) {
		u.is_b = w < units_b;
		const ChainSide &sd = u.is_b ? B : A;
		const long v = u.is_b ? w : w - units_b;
		u.r = sd.begin + (int)(v / nb);
		u.gb = (int)(v % nb);
		u.own = sd.id[u.r];
		u.beg = sd.ptr[u.r];
		u.m = sd.ptr[u.r + 1] - u.beg;
	};
	auto first_block = [&](Unit &u) {
		const ChainSide &sd = u.is_b ? B : A;
		const int n = min(lane, u.m - 1);
		u.idx0 = sd.idx[u.beg + n];
		u.e0 = sd.e[u.beg + n];
	};

	long w = take();
	Unit cur, nxt;
	if (w < units_total) {
		describe(w, cur);
		first_block(cur);
	}
	while (w < units_total) {
		const long wn = take();
		const ChainSide &sd = cur.is_b ? B : A;
		const int g0 = cur.gb * GW;
		const int k0 = g0 * 32 + lane;
		double acc[GW];
#pragma unroll
		for (int q = 0; q < GW; q++) {
			const int k = k0 + q * 32;
			acc[q] = (sd.init && k < F) ? sd.init[(size_t)(cur.own - sd.own_off) * ld + k] : 0.0;
		}
		const double *srcb = sd.src + k0 - (size_t)sd.src_off * ld;
		const bool ok_last = k0 + (GW - 1) * 32 < F;      /* this lane's element of the block's last group exists */
		if (wn < units_total)
			describe(wn, nxt);
		int bi = cur.idx0;
		double be = cur.e0;
		for (int b0 = 0; b0 < cur.m; b0 += 32) {
			const int mb = min(32, cur.m - b0);
			__syncwarp();                      /* the slot is no longer read */
			s_idx[warp][lane] = bi;
			s_e[warp][lane] = be;
			__syncwarp();
			if (b0 + 32 < cur.m) {             /* next block of this chain */
				const int n = min(b0 + 32 + lane, cur.m - 1);
				bi = sd.idx[cur.beg + n];
				be = sd.e[cur.beg + n];
			} else if (wn < units_total) {
				first_block(nxt);              /* first block of the next unit */
			}
			for (int s0 = 0; s0 < mb; s0 += ACC_SB) {
				const int m = min(ACC_SB, mb - s0);
				const double *rp[ACC_SB];
				double e[ACC_SB];
				acc_prep(s_idx[warp], s_e[warp], s0, srcb, ld, rp, e);
				double x[GW][ACC_SB];
				AccGroups<GW, 0>::load(x, rp, ok_last);
				AccGroups<GW, 0>::add(acc, x, e, m);
			}
		}
#pragma unroll
		for (int q = 0; q < GW; q++) {
			const int k = k0 + q * 32;
			if (k < F)
				chain_store(sd, cur.own, ld, k, acc[q]);
		}
		w = wn;
		cur = nxt;
	}
}

template <int GW>
__global__ void __launch_bounds__(ACC_WARPS * 32, GW <= 2 ? 8 : GW <= 4 ? 6 : ACC_CTAS)
k_acc(const ChainSide A, const ChainSide B, long units_b, long units_total, int F, int ld,
      unsigned long long *ticket)
{
	acc_body<GW>(A, B, units_b, units_total, F, ld, ticket);
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
