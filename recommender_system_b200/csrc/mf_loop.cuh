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
 *               its nonzeros are brought to shared memory by one cp.async.bulk
 *               each (mbarrier transaction count), e_n is computed with one
 *               lane per nonzero walking k sequentially (the dot of
 *               mat2d.c:133-137), then L_new[i][k] is accumulated with lanes
 *               over features walking the nonzeros sequentially.  Rows longer
 *               than a threshold get a whole CTA: all warps compute dots, then
 *               the sequential accumulation streams the R rows through the
 *               shared-memory ring below.  Writes e_n at its by-item position.
 *   k_update_R  by-item pass.  A CTA owns (item, block of <= 256 features),
 *               thread = feature; the L rows of the item's nonzeros stream
 *               through a ring of bulk copies kept several stages ahead of the
 *               strictly sequential accumulation (file order of the column).
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

constexpr int RING_MAX_STAGES = 8;
constexpr int BAR_BYTES = 128;           /* barrier block at the start of dynamic shared memory */

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

/*
 * acc = ((acc + ev[0]*src[idx[0]][k]) + ev[1]*src[idx[1]][k]) + ...   for n < m, in order.
 *
 * Called by every thread of a CTA of TPB threads; thread `tid` owns feature
 * `tid` of the `seg` doubles copied from each row (src already points at the
 * first of them).  Rows move global -> shared in stages of RB rows, one
 * cp.async.bulk per row issued by the lanes of warp 0; a stage is reused only
 * after the whole CTA passed the barrier that ends its consumption.  `sbase`
 * counts the stages ever pushed through these barriers (phase bookkeeping
 * across calls).
 */
template <int TPB>
__device__ __forceinline__ double chain_pipeline(double acc, bool active, int tid,
                                                 const int *__restrict__ idx,
                                                 const double *__restrict__ ev, int m,
                                                 const double *__restrict__ src, int ld, int row_off,
                                                 int seg, double *ring, double *es, uint64_t *bars,
                                                 int RB, int NST, uint32_t &sbase)
{
	const int n_stages = (m + RB - 1) / RB;
	const uint32_t seg_bytes = (uint32_t)seg * 8u;
	auto issue = [&](int s) {
		if (tid < 32) {
			const uint32_t g = sbase + (uint32_t)s;
			const int buf = (int)(g % (uint32_t)NST);
			const int n = s * RB + tid;
			if (tid < RB && n < m) {
				const int i = idx[n] - row_off;
				es[buf * RB + tid] = ev[n];
				mbar_arrive_expect_tx(&bars[buf], seg_bytes);
				bulk_g2s(ring + (size_t)(buf * RB + tid) * seg, src + (size_t)i * ld, seg_bytes, &bars[buf]);
			} else {
				mbar_arrive(&bars[buf]);
			}
		}
	};
	for (int s = 0; s < NST && s < n_stages; s++)
		issue(s);
	for (int s = 0; s < n_stages; s++) {
		const uint32_t g = sbase + (uint32_t)s;
		const int buf = (int)(g % (uint32_t)NST);
		mbar_wait(&bars[buf], (g / (uint32_t)NST) & 1u);
		const int rows = min(RB, m - s * RB);
		if (active) {
			const double *row = ring + (size_t)buf * RB * seg + tid;
			const double *e = es + buf * RB;
#pragma unroll 4
			for (int r = 0; r < rows; r++)
				acc = __dadd_rn(acc, __dmul_rn(e[r], row[(size_t)r * seg]));
		}
		if (TPB == 32)
			__syncwarp();
		else
			__syncthreads();
		if (s + NST < n_stages)
			issue(s + NST);
	}
	sbase += (uint32_t)n_stages;
	return acc;
}

/* Lout[k] = ((init[k] + es[0]*Rs[0][k]) + es[1]*Rs[1][k]) + ...  for the NG feature
 * groups k = kb + g*32 + lane of this lane (matFact.c:48-49 order within the row). */
template <int NG>
__device__ __forceinline__ void row_update(const double *init, double *Lout, const double *Rs,
                                           const double *es, int cnt, int S, int kb, int F, int lane)
{
	double acc[NG];
	const bool last_ok = kb + (NG - 1) * 32 + lane < F;   /* only the last group can be partial */
#pragma unroll
	for (int g = 0; g < NG; g++)
		acc[g] = (g < NG - 1 || last_ok) ? init[kb + g * 32 + lane] : 0.0;
	const double *col = Rs + kb + lane;
	for (int n = 0; n < cnt; n++) {
		const double e = es[n];
		const double *x = col + (size_t)n * S;
#pragma unroll
		for (int g = 0; g < NG; g++)
			if (g < NG - 1 || last_ok)
				acc[g] = __dadd_rn(acc[g], __dmul_rn(e, x[g * 32]));
	}
#pragma unroll
	for (int g = 0; g < NG; g++)
		if (g < NG - 1 || last_ok)
			Lout[kb + g * 32 + lane] = acc[g];
}

/*
 * By-user pass.  Dynamic shared memory: [BAR_BYTES of mbarriers][WARPS x (Rs[CH*S] | Ls[S] | es[32])].
 * bars[w] (count 1) tracks warp w's row staging; bars[WARPS + s] (count 32) are the ring
 * stages of the whole-CTA rows, whose ring re-uses the staging area once the dots are done.
 */
template <int WARPS>
__global__ void __launch_bounds__(WARPS * 32)
k_update_L(const LoopArgs p, const double *__restrict__ Lold, double *__restrict__ Lnew,
           const double *__restrict__ Rold)
{
	extern __shared__ __align__(128) unsigned char smem_raw[];
	uint64_t *bars = reinterpret_cast<uint64_t *>(smem_raw);
	double *data = reinterpret_cast<double *>(smem_raw + BAR_BYTES);
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const int F = p.F, ld = p.ld, S = p.S, CH = p.CH;
	const uint32_t row_bytes = (uint32_t)ld * 8u;
	const size_t per_warp = (size_t)CH * S + S + 32;
	double *Rs = data + per_warp * warp;
	double *Ls = Rs + (size_t)CH * S;
	double *es = Ls + S;
	uint64_t *bar = &bars[warp];

	const bool is_long = (int)blockIdx.x < p.n_long;
	const int r = is_long ? (int)blockIdx.x : p.n_long + ((int)blockIdx.x - p.n_long) * WARPS + warp;
	if (lane == 0) {
		mbar_init(bar, 1);
		if (is_long && warp == 0)
			for (int s = 0; s < p.nst_long; s++)
				mbar_init(&bars[WARPS + s], 32);
		mbar_fence_init();
	}
	if (is_long)
		__syncthreads();
	else
		__syncwarp();
	if (r >= p.n_rows)
		return;                       /* whole warp leaves; short-path blocks never use a block barrier */
	const int i = p.row_id[r] - p.user_lo;
	const int base = p.row_ptr[r];
	const int m = p.row_ptr[r + 1] - base;
	const double *Lrow = Lold + (size_t)i * ld;
	double *Lout = Lnew + (size_t)i * ld;

	uint32_t phase = 0;
	bool first = true;
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
		/* stage: one bulk copy per R row (+ the user's own L row the first time) */
		if (lane == 0) {
			mbar_arrive_expect_tx(bar, row_bytes * (uint32_t)(cnt + (first ? 1 : 0)));
			if (first)
				bulk_g2s(Ls, Lrow, row_bytes, bar);
		}
		if (lane < cnt)
			bulk_g2s(Rs + (size_t)lane * S, Rold + (size_t)j * ld, row_bytes, bar);
		mbar_wait(bar, phase);
		phase ^= 1u;
		first = false;

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
			/* L_new[i][k] chain over this chunk's nonzeros; the feature groups of a lane
			 * are independent chains and run interleaved */
			for (int kb = 0; kb < F; kb += 256) {
				const double *init = (c0 == 0) ? Ls : Lout;
				switch (min(8, (F - kb + 31) >> 5)) {
				case 1: row_update<1>(init, Lout, Rs, es, cnt, S, kb, F, lane); break;
				case 2: row_update<2>(init, Lout, Rs, es, cnt, S, kb, F, lane); break;
				case 3: row_update<3>(init, Lout, Rs, es, cnt, S, kb, F, lane); break;
				case 4: row_update<4>(init, Lout, Rs, es, cnt, S, kb, F, lane); break;
				case 5: row_update<5>(init, Lout, Rs, es, cnt, S, kb, F, lane); break;
				case 6: row_update<6>(init, Lout, Rs, es, cnt, S, kb, F, lane); break;
				case 7: row_update<7>(init, Lout, Rs, es, cnt, S, kb, F, lane); break;
				default: row_update<8>(init, Lout, Rs, es, cnt, S, kb, F, lane); break;
				}
			}
		}
		__syncwarp();   /* everyone is done with Rs / es before the next chunk overwrites them */
	}
	if (!is_long)
		return;

	/* whole-CTA row: every e_n is in e_csr; thread = feature, R rows stream through the ring */
	__syncthreads();
	uint32_t sbase = 0;
	const int RB = p.rb_long, NST = p.nst_long;
	double *ring = data;
	for (int fb = 0; fb < F; fb += WARPS * 32) {
		const int seg = min(WARPS * 32, ld - fb);
		double *ring_es = ring + (size_t)NST * RB * seg;
		const int k = fb + (int)threadIdx.x;
		const bool active = k < F;
		double acc = active ? Lrow[k] : 0.0;
		acc = chain_pipeline<WARPS * 32>(acc, active, (int)threadIdx.x, p.col_idx + base, p.e_csr + base, m,
		                                 Rold + fb, ld, 0, seg, ring, ring_es, &bars[WARPS], RB, NST, sbase);
		if (active)
			Lout[k] = acc;
	}
}

/*
 * By-item pass.  CTA = (stored column c, feature block fbi); thread = feature.
 * Dynamic shared memory: [BAR_BYTES of mbarriers][ring NST*RB rows of seg doubles][es NST*RB].
 * single-GPU: Rout = R_new, starts from R_old[j][k]        (matFact.c:50-51)
 * N shards:   Rout = G (rows by item_slot), starts from 0  (matFact-mpi.c:187-205)
 */
template <int TPB>
__global__ void __launch_bounds__(TPB)
k_update_R(const LoopArgs p, const double *__restrict__ Lold, const double *__restrict__ Rold,
           double *__restrict__ Rout)
{
	extern __shared__ __align__(128) unsigned char smem_raw[];
	uint64_t *bars = reinterpret_cast<uint64_t *>(smem_raw);
	double *ring = reinterpret_cast<double *>(smem_raw + BAR_BYTES);
	const int tid = threadIdx.x;
	const int c = (int)(blockIdx.x / (unsigned)p.nfb), fb = (int)(blockIdx.x % (unsigned)p.nfb) * TPB;
	const int RB = p.rb, NST = p.nst;
	if (tid == 0) {
		for (int s = 0; s < NST; s++)
			mbar_init(&bars[s], 32);
		mbar_fence_init();
	}
	if (TPB == 32)
		__syncwarp();
	else
		__syncthreads();
	const int ld = p.ld;
	const int seg = min(TPB, ld - fb);
	double *es = ring + (size_t)NST * RB * seg;
	const int j = p.col_id[c];
	const int b = p.col_ptr[c], m = p.col_ptr[c + 1] - b;
	const int k = fb + tid;
	const bool active = k < p.F;
	double acc = (p.item_slot || !active) ? 0.0 : Rold[(size_t)j * ld + k];
	uint32_t sbase = 0;
	acc = chain_pipeline<TPB>(acc, active, tid, p.row_idx + b, p.e_csc + b, m, Lold + fb, ld, p.user_lo,
	                          seg, ring, es, bars, RB, NST, sbase);
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
