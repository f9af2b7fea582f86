/*
 * csrc/mf_iter.cu -- the only translation unit that instantiates and launches
 * kernels: the iteration loop (matFact.c:36-53), the masked arg-max
 * (matFact.c:55 + :10-27) and the one-shot entries of the C ABI.
 *
 * One iteration on one shard is two launches on the double-buffered factors:
 *     k_update_L(L[c] -> L[1-c], R[c])      writes e_n
 *     k_update_R(L[c], R[c] -> R[1-c])      reads  e_n
 * With N shards the second launch writes the partial gradient G instead and
 * is followed by the exchange:
 *   - one process, N GPUs: every GPU reads all G over peer memory (NVLink) in
 *     k_apply_R_peer and adds them in shard order, so all replicas of R are
 *     bit-identical and equal to the N-shard ordering of the oracle;
 *   - one process per GPU (rank mode): ncclAllReduce(G) + k_apply_R.
 * Sequences of iterations are captured in a CUDA graph (launch latency is the
 * first bound on the small inputs: one iteration of instML100k is ~10 us).
 */
#include "mf_internal.h"
#include "mf_loop.cuh"
#include "mf_argmax.cuh"

#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <thread>
#include <vector>

namespace mf {

struct PeerG {
	const double *g[MF_MAX_SHARDS];
	int n;
};

/* one process, N GPUs: R_new = R_old + ((G_0 + G_1) + ...), G_h read over NVLink */
__global__ void __launch_bounds__(256)
k_apply_R_peer(int n_touched, int F, int ld, const int *__restrict__ touched_item,
               const double *__restrict__ Rold, const PeerG pg, double *__restrict__ Rnew)
{
	const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
	const long total = (long)n_touched * F;
	if (t >= total)
		return;
	const int s = (int)(t / F), k = (int)(t % F);
	const size_t gi = (size_t)s * ld + k;
	double sum = pg.g[0][gi];
	for (int h = 1; h < pg.n; h++)
		sum = __dadd_rn(sum, pg.g[h][gi]);
	const size_t o = (size_t)touched_item[s] * ld + k;
	Rnew[o] = __dadd_rn(Rold[o], sum);
}

/* ---- rank mode over peer memory: the same reduce + apply, synchronised by flags ----------
 * Sync block of a rank (unsigned long long[8], mapped by every other rank):
 *   done[b]      how many times the rank has finished writing its G[b]
 *   consumed[b]  how many times the rank has reduced the G[b] of everybody (local use)
 * k_xchg_signal runs after the kernels that write G[b]; k_apply_R_xchg waits until every
 * rank's done[b] has reached consumed[b] + 1, then reads all G[b] over NVLink and adds them
 * in rank order (bit-identical on every rank, and to the oracle's N-shard order); its last
 * CTA closes the round (consumed[b] += 1).  G[b] is rewritten two iterations later, after this
 * rank has seen done[1-b] of the next iteration from everybody, which they signal only
 * after their own reduce of G[b] -- nobody can still be reading it. */
struct PeerSync {
	const unsigned long long *sync[MF_MAX_SHARDS];
	int n;
};

/* wait until the word another rank writes has reached `want`; bounded by time (mf_async.cuh) */
__device__ __forceinline__ void spin_until(const volatile unsigned long long *w, unsigned long long want,
                                           const WaitCtl &ctl, int who)
{
	unsigned spins = 0;
	unsigned long long t0 = 0;
	while (*w < want)
		if (!wait_more(ctl, spins, t0, who))
			return;
}

__global__ void k_xchg_signal(unsigned long long *done)
{
	__threadfence_system();              /* G[b], written by the kernels before this one */
	*(volatile unsigned long long *)done = *done + 1;
	__threadfence_system();
}

__global__ void __launch_bounds__(256)
k_apply_R_xchg(int n_touched, int F, int ld, const int *__restrict__ touched_item,
               const double *__restrict__ Rold, const PeerG pg, const PeerSync ps, int b,
               unsigned long long *sync_own, double *__restrict__ Rnew, unsigned long long timeout_ns)
{
	/* sync_own: done[0..1], consumed[2..3], CTAs finished [4] */
	__shared__ bool last_cta;
	if (threadIdx.x < ps.n) {
		const unsigned long long want = sync_own[2 + b] + 1;
		spin_until(ps.sync[threadIdx.x] + b, want, WaitCtl{sync_own + MF_SYNC_ERR, timeout_ns}, threadIdx.x);
		__threadfence_system();
	}
	__syncthreads();
	const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
	const long total = (long)n_touched * F;
	if (t < total) {
		const int s = (int)(t / F), k = (int)(t % F);
		const size_t gi = (size_t)s * ld + k;
		double sum = __ldcv(pg.g[0] + gi);
		for (int h = 1; h < pg.n; h++)
			sum = __dadd_rn(sum, __ldcv(pg.g[h] + gi));
		const size_t o = (size_t)touched_item[s] * ld + k;
		Rnew[o] = __dadd_rn(Rold[o], sum);
	}
	/* the last CTA to get here closes the round: every CTA has read consumed[b] by then */
	__syncthreads();
	if (threadIdx.x == 0) {
		__threadfence();
		last_cta = atomicAdd(sync_own + 4, 1ull) == (unsigned long long)gridDim.x - 1;
	}
	__syncthreads();
	if (last_cta && threadIdx.x == 0) {
		sync_own[4] = 0;
		sync_own[2 + b] = sync_own[2 + b] + 1;
	}
}

/* ---- rank mode, push exchange -------------------------------------------------------------
 * The touched items are dealt to the ranks in blocks of `per` slots.  During the iteration the
 * R-side chains store every finished gradient row into the receive buffer of the rank that
 * owns the item (ChainSide::peer_out, mf_loop.cuh) -- the transfer rides under the chain
 * kernels.  Then, per iteration and rank:
 *   k_push_done     after the chain kernels: tell every rank "my contributions are with you"
 *                   (doing this from the last chain CTA instead was measured: a system-scope
 *                   fence at the end of every chain CTA costs 14 us per iteration at two ranks)
 *   k_reduce_bcast  wait for that word from every rank; for the own block of items add the N
 *                   contributions in rank order (bit-identical everywhere and to the oracle's
 *                   N-shard order), add R_old, store the new rows into every replica of R_new;
 *                   the last CTA tells every rank "my rows of R_new are with you" and waits for
 *                   that word from every rank: R_new is then complete in this replica
 * The receive buffer needs no double buffering: a rank stores into it again only after it has
 * seen everybody's `ready` of the previous iteration, which follows everybody's reduce.  The
 * words are iteration counts (never reset), each rank keeps its own count in `epoch`. */
struct PushPeers {
	double *Rnew[MF_MAX_SHARDS];
	unsigned long long *sync[MF_MAX_SHARDS];
	int n, me;
	unsigned long long timeout_ns;
};

__global__ void k_push_done(const PushPeers pp, const unsigned long long *sync_own)
{
	if ((int)threadIdx.x < pp.n) {
		__threadfence_system();          /* the rows stored by the kernels before this one */
		*(volatile unsigned long long *)(pp.sync[threadIdx.x] + MF_SYNC_DONE + pp.me) = sync_own[MF_SYNC_EPOCH] + 1;
	}
}

__global__ void __launch_bounds__(256)
k_reduce_bcast(const PushPeers pp, int per, int cnt, int ld, const int *__restrict__ own_item,
               const double *__restrict__ Rold, const double *recv, unsigned long long *sync_own)
{
	__shared__ bool last_cta;
	const unsigned long long want = sync_own[MF_SYNC_EPOCH] + 1;
	if ((int)threadIdx.x < pp.n) {
		spin_until(sync_own + MF_SYNC_DONE + threadIdx.x, want, WaitCtl{sync_own + MF_SYNC_ERR, pp.timeout_ns}, threadIdx.x);
		__threadfence_system();
	}
	__syncthreads();
	const int ld2 = ld >> 1;
	const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
	if (t < (long)cnt * ld2) {
		const int q = (int)(t / ld2), kp = (int)(t % ld2);
		const double *g = recv + (size_t)q * ld + 2 * kp;
		const size_t blk = (size_t)per * ld;
		double2 sum = __ldcv(reinterpret_cast<const double2 *>(g));
		for (int h = 1; h < pp.n; h++) {
			const double2 x = __ldcv(reinterpret_cast<const double2 *>(g + h * blk));
			sum.x = __dadd_rn(sum.x, x.x);
			sum.y = __dadd_rn(sum.y, x.y);
		}
		const size_t o = (size_t)own_item[q] * ld + 2 * kp;
		const double2 r = *reinterpret_cast<const double2 *>(Rold + o);
		double2 v;
		v.x = __dadd_rn(r.x, sum.x);
		v.y = __dadd_rn(r.y, sum.y);
		for (int d = 0; d < pp.n; d++) {         /* start with the neighbour: spread the links */
			int h = pp.me + 1 + d;
			if (h >= pp.n)
				h -= pp.n;
			*reinterpret_cast<double2 *>(pp.Rnew[h] + o) = v;
		}
	}
	/* the last CTA to get here publishes: every CTA has stored its rows and read `epoch` */
	__syncthreads();
	if (threadIdx.x == 0) {
		__threadfence_system();
		last_cta = atomicAdd(sync_own + MF_SYNC_CTAS, 1ull) == (unsigned long long)gridDim.x - 1;
	}
	__syncthreads();
	if (last_cta) {
		if ((int)threadIdx.x < pp.n) {
			__threadfence_system();
			*(volatile unsigned long long *)(pp.sync[threadIdx.x] + MF_SYNC_READY + pp.me) = want;
			/* ... and stays until everybody's rows are here: when this launch has finished,
			 * R_new is complete in this replica */
			spin_until(sync_own + MF_SYNC_READY + threadIdx.x, want, WaitCtl{sync_own + MF_SYNC_ERR, pp.timeout_ns}, threadIdx.x);
			__threadfence_system();
		}
		if (threadIdx.x == 0) {
			sync_own[MF_SYNC_CTAS] = 0;
			sync_own[MF_SYNC_EPOCH] = want;
		}
	}
}

/* ---- low-latency exchange (N shards: one process per GPU, or one process driving N GPUs) -----
 * The same data flow as the push exchange, but every element travels as a (value, tag) pair in
 * one 16-byte store (st_pair / ld_pair_when, mf_async.cuh), tag = number of the iteration, and
 * the readers poll the element itself: no fence, no flag word, no launch that only signals.
 *   chain kernels   store (gradient value, tag) into the receive buffer of the reducing rank
 *   k_xchg_ll       reduce part, one thread per element of the own block of items: take the
 *                   contributions in rank order as they arrive (ranks without a nonzero in
 *                   the item contribute +0.0 and are not waited for: `contrib`), add R_old, store
 *                   R_new here and as (value, tag) into every other rank;
 *                   unpack part, one thread per element of the other ranks' blocks: take the pair
 *                   when it arrives and store the value into R_new.
 * Progress does not depend on the order in which the hardware starts CTAs: the grid is never
 * larger than what is resident at once and every CTA does its share of the reduce part before
 * its share of the unpack part.  The reduce part waits only for chain kernels of other ranks
 * (earlier in their streams), the unpack part only for reduce parts: no cycle.
 * The pair buffers are double (ChainSide::ll_half): iteration t uses half t & 1.  A rank can
 * be at most one iteration ahead of another (it needs everybody's rows of R_new(t) before its
 * chain kernels of t + 1 can start), so a half is rewritten only after everyone has read it --
 * a rank that has no nonzero in an item and therefore is not waited for included. */
struct LLArgs {
	double *stage[MF_MAX_SHARDS];     /* pair buffer for R_new rows of every rank (own: unused) */
	const double *recv;               /* own pair buffer for contributions: [rank][slot in block][ld] */
	const double *stage_own;
	const double *Rold;
	double *Rnew;
	const int *touched_item, *contrib;
	unsigned long long *sync_own;
	unsigned long long timeout_ns;
	long half;                        /* doubles per parity half of a pair buffer */
	int n, me, per, n_touched, lo, cnt, F, ld;
};

__global__ void __launch_bounds__(256)
k_xchg_ll(const LLArgs a)
{
	__shared__ bool last_cta;
	const unsigned long long want = a.sync_own[MF_SYNC_EPOCH] + 1;
	const long par = (long)(want & 1) * a.half;
	const WaitCtl ctl{a.sync_own + MF_SYNC_ERR, a.timeout_ns};
	const long stride = (long)gridDim.x * blockDim.x;
	const long t0 = (long)blockIdx.x * blockDim.x + threadIdx.x;
	const size_t blk = (size_t)a.per * a.ld;
	const double *recv = a.recv + par;
	for (long t = t0; t < (long)a.cnt * a.F; t += stride) {
		const int q = (int)(t / a.F), k = (int)(t % a.F);
		const int slot = a.lo + q, who = a.contrib[slot];
		const double *g = recv + 2 * ((size_t)q * a.ld + k);
		double sum = (who & 1) ? ld_pair_when(g, want, ctl, 0) : 0.0;
		for (int h = 1; h < a.n; h++)
			sum = __dadd_rn(sum, ((who >> h) & 1) ? ld_pair_when(g + 2 * h * blk, want, ctl, h) : 0.0);
		const size_t o = (size_t)a.touched_item[slot] * a.ld + k;
		const double v = __dadd_rn(a.Rold[o], sum);
		a.Rnew[o] = v;
		const size_t so = 2 * ((size_t)slot * a.ld + k);
		for (int d = 1; d < a.n; d++) {              /* start with the neighbour: spread the links */
			int h = a.me + d;
			if (h >= a.n)
				h -= a.n;
			st_pair(a.stage[h] + par + so, v, want);
		}
	}
	const double *stage_own = a.stage_own + par;
	for (long u = t0; u < (long)(a.n_touched - a.cnt) * a.F; u += stride) {
		int slot = (int)(u / a.F);
		const int k = (int)(u % a.F);
		if (slot >= a.lo)
			slot += a.cnt;                            /* skip the own block */
		const double v = ld_pair_when(stage_own + 2 * ((size_t)slot * a.ld + k), want, ctl, slot / a.per);
		a.Rnew[(size_t)a.touched_item[slot] * a.ld + k] = v;
	}
	/* the last CTA to get here closes the iteration */
	__syncthreads();
	if (threadIdx.x == 0) {
		__threadfence();
		last_cta = atomicAdd(a.sync_own + MF_SYNC_CTAS, 1ull) == (unsigned long long)gridDim.x - 1;
	}
	__syncthreads();
	if (last_cta && threadIdx.x == 0) {
		a.sync_own[MF_SYNC_CTAS] = 0;
		a.sync_own[MF_SYNC_EPOCH] = want;
	}
}

/* ---- the same exchange inside k_acc (split mode; acc_body / NoTail in mf_loop.cuh) ----------
 * Reduce unit: 32 elements of the own block, a lane each; unpack unit: 128 elements of the other
 * blocks, four per lane.  A lane first looks at up to four pairs at once and only then waits for
 * the ones that were not there: a warp of k_acc is worth ~14 KB of gathers in flight, and it
 * should be back at them quickly.  The arithmetic is k_xchg_ll's, element for element. */
constexpr int LL_RED_ELEMS = 32, LL_UNP_ELEMS = 128;

__device__ __noinline__ void ll_reduce_unit(const LLArgs *ap, long u, int lane)
{
	const LLArgs &a = *ap;
	const unsigned long long want = *(volatile unsigned long long *)(a.sync_own + MF_SYNC_EPOCH) + 1;
	const long par = (long)(want & 1) * a.half;
	const WaitCtl ctl{a.sync_own + MF_SYNC_ERR, a.timeout_ns};
	const long t = u * LL_RED_ELEMS + lane;
	if (t >= (long)a.cnt * a.F)
		return;
	const int q = (int)(t / a.F), k = (int)(t % a.F);
	const int slot = a.lo + q, who = a.contrib[slot];
	const size_t blk = (size_t)a.per * a.ld;
	const double *g = a.recv + par + 2 * ((size_t)q * a.ld + k);
	const size_t o = (size_t)a.touched_item[slot] * a.ld + k;
	const double r_old = a.Rold[o];
	double sum = 0.0;
	for (int h0 = 0; h0 < a.n; h0 += 4) {
		unsigned long long v[4], tg[4];
#pragma unroll
		for (int j = 0; j < 4; j++) {
			const int h = h0 + j;
			tg[j] = 0;                       /* never a tag: the first exchange waits for 1 */
			if (h < a.n && ((who >> h) & 1))
				ld_pair_try(g + 2 * h * blk, v[j], tg[j]);
		}
#pragma unroll
		for (int j = 0; j < 4; j++) {
			const int h = h0 + j;
			if (h >= a.n)
				break;
			double c = 0.0;
			if ((who >> h) & 1)
				c = tg[j] == want ? __longlong_as_double((long long)v[j]) : ld_pair_when(g + 2 * h * blk, want, ctl, h);
			sum = h == 0 ? c : __dadd_rn(sum, c);
		}
	}
	const double val = __dadd_rn(r_old, sum);
	a.Rnew[o] = val;
	const size_t so = 2 * ((size_t)slot * a.ld + k);
	for (int d = 1; d < a.n; d++) {              /* start with the neighbour: spread the links */
		int h = a.me + d;
		if (h >= a.n)
			h -= a.n;
		st_pair(a.stage[h] + par + so, val, want);
	}
}

__device__ __noinline__ void ll_unpack_unit(const LLArgs *ap, long u, int lane)
{
	const LLArgs &a = *ap;
	const unsigned long long want = *(volatile unsigned long long *)(a.sync_own + MF_SYNC_EPOCH) + 1;
	const long par = (long)(want & 1) * a.half;
	const WaitCtl ctl{a.sync_own + MF_SYNC_ERR, a.timeout_ns};
	const long total = (long)(a.n_touched - a.cnt) * a.F;
	const double *stage_own = a.stage_own + par;
	unsigned long long v[4], tg[4];
	int slot[4], k[4];
#pragma unroll
	for (int j = 0; j < 4; j++) {
		const long e = u * LL_UNP_ELEMS + j * 32 + lane;
		tg[j] = 0;
		slot[j] = -1;
		if (e < total) {
			slot[j] = (int)(e / a.F);
			k[j] = (int)(e % a.F);
			if (slot[j] >= a.lo)
				slot[j] += a.cnt;                     /* skip the own block */
			ld_pair_try(stage_own + 2 * ((size_t)slot[j] * a.ld + k[j]), v[j], tg[j]);
		}
	}
#pragma unroll
	for (int j = 0; j < 4; j++) {
		if (slot[j] < 0)
			continue;
		const double val = tg[j] == want ? __longlong_as_double((long long)v[j])
		                                 : ld_pair_when(stage_own + 2 * ((size_t)slot[j] * a.ld + k[j]), want, ctl, slot[j] / a.per);
		a.Rnew[(size_t)a.touched_item[slot[j]] * a.ld + k[j]] = val;
	}
}

/* k_acc with the exchange in it (N shards, split mode), in four phases with a ticket counter each:
 *   the columns | reduce units | the rows | unpack units
 * The own block of items is reduced as soon as the columns are through and travels while the rows
 * are being accumulated; the other blocks are taken at the end, when they have long arrived.
 * A warp takes reduce units only when it has no column in hand, so a column that somebody waits
 * for is always with a warp that is working on it, whatever the hardware's order of starting
 * CTAs: no cycle, as in k_xchg_ll.  The last CTA to finish closes the iteration. */
template <int GW>
__global__ void __launch_bounds__(ACC_WARPS * 32, GW <= 2 ? 8 : GW <= 4 ? 6 : ACC_CTAS)
k_acc_ll(const ChainSide A, const ChainSide B, long units_b, long units_a, int F, int ld,
         unsigned long long *ticket, const LLArgs *xa, long r_units, long u_units)
{
	const int lane = threadIdx.x & 31;
	auto take = [&](unsigned long long *t) {
		unsigned long long v = 0;
		if (lane == 0)
			v = atomicAdd(t, 1ull);
		return (long)__shfl_sync(FULL, v, 0);
	};
	acc_body<GW>(A, B, units_b, units_b, F, ld, ticket);
	for (long u = take(ticket + 1); u < r_units; u = take(ticket + 1))
		ll_reduce_unit(xa, u, lane);
	acc_body<GW>(A, B, 0, units_a, F, ld, ticket + 2);
	for (long u = take(ticket + 3); u < u_units; u = take(ticket + 3))
		ll_unpack_unit(xa, u, lane);
	__syncthreads();
	if (threadIdx.x == 0) {
		__threadfence();
		volatile unsigned long long *sync = xa->sync_own;
		const unsigned long long want = sync[MF_SYNC_EPOCH] + 1;
		if (atomicAdd(xa->sync_own + MF_SYNC_CTAS, 1ull) == (unsigned long long)gridDim.x - 1) {
			sync[MF_SYNC_CTAS] = 0;
			sync[MF_SYNC_EPOCH] = want;
		}
	}
}

/* profiling only: this rank's contributions again, with the tag of the next exchange and no
 * chain kernels before it (the values do not matter: the exchange writes the output buffer) */
__global__ void __launch_bounds__(256)
k_ll_repush(const LLArgs a, const PeerG recv_of)
{
	const unsigned long long want = a.sync_own[MF_SYNC_EPOCH] + 1;
	const long par = (long)(want & 1) * a.half;
	const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= (long)a.n_touched * a.F)
		return;
	const int slot = (int)(t / a.F), k = (int)(t % a.F);
	if (!((a.contrib[slot] >> a.me) & 1))
		return;
	const size_t row = (size_t)a.me * a.per + slot % a.per;
	st_pair(const_cast<double *>(recv_of.g[slot / a.per]) + par + 2 * (row * a.ld + k), 0.0, want);
}

} // namespace mf

template <int NG>
static int prepare_res(size_t smem)
{
	CK(cudaFuncSetAttribute(mf::k_update_L_res<NG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
	/* the resident rows decide how many warps an SM holds: take all the shared memory there is */
	CK(cudaFuncSetAttribute(mf::k_update_L_res<NG>, cudaFuncAttributePreferredSharedMemoryCarveout,
	                        (int)cudaSharedmemCarveoutMaxShared));
	return MFGPU_OK;
}

/* k_dot_units<W, NT> for the configured (W, NT): its function pointer */
typedef void (*dot_units_fn)(const mf::LoopArgs, const int *, int, const double *, const double *);
static dot_units_fn dot_units_kernel(int w, int nt)
{
#define MF_ROW(W) \
	case W: switch (nt) { case 2: return mf::k_dot_units<W, 2>; case 3: return mf::k_dot_units<W, 3>; \
	                      case 4: return mf::k_dot_units<W, 4>; case 5: return mf::k_dot_units<W, 5>; \
	                      default: return mf::k_dot_units<W, 6>; }
	switch (w) {
		MF_ROW(4) MF_ROW(5) MF_ROW(6)
	default:
		switch (nt) { case 2: return mf::k_dot_units<8, 2>; case 3: return mf::k_dot_units<8, 3>;
		              case 4: return mf::k_dot_units<8, 4>; case 5: return mf::k_dot_units<8, 5>;
		              default: return mf::k_dot_units<8, 6>; }
	}
#undef MF_ROW
}

/* k_acc<GW>: its function pointer */
typedef void (*acc_fn)(const mf::ChainSide, const mf::ChainSide, long, long, int, int, unsigned long long *);
static acc_fn acc_kernel(int gw)
{
	switch (gw) {
	case 1: return mf::k_acc<1>;
	case 2: return mf::k_acc<2>;
	case 3: return mf::k_acc<3>;
	case 4: return mf::k_acc<4>;
	case 5: return mf::k_acc<5>;
	case 6: return mf::k_acc<6>;
	case 7: return mf::k_acc<7>;
	default: return mf::k_acc<8>;
	}
}

/* With CUDA's lazy module loading a kernel's code is loaded at its first launch (or capture):
 * milliseconds per kernel, which a program that runs once (the CLI) would pay inside its loop.
 * Asking for the attributes of the kernels a plan is going to use loads them at plan creation. */
int mf_kernels_preload(int F, int split, int nshards)
{
	cudaFuncAttributes fa;
	const int gw = std::min((F + 31) / 32, 8);
	if (split)
		CK(cudaFuncGetAttributes(&fa, acc_kernel(gw)));
	CK(cudaFuncGetAttributes(&fa, mf::k_chain_ws));
	CK(cudaFuncGetAttributes(&fa, mf::k_chain<8, mf::CHAIN_BULK_RING>));
	CK(cudaFuncGetAttributes(&fa, mf::k_masked_argmax));
	CK(cudaFuncGetAttributes(&fa, mf::k_argmax_final));
	if (nshards > 1) {
		CK(cudaFuncGetAttributes(&fa, mf::k_xchg_ll));
		CK(cudaFuncGetAttributes(&fa, mf::k_argmax_block));
	}
	return MFGPU_OK;
}

int mf_kernels_prepare(int warps, size_t k1_smem, int res_ng, size_t res_smem, size_t pair_smem, int pair_warps, int pair_nt)
{
	const auto attr = cudaFuncAttributeMaxDynamicSharedMemorySize;
	CK(cudaFuncSetAttribute(mf::k_chain_ws, attr, mf::WS_SMEM));
	switch (res_ng) {
	case 1: TRY(prepare_res<1>(res_smem)); break;
	case 2: TRY(prepare_res<2>(res_smem)); break;
	case 3: TRY(prepare_res<3>(res_smem)); break;
	case 4: TRY(prepare_res<4>(res_smem)); break;
	case 5: TRY(prepare_res<5>(res_smem)); break;
	case 6: TRY(prepare_res<6>(res_smem)); break;
	case 7: TRY(prepare_res<7>(res_smem)); break;
	case 8: TRY(prepare_res<8>(res_smem)); break;
	default: break;
	}
	if (pair_smem) {
		CK(cudaFuncSetAttribute(dot_units_kernel(pair_warps, pair_nt), attr, (int)pair_smem));
		CK(cudaFuncSetAttribute(dot_units_kernel(pair_warps, pair_nt), cudaFuncAttributePreferredSharedMemoryCarveout,
		                        (int)cudaSharedmemCarveoutMaxShared));
	}
	if (warps == 8) {
		CK(cudaFuncSetAttribute(mf::k_update_L<8, 1, 2>, attr, (int)k1_smem));
		CK(cudaFuncSetAttribute(mf::k_update_L<8, 3, 2>, attr, (int)k1_smem));
		CK(cudaFuncSetAttribute(mf::k_update_L<8, 1, 4>, attr, (int)k1_smem));
		CK(cudaFuncSetAttribute(mf::k_update_L<8, 3, 4>, attr, (int)k1_smem));
		CK(cudaFuncSetAttribute(mf::k_dot<8, 2>, attr, (int)k1_smem));
		CK(cudaFuncSetAttribute(mf::k_dot<8, 4>, attr, (int)k1_smem));
		CK(cudaFuncSetAttribute(mf::k_chain<8, mf::CHAIN_DEEP_RING>, attr,
		                        8 * mf::ring_doubles(mf::RING_DEEP_ROWS, mf::RING_DEEP_STAGES) * (int)sizeof(double) + 128));
		CK(cudaFuncSetAttribute(mf::k_chain<8, mf::CHAIN_BULK_RING>, attr,
		                        8 * mf::ring_doubles(mf::RING_BULK_ROWS, mf::RING_BULK_STAGES) * (int)sizeof(double) + 128));
	} else {
		CK(cudaFuncSetAttribute(mf::k_update_L<1, 1, 2>, attr, (int)k1_smem));
	}
	return MFGPU_OK;
}

static int local_shards(const mfgpu_plan *pl) { return pl->comm ? 1 : pl->nshards; }

/* launches issued so far (mfgpu_plan_launch_count); launches recorded into a graph under capture
 * are not counted here but when the graph is launched (several threads may be capturing) */
static thread_local bool t_capturing = false;
static inline void count_launches(mfgpu_plan *pl, int n = 1)
{
	if (!t_capturing)
		pl->launches += n;
}

/* ---- launches ------------------------------------------------------------------ */

static mf::ChainSide side_L(const mfgpu_plan *, const mf_shard *s, int c, int begin, int end)
{
	mf::ChainSide d;
	d.begin = begin; d.end = end;
	d.ptr = s->d_row_ptr; d.idx = s->d_col_idx; d.id = s->d_row_id;
	d.e = s->d_e_csr;
	d.init = s->L[c]; d.src = s->R[c]; d.out = s->L[1 - c];
	d.own_off = s->user_lo; d.src_off = 0;
	d.out_slot = nullptr;
	d.out_owner = nullptr;
	d.ll_epoch = nullptr;
	d.ll_half = 0;
	return d;
}

static mf::ChainSide side_R(const mfgpu_plan *pl, const mf_shard *s, int c, int begin, int end, double *rout)
{
	mf::ChainSide d;
	d.begin = begin; d.end = end;
	d.ptr = s->d_col_ptr; d.idx = s->d_row_idx; d.id = s->d_col_id;
	d.e = s->d_e_csc;
	d.init = pl->nshards > 1 ? nullptr : s->R[c];    /* N shards: the gradient G starts from 0.0 */
	d.src = s->L[c]; d.out = rout;
	d.own_off = 0; d.src_off = s->user_lo;
	d.out_owner = nullptr;
	d.out_slot = pl->nshards > 1 ? s->d_item_slot : nullptr;
	d.ll_epoch = nullptr;
	d.ll_half = 0;
	if (pl->xchg_ipc >= 2 && rout == nullptr) {
		/* push / low-latency exchange: rows go to the reducing rank's receive buffer (its G[0]) */
		d.out_slot = s->d_push_row;
		d.out_owner = s->d_push_owner;
		for (int h = 0; h < pl->nshards; h++)
			d.peer_out[h] = pl->peer_G[0][h];
		if (pl->xchg_ipc == 3) {
			d.ll_epoch = s->d_sync + MF_SYNC_EPOCH;
			d.ll_half = pl->ll_half;
		}
	}
	return d;
}

enum { K_FUSED = 1, K_DOT = 2, K_BULK = 4, K_DEEP = 8, K_ALL = 15 };

static int launch_fused(mfgpu_plan *pl, mf_shard *s, int c, cudaStream_t st)
{
	if (s->n_units > 0) {
		/* persistent warps; as many CTAs as fit an SM (at most 16 warps: 128 registers) */
		int n_sm = 148;
		CK(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, s->dev));
		const int w = s->pair_warps;
		const int per_sm = std::max(1, std::min((w == 4 ? 4 * MF_DOT_MINB : 16) / w, (int)(227 * 1024 / (s->pair_smem + 1024))));
		const unsigned blocks = (unsigned)std::min((s->n_units + w - 1) / w, n_sm * per_sm);
		dot_units_kernel(w, s->pair_nt)<<<blocks, 32 * w, s->pair_smem, st>>>(s->la, s->d_units, s->n_units, s->L[c], s->R[c]);
		CK(cudaGetLastError());
		count_launches(pl);
	}
	if (s->k1_blocks <= 0)
		return MFGPU_OK;
#define MF_RES(NG) \
	case NG: mf::k_update_L_res<NG><<<s->k1_blocks, 32 * mf::RES_WARPS, s->res_smem, st>>>( \
		s->la, s->L[c], s->L[1 - c], s->R[c]); break;
	if (s->res_ng) {
		switch (s->res_ng) {
			MF_RES(1) MF_RES(2) MF_RES(3) MF_RES(4) MF_RES(5) MF_RES(6) MF_RES(7) MF_RES(8)
		default: return mf_fail(MFGPU_EINVAL, "resident kernel: bad group count %d", s->res_ng);
		}
	} else if (s->warps == 8) {
		const bool deep_acc = s->acc_rows > 1, deep_dot = s->la.NT > 2;
		mf::LoopArgs la = s->la;
		la.n_rows = s->n_mid;             /* the shorter rows went to the paired launch above */
#define MF_UPD(P, NT) mf::k_update_L<8, P, NT><<<s->k1_blocks, 256, s->k1_smem, st>>>(la, s->L[c], s->L[1 - c], s->R[c])
		if (deep_acc && deep_dot) MF_UPD(3, 4);
		else if (deep_acc) MF_UPD(3, 2);
		else if (deep_dot) MF_UPD(1, 4);
		else MF_UPD(1, 2);
#undef MF_UPD
	} else
		mf::k_update_L<1, 1, 2><<<s->k1_blocks, 32, s->k1_smem, st>>>(s->la, s->L[c], s->L[1 - c], s->R[c]);
#undef MF_RES
	CK(cudaGetLastError());
	count_launches(pl);
	return MFGPU_OK;
}

static int launch_dot(mfgpu_plan *pl, mf_shard *s, int c, cudaStream_t st)
{
	if (s->dot_blocks <= 0)
		return MFGPU_OK;
	if (s->la.NT > 2)
		mf::k_dot<8, 4><<<s->dot_blocks, 256, s->k1_smem, st>>>(s->la, s->L[c], s->R[c]);
	else
		mf::k_dot<8, 2><<<s->dot_blocks, 256, s->k1_smem, st>>>(s->la, s->L[c], s->R[c]);
	CK(cudaGetLastError());
	count_launches(pl);
	return MFGPU_OK;
}

static bool has_deep(const mf_shard *s) { return s->la.n_deep_rows > 0 || s->la.n_deep_cols > 0; }

static mf::LLArgs ll_args(const mfgpu_plan *pl, const mf_shard *s, int c);

/* N shards, low-latency exchange, split mode: the exchange goes into k_acc (k_acc_ll) -- unless
 * some of the shard's chains run in another launch beside it (k_chain_ws for the deepest): a
 * reduce unit would wait for their contributions, and nothing promises that the hardware runs
 * the two launches at the same time (two kernels that want different splits of an SM's
 * L1 / shared memory do not share it: measured, the wait bound fired). */
static bool ll_fusable(const mfgpu_plan *pl, const mf_shard *s)
{
	return pl->nshards > 1 && pl->xchg_ipc == 3 && s->split && pl->opt_xchg_fused != 0 &&
	       !(s->dot_blocks > 0 || has_deep(s)) && (long)pl->n_touched * pl->F > 0;
}

template <int GW>
static void launch_acc_ll_gw(const mf::ChainSide &A, const mf::ChainSide &B, long ub, long ut, int F, int ld,
                             unsigned long long *ticket, int n_sm, const mf::LLArgs *xa, long ur, long uu, cudaStream_t st)
{
	const long ctas = std::min<long>((ut + ur + uu + mf::ACC_WARPS - 1) / mf::ACC_WARPS,
	                                 (long)n_sm * (GW <= 2 ? 8 : GW <= 4 ? 6 : mf::ACC_CTAS));
	mf::k_acc_ll<GW><<<(unsigned)ctas, mf::ACC_WARPS * 32, 0, st>>>(A, B, ub, ut - ub, F, ld, ticket, xa, ur, uu);
}

template <int GW>
static void launch_acc_gw(const mf::ChainSide &A, const mf::ChainSide &B, long ub, long ut, int F, int ld,
                          unsigned long long *ticket, int n_sm, cudaStream_t st)
{
	const long ctas = std::min<long>((ut + mf::ACC_WARPS - 1) / mf::ACC_WARPS, (long)n_sm * (GW <= 2 ? 8 : GW <= 4 ? 6 : mf::ACC_CTAS));
	mf::k_acc<GW><<<(unsigned)ctas, mf::ACC_WARPS * 32, 0, st>>>(A, B, ub, ut, F, ld, ticket);
}

/* split mode: every accumulation that is not a deep chain, straight from L2 (k_acc) */
static int launch_acc(mfgpu_plan *pl, mf_shard *s, int c, double *rout, cudaStream_t st)
{
	const mf::LoopArgs &a = s->la;
	const mf::ChainSide A = side_L(pl, s, c, a.n_deep_rows, a.n_rows);
	const mf::ChainSide B = side_R(pl, s, c, a.n_deep_cols, a.n_cols, rout);
	const int ng = (a.F + 31) / 32;
	const int gw = pl->opt_acc_gw > 0 ? std::min(pl->opt_acc_gw, std::min(ng, 8)) : std::min(ng, 8), nb = (ng + gw - 1) / gw;
	const long ub = (long)(B.end - B.begin) * nb, ut = ub + (long)(A.end - A.begin) * nb;
	const bool fuse = rout == nullptr && ll_fusable(pl, s);
	if (ut <= 0 && !fuse)
		return MFGPU_OK;
	int n_sm = 148;
	CK(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, s->dev));
	CK(cudaMemsetAsync(s->d_ticket, 0, (fuse ? 4 : 1) * sizeof(unsigned long long), st));
	if (fuse) {
		/* the exchange's arguments are read from global memory (s->d_ll_args, one copy per factor
		 * parity, written by ll_args_upload before the first iteration): as a kernel parameter
		 * their address stayed in registers across the accumulation loop, which has none to spare */
		if (!s->d_ll_args)
			return mf_fail(MFGPU_EINVAL, "k_acc_ll without its arguments (internal error)");
		const mf::LLArgs *xa = (const mf::LLArgs *)s->d_ll_args + c;
		const mf::LLArgs xh = ll_args(pl, s, c);
		const long ur = ((long)xh.cnt * xh.F + mf::LL_RED_ELEMS - 1) / mf::LL_RED_ELEMS;
		const long uu = ((long)(xh.n_touched - xh.cnt) * xh.F + mf::LL_UNP_ELEMS - 1) / mf::LL_UNP_ELEMS;
#define MF_ACC_LL(G) case G: launch_acc_ll_gw<G>(A, B, ub, std::max(ut, 0L), a.F, a.ld, s->d_ticket, n_sm, xa, ur, uu, st); break;
		switch (gw) {
			MF_ACC_LL(1) MF_ACC_LL(2) MF_ACC_LL(3) MF_ACC_LL(4) MF_ACC_LL(5) MF_ACC_LL(6) MF_ACC_LL(7)
		default: launch_acc_ll_gw<8>(A, B, ub, std::max(ut, 0L), a.F, a.ld, s->d_ticket, n_sm, xa, ur, uu, st); break;
		}
#undef MF_ACC_LL
		CK(cudaGetLastError());
		count_launches(pl);
		s->ll_fused = true;          /* enqueue_iteration: no exchange launch after this one */
		return MFGPU_OK;
	}
	switch (gw) {
	case 1: launch_acc_gw<1>(A, B, ub, ut, a.F, a.ld, s->d_ticket, n_sm, st); break;
	case 2: launch_acc_gw<2>(A, B, ub, ut, a.F, a.ld, s->d_ticket, n_sm, st); break;
	case 3: launch_acc_gw<3>(A, B, ub, ut, a.F, a.ld, s->d_ticket, n_sm, st); break;
	case 4: launch_acc_gw<4>(A, B, ub, ut, a.F, a.ld, s->d_ticket, n_sm, st); break;
	case 5: launch_acc_gw<5>(A, B, ub, ut, a.F, a.ld, s->d_ticket, n_sm, st); break;
	case 6: launch_acc_gw<6>(A, B, ub, ut, a.F, a.ld, s->d_ticket, n_sm, st); break;
	case 7: launch_acc_gw<7>(A, B, ub, ut, a.F, a.ld, s->d_ticket, n_sm, st); break;
	default: launch_acc_gw<8>(A, B, ub, ut, a.F, a.ld, s->d_ticket, n_sm, st); break;
	}
	CK(cudaGetLastError());
	count_launches(pl);
	return MFGPU_OK;
}

static int launch_chain(mfgpu_plan *pl, mf_shard *s, int c, double *rout, bool deep, cudaStream_t st)
{
	if (!deep && s->split)
		return launch_acc(pl, s, c, rout, st);
	const mf::LoopArgs &a = s->la;
	const mf::ChainSide A = deep ? side_L(pl, s, c, 0, a.n_deep_rows) : side_L(pl, s, c, a.n_deep_rows, a.n_long);
	const mf::ChainSide B = deep ? side_R(pl, s, c, 0, a.n_deep_cols, rout)
	                             : side_R(pl, s, c, a.n_deep_cols, a.n_cols, rout);
	const long ng = (a.F + 31) / 32;
	const long wa = (long)(A.end - A.begin) * ng, wb = (long)(B.end - B.begin) * ng;
	const long blocks = (wa + wb + 7) / 8;
	if (blocks <= 0)
		return MFGPU_OK;
	const size_t ring_deep = (size_t)8 * mf::ring_doubles(mf::RING_DEEP_ROWS, mf::RING_DEEP_STAGES) * sizeof(double) + 128;
	const size_t ring_bulk = (size_t)8 * mf::ring_doubles(mf::RING_BULK_ROWS, mf::RING_BULK_STAGES) * sizeof(double) + 128;
	const bool ring = pl->opt_chain_ring != 0;
	if (deep && pl->opt_chain_ws)
		mf::k_chain_ws<<<(unsigned)(wa + wb), mf::WS_THREADS, mf::WS_SMEM, st>>>(A, B, wa, a.F, a.ld);
	else if (deep && ring)
		mf::k_chain<8, mf::CHAIN_DEEP_RING><<<(unsigned)blocks, 256, ring_deep, st>>>(A, B, wa, a.F, a.ld);
	else if (deep)
		mf::k_chain<8, mf::CHAIN_DEEP_REGS><<<(unsigned)blocks, 256, 0, st>>>(A, B, wa, a.F, a.ld);
	else if (ring)
		mf::k_chain<8, mf::CHAIN_BULK_RING><<<(unsigned)blocks, 256, ring_bulk, st>>>(A, B, wa, a.F, a.ld);
	else
		mf::k_chain<8, mf::CHAIN_BULK_REGS><<<(unsigned)blocks, 256, 0, st>>>(A, B, wa, a.F, a.ld);
	CK(cudaGetLastError());
	count_launches(pl);
	return MFGPU_OK;
}

/* The kernels of one iteration on one shard; R-side results go to rout.
 *   st : k_update_L(short rows) ------------\                /-> k_chain<bulk> --\
 *   st2: k_dot(long rows) ------------------- both done ----+--> k_chain<deep> ---+-> st
 * (k_chain needs every e_n: the short rows' from k_update_L, the long rows' from k_dot) */
static int enqueue_loop_kernels(mfgpu_plan *pl, mf_shard *s, int c, double *rout)
{
	const bool second = s->dot_blocks > 0 || has_deep(s);
	if (!second) {
		TRY(launch_fused(pl, s, c, s->st));
		return launch_chain(pl, s, c, rout, false, s->st);
	}
	CK(cudaEventRecord(s->ev_fork, s->st));
	CK(cudaStreamWaitEvent(s->st2, s->ev_fork, 0));
	TRY(launch_fused(pl, s, c, s->st));
	CK(cudaEventRecord(s->ev_a, s->st));
	TRY(launch_dot(pl, s, c, s->st2));
	CK(cudaEventRecord(s->ev_b, s->st2));
	CK(cudaStreamWaitEvent(s->st, s->ev_b, 0));
	TRY(launch_chain(pl, s, c, rout, false, s->st));
	CK(cudaStreamWaitEvent(s->st2, s->ev_a, 0));
	TRY(launch_chain(pl, s, c, rout, true, s->st2));
	CK(cudaEventRecord(s->ev_join, s->st2));
	CK(cudaStreamWaitEvent(s->st, s->ev_join, 0));
	return MFGPU_OK;
}

/* one iteration of a 1-shard plan, or of one rank in rank mode */
/* rank mode over peer memory: signal G[b], reduce everybody's G[b] and apply, close the round */
static int enqueue_xchg(mfgpu_plan *pl, mf_shard *s, int c, int b)
{
	mf::PeerG pg;
	mf::PeerSync ps;
	pg.n = ps.n = pl->nshards;
	for (int h = 0; h < pg.n; h++) {
		pg.g[h] = pl->peer_G[b][h];
		ps.sync[h] = pl->peer_sync[h];
	}
	const long total = (long)pl->n_touched * pl->F;
	mf::k_xchg_signal<<<1, 1, 0, s->st>>>(s->d_sync + b);
	mf::k_apply_R_xchg<<<(unsigned)((total + 255) / 256), 256, 0, s->st>>>(
	    pl->n_touched, pl->F, pl->ld, s->d_touched_item, s->R[c], pg, ps, b, s->d_sync, s->R[1 - c], pl->comm_timeout_ns);
	CK(cudaGetLastError());
	count_launches(pl, 2);
	return MFGPU_OK;
}

/* rank mode, push exchange: publish the contributions, reduce + broadcast the own block of
 * items, wait until every block has arrived (kernels above) */
static int enqueue_push(mfgpu_plan *pl, mf_shard *s, int c)
{
	mf::PushPeers pp;
	pp.n = pl->nshards;
	pp.me = s->me;
	pp.timeout_ns = pl->comm_timeout_ns;
	for (int h = 0; h < pp.n; h++) {
		pp.Rnew[h] = pl->peer_R[1 - c][h];
		pp.sync[h] = pl->peer_sync[h];
	}
	const int per = pl->slots_per_rank;
	const int lo = std::min(pl->n_touched, pp.me * per), cnt = std::min(per, pl->n_touched - lo);
	const long pairs = (long)cnt * (pl->ld / 2);
	mf::k_push_done<<<1, 32, 0, s->st>>>(pp, s->d_sync);
	/* a rank without items of its own still takes part: it publishes `ready` */
	mf::k_reduce_bcast<<<(unsigned)std::max(1L, (pairs + 255) / 256), 256, 0, s->st>>>(
	    pp, per, cnt, pl->ld, s->d_touched_item + lo, s->R[c], s->G[0], s->d_sync);
	CK(cudaGetLastError());
	count_launches(pl, 2);
	return MFGPU_OK;
}

/* rank mode, low-latency exchange: one launch (k_xchg_ll above) */
static mf::LLArgs ll_args(const mfgpu_plan *pl, const mf_shard *s, int c)
{
	mf::LLArgs a;
	a.n = pl->nshards;
	a.me = s->me;
	a.timeout_ns = pl->comm_timeout_ns;
	a.half = pl->ll_half;
	for (int h = 0; h < a.n; h++)
		a.stage[h] = pl->peer_G[1][h];
	a.recv = s->G[0];
	a.stage_own = s->G[1];
	a.Rold = s->R[c];
	a.Rnew = s->R[1 - c];
	a.touched_item = s->d_touched_item;
	a.contrib = s->d_contrib;
	a.sync_own = s->d_sync;
	a.per = pl->slots_per_rank;
	a.n_touched = pl->n_touched;
	a.lo = std::min(pl->n_touched, a.me * a.per);
	a.cnt = std::min(a.per, pl->n_touched - a.lo);
	a.F = pl->F;
	a.ld = pl->ld;
	return a;
}

static int enqueue_ll(mfgpu_plan *pl, mf_shard *s, int c)
{
	const mf::LLArgs a = ll_args(pl, s, c);
	/* never more CTAs than are resident at once (148 SMs x 8 CTAs of 256 threads): see k_xchg_ll */
	const long work = std::max((long)a.cnt, (long)(a.n_touched - a.cnt)) * a.F;
	int n_sm = 148;
	CK(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, s->dev));
	mf::k_xchg_ll<<<(unsigned)std::max(1L, std::min((work + 255) / 256, (long)n_sm * 4)), 256, 0, s->st>>>(a);
	CK(cudaGetLastError());
	count_launches(pl);
	return MFGPU_OK;
}

/* device copies of the exchange's arguments for k_acc_ll, both factor parities (not inside a capture) */
static int ll_args_upload(mfgpu_plan *pl, mf_shard *s)
{
	if (s->d_ll_args || !ll_fusable(pl, s))
		return MFGPU_OK;
	const mf::LLArgs both[2] = {ll_args(pl, s, 0), ll_args(pl, s, 1)};
	CK(cudaMalloc(&s->d_ll_args, sizeof both));
	CK(cudaMemcpy(s->d_ll_args, both, sizeof both, cudaMemcpyHostToDevice));
	return MFGPU_OK;
}

static int enqueue_iteration(mfgpu_plan *pl, mf_shard *s, int c)
{
	if (pl->nshards == 1)
		return enqueue_loop_kernels(pl, s, c, s->R[1 - c]);
	if (pl->xchg_ipc >= 2) {
		s->ll_fused = false;
		TRY(enqueue_loop_kernels(pl, s, c, nullptr));
		if (pl->xchg_ipc == 3)
			return s->ll_fused ? MFGPU_OK : enqueue_ll(pl, s, c);
		return enqueue_push(pl, s, c);
	}
	/* rank mode: the gradient goes to G[c] for the peer-memory exchange (double buffered,
	 * see k_apply_R_xchg), to G[0] for NCCL */
	TRY(enqueue_loop_kernels(pl, s, c, pl->xchg_ipc ? s->G[c] : s->G[0]));
	const size_t count = (size_t)pl->n_touched * pl->ld;
	if (count == 0)
		return MFGPU_OK;
	if (pl->xchg_ipc)
		return enqueue_xchg(pl, s, c, c);
	if (!pl->comm)
		return mf_fail(MFGPU_EINVAL, "N shards without an exchange (internal error)");
	/* out of place: the columns this rank never writes must stay 0.0 in G[0] */
	NCK(g_nccl.AllReduce(s->G[0], s->G[1], count, /*ncclDouble*/ 8, /*ncclSum*/ 0, pl->comm->nccl, s->st));
	const long total = (long)pl->n_touched * pl->F;
	mf::k_apply_R<<<(unsigned)((total + 255) / 256), 256, 0, s->st>>>(pl->n_touched, pl->F, pl->ld,
	                                                                 s->d_touched_item, s->R[c], s->G[1],
	                                                                 s->R[1 - c]);
	CK(cudaGetLastError());
	count_launches(pl);
	return MFGPU_OK;
}

/* One iteration of a one-process N-GPU plan with the event-ordered pull exchange (option
 * exchange = 2; the default is the low-latency exchange through enqueue_iteration).  The G buffer
 * follows the factor parity c, which persists from one mfgpu_plan_iterate call to the next: G[c]
 * is rewritten two iterations after the peers read it, and every shard's stream has by then
 * waited for the events of the iteration in between, whatever the split into calls. */
static int enqueue_iteration_peer(mfgpu_plan *pl, int c)
{
	const int gb = c;
	for (int g = 0; g < pl->nshards; g++) {
		mf_shard *s = &pl->sh[g];
		CK(cudaSetDevice(s->dev));
		TRY(enqueue_loop_kernels(pl, s, c, s->G[gb]));
		CK(cudaEventRecord(s->ev_g[gb], s->st));
	}
	const long total = (long)pl->n_touched * pl->F;
	if (total == 0)
		return MFGPU_OK;
	mf::PeerG pg;
	pg.n = pl->nshards;
	for (int h = 0; h < pl->nshards; h++)
		pg.g[h] = pl->sh[h].G[gb];
	for (int g = 0; g < pl->nshards; g++) {
		mf_shard *s = &pl->sh[g];
		CK(cudaSetDevice(s->dev));
		for (int h = 0; h < pl->nshards; h++)
			if (h != g)
				CK(cudaStreamWaitEvent(s->st, pl->sh[h].ev_g[gb], 0));
		mf::k_apply_R_peer<<<(unsigned)((total + 255) / 256), 256, 0, s->st>>>(
		    pl->n_touched, pl->F, pl->ld, s->d_touched_item, s->R[c], pg, s->R[1 - c]);
		CK(cudaGetLastError());
		count_launches(pl);
	}
	return MFGPU_OK;
}

/* graph of `n` (even) iterations starting at parity c, single shard or rank mode */
static int build_graph(mfgpu_plan *pl, mf_shard *s, int c, int n)
{
	cudaGraph_t graph = nullptr;
	CK(cudaStreamBeginCapture(s->st, cudaStreamCaptureModeThreadLocal));
	int rc = MFGPU_OK;
	t_capturing = true;          /* captured, not launched */
	for (int it = 0; it < n && rc == MFGPU_OK; it++)
		rc = enqueue_iteration(pl, s, (c + it) & 1);
	t_capturing = false;
	cudaError_t e = cudaStreamEndCapture(s->st, &graph);
	if (rc != MFGPU_OK) {
		if (graph) cudaGraphDestroy(graph);
		return rc;
	}
	if (e != cudaSuccess)
		return mf_fail(MFGPU_ECUDA, "cudaStreamEndCapture: %s", cudaGetErrorString(e));
	e = cudaGraphInstantiate(&s->gexec[c], graph, 0);
	cudaGraphDestroy(graph);
	if (e != cudaSuccess)
		return mf_fail(MFGPU_ECUDA, "cudaGraphInstantiate: %s", cudaGetErrorString(e));
	return MFGPU_OK;
}

static long long launches_per_iteration(const mfgpu_plan *pl, const mf_shard *s)
{
	const mf::LoopArgs &a = s->la;
	long long n = (s->k1_blocks > 0) + (s->n_units > 0) + (s->dot_blocks > 0);
	const bool bulk = a.n_long - a.n_deep_rows > 0 || a.n_cols - a.n_deep_cols > 0;
	n += bulk;
	n += (a.n_deep_rows > 0 || a.n_deep_cols > 0);
	if (pl->nshards > 1 && (long)pl->n_touched * pl->F > 0)
		n += pl->xchg_ipc == 3 ? (ll_fusable(pl, s) && bulk ? 0 : 1) : pl->xchg_ipc ? 2 : 1;
	return n;
}

extern "C" int mfgpu_plan_iterate(mfgpu_plan *pl, int iters)
{
	if (!pl || iters < 0)
		return mf_fail(MFGPU_EINVAL, "bad argument");
	if (!pl->have_factors)
		return mf_fail(MFGPU_EINVAL, "mfgpu_plan_set_factors has not been called");
	const int nl = local_shards(pl);
	for (int g = 0; g < nl; g++) {
		CK(cudaSetDevice(pl->sh[g].dev));
		CK(cudaEventRecord(pl->sh[g].ev0, pl->sh[g].st));
		TRY(ll_args_upload(pl, &pl->sh[g]));
	}
	int c = pl->cur;
	if (pl->F == 0 || pl->nnz == 0) {
		/* nothing changes (matFact.c:41 loop is empty / rows have no columns) */
	} else if (nl > 1 && pl->peer_events) {
		for (int t = 0; t < iters; t++, c ^= 1)
			TRY(enqueue_iteration_peer(pl, c));
	} else {
		/* one shard, one rank of a group, or one process driving N GPUs with the low-latency
		 * exchange: every local shard replays its own graph of gi iterations.  The graphs of all
		 * shards are built before any of them is launched: a shard's exchange waits for its peers,
		 * which must not be held up by a graph instantiation on the host. */
		int left = iters;
		/* iterations per graph: the option's (16) -- or, for the first graph of a plan, the even
		 * length between that and twice that which leaves the fewest iterations of this call to be
		 * enqueued one by one (100 iterations: 5 graphs of 20 instead of 6 of 16 + 4 single) */
		int gi = pl->graph_len ? pl->graph_len : pl->opt_graph_iters;
		if (!pl->graph_len && pl->opt_graph) {
			/* short runs get shorter graphs (capture and instantiation cost ~50 us per iteration
			 * and shard; a run of 100 iterations is better off with 10 launches of 10) */
			const int lo = (pl->opt_graph_iters == 16 && iters < 256) ? 8 : pl->opt_graph_iters;
			if (iters >= lo) {
				int best = lo;
				for (int g = lo; g <= 2 * lo && g <= iters; g += 2)
					if (iters % g < iters % best || (iters % g == iters % best && g > best))
						best = g;
				gi = best;
			}
		}
		if (pl->opt_graph && left >= gi) {
			pl->graph_len = gi;
			/* one host thread per shard that still lacks its graph: capture and instantiation
			 * take ~0.3 ms each, and the program that runs once (the CLI) pays them inside its loop */
			TraceClock tc;
			int rcs[MF_MAX_SHARDS] = {0};
			char msgs[MF_MAX_SHARDS][256];
			std::vector<std::thread> th;
			for (int g = 0; g < nl; g++) {
				mf_shard *s = &pl->sh[g];
				if (s->gexec[c])
					continue;
				auto work = [pl, s, c, gi, g, &rcs, &msgs]() {
					rcs[g] = cudaSetDevice(s->dev) == cudaSuccess ? build_graph(pl, s, c, gi)
					                                              : mf_fail(MFGPU_ECUDA, "cudaSetDevice failed");
					snprintf(msgs[g], sizeof msgs[g], "%s", rcs[g] ? mfgpu_last_error() : "");
				};
				if (nl > 1)
					th.emplace_back(work);
				else
					work();
			}
			for (auto &t : th)
				t.join();
			for (int g = 0; g < nl; g++)
				if (rcs[g] != MFGPU_OK)
					return mf_fail(rcs[g], "%s", msgs[g]);
			tc.lap("iterate: graph build");
			for (; left >= gi; left -= gi)
				for (int g = 0; g < nl; g++) {
					mf_shard *s = &pl->sh[g];
					CK(cudaSetDevice(s->dev));
					CK(cudaGraphLaunch(s->gexec[c], s->st));
					pl->launches += launches_per_iteration(pl, s) * gi;
				}
		}
		for (; left > 0; left--, c ^= 1)
			for (int g = 0; g < nl; g++) {
				CK(cudaSetDevice(pl->sh[g].dev));
				TRY(enqueue_iteration(pl, &pl->sh[g], c));
			}
	}
	pl->cur = c;
	for (int g = 0; g < nl; g++) {
		CK(cudaSetDevice(pl->sh[g].dev));
		CK(cudaEventRecord(pl->sh[g].ev1, pl->sh[g].st));
	}
	return MFGPU_OK;
}

/* After a sync: has a device-side wait for a peer run out of time (mf_async.cuh)?  The word is
 * cleared, so that one late peer is reported once. */
int mf_check_comm_error(mfgpu_plan *pl)
{
	if (pl->nshards < 2)
		return MFGPU_OK;
	for (int g = 0; g < local_shards(pl); g++) {
		mf_shard *s = &pl->sh[g];
		if (!s->d_sync)
			continue;
		unsigned long long word = 0;
		CK(cudaSetDevice(s->dev));
		CK(cudaMemcpy(&word, s->d_sync + MF_SYNC_ERR, sizeof word, cudaMemcpyDeviceToHost));
		if (word) {
			CK(cudaMemset(s->d_sync + MF_SYNC_ERR, 0, sizeof word));
			return mf_fail(MFGPU_ECOMM, "shard %d of %d waited more than %.0f ms for shard %d during the exchange "
			                            "of the R gradient (a peer stopped, or ranks issue different calls); "
			                            "the factors are no longer valid",
			               s->me, pl->nshards, pl->comm_timeout_ns / 1e6, (int)word - 1);
		}
	}
	return MFGPU_OK;
}

extern "C" int mfgpu_plan_last_loop_ms(mfgpu_plan *pl, float *ms)
{
	if (!pl || !ms)
		return mf_fail(MFGPU_EINVAL, "null argument");
	float best = 0.f;
	for (int g = 0; g < local_shards(pl); g++) {
		mf_shard *s = &pl->sh[g];
		CK(cudaSetDevice(s->dev));
		CK(cudaEventSynchronize(s->ev1));
		float t = 0.f;
		CK(cudaEventElapsedTime(&t, s->ev0, s->ev1));
		best = std::max(best, t);
	}
	*ms = pl->loop_ms = best;
	return mf_check_comm_error(pl);
}

/* Per-kernel device times for the roofline report of bench.py: each kernel `reps` times back
 * to back on the plan's stream.  ms[0] k_update_L (short rows), [1] k_dot (long rows),
 * [2] k_chain<bulk>, [3] k_chain<deep>, [4] exchange + apply (rank mode). */
extern "C" int mfgpu_plan_profile_kernels(mfgpu_plan *pl, int reps, float ms[5])
{
	if (!pl || !ms || reps < 1)
		return mf_fail(MFGPU_EINVAL, "bad argument");
	if (!pl->have_factors)
		return mf_fail(MFGPU_EINVAL, "mfgpu_plan_set_factors has not been called");
	for (int k = 0; k < 5; k++)
		ms[k] = 0.f;
	if (pl->F == 0 || pl->nnz == 0)
		return MFGPU_OK;
	mf_shard *s = &pl->sh[0];
	const int c = pl->cur;
	const long long before = pl->launches;
	CK(cudaSetDevice(s->dev));
	double *rout = pl->nshards == 1 ? s->R[1 - c] : s->G[0];
	if (pl->xchg_ipc >= 2) {
		/* the chain launches below must not store into the other ranks outside an iteration */
		if (!s->G_scratch)
			CK(cudaMalloc((void **)&s->G_scratch, ((size_t)pl->n_touched * pl->ld + 2) * sizeof(double)));
		rout = s->G_scratch;
	}
	if (pl->xchg_ipc && pl->comm) {
		/* the launches below write G[0] outside the exchange protocol: wait until no peer
		 * can still be reducing it (a one-word all-reduce after everybody's last iteration) */
		CK(cudaStreamSynchronize(s->st));
		NCK(g_nccl.AllReduce((int *)s->d_sync + 14, (int *)s->d_sync + 14, 1, 2, 0, pl->comm->nccl, s->st));
		CK(cudaStreamSynchronize(s->st));
	}
	for (int phase = 0; phase < 5; phase++) {
		if (phase == 4 && !(pl->comm && (size_t)pl->n_touched * pl->ld > 0))
			break;
		CK(cudaEventRecord(s->ev0, s->st));
		for (int r = 0; r < reps; r++) {
			if (phase == 0)
				TRY(launch_fused(pl, s, c, s->st));
			else if (phase == 1)
				TRY(launch_dot(pl, s, c, s->st));
			else if (phase == 2)
				TRY(launch_chain(pl, s, c, rout, false, s->st));
			else if (phase == 3)
				TRY(launch_chain(pl, s, c, rout, true, s->st));
			else if (pl->xchg_ipc == 3) {
				const mf::LLArgs a = ll_args(pl, s, c);
				mf::PeerG recv_of;
				recv_of.n = a.n;
				for (int h = 0; h < a.n; h++)
					recv_of.g[h] = pl->peer_G[0][h];
				mf::k_ll_repush<<<(unsigned)(((long)a.n_touched * a.F + 255) / 256), 256, 0, s->st>>>(a, recv_of);
				TRY(enqueue_ll(pl, s, c));
			} else if (pl->xchg_ipc == 2) {
				TRY(enqueue_push(pl, s, c));
			} else if (pl->xchg_ipc) {
				TRY(enqueue_xchg(pl, s, c, 0));
			} else {
				NCK(g_nccl.AllReduce(s->G[0], s->G[1], (size_t)pl->n_touched * pl->ld, 8, 0,
				                     pl->comm->nccl, s->st));
				const long total = (long)pl->n_touched * pl->F;
				mf::k_apply_R<<<(unsigned)((total + 255) / 256), 256, 0, s->st>>>(
				    pl->n_touched, pl->F, pl->ld, s->d_touched_item, s->R[c], s->G[1], s->R[1 - c]);
				CK(cudaGetLastError());
			}
		}
		CK(cudaEventRecord(s->ev1, s->st));
		CK(cudaEventSynchronize(s->ev1));
		float t = 0.f;
		CK(cudaEventElapsedTime(&t, s->ev0, s->ev1));
		ms[phase] = t / reps;
	}
	pl->launches = before;     /* diagnostic launches are not part of any timed step */
	return MFGPU_OK;
}

/* ---- masked arg-max -------------------------------------------------------------- */

/* host twin of mf::better (mf_argmax.cuh): NaN never wins; ties go to the lower item */
static bool better_host(double v, int j, double bv, int bj)
{
	return (v == v) && (bj < 0 || v > bv || (v == bv && j < bj));
}

/* Item-sharded masked arg-max (one process, N GPUs; few users, many items -- the shape of
 * samples/inst1000-1e6-1000-1-3.in, where the arg-max is the whole run time): GPU g scans ITS
 * block of items for ALL users and the per-user results of the blocks are merged with
 * "(value greater) else (index lower)", the serial tie rule of matFact.c:13-25 -- the reference's
 * own distributed version does the same with MPI_Reduce and max_cmp (matFact-mpi.c:23-28,76-98),
 * except that its strict > keeps the incoming entry, so it is not the serial rule on ties.
 * Every GPU needs all of L for that (users x ld doubles, small by assumption): the rows are
 * copied from their owners over NVLink.  Blocks are whole 64-item tiles. */
static int recommend_by_items(mfgpu_plan *pl, int *best)
{
	const int N = pl->nshards, U = pl->users, c = pl->cur;
	const int n_tiles = (pl->items + mf::AM_TJ - 1) / mf::AM_TJ;
	const int tiles_y = (U + mf::AM_TI - 1) / mf::AM_TI;
	for (int g = 0; g < N; g++) {                 /* everything enqueued so far, L rows included, is done */
		CK(cudaSetDevice(pl->sh[g].dev));
		CK(cudaStreamSynchronize(pl->sh[g].st));
	}
	/* the rated-item mask of all users: the shards hold consecutive user ranges */
	std::vector<int> mptr((size_t)U + 1, 0), mcol, ff((size_t)U, -1);
	for (int g = 0; g < N; g++) {
		const mf_shard *s = &pl->sh[g];
		for (int i = s->user_lo; i < s->user_hi; i++) {
			const int a = s->pat.mask_ptr[i], b = s->pat.mask_ptr[i + 1];
			mcol.insert(mcol.end(), s->pat.mask_col + a, s->pat.mask_col + b);
			mptr[(size_t)i + 1] = (int)mcol.size();
			ff[(size_t)i] = s->pat.first_free[i];
		}
	}
	std::vector<double> hv((size_t)N * U);
	std::vector<int> hj((size_t)N * U), hn((size_t)N * U);
	for (int g = 0; g < N; g++) {
		mf_shard *s = &pl->sh[g];
		CK(cudaSetDevice(s->dev));
		CK(cudaEventRecord(s->ev0, s->st));
		if (!s->am_L) {
			s->am_tile0 = (int)((long)n_tiles * g / N);
			s->am_tile1 = (int)((long)n_tiles * (g + 1) / N);
			const int tiles = s->am_tile1 - s->am_tile0;
			s->am2_splits = std::max(1, std::min(std::max(1, tiles), (148 * 2 * 8 + tiles_y - 1) / std::max(1, tiles_y)));
			CK(cudaMallocAsync((void **)&s->am_L, ((size_t)U * pl->ld + 34) * sizeof(double), s->st));
			CK(cudaMallocAsync((void **)&s->am_mask_ptr, ((size_t)U + 1) * sizeof(int), s->st));
			CK(cudaMallocAsync((void **)&s->am_mask_col, (mcol.size() + 1) * sizeof(int), s->st));
			CK(cudaMallocAsync((void **)&s->am_first_free, ((size_t)U + 1) * sizeof(int), s->st));
			CK(cudaMallocAsync((void **)&s->am_part_v, ((size_t)U * s->am2_splits + 1) * sizeof(double), s->st));
			CK(cudaMallocAsync((void **)&s->am_part_j, ((size_t)U * s->am2_splits + 1) * sizeof(int), s->st));
			CK(cudaMallocAsync((void **)&s->am_nan, ((size_t)U + 1) * sizeof(int), s->st));
			CK(cudaMallocAsync((void **)&s->am_out_v, ((size_t)U + 1) * sizeof(double), s->st));
			CK(cudaMallocAsync((void **)&s->am_out_j, ((size_t)U + 1) * sizeof(int), s->st));
			CK(cudaMemcpyAsync(s->am_mask_ptr, mptr.data(), mptr.size() * sizeof(int), cudaMemcpyHostToDevice, s->st));
			if (!mcol.empty())
				CK(cudaMemcpyAsync(s->am_mask_col, mcol.data(), mcol.size() * sizeof(int), cudaMemcpyHostToDevice, s->st));
			CK(cudaMemcpyAsync(s->am_first_free, ff.data(), ff.size() * sizeof(int), cudaMemcpyHostToDevice, s->st));
		}
		for (int h = 0; h < N; h++) {
			const mf_shard *o = &pl->sh[h];
			const size_t rows = (size_t)(o->user_hi - o->user_lo);
			if (rows)
				CK(cudaMemcpyPeerAsync(s->am_L + (size_t)o->user_lo * pl->ld, s->dev, o->L[c], o->dev,
				                       rows * pl->ld * sizeof(double), s->st));
		}
		CK(cudaMemsetAsync(s->am_nan, 0, (size_t)U * sizeof(int), s->st));
		mf::ArgmaxArgs a;
		a.users_lo = 0;
		a.users_hi = U;
		a.l_row0 = 0;
		a.items = pl->items;
		a.F = pl->F;
		a.ld = pl->ld;
		a.n_splits = s->am2_splits;
		a.tile_begin = s->am_tile0;
		a.tile_end = s->am_tile1;
		a.tiles_per_split = (s->am_tile1 - s->am_tile0 + s->am2_splits - 1) / s->am2_splits;
		a.mask_ptr = s->am_mask_ptr;
		a.mask_col = s->am_mask_col;
		a.first_free = s->am_first_free;
		a.part_v = s->am_part_v;
		a.part_j = s->am_part_j;
		a.nan_first = s->am_nan;
		dim3 grid((unsigned)s->am2_splits, (unsigned)tiles_y);
		mf::k_masked_argmax<<<grid, 256, 0, s->st>>>(a, s->am_L, s->R[c]);
		CK(cudaGetLastError());
		mf::k_argmax_block<<<(U + 255) / 256, 256, 0, s->st>>>(a, s->am_out_v, s->am_out_j);
		CK(cudaGetLastError());
		count_launches(pl, 2);
		CK(cudaMemcpyAsync(hv.data() + (size_t)g * U, s->am_out_v, (size_t)U * sizeof(double), cudaMemcpyDeviceToHost, s->st));
		CK(cudaMemcpyAsync(hj.data() + (size_t)g * U, s->am_out_j, (size_t)U * sizeof(int), cudaMemcpyDeviceToHost, s->st));
		CK(cudaMemcpyAsync(hn.data() + (size_t)g * U, s->am_nan, (size_t)U * sizeof(int), cudaMemcpyDeviceToHost, s->st));
		CK(cudaEventRecord(s->ev1, s->st));
	}
	float worst = 0.f;
	for (int g = 0; g < N; g++) {
		mf_shard *s = &pl->sh[g];
		CK(cudaSetDevice(s->dev));
		CK(cudaStreamSynchronize(s->st));
		float t = 0.f;
		CK(cudaEventElapsedTime(&t, s->ev0, s->ev1));
		worst = std::max(worst, t);
	}
	pl->rec_ms = worst;
	/* merge: blocks in ascending item order; a NaN at the first unrated item is never displaced */
	for (int i = 0; i < U; i++) {
		double v = 0.0;
		int j = -1, nan_first = 0;
		for (int g = 0; g < N; g++) {
			const size_t o = (size_t)g * U + i;
			if (hj[o] >= 0 && better_host(hv[o], hj[o], v, j)) {
				v = hv[o];
				j = hj[o];
			}
			nan_first |= hn[o];
		}
		if (ff[(size_t)i] >= 0 && nan_first)
			j = ff[(size_t)i];
		best[i] = j;
	}
	return MFGPU_OK;
}

extern "C" int mfgpu_plan_recommend(mfgpu_plan *pl, int *best)
{
	if (!pl || !best)
		return mf_fail(MFGPU_EINVAL, "null argument");
	if (!pl->have_factors)
		return mf_fail(MFGPU_EINVAL, "mfgpu_plan_set_factors has not been called");
	const int nl = local_shards(pl);
	const int c = pl->cur;
	const bool on_device = pl->F > 0 && pl->items > 0;
	if (nl > 1 && on_device && pl->users > 0 &&
	    (pl->opt_argmax_items > 0 || (pl->opt_argmax_items < 0 && (long)pl->items >= 8L * pl->users)))
		return recommend_by_items(pl, best);
	for (int g = 0; g < nl; g++) {
		mf_shard *s = &pl->sh[g];
		const int nu = s->user_hi - s->user_lo;
		CK(cudaSetDevice(s->dev));
		CK(cudaEventRecord(s->ev0, s->st));
		if (on_device && nu > 0) {
			CK(cudaMemsetAsync(s->nan_first, 0, (size_t)nu * sizeof(int), s->st));
			const int n_tiles = (pl->items + mf::AM_TJ - 1) / mf::AM_TJ;
			mf::ArgmaxArgs a;
			a.l_row0 = s->user_lo;
			a.items = pl->items;
			a.F = pl->F;
			a.ld = pl->ld;
			a.n_splits = s->am_splits;
			a.tiles_per_split = (n_tiles + s->am_splits - 1) / s->am_splits;
			a.tile_begin = 0;
			a.tile_end = n_tiles;
			a.mask_ptr = s->d_mask_ptr;
			a.mask_col = s->d_mask_col;
			a.first_free = s->d_first_free;
			const int chunk = 65535 * mf::AM_TI;      /* grid.y limit */
			for (int lo = s->user_lo; lo < s->user_hi; lo += chunk) {
				a.users_lo = lo;
				a.users_hi = std::min(s->user_hi, lo + chunk);
				const size_t off = (size_t)(lo - s->user_lo);
				a.part_v = s->part_v + off * s->am_splits;
				a.part_j = s->part_j + off * s->am_splits;
				a.nan_first = s->nan_first + off;
				const int cnt = a.users_hi - a.users_lo;
				dim3 grid((unsigned)s->am_splits, (unsigned)((cnt + mf::AM_TI - 1) / mf::AM_TI));
				mf::k_masked_argmax<<<grid, 256, 0, s->st>>>(a, s->L[c], s->R[c]);
				CK(cudaGetLastError());
				mf::k_argmax_final<<<(cnt + 255) / 256, 256, 0, s->st>>>(a, s->d_best + off);
				CK(cudaGetLastError());
				count_launches(pl, 2);
			}
			CK(cudaMemcpyAsync(best + s->user_lo, s->d_best, (size_t)nu * sizeof(int),
			                   cudaMemcpyDeviceToHost, s->st));
		} else {
			/* B has no columns, or is all +0.0 (features == 0): print_output keeps
			 * the first item its cursor does not skip (matFact.c:17) */
			for (int i = s->user_lo; i < s->user_hi; i++)
				best[i] = pl->items > 0 ? s->pat.first_free[i] : -1;
		}
		CK(cudaEventRecord(s->ev1, s->st));
	}
	float worst = 0.f;
	for (int g = 0; g < nl; g++) {
		mf_shard *s = &pl->sh[g];
		CK(cudaSetDevice(s->dev));
		CK(cudaStreamSynchronize(s->st));
		float t = 0.f;
		CK(cudaEventElapsedTime(&t, s->ev0, s->ev1));
		worst = std::max(worst, t);
	}
	pl->rec_ms = worst;
	return MFGPU_OK;
}

extern "C" int mfgpu_plan_last_recommend_ms(mfgpu_plan *pl, float *ms)
{
	if (!pl || !ms)
		return mf_fail(MFGPU_EINVAL, "null argument");
	*ms = pl->rec_ms;
	return MFGPU_OK;
}

/* ---- one-shot entries with host buffers ------------------------------------------- */

static int run_all(int users, int items, int features, int nnz, const mfgpu_entry *entries,
                   int iters, double alpha, double *L, double *R, int *best, int ngpus,
                   bool factorize, bool writeback)
{
	if (!L || !R)
		return mf_fail(MFGPU_EINVAL, "null factor pointer");
	mfgpu_plan *pl = nullptr;
	TRY(mfgpu_plan_create(&pl, users, items, features, nnz, entries, alpha, ngpus));
	int rc = mfgpu_plan_set_factors(pl, L, R);
	if (rc == MFGPU_OK && factorize)
		rc = mfgpu_plan_iterate(pl, iters);
	if (rc == MFGPU_OK && writeback && best)
		rc = mfgpu_plan_get_results(pl, L, R, best);
	else if (rc == MFGPU_OK && writeback)
		rc = mfgpu_plan_get_factors(pl, L, R);
	else if (rc == MFGPU_OK && best)
		rc = mfgpu_plan_recommend(pl, best);
	if (rc == MFGPU_OK)
		rc = mfgpu_plan_sync(pl);
	char keep[512];
	snprintf(keep, sizeof keep, "%s", mfgpu_last_error());
	mfgpu_plan_destroy(pl);
	return rc == MFGPU_OK ? rc : mf_fail(rc, "%s", keep);
}

extern "C" int mfgpu_factorize(int users, int items, int features, int nnz,
                               const mfgpu_entry *entries, int iters, double alpha,
                               double *L, double *R, int ngpus)
{
	if (iters < 0)
		return mf_fail(MFGPU_EINVAL, "negative iteration count");
	return run_all(users, items, features, nnz, entries, iters, alpha, L, R, nullptr, ngpus, true, true);
}

extern "C" int mfgpu_recommend(int users, int items, int features, int nnz,
                               const mfgpu_entry *entries, const double *L, const double *R,
                               int *best, int ngpus)
{
	if (!best)
		return mf_fail(MFGPU_EINVAL, "null output pointer");
	return run_all(users, items, features, nnz, entries, 0, 0.0, const_cast<double *>(L),
	               const_cast<double *>(R), best, ngpus, false, false);
}

extern "C" int mfgpu_run(int users, int items, int features, int nnz,
                         const mfgpu_entry *entries, int iters, double alpha,
                         double *L, double *R, int *best, int ngpus)
{
	if (iters < 0 || !best)
		return mf_fail(MFGPU_EINVAL, "bad argument");
	return run_all(users, items, features, nnz, entries, iters, alpha, L, R, best, ngpus, true, true);
}

/* ---- masked arg-max without any allocation (mfgpu_recommend_noalloc) -----------------------
 *
 * No counterpart in the reference; added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  The same two kernels as mfgpu_plan_recommend, with their scratch (the
 * per-split partial maxima part_v / part_j and the nan_first flags) in a file-scope device
 * array instead of pool memory.  Every GPU that loads the library holds the array until the
 * process exits; every call shares it, so calls that overlap on one GPU race.  In host code
 * the array's name is a host-side stand-in, not a device address: the address the kernels get
 * comes from cudaGetSymbolAddress, once per device, with that device current. */
static __device__ __align__(256) unsigned char g_noalloc_scratch[MFGPU_NOALLOC_SCRATCH_BYTES];
static void *g_noalloc_ptr[64];

extern "C" int mfgpu_recommend_noalloc(int users, int items, int features, int ld,
                                       const double *L, const double *R, const int *mask_ptr,
                                       const int *mask_col, const int *first_free, int *best,
                                       void *stream)
{
	if (users < 0 || items < 0 || features < 0)
		return mf_fail(MFGPU_EINVAL, "mfgpu_recommend_noalloc: negative size");
	if (users == 0)
		return MFGPU_OK;
	if (!best || !mask_ptr || !first_free)
		return mf_fail(MFGPU_EINVAL, "mfgpu_recommend_noalloc: null pointer");
	if (features > 0 && items > 0) {
		if (!L || !R)
			return mf_fail(MFGPU_EINVAL, "mfgpu_recommend_noalloc: null factor pointer");
		if (ld < features || (ld & 1) || ((uintptr_t)L & 15) || ((uintptr_t)R & 15))
			return mf_fail(MFGPU_EINVAL, "mfgpu_recommend_noalloc: rows must be 16-byte aligned "
			                             "(ld even and >= features, L and R 16-byte aligned)");
	}
	/* scratch: nan_first[users] ints | part_j[users*splits] ints | part_v[users*splits] doubles */
	const size_t cap = MFGPU_NOALLOC_SCRATCH_BYTES;
	const size_t fixed = 64 + (size_t)users * sizeof(int);
	const size_t per_split = (size_t)users * (sizeof(int) + sizeof(double));
	if (fixed + per_split > cap)
		return mf_fail(MFGPU_EINVAL, "mfgpu_recommend_noalloc: %d users need %zu bytes of scratch, the "
		                             "built-in array has %zu (at most %zu users)", users, fixed + per_split,
		               cap, (cap - 64) / (2 * sizeof(int) + sizeof(double)));
	int dev = 0;
	CK(cudaGetDevice(&dev));
	if (dev < 0 || dev >= 64)
		return mf_fail(MFGPU_ENODEV, "mfgpu_recommend_noalloc: device ordinal %d out of range", dev);
	int major = 0;
	CK(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
	if (major != 10)
		return mf_fail(MFGPU_ENODEV, "mfgpu_recommend_noalloc: the current device is not an sm_100 (B200) device");
	if (!g_noalloc_ptr[dev]) {
		void *q = nullptr;
		CK(cudaGetSymbolAddress(&q, g_noalloc_scratch));
		g_noalloc_ptr[dev] = q;
	}
	cudaStream_t st = (cudaStream_t)stream;
	if (features == 0 || items == 0) {
		/* B has no columns, or is all +0.0: print_output keeps the first item its cursor does not
		 * skip (matFact.c:17); first_free already holds it (-1 when there is none) */
		if (items == 0)
			CK(cudaMemsetAsync(best, 0xff, (size_t)users * sizeof(int), st));
		else
			CK(cudaMemcpyAsync(best, first_free, (size_t)users * sizeof(int), cudaMemcpyDeviceToDevice, st));
		return MFGPU_OK;
	}
	const int tiles_y = (users + mf::AM_TI - 1) / mf::AM_TI;
	if (tiles_y > 65535)
		return mf_fail(MFGPU_EINVAL, "mfgpu_recommend_noalloc: too many users");
	const int n_tiles = (items + mf::AM_TJ - 1) / mf::AM_TJ;
	int splits = (148 * 2 * 8 + tiles_y - 1) / tiles_y;            /* as shard_create (mf_plan.cu) */
	splits = std::max(1, std::min(splits, n_tiles));
	splits = (int)std::min<size_t>((size_t)splits, (cap - fixed) / per_split);
	unsigned char *base = (unsigned char *)g_noalloc_ptr[dev];
	mf::ArgmaxArgs a;
	a.users_lo = 0;
	a.users_hi = users;
	a.l_row0 = 0;
	a.items = items;
	a.F = features;
	a.ld = ld;
	a.n_splits = splits;
	a.tiles_per_split = (n_tiles + splits - 1) / splits;
	a.tile_begin = 0;
	a.tile_end = n_tiles;
	a.mask_ptr = mask_ptr;
	a.mask_col = mask_col;
	a.first_free = first_free;
	a.nan_first = (int *)base;
	a.part_j = a.nan_first + users;
	a.part_v = (double *)(base + (((size_t)users * (1 + (size_t)splits) * sizeof(int) + 63) & ~(size_t)63));
	/* the array keeps the last call's contents: the flags must start at zero, the partial
	 * maxima are written in full by every CTA before k_argmax_final reads them */
	CK(cudaMemsetAsync(a.nan_first, 0, (size_t)users * sizeof(int), st));
	dim3 grid((unsigned)splits, (unsigned)tiles_y);
	mf::k_masked_argmax<<<grid, 256, 0, st>>>(a, L, R);
	CK(cudaGetLastError());
	mf::k_argmax_final<<<(users + 255) / 256, 256, 0, st>>>(a, best);
	CK(cudaGetLastError());
	return MFGPU_OK;
}
