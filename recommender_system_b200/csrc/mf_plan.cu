/*
 * csrc/mf_plan.cu -- plan life cycle: pattern upload, factor buffers, launch geometry.
 *
 * Data layout in HBM (per shard = per GPU):
 *   L[2]   (users of the shard) x ld doubles, row-major, double buffered
 *   R[2]   items x ld doubles, row-major (the reference's transposed R,
 *          matFact.c:117-120), double buffered, replicated on every shard
 *   by-user grouping  row_id / row_ptr / col_idx / val / epos   (longest row first)
 *   by-item grouping  col_id / col_ptr / row_idx                (longest column first)
 *   e_csc  nnz doubles, e_n at by-item positions; e_csr the same at by-user positions
 *   G[2]   (N shards) touched items x ld doubles: partial R gradient
 * ld = features rounded up to even so every row starts 16-byte aligned.
 */
#include "mf_internal.h"
#include "../host/mf_ingest.h"

#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <new>
#include <vector>

/* Device buffers come from the device's default memory pool in stream order, and the pool keeps
 * what is freed (release threshold = everything): a process that factorizes one matrix after
 * another pays cudaMalloc/cudaFree (milliseconds per 100 MB) once, not per plan.  Buffers that
 * peers or NCCL touch (G) stay plain cudaMalloc. */
static int pool_prepare(int dev)
{
	static bool done[64];
	if (dev < 0 || dev >= 64 || done[dev])
		return MFGPU_OK;
	cudaMemPool_t pool;
	CK(cudaDeviceGetDefaultMemPool(&pool, dev));
	unsigned long long keep = ~0ull;
	CK(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
	done[dev] = true;
	return MFGPU_OK;
}

static int dev_alloc(void **p, size_t bytes, cudaStream_t st)
{
	CK(cudaMallocAsync(p, bytes ? bytes : 1, st));
	return MFGPU_OK;
}

static void dev_free(void *p, cudaStream_t st)
{
	if (p)
		cudaFreeAsync(p, st);
}

/* Page-locked host workspaces for the patterns, kept from one plan to the next (a handful,
 * at most 1 GB in total): cudaHostAlloc costs milliseconds, first-touch page faults of a
 * fresh 20 MB pattern cost more than building it, and uploads from pageable memory are not
 * DMA.  A workspace that cannot be had is not an error: the pattern then uses malloc. */
#include <mutex>
namespace {
struct HostWs { void *p; size_t bytes; };
std::mutex g_ws_mutex;
std::vector<HostWs> g_ws_free;

void *ws_acquire(size_t need, size_t *got)
{
	{
		std::lock_guard<std::mutex> lk(g_ws_mutex);
		int best = -1;
		for (int i = 0; i < (int)g_ws_free.size(); i++)
			if (g_ws_free[i].bytes >= need && (best < 0 || g_ws_free[i].bytes < g_ws_free[best].bytes))
				best = i;
		if (best >= 0) {
			HostWs w = g_ws_free[best];
			g_ws_free.erase(g_ws_free.begin() + best);
			*got = w.bytes;
			return w.p;
		}
	}
	void *p = nullptr;
	const size_t bytes = need + need / 8 + 4096;
	if (cudaHostAlloc(&p, bytes, cudaHostAllocPortable) != cudaSuccess) {
		cudaGetLastError();
		return nullptr;
	}
	*got = bytes;
	return p;
}

void ws_release(void *p, size_t bytes)
{
	if (!p)
		return;
	{
		std::lock_guard<std::mutex> lk(g_ws_mutex);
		size_t total = bytes;
		for (const HostWs &w : g_ws_free)
			total += w.bytes;
		if (g_ws_free.size() < 16 && total <= ((size_t)1 << 30)) {
			g_ws_free.push_back({p, bytes});
			return;
		}
	}
	cudaFreeHost(p);
}
} // namespace

/* rows of `width` doubles from a matrix of row stride sld to one of row stride dld (device) */
__global__ void __launch_bounds__(256)
k_pack_rows(double *__restrict__ dst, size_t dld, const double *__restrict__ src, size_t sld, size_t width, size_t total)
{
	for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
		const size_t r = t / width, k = t - r * width;
		dst[r * dld + k] = src[r * sld + k];
	}
}

static int pack_rows(double *dst, size_t dld, const double *src, size_t sld, size_t width, size_t rows, cudaStream_t st)
{
	const size_t total = width * rows;
	if (!total)
		return MFGPU_OK;
	k_pack_rows<<<(unsigned)std::min<size_t>((total + 255) / 256, 148 * 32), 256, 0, st>>>(dst, dld, src, sld, width, total);
	CK(cudaGetLastError());
	return MFGPU_OK;
}

/* Host rows of F doubles -> device rows of stride ld, through `scratch` (a device buffer of the
 * same size, whose content is dead): one contiguous copy + a kernel instead of a strided copy
 * (cudaMemcpy2DAsync of 1600-byte rows runs at a third of the contiguous rate). */
static int upload_rows(double *dst, double *scratch, size_t ld, const double *host, size_t F, size_t rows, cudaStream_t st)
{
	if (ld == F) {
		CK(cudaMemcpyAsync(dst, host, rows * F * sizeof(double), cudaMemcpyHostToDevice, st));
		return MFGPU_OK;
	}
	CK(cudaMemcpyAsync(scratch, host, rows * F * sizeof(double), cudaMemcpyHostToDevice, st));
	return pack_rows(dst, ld, scratch, F, F, rows, st);
}

/* ... and back: packed into `scratch`, one contiguous copy to the host, and `scratch` restored as
 * a copy of `src` (it is the other buffer of the pair: rows that are never rewritten must be
 * equal in both, and any valid content will do for the rest). */
static int download_rows(double *host, const double *src, double *scratch, size_t ld, size_t F, size_t rows, cudaStream_t st)
{
	if (ld == F) {
		CK(cudaMemcpyAsync(host, src, rows * F * sizeof(double), cudaMemcpyDeviceToHost, st));
		return MFGPU_OK;
	}
	TRY(pack_rows(scratch, F, src, ld, F, rows, st));
	CK(cudaMemcpyAsync(host, scratch, rows * F * sizeof(double), cudaMemcpyDeviceToHost, st));
	CK(cudaMemcpyAsync(scratch, src, rows * ld * sizeof(double), cudaMemcpyDeviceToDevice, st));
	return MFGPU_OK;
}

template <typename T>
static int upload(T **dst, const T *src, size_t n, cudaStream_t st)
{
	TRY(dev_alloc((void **)dst, n * sizeof(T), st));
	if (n)
		CK(cudaMemcpyAsync(*dst, src, n * sizeof(T), cudaMemcpyHostToDevice, st));
	return MFGPU_OK;
}

/* Doubles of one parity half of a G buffer.  One shard per row of G: a row of ld doubles per
 * touched item.  N shards: the receive buffers of the exchanges -- nshards blocks (one per
 * contributing shard) of slots_per_rank rows, twice that for the (value, tag) pairs of the
 * low-latency exchange (G[0]: contributions received, G[1]: rows of R_new received). */
static size_t mf_g_half(const mfgpu_plan *pl)
{
	size_t rows = (size_t)pl->n_touched;
	if (pl->nshards > 1)
		rows = 2 * std::max(rows, (size_t)pl->nshards * pl->slots_per_rank);
	return rows * pl->ld;
}

/* Bytes of one G buffer: two halves, used alternately by the low-latency exchange (k_xchg_ll). */
static size_t mf_g_bytes(const mfgpu_plan *pl)
{
	return (2 * mf_g_half(pl) + 2) * sizeof(double);
}

/* Launch geometry of k_update_L / k_update_R for one shard (see mf_loop.cuh). */
int mf_configure_shard(mfgpu_plan *pl, mf_shard *s)
{
	const int F = pl->F;
	int max_smem = 0;
	CK(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, s->dev));
	/* nonzeros per chunk = lanes of the dot phase: the whole row when it has <= 32 nonzeros */
	int CH = std::max(1, std::min(32, s->pat.max_row_len));
	if (pl->opt_ch > 0)
		CH = std::max(1, std::min(32, pl->opt_ch));
	/* Depth of the two pipelines of a k_update_L warp: NT tiles in the dot-phase ring (2 or 4),
	 * acc_rows rows in flight ahead of the addition in the accumulation phase (1 or 3).  The
	 * shallow ones are the default at every size: measured on B200 at F = 200, the deep variants
	 * (fewer warps per SM: 10.5 KB of tiles, 128 registers) lose both on the full inst60000 shape
	 * (273 -> 327 / 347 us) and on one eighth of it, the shard of one rank of eight (41 -> 49 /
	 * 51 us): resident warps hide latency better than a deeper pipeline per warp.  Options
	 * dot_stages / acc_rows select them for experiments. */
	int NT = pl->opt_dot_stages > 2 ? 4 : 2;
	s->acc_rows = pl->opt_acc_rows > 1 ? 3 : 1;
	auto warp_doubles = [&](int ch) { return ((size_t)NT * ch * 16 + pl->ld + 32 + 15) & ~(size_t)15; };   /* tiles of [ch][16], mf_loop.cuh */
	int warps = 8;
	if (8 * warp_doubles(CH) * sizeof(double) + 128 > (size_t)max_smem)
		NT = 2;
	if (8 * warp_doubles(CH) * sizeof(double) + 128 > (size_t)max_smem) {
		warps = 1;      /* one warp per CTA: k_update_L<1, 1, 2> */
		s->acc_rows = 1;
	}
	if (warps * warp_doubles(CH) * sizeof(double) + 128 > (size_t)max_smem)
		return mf_fail(MFGPU_EINVAL, "features=%d does not fit shared memory (%d bytes)", F, max_smem);
	s->warps = warps;
	s->k1_smem = warps * warp_doubles(CH) * sizeof(double) + 128;      /* + slack to start on a 128-byte line */
	/* Row classes.  A warp of k_update_L spends ~100 cycles per nonzero, so a row much longer
	 * than the per-warp share of the shard (nnz / (8 warps x SMs)) would outlast the rest of
	 * the launch: such rows have their dots spread over warps by k_dot.  Chains longer than
	 * ~nnz / (8 x SMs) go to k_chain_ws (one CTA of specialised warps per chain and feature
	 * group: ~20 cycles per nonzero against ~70 for a warp of k_chain, but three warps and
	 * 52 KB of shared memory per chain), the rest to k_chain. */
	int n_sm = 148;
	CK(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, s->dev));
	n_sm = std::max(1, n_sm);
	const mf_pattern &pt = s->pat;
	const int share = (int)std::min<long>(1 << 30, (long)pt.nnz / (16L * n_sm));
	/* split mode (option "pair", the default when eight dot warps fit): the dots of rows of at most
	 * 32 nonzeros by k_dot_pair, of longer rows by k_dot, every accumulation by k_acc */
	int pair_warps = pl->opt_pair_warps;
	if (pair_warps != 4 && pair_warps != 5 && pair_warps != 6 && pair_warps != 8)
		pair_warps = 8;
	auto pair_bytes = [&](int nt) { return pair_warps * ((size_t)nt * 34 * 16 + 144) * sizeof(double) + 128; };   /* mf::dot_units_doubles + alignment slack */
	/* default: where it was measured ahead of the fused kernels (us per iteration, split / fused):
	 * inst60000 shape 316 / 401, one eighth of it 53 / 61, one eighth of the ML1M shape 52 / 59;
	 * the whole ML1M shape 252 / 248 -- wide factors always, F >= 64 on shards below 6e5 nonzeros;
	 * not for rows of a few nonzeros (two rows to a warp leave its lanes idle: inst50000-5000-100-2-5
	 * 113 / 97, inst1e6-100-700-1-3 1331 / 1114): at least 8 nonzeros per row on average */
	const int want_split = pl->opt_pair >= 0 ? pl->opt_pair
	                       : ((F >= 128 || (F >= 64 && pt.nnz < 600000)) && (long)pt.nnz >= 8L * pt.n_rows);
	const bool split = warps == 8 && want_split && pair_bytes(2) <= (size_t)max_smem;
	/* the ring of a k_dot_pair warp is what it has in flight from L2: as deep as one CTA of
	 * eight warps per SM can hold (measured latency under load ~2000 cycles: ~100 KB in flight
	 * per SM for 50 B/clk) */
	int pair_nt = 2;
	if (split) {
		pair_nt = pl->opt_pair_stages >= 2 ? std::min(6, pl->opt_pair_stages) : 6;
		pair_nt = std::min(pair_nt, (F + 15) / 16 + 1);      /* the ring runs at most one unit ahead */
		while (pair_nt > 2 && pair_bytes(pair_nt) > (size_t)max_smem)
			pair_nt--;
	}
	const int long_thresh = split ? 32 : pl->opt_long_row > 0 ? pl->opt_long_row : std::max(2 * CH, share);
	const int deep_thresh = pl->opt_long_col > 0 ? pl->opt_long_col
	                                             : std::max(64, (int)std::min<long>(1 << 30, (long)pt.nnz / (8L * n_sm)));
	int n_long = 0, n_deep_rows = 0, n_deep_cols = 0;
	if (warps > 1)
		while (n_long < pt.n_rows && pt.row_ptr[n_long + 1] - pt.row_ptr[n_long] > long_thresh)
			n_long++;
	while (n_deep_rows < n_long && pt.row_ptr[n_deep_rows + 1] - pt.row_ptr[n_deep_rows] > deep_thresh)
		n_deep_rows++;
	while (n_deep_cols < pt.n_cols && pt.col_ptr[n_deep_cols + 1] - pt.col_ptr[n_deep_cols] > deep_thresh)
		n_deep_cols++;

	/* Short rows with their gathered R rows resident in shared memory (k_update_L_res): rows
	 * RST doubles apart (RST/2 odd), CR rows per chunk = the longest short row (at most 32).
	 * It reads every R row once instead of twice, but a warp then owns CR x RST x 8 bytes of
	 * shared memory: measured on B200 it wins while eight or more warps fit an SM (F = 100:
	 * 107 us against 140 us on the MovieLens-1M shape) and loses to the tiled kernel when
	 * fewer do (F = 200: 311 us against 276 us on inst60000).  resident=2 forces it. */
	int CR = 0, RST = pl->ld + (((pl->ld / 2) & 1) ? 0 : 2);
	s->res_ng = 0;
	s->res_smem = 0;
	if (warps > 1 && pl->opt_resident && !split && F <= 256 && n_long < pt.n_rows) {
		const int longest = pt.row_ptr[n_long + 1] - pt.row_ptr[n_long];
		CR = std::min(32, longest);
		if (pl->opt_res_rows > 0)
			CR = std::min(32, pl->opt_res_rows);
		const size_t per_warp = ((size_t)CR * RST + pl->ld + 32) * sizeof(double);
		const bool eight_fit = 8 * per_warp <= (size_t)max_smem;
		if ((eight_fit || pl->opt_resident > 1) && 2 * per_warp <= (size_t)max_smem) {
			s->res_ng = (F + 31) / 32;
			s->res_smem = 2 * per_warp;          /* mf::RES_WARPS warps per CTA */
		}
	}
	/* Rows of at most 32 nonzeros two to a warp (k_update_L_pair, mf_loop.cuh): the longest
	 * unpaired row with the shortest one while their lengths add up to at most 32.  Rows are
	 * stored longest first, so [n_mid, n_rows) are the rows of at most 32 nonzeros. */
	std::vector<int> units;        /* 8 ints per unit: {ia, basea, la, ib, baseb, lb, 0, 0} */
	s->n_units = 0;
	s->n_mid = pt.n_rows;
	s->pair_smem = 0;
	s->split = split;
	if (split) {
		s->pair_smem = pair_bytes(pair_nt);
		s->pair_nt = pair_nt;
		s->pair_warps = pair_warps;
		s->n_mid = n_long;
		units.reserve(8 * ((size_t)(n_long ? pt.row_ptr[n_long] : 0) / 32 + (size_t)pt.n_rows + 1));
		auto unit = [&](int ra, int basea, int la, int rb) {
			const int ib = rb >= 0 ? pt.row_id[rb] - s->user_lo : 0;
			const int baseb = rb >= 0 ? pt.row_ptr[rb] : 0, lb = rb >= 0 ? pt.row_ptr[rb + 1] - baseb : 0;
			const int d[8] = {pt.row_id[ra] - s->user_lo, basea, la, ib, baseb, lb, 0, 0};
			units.insert(units.end(), d, d + 8);
		};
		/* rows longer than 32 (the rows [0, n_long), long_thresh = 32): chunks of 32 */
		for (int r = 0; r < n_long; r++)
			for (int pos = pt.row_ptr[r]; pos < pt.row_ptr[r + 1]; pos += 32)
				unit(r, pos, std::min(32, pt.row_ptr[r + 1] - pos), -1);
		/* the others two to a warp: the longest unpaired row with the shortest one while their
		 * lengths add up to at most 32 (rows are stored longest first) */
		int lo = n_long, hi = pt.n_rows - 1;
		while (lo <= hi) {
			const int la = pt.row_ptr[lo + 1] - pt.row_ptr[lo];
			const bool two = lo < hi && la + (pt.row_ptr[hi + 1] - pt.row_ptr[hi]) <= 32;
			unit(lo, pt.row_ptr[lo], la, two ? hi : -1);
			lo++;
			if (two)
				hi--;
		}
		s->n_units = (int)(units.size() / 8);
	}
	TRY(mf_kernels_prepare(warps, s->k1_smem, s->res_ng, s->res_smem, s->pair_smem, s->pair_warps, s->pair_nt));
	TRY(mf_kernels_preload(F, s->split, pl->nshards));

	/* chunks of the long rows for k_dot */
	std::vector<int> chunk_row, chunk_pos;
	for (int r = 0; r < (split ? 0 : n_long); r++)     /* split mode: the chunks are units of k_dot_units */
		for (int pos = pt.row_ptr[r]; pos < pt.row_ptr[r + 1]; pos += CH) {
			chunk_row.push_back(r);
			chunk_pos.push_back(pos);
		}
	dev_free(s->d_chunk_row, s->st);
	dev_free(s->d_chunk_pos, s->st);
	dev_free(s->d_units, s->st);
	s->d_chunk_row = s->d_chunk_pos = s->d_units = nullptr;
	TRY(upload(&s->d_units, units.data(), units.size(), s->st));
	TRY(upload(&s->d_chunk_row, chunk_row.data(), chunk_row.size(), s->st));
	TRY(upload(&s->d_chunk_pos, chunk_pos.data(), chunk_pos.size(), s->st));
	CK(cudaStreamSynchronize(s->st));     /* the vectors die with this scope */

	mf::LoopArgs &a = s->la;
	a.F = F;
	a.ld = pl->ld;
	a.CH = CH;
	a.NT = NT;
	a.CR = CR;
	a.RST = RST;
	a.n_rows = pt.n_rows;
	a.n_long = n_long;
	a.n_deep_rows = n_deep_rows;
	a.n_cols = pt.n_cols;
	a.n_deep_cols = n_deep_cols;
	a.n_chunks = (int)chunk_row.size();
	a.user_lo = s->user_lo;
	a.cs_store = pl->opt_cs_store;
	a.alpha2 = pl->alpha * 2;     /* matFact.c:45: alpha * 2 * (...) associates left, exact */
	a.row_id = s->d_row_id;
	a.row_ptr = s->d_row_ptr;
	a.col_idx = s->d_col_idx;
	a.epos = s->d_epos;
	a.val = s->d_val;
	a.col_id = s->d_col_id;
	a.col_ptr = s->d_col_ptr;
	a.row_idx = s->d_row_idx;
	a.chunk_row = s->d_chunk_row;
	a.chunk_pos = s->d_chunk_pos;
	a.item_slot = pl->nshards > 1 ? s->d_item_slot : nullptr;
	a.e_csc = s->d_e_csc;
	a.e_csr = s->d_e_csr;
	s->k1_blocks = (s->n_mid - n_long + warps - 1) / warps;      /* rows [n_mid, n_rows) go to the paired launch */
	if (s->res_ng) {
		/* persistent warps: as many CTAs as are resident at once, rows dealt round robin */
		int smem_sm = 228 * 1024;
		CK(cudaDeviceGetAttribute(&smem_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, s->dev));
		const int per_sm = (int)std::max<size_t>(1, std::min<size_t>(32, smem_sm / (s->res_smem + 1024)));
		s->k1_blocks = std::min((a.n_rows - n_long + 1) / 2, n_sm * per_sm);
	}
	s->dot_blocks = (a.n_chunks + 7) / 8;

	/* algorithmic (compulsory) bytes of one launch of each kernel, DESIGN.md section 4 */
	{
		const long long Fb = 8LL * F;
		const long long nnz_long = n_long ? pt.row_ptr[n_long] : 0, nnz_short = pt.nnz - nnz_long;
		const long long nnz_deep_r = n_deep_rows ? pt.row_ptr[n_deep_rows] : 0;
		const long long nnz_deep_c = n_deep_cols ? pt.col_ptr[n_deep_cols] : 0;
		auto touched = [&](long long nnz_part, long long other_total) { return std::min(nnz_part, other_total); };
		s->kbytes[0] = 2 * Fb * (pt.n_rows - n_long) + Fb * touched(nnz_short, pt.n_cols) + 24 * nnz_short;
		s->kbytes[1] = Fb * n_long + Fb * touched(nnz_long, pt.n_cols) + 32 * nnz_long;
		s->kbytes[2] = 2 * Fb * (n_long - n_deep_rows) + Fb * touched(nnz_long - nnz_deep_r, pt.n_cols) +
		               12 * (nnz_long - nnz_deep_r) +
		               2 * Fb * (pt.n_cols - n_deep_cols) + Fb * touched(pt.nnz - nnz_deep_c, pt.n_rows) +
		               12 * (pt.nnz - nnz_deep_c);
		s->kbytes[3] = 2 * Fb * n_deep_rows + Fb * touched(nnz_deep_r, pt.n_cols) + 12 * nnz_deep_r +
		               2 * Fb * n_deep_cols + Fb * touched(nnz_deep_c, pt.n_rows) + 12 * nnz_deep_c;
		s->kbytes[4] = pl->nshards > 1 ? 3 * Fb * pl->n_touched : 0;
		/* bytes the launch requests from L2 (the diagnostic model of SURVEY.md 8d): the own rows
		 * as above, but a gathered row of the other factor for EVERY nonzero and pass -- twice
		 * per nonzero in the tiled k_update_L (dot, accumulation), once in the resident one */
		const long long passes = s->res_ng ? 1 : 2;
		s->gbytes[0] = 2 * Fb * (pt.n_rows - n_long) + passes * Fb * nnz_short + 24 * nnz_short;
		s->gbytes[1] = Fb * n_long + Fb * nnz_long + 32 * nnz_long;
		s->gbytes[2] = 2 * Fb * (n_long - n_deep_rows) + Fb * (nnz_long - nnz_deep_r) + 12 * (nnz_long - nnz_deep_r) +
		               2 * Fb * (pt.n_cols - n_deep_cols) + Fb * (pt.nnz - nnz_deep_c) + 12 * (pt.nnz - nnz_deep_c);
		s->gbytes[3] = 2 * Fb * n_deep_rows + Fb * nnz_deep_r + 12 * nnz_deep_r +
		               2 * Fb * n_deep_cols + Fb * nnz_deep_c + 12 * nnz_deep_c;
		s->gbytes[4] = s->kbytes[4];
	}
	for (int g = 0; g < 2; g++)
		if (s->gexec[g]) {
			cudaGraphExecDestroy(s->gexec[g]);
			s->gexec[g] = nullptr;
		}
	pl->graph_len = 0;               /* the next graph may choose its length again */
	return MFGPU_OK;
}

/* ---- rank mode: the communicator's exchange arena (mf_internal.h) ---------------------- */

/* Collective: unmap the peers' buffers and free the own ones.  Nobody may do that while a peer's
 * kernels can still touch them, hence the one-word all-reduce as a barrier first. */
void mf_arena_reset(mfgpu_comm *c)
{
	mf_xchg_arena *ar = &c->arena;
	if (!ar->G[0] && !ar->sync)
		return;
	cudaSetDevice(c->device);
	cudaDeviceSynchronize();
	if (ar->sync && g_nccl.handle && c->nccl &&
	    g_nccl.AllReduce((int *)ar->sync + 14, (int *)ar->sync + 14, 1, 2, 0, c->nccl, 0) == 0)
		cudaDeviceSynchronize();
	for (int k = 0; k < ar->n_ipc_mapped; k++)
		cudaIpcCloseMemHandle(ar->ipc_mapped[k]);
	for (int b = 0; b < 2; b++) {
		cudaFree(ar->G[b]);
		cudaFree(ar->R[b]);
	}
	cudaFree(ar->sync);
	cudaGetLastError();
	memset(ar, 0, sizeof *ar);
}

/* Buffers of at least the given sizes for a new plan of this communicator: the arena's when no
 * live plan holds it (re-created, collectively, when it is too small), else nothing (the plan
 * allocates its own). */
static int arena_borrow(mfgpu_comm *c, size_t gbytes, size_t rbytes, mf_xchg_arena **out)
{
	mf_xchg_arena *ar = &c->arena;
	*out = nullptr;
	if (ar->in_use)
		return MFGPU_OK;
	if (!ar->G[0] || ar->g_bytes < gbytes || ar->r_bytes < rbytes) {
		mf_arena_reset(c);
		for (int b = 0; b < 2; b++) {
			CK(cudaMalloc((void **)&ar->G[b], gbytes));
			CK(cudaMalloc((void **)&ar->R[b], rbytes));
		}
		CK(cudaMalloc((void **)&ar->sync, MF_SYNC_WORDS * sizeof(unsigned long long)));
		CK(cudaMemset(ar->sync, 0, MF_SYNC_WORDS * sizeof(unsigned long long)));
		ar->g_bytes = gbytes;
		ar->r_bytes = rbytes;
	}
	ar->in_use = 1;
	*out = ar;
	return MFGPU_OK;
}

static int shard_create(mfgpu_plan *pl, mf_shard *s, int dev, const mfgpu_entry *entries,
                        const int *item_slot, const int *touched_item, const int *contrib)
{
	s->dev = dev;
	CK(cudaSetDevice(dev));
	TRY(pool_prepare(dev));
	CK(cudaStreamCreateWithFlags(&s->st, cudaStreamNonBlocking));
	CK(cudaStreamCreateWithFlags(&s->st2, cudaStreamNonBlocking));
	CK(cudaEventCreateWithFlags(&s->ev_fork, cudaEventDisableTiming));
	CK(cudaEventCreateWithFlags(&s->ev_join, cudaEventDisableTiming));
	CK(cudaEventCreateWithFlags(&s->ev_a, cudaEventDisableTiming));
	CK(cudaEventCreateWithFlags(&s->ev_b, cudaEventDisableTiming));
	CK(cudaEventCreate(&s->ev0));
	CK(cudaEventCreate(&s->ev1));
	CK(cudaEventCreate(&s->ev_t0));
	CK(cudaEventCreate(&s->ev_t1));
	for (int g = 0; g < 2; g++)
		CK(cudaEventCreateWithFlags(&s->ev_g[g], cudaEventDisableTiming));
	TraceClock tc;
	s->host_ws = ws_acquire(mf_pattern_ws_bytes(pl->users, pl->items, pl->nnz, s->user_lo, s->user_hi),
	                        &s->host_ws_bytes);
	tc.lap("create: host workspace");
	int rc = mf_pattern_build_scan(&s->pat, pl->users, pl->items, pl->nnz, entries, s->user_lo, s->user_hi,
	                               s->host_ws, s->host_ws_bytes, pl->scan);
	tc.lap("create: pattern build (host)");
	if (rc != MFGPU_OK)
		return mf_fail(rc, rc == MFGPU_EINVAL ? "an entry lies outside the users x items matrix"
		                                      : "out of host memory building the pattern");
	const mf_pattern &p = s->pat;
	TRY(upload(&s->d_row_id, p.row_id, (size_t)p.n_rows, s->st));
	TRY(upload(&s->d_row_ptr, p.row_ptr, (size_t)p.n_rows + 1, s->st));
	TRY(upload(&s->d_col_idx, p.col_idx, (size_t)p.nnz, s->st));
	TRY(upload(&s->d_epos, p.epos, (size_t)p.nnz, s->st));
	TRY(upload(&s->d_val, p.val, (size_t)p.nnz, s->st));
	TRY(upload(&s->d_col_id, p.col_id, (size_t)p.n_cols, s->st));
	TRY(upload(&s->d_col_ptr, p.col_ptr, (size_t)p.n_cols + 1, s->st));
	TRY(upload(&s->d_row_idx, p.row_idx, (size_t)p.nnz, s->st));
	TRY(upload(&s->d_mask_ptr, p.mask_ptr, (size_t)pl->users + 1, s->st));
	TRY(upload(&s->d_mask_col, p.mask_col, (size_t)p.mask_ptr[pl->users], s->st));
	TRY(upload(&s->d_first_free, p.first_free, (size_t)pl->users, s->st));
	tc.lap("create: pattern upload");
	TRY(dev_alloc((void **)&s->d_e_csc, ((size_t)p.nnz + 1) * sizeof(double), s->st));
	TRY(dev_alloc((void **)&s->d_e_csr, ((size_t)p.nnz + 1) * sizeof(double), s->st));
	TRY(dev_alloc((void **)&s->d_ticket, 64, s->st));

	const size_t nu = (size_t)(s->user_hi - s->user_lo);
	/* 34 spare doubles: k_acc lets the lanes past F of the last feature group read what follows the row */
	const size_t lbytes = (nu * pl->ld + 34) * sizeof(double);
	const size_t rbytes = ((size_t)pl->items * pl->ld + 34) * sizeof(double);
	/* rank mode: peers store their rows of R_new straight into this replica (push exchange), so
	 * R must be exportable over CUDA IPC, which pool memory is not */
	s->r_plain = pl->nshards > 1;
	if (pl->comm && pl->nshards > 1) {
		TRY(arena_borrow(pl->comm, mf_g_bytes(pl), rbytes, &pl->arena));
		s->borrowed = pl->arena != nullptr;
	}
	for (int b = 0; b < 2; b++) {
		TRY(dev_alloc((void **)&s->L[b], lbytes, s->st));
		if (s->borrowed)
			s->R[b] = pl->arena->R[b];
		else if (s->r_plain)
			CK(cudaMalloc((void **)&s->R[b], rbytes));
		else
			TRY(dev_alloc((void **)&s->R[b], rbytes, s->st));
		CK(cudaMemsetAsync(s->L[b], 0, lbytes, s->st));
		CK(cudaMemsetAsync(s->R[b], 0, rbytes, s->st));
	}
	if (pl->nshards > 1) {
		TRY(upload(&s->d_item_slot, item_slot, (size_t)pl->items, s->st));
		TRY(upload(&s->d_touched_item, touched_item, (size_t)pl->n_touched, s->st));
		if (contrib) {
			/* push exchange: touched item of slot t is reduced by shard t / slots_per_rank; this
			 * shard's contribution goes to row me * slots_per_rank + (t mod slots_per_rank) of
			 * that shard's receive buffer */
			const int per = pl->slots_per_rank, me = s->me;
			std::vector<int> owner((size_t)std::max(1, pl->items), 0), prow((size_t)std::max(1, pl->items), 0);
			for (int j = 0; j < pl->items; j++)
				if (item_slot[j] >= 0) {
					owner[j] = item_slot[j] / per;
					prow[j] = me * per + item_slot[j] % per;
				}
			TRY(upload(&s->d_push_owner, owner.data(), (size_t)pl->items, s->st));
			TRY(upload(&s->d_push_row, prow.data(), (size_t)pl->items, s->st));
			TRY(upload(&s->d_contrib, contrib, (size_t)pl->n_touched, s->st));
			CK(cudaStreamSynchronize(s->st));     /* the vectors die with this scope */
		}
		/* G: one row per touched item; as the receive buffer of the push exchange, nshards
		 * blocks of slots_per_rank rows (one block per contributing rank) */
		const size_t gbytes = mf_g_bytes(pl);
		for (int b = 0; b < 2; b++) {
			if (s->borrowed)
				s->G[b] = pl->arena->G[b];
			else
				CK(cudaMalloc((void **)&s->G[b], gbytes));
			/* columns without a local nonzero are never written: they stay 0.0 */
			CK(cudaMemsetAsync(s->G[b], 0, gbytes, s->st));
		}
	}
	/* arg-max scratch */
	const int tiles_y = (int)((nu + mf::AM_TI - 1) / mf::AM_TI);
	const int n_tiles = (pl->items + mf::AM_TJ - 1) / mf::AM_TJ;
	/* two CTAs are resident per SM (registers): aim at eight waves of them, so that the last,
	 * partly filled wave is a small part of the launch (three waves left 20 % of the SMs idle) */
	int splits = tiles_y > 0 ? (148 * 2 * 8 + tiles_y - 1) / tiles_y : 1;
	splits = std::max(1, std::min(splits, std::max(1, n_tiles)));
	s->am_splits = splits;
	TRY(dev_alloc((void **)&s->part_v, (nu * splits + 1) * sizeof(double), s->st));
	TRY(dev_alloc((void **)&s->part_j, (nu * splits + 1) * sizeof(int), s->st));
	TRY(dev_alloc((void **)&s->nan_first, (nu + 1) * sizeof(int), s->st));
	TRY(dev_alloc((void **)&s->d_best, (nu + 1) * sizeof(int), s->st));
	tc.lap("create: factor buffers");
	TRY(mf_configure_shard(pl, s));
	tc.lap("create: configure");
	CK(cudaStreamSynchronize(s->st));
	tc.lap("create: sync");
	return MFGPU_OK;
}

static void shard_destroy(mf_shard *s, bool own_sync)
{
	if (s->dev >= 0)
		cudaSetDevice(s->dev);
	if (own_sync && s->d_sync) {
		if (s->st) cudaStreamSynchronize(s->st);
		cudaFree(s->d_sync);
		s->d_sync = nullptr;
	}
	TraceClock tc;
	if (s->st) cudaStreamSynchronize(s->st);
	if (s->st2) cudaStreamSynchronize(s->st2);
	for (int g = 0; g < 2; g++)
		if (s->gexec[g]) cudaGraphExecDestroy(s->gexec[g]);
	tc.lap("destroy: graphs");
	for (int g = 0; g < 2; g++) {
		if (s->ev_g[g]) cudaEventDestroy(s->ev_g[g]);
		if (g == 0) cudaFree(s->G_scratch);
		dev_free(s->L[g], s->st);
		if (s->borrowed)
			;                                   /* the communicator's arena keeps G and R */
		else if (s->r_plain) {
			cudaFree(s->R[g]);
			cudaFree(s->G[g]);
		} else {
			dev_free(s->R[g], s->st);
			cudaFree(s->G[g]);
		}
	}
	tc.lap("destroy: factor buffers");
	void *ptrs[] = {s->d_row_id, s->d_row_ptr, s->d_col_idx, s->d_epos, s->d_col_id, s->d_col_ptr,
	                s->d_row_idx, s->d_val, s->d_e_csc, s->d_e_csr, s->d_mask_ptr, s->d_mask_col,
	                s->d_first_free, s->d_item_slot, s->d_touched_item, s->d_push_owner, s->d_push_row, s->d_contrib, s->part_v, s->part_j,
	                s->nan_first, s->d_best, s->L_keep, s->R_keep, s->d_chunk_row, s->d_chunk_pos, s->d_units, s->d_ticket,
	                s->am_L, s->am_mask_ptr, s->am_mask_col, s->am_first_free, s->am_part_v, s->am_out_v, s->am_part_j, s->am_nan, s->am_out_j};
	for (void *q : ptrs)
		dev_free(q, s->st);
	if (s->d_ll_args) cudaFree(s->d_ll_args);
	tc.lap("destroy: other buffers");
	if (s->ev0) cudaEventDestroy(s->ev0);
	if (s->ev1) cudaEventDestroy(s->ev1);
	if (s->ev_t0) cudaEventDestroy(s->ev_t0);
	if (s->ev_t1) cudaEventDestroy(s->ev_t1);
	if (s->ev_fork) cudaEventDestroy(s->ev_fork);
	if (s->ev_join) cudaEventDestroy(s->ev_join);
	if (s->ev_a) cudaEventDestroy(s->ev_a);
	if (s->ev_b) cudaEventDestroy(s->ev_b);
	if (s->st2) cudaStreamDestroy(s->st2);
	if (s->st) cudaStreamDestroy(s->st);
	tc.lap("destroy: events, streams");
	mf_pattern_free(&s->pat);
	ws_release(s->host_ws, s->host_ws_bytes);
	s->host_ws = nullptr;
	tc.lap("destroy: host pattern");
	cudaGetLastError();
}

/* ---- rank mode: exchange over peer memory ---------------------------------------------------
 * Collective over the ranks of the communicator.  Every rank publishes CUDA IPC handles of its
 * two G buffers and of a small flag block, and maps those of the others; the handles travel
 * through one ncclAllGather.  If any rank cannot export or map (no peer access, IPC disabled),
 * all ranks fall back to ncclAllReduce together. */
struct XchgHandles {
	cudaIpcMemHandle_t g[2], r[2], sync;
	int ok, pad[15];
};

static bool xchg_all_ok(mfgpu_plan *pl, mf_shard *s, int mine, int *d_scratch)
{
	/* logical AND over the ranks: sum of (1 - ok) must be 0 */
	int bad = mine ? 0 : 1, total = 1;
	if (cudaMemcpyAsync(d_scratch, &bad, sizeof bad, cudaMemcpyHostToDevice, s->st) != cudaSuccess)
		return false;
	if (g_nccl.AllReduce(d_scratch, d_scratch, 1, /*ncclInt32*/ 2, /*ncclSum*/ 0, pl->comm->nccl, s->st) != 0)
		return false;
	if (cudaMemcpyAsync(&total, d_scratch, sizeof total, cudaMemcpyDeviceToHost, s->st) != cudaSuccess)
		return false;
	if (cudaStreamSynchronize(s->st) != cudaSuccess)
		return false;
	return total == 0;
}

static int xchg_setup_ranks(mfgpu_plan *pl)
{
	pl->xchg_ipc = 0;
	if (!pl->comm || pl->nshards < 2 || (size_t)pl->n_touched * pl->ld == 0)
		return MFGPU_OK;
	mf_shard *s = &pl->sh[0];
	const int n = pl->comm->nranks, me = pl->comm->rank;
	CK(cudaSetDevice(s->dev));
	/* every transport relies on "rows of G nobody writes are 0.0", and each lays G out in its
	 * own way: start clean (the all-reduce of xchg_all_ok below is the barrier after this) */
	for (int b = 0; b < 2; b++)
		CK(cudaMemsetAsync(s->G[b], 0, mf_g_bytes(pl), s->st));
	if (pl->opt_exchange == 0)
		return MFGPU_OK;
	mf_xchg_arena *ar = pl->arena;
	if (ar && ar->mapped < 0)
		return MFGPU_OK;                 /* IPC was found unavailable before: ncclAllReduce */
	if (ar)
		pl->d_sync = ar->sync;
	else
		CK(cudaMalloc((void **)&pl->d_sync, MF_SYNC_WORDS * sizeof(unsigned long long)));
	CK(cudaMemsetAsync(pl->d_sync, 0, MF_SYNC_WORDS * sizeof(unsigned long long), s->st));
	if (ar && ar->mapped > 0) {
		/* the peers' buffers are mapped already: only the barrier after the clean-up above */
		for (int h = 0; h < n; h++) {
			for (int b = 0; b < 2; b++) {
				pl->peer_G[b][h] = ar->peer_G[b][h];
				pl->peer_R[b][h] = ar->peer_R[b][h];
			}
			pl->peer_sync[h] = ar->peer_sync[h];
		}
		if (!xchg_all_ok(pl, s, 1, (int *)pl->d_sync + 14))
			return mf_fail(MFGPU_ECOMM, "exchange set-up: barrier failed");
		pl->xchg_ipc = pl->opt_exchange == 2 ? 1 : pl->opt_exchange == 3 ? 2 : 3;
		return MFGPU_OK;
	}

	XchgHandles mine;
	memset(&mine, 0, sizeof mine);
	mine.ok = cudaIpcGetMemHandle(&mine.g[0], s->G[0]) == cudaSuccess &&
	          cudaIpcGetMemHandle(&mine.g[1], s->G[1]) == cudaSuccess &&
	          cudaIpcGetMemHandle(&mine.r[0], s->R[0]) == cudaSuccess &&
	          cudaIpcGetMemHandle(&mine.r[1], s->R[1]) == cudaSuccess &&
	          cudaIpcGetMemHandle(&mine.sync, pl->d_sync) == cudaSuccess;
	cudaGetLastError();

	XchgHandles *d_all = nullptr;
	std::vector<XchgHandles> all((size_t)n);
	CK(cudaMalloc((void **)&d_all, sizeof(XchgHandles) * (size_t)(n + 1)));
	CK(cudaMemcpyAsync(d_all + n, &mine, sizeof mine, cudaMemcpyHostToDevice, s->st));
	NCK(g_nccl.AllGather(d_all + n, d_all, sizeof(XchgHandles), /*ncclInt8*/ 0, pl->comm->nccl, s->st));
	CK(cudaMemcpyAsync(all.data(), d_all, sizeof(XchgHandles) * (size_t)n, cudaMemcpyDeviceToHost, s->st));
	CK(cudaStreamSynchronize(s->st));

	int ok = 1;
	for (int h = 0; h < n; h++)
		ok &= all[h].ok;
	pl->n_ipc_mapped = 0;
	for (int h = 0; h < n && ok; h++) {
		if (h == me) {
			for (int b = 0; b < 2; b++) {
				pl->peer_G[b][h] = s->G[b];
				pl->peer_R[b][h] = s->R[b];
			}
			pl->peer_sync[h] = pl->d_sync;
			continue;
		}
		void *q[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
		const cudaIpcMemHandle_t *hd[5] = {&all[h].g[0], &all[h].g[1], &all[h].r[0], &all[h].r[1], &all[h].sync};
		for (int k = 0; k < 5 && ok; k++) {
			if (cudaIpcOpenMemHandle(&q[k], *hd[k], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
				cudaGetLastError();
				ok = 0;
			} else {
				pl->ipc_mapped[pl->n_ipc_mapped++] = q[k];
			}
		}
		pl->peer_G[0][h] = (double *)q[0];
		pl->peer_G[1][h] = (double *)q[1];
		pl->peer_R[0][h] = (double *)q[2];
		pl->peer_R[1][h] = (double *)q[3];
		pl->peer_sync[h] = (unsigned long long *)q[4];
	}
	/* everybody, or nobody */
	const bool all_ok = xchg_all_ok(pl, s, ok, (int *)d_all);
	cudaFree(d_all);
	if (!all_ok) {
		for (int k = 0; k < pl->n_ipc_mapped; k++)
			cudaIpcCloseMemHandle(pl->ipc_mapped[k]);
		pl->n_ipc_mapped = 0;
		cudaGetLastError();
		if (ar) {
			ar->mapped = -1;
			pl->d_sync = nullptr;
		}
		if (pl->opt_exchange > 0 && getenv("MFGPU_TRACE"))
			fprintf(stderr, "[mfgpu] peer-memory exchange unavailable, using ncclAllReduce\n");
		return MFGPU_OK;
	}
	if (ar) {
		/* the mappings now belong to the arena and outlive this plan */
		for (int h = 0; h < n; h++) {
			for (int b = 0; b < 2; b++) {
				ar->peer_G[b][h] = pl->peer_G[b][h];
				ar->peer_R[b][h] = pl->peer_R[b][h];
			}
			ar->peer_sync[h] = pl->peer_sync[h];
		}
		memcpy(ar->ipc_mapped, pl->ipc_mapped, sizeof(void *) * (size_t)pl->n_ipc_mapped);
		ar->n_ipc_mapped = pl->n_ipc_mapped;
		pl->n_ipc_mapped = 0;
		ar->mapped = 1;
	}
	/* option exchange: 2 = pull kernel, 3 = push with fences and flags, else low latency */
	pl->xchg_ipc = pl->opt_exchange == 2 ? 1 : pl->opt_exchange == 3 ? 2 : 3;
	return MFGPU_OK;
}

int mf_xchg_setup(mfgpu_plan *pl)
{
	const int rc = xchg_setup_ranks(pl);
	if (pl->comm)
		pl->sh[0].d_sync = pl->d_sync;       /* the rank's one shard waits on the rank's sync block */
	return rc;
}

/* Collective: nobody may unmap or free while a peer's kernels can still read its G. */
void mf_xchg_teardown(mfgpu_plan *pl)
{
	if (!pl->comm || !pl->d_sync)
		return;
	pl->sh[0].d_sync = nullptr;
	mf_shard *s = &pl->sh[0];
	if (s->dev >= 0)
		cudaSetDevice(s->dev);
	if (pl->xchg_ipc && s->st) {
		cudaStreamSynchronize(s->st);
		if (s->st2) cudaStreamSynchronize(s->st2);
		/* a one-word all-reduce as the barrier: a peer joins it only after its kernels */
		if (g_nccl.AllReduce((int *)pl->d_sync + 14, (int *)pl->d_sync + 14, 1, 2, 0, pl->comm->nccl, s->st) == 0)
			cudaStreamSynchronize(s->st);
	}
	for (int k = 0; k < pl->n_ipc_mapped; k++)
		cudaIpcCloseMemHandle(pl->ipc_mapped[k]);
	pl->n_ipc_mapped = 0;
	if (!pl->arena)
		cudaFree(pl->d_sync);            /* else the flag block is the arena's */
	pl->d_sync = nullptr;
	pl->xchg_ipc = 0;
	cudaGetLastError();
}

/* One process, N GPUs: the shards are each other's peers directly (cudaDeviceEnablePeerAccess;
 * no IPC): every shard gets a sync block, and the peer tables of the plan point at the shards'
 * own buffers.  Default: the low-latency exchange, exactly as between processes; option
 * exchange = 2: the event-ordered pull (k_apply_R_peer). */
static int mf_xchg_setup_local(mfgpu_plan *pl)
{
	for (int g = 0; g < pl->nshards; g++) {
		mf_shard *s = &pl->sh[g];
		CK(cudaSetDevice(s->dev));
		if (!s->d_sync)
			CK(cudaMalloc((void **)&s->d_sync, MF_SYNC_WORDS * sizeof(unsigned long long)));
		CK(cudaMemsetAsync(s->d_sync, 0, MF_SYNC_WORDS * sizeof(unsigned long long), s->st));
		for (int b = 0; b < 2; b++) {
			CK(cudaMemsetAsync(s->G[b], 0, mf_g_bytes(pl), s->st));
			pl->peer_G[b][g] = s->G[b];
			pl->peer_R[b][g] = s->R[b];
		}
		pl->peer_sync[g] = s->d_sync;
		CK(cudaStreamSynchronize(s->st));
	}
	pl->peer_events = pl->opt_exchange == 2;
	pl->xchg_ipc = pl->peer_events ? 0 : 3;
	return MFGPU_OK;
}

static int plan_create(mfgpu_plan **out, int users, int items, int features, int nnz,
                       const mfgpu_entry *entries, double alpha, int ngpus, mfgpu_comm *comm)
{
	if (!out)
		return mf_fail(MFGPU_EINVAL, "null plan pointer");
	*out = nullptr;
	if (users < 0 || items < 0 || features < 0 || nnz < 0 || (nnz > 0 && !entries))
		return mf_fail(MFGPU_EINVAL, "negative size or missing entries");
	if ((double)items * ((features + 1) & ~1) >= 2147483647.0 * 4 ||
	    (double)users * ((features + 1) & ~1) >= 2147483647.0 * 4)
		return mf_fail(MFGPU_EINVAL, "factor matrix too large");
	int devs[MF_MAX_SHARDS];
	const int avail = mf_usable_devices(devs, MF_MAX_SHARDS);
	if (avail == 0)
		return mf_fail(MFGPU_ENODEV, "no sm_100 (B200) device visible; this library has no CPU path");
	int nshards = comm ? comm->nranks : (ngpus <= 0 ? avail : ngpus);
	if (!comm && nshards > avail)
		return mf_fail(MFGPU_ENODEV, "%d GPUs requested, %d usable", nshards, avail);
	if (nshards > MF_MAX_SHARDS)
		return mf_fail(MFGPU_EINVAL, "at most %d shards", MF_MAX_SHARDS);

	mfgpu_plan *pl = new (std::nothrow) mfgpu_plan();
	if (!pl)
		return mf_fail(MFGPU_ENOMEM, "out of host memory");
	for (int i = 0; i < MF_MAX_SHARDS; i++)
		pl->sh[i].dev = -1;
	pl->users = users;
	pl->items = items;
	pl->F = features;
	pl->ld = std::max(2, (features + 1) & ~1);
	{
		/* Rows on whole 128-byte lines when that costs at most 1/16 more bytes (F = 200 -> 208,
		 * 1000 -> 1008; not 100 -> 112): a 16-feature tile of a gathered row is then one line
		 * instead of the halves of two (ncu: the dot kernel pulled 2.2 GB through L2 -> SM for the
		 * 1.45 GB it needed; 169 -> 154 us per iteration on the inst60000 shape).  The host copies
		 * stay contiguous: rows are packed / padded by a kernel on the device (pack_rows). */
		const char *env = getenv("MFGPU_LD_ALIGN");      /* 0 / 1: never; default 16 doubles */
		const int al = env ? atoi(env) : 16;
		const int padded = al > 1 ? (features + al - 1) / al * al : pl->ld;
		if (features >= 64 && padded - features <= features / 16)
			pl->ld = padded;
	}
	pl->nnz = nnz;
	pl->alpha = alpha;
	pl->nshards = nshards;
	pl->comm = comm;
	pl->opt_graph = 1;
	pl->opt_graph_iters = 16;
	pl->opt_chain_ring = 1;
	pl->opt_resident = 1;
	pl->opt_res_rows = 0;
	pl->opt_chain_ws = 1;
	pl->opt_xchg_fused = 0;      /* 1: the low-latency exchange inside k_acc (k_acc_ll; measured slower) */
	if (const char *env = getenv("MFGPU_XCHG_FUSED"))
		pl->opt_xchg_fused = atoi(env);
	pl->opt_argmax_items = -1;   /* item-sharded arg-max when users << items (mfgpu_plan_recommend) */
	pl->opt_pair = -1;           /* split mode by shape (mf_configure_shard) */
	pl->opt_pair_warps = 4;
	pl->opt_pair_stages = 2;
	pl->opt_exchange = -1;       /* peer memory when it can be had */

	TraceClock tc;
	/* one pass over the whole entry list (host/mf_prep.c: mf_scan_entries): the cut, the touched
	 * items, the contributors of every item, and -- when the rows never decrease, as in every
	 * file of the reference -- where each user's entries start, so that a shard builds its
	 * pattern from its own entries only */
	int rc = MFGPU_OK;
	int *item_slot = (int *)malloc(sizeof(int) * (size_t)(items ? items : 1));
	int *row_first = (int *)malloc(sizeof(int) * ((size_t)users + 1));
	int *touched = nullptr, *contrib = nullptr;
	mf_scan scan;
	memset(&scan, 0, sizeof scan);
	if (nshards > 1)
		contrib = (int *)calloc((size_t)(items ? items : 1), sizeof(int));
	if (!item_slot || !row_first || (nshards > 1 && !contrib))
		rc = MFGPU_ENOMEM;
	else {
		int per = 1;
		const int cnt = mf_scan_entries(users, items, nnz, entries, nshards, pl->cut, item_slot, row_first, &scan,
		                                contrib, &per);
		if (cnt < 0)
			rc = cnt;
		else if (nshards > 1) {
			pl->n_touched = cnt;
			pl->slots_per_rank = per;
			touched = (int *)malloc(sizeof(int) * (size_t)(cnt ? cnt : 1));
			if (!touched)
				rc = MFGPU_ENOMEM;
			else
				for (int j = 0; j < items; j++)
					if (item_slot[j] >= 0)
						touched[item_slot[j]] = j;
		}
	}
	pl->scan = &scan;
	if (rc != MFGPU_OK)
		mf_fail(rc, rc == MFGPU_EINVAL ? "an entry lies outside the users x items matrix"
		                               : "out of host memory");
	tc.lap("create: user split, items");
	pl->ll_half = (long)mf_g_half(pl);
	{
		const char *env = getenv("MFGPU_COMM_TIMEOUT_MS");
		const double ms = env && atof(env) > 0 ? atof(env) : 30000.0;
		pl->comm_timeout_ns = (unsigned long long)(ms * 1e6);
	}
	if (rc == MFGPU_OK) {
		if (comm) {
			/* rank mode: this process holds one shard, the users of its rank */
			pl->sh[0].user_lo = pl->cut[comm->rank];
			pl->sh[0].user_hi = pl->cut[comm->rank + 1];
			pl->sh[0].me = comm->rank;
			rc = shard_create(pl, &pl->sh[0], comm->device, entries, item_slot, touched, contrib);
			if (rc == MFGPU_OK)
				rc = mf_xchg_setup(pl);
		} else {
			for (int g = 0; g < nshards && rc == MFGPU_OK; g++) {
				pl->sh[g].user_lo = pl->cut[g];
				pl->sh[g].user_hi = pl->cut[g + 1];
				pl->sh[g].me = g;
				rc = shard_create(pl, &pl->sh[g], devs[g], entries, item_slot, touched, contrib);
			}
			/* the peer exchange reads every other shard's G directly */
			for (int g = 0; g < nshards && rc == MFGPU_OK && nshards > 1; g++) {
				cudaSetDevice(devs[g]);
				for (int h = 0; h < nshards; h++) {
					if (h == g) continue;
					int ok = 0;
					cudaDeviceCanAccessPeer(&ok, devs[g], devs[h]);
					if (!ok) {
						rc = mf_fail(MFGPU_ECUDA, "no peer access between devices %d and %d", devs[g], devs[h]);
						break;
					}
					cudaError_t e = cudaDeviceEnablePeerAccess(devs[h], 0);
					if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
						rc = mf_fail(MFGPU_ECUDA, "cudaDeviceEnablePeerAccess: %s", cudaGetErrorString(e));
						break;
					}
					cudaGetLastError();
				}
			}
			if (rc == MFGPU_OK && nshards > 1)
				rc = mf_xchg_setup_local(pl);
		}
	}
	pl->scan = nullptr;
	free(item_slot);
	free(row_first);
	free(touched);
	free(contrib);
	if (rc != MFGPU_OK) {
		char keep[512];
		snprintf(keep, sizeof keep, "%s", mfgpu_last_error());
		mfgpu_plan_destroy(pl);
		return mf_fail(rc, "%s", keep);
	}
	*out = pl;
	return MFGPU_OK;
}

extern "C" int mfgpu_plan_create(mfgpu_plan **out, int users, int items, int features, int nnz,
                                 const mfgpu_entry *entries, double alpha, int ngpus)
{
	return plan_create(out, users, items, features, nnz, entries, alpha, ngpus, nullptr);
}

extern "C" int mfgpu_plan_create_rank(mfgpu_plan **out, int users, int items, int features, int nnz,
                                      const mfgpu_entry *entries, double alpha, mfgpu_comm *comm)
{
	if (!comm)
		return mf_fail(MFGPU_EINVAL, "null communicator");
	return plan_create(out, users, items, features, nnz, entries, alpha, comm->nranks, comm);
}

extern "C" void mfgpu_plan_destroy(mfgpu_plan *pl)
{
	if (!pl)
		return;
	const int n = pl->comm ? 1 : pl->nshards;
	mf_xchg_teardown(pl);            /* rank mode: collective, see there */
	if (!pl->comm)
		for (int g = 0; g < n; g++)
			if (pl->sh[g].dev >= 0 && pl->sh[g].d_sync) {
				cudaSetDevice(pl->sh[g].dev);
				if (pl->sh[g].st) cudaStreamSynchronize(pl->sh[g].st);
			}
	for (int g = 0; g < n; g++)
		if (pl->sh[g].dev >= 0)
			shard_destroy(&pl->sh[g], pl->comm == nullptr);
	if (pl->arena)
		pl->arena->in_use = 0;       /* the next plan of this communicator may borrow it */
	delete pl;
}

static int local_shards(const mfgpu_plan *pl) { return pl->comm ? 1 : pl->nshards; }

extern "C" int mfgpu_plan_set_factors(mfgpu_plan *pl, const double *L, const double *R)
{
	if (!pl || !L || !R)
		return mf_fail(MFGPU_EINVAL, "null argument");
	const size_t F = (size_t)pl->F, ld = (size_t)pl->ld;
	for (int g = 0; g < local_shards(pl); g++) {
		mf_shard *s = &pl->sh[g];
		CK(cudaSetDevice(s->dev));
		const size_t nu = (size_t)(s->user_hi - s->user_lo);
		if (F && nu)
			TRY(upload_rows(s->L[0], s->L[1], ld, L + (size_t)s->user_lo * F, F, nu, s->st));
		if (F && pl->items)
			TRY(upload_rows(s->R[0], s->R[1], ld, R, F, (size_t)pl->items, s->st));
		/* rows without a nonzero are never rewritten, so both buffers start equal */
		CK(cudaMemcpyAsync(s->L[1], s->L[0], nu * ld * 8, cudaMemcpyDeviceToDevice, s->st));
		CK(cudaMemcpyAsync(s->R[1], s->R[0], (size_t)pl->items * ld * 8, cudaMemcpyDeviceToDevice, s->st));
	}
	for (int g = 0; g < local_shards(pl); g++)
		CK(cudaStreamSynchronize(pl->sh[g].st));
	pl->cur = 0;
	pl->have_factors = 1;
	return MFGPU_OK;
}

/* ---- initial factors generated straight into the device buffers ------------------------------
 *
 * mat2d_random_fill_LR + mat2d_transpose (mat2d.c:61-72,115-124) without the host matrices:
 * the calling thread produces the random() stream block by block into page-locked staging
 * (host/mf_ingest.c: mf_rand_block), each block crosses as 4-byte draws and a kernel turns it
 * into ((double)r / RAND_MAX) / F -- two IEEE divisions, the reference's expression -- at the
 * element the draw belongs to: draw d < U*F is L[d / F][d % F], draw U*F + e is R[e % I][e / I]
 * (R is drawn feature by feature and used transposed).  Block b+1 is generated while block b
 * is copied and converted. */
#define MF_INIT_BLOCK ((size_t)1 << 22)

__global__ void __launch_bounds__(256)
k_init_factors(const uint32_t *__restrict__ draws, unsigned long long d0, unsigned count, unsigned long long nl,
               int F, int I, int ld, int user_lo, int user_hi, double *__restrict__ L, double *__restrict__ R)
{
	const unsigned t = blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= count)
		return;
	const unsigned long long d = d0 + t;
	const double v = __ddiv_rn(__ddiv_rn((double)draws[t], 2147483647.0), (double)F);
	if (d < nl) {
		const unsigned long long i = d / (unsigned)F;
		const int k = (int)(d - i * (unsigned)F);
		if (i >= (unsigned long long)user_lo && i < (unsigned long long)user_hi)
			L[(size_t)(i - (unsigned long long)user_lo) * ld + k] = v;
	} else {
		const unsigned long long e = d - nl, k = e / (unsigned)I, j = e - k * (unsigned)I;
		R[(size_t)j * ld + k] = v;
	}
}

struct InitStaging {
	uint32_t *host[2] = {nullptr, nullptr};
	uint32_t *dev[MF_MAX_SHARDS][2] = {};
	cudaEvent_t done[MF_MAX_SHARDS][2] = {};
	int devs[MF_MAX_SHARDS] = {};
	int n = 0;
	~InitStaging()
	{
		for (int g = 0; g < n; g++) {
			cudaSetDevice(devs[g]);
			for (int b = 0; b < 2; b++) {
				if (done[g][b])
					cudaEventDestroy(done[g][b]);
				if (dev[g][b])
					cudaFree(dev[g][b]);
			}
		}
		for (int b = 0; b < 2; b++)
			if (host[b])
				cudaFreeHost(host[b]);
	}
};

extern "C" int mfgpu_plan_init_factors(mfgpu_plan *pl)
{
	if (!pl)
		return mf_fail(MFGPU_EINVAL, "null plan");
	const unsigned long long nl = (unsigned long long)pl->users * pl->F, nr = (unsigned long long)pl->items * pl->F;
	const unsigned long long total = nl + nr;
	const int ns = local_shards(pl);
	mf_rand gen;
	mf_rand_seed(&gen, 0);                               /* mat2d.c:62: srandom(0) */
	if (total) {
		InitStaging sg;
		const size_t cap = (size_t)std::min<unsigned long long>(total, MF_INIT_BLOCK);
		for (int b = 0; b < 2; b++)
			CK(cudaHostAlloc((void **)&sg.host[b], cap * sizeof(uint32_t), cudaHostAllocDefault));
		for (int g = 0; g < ns; g++) {
			CK(cudaSetDevice(pl->sh[g].dev));
			sg.devs[g] = pl->sh[g].dev;
			sg.n = g + 1;
			for (int b = 0; b < 2; b++) {
				CK(cudaMalloc((void **)&sg.dev[g][b], cap * sizeof(uint32_t)));
				CK(cudaEventCreateWithFlags(&sg.done[g][b], cudaEventDisableTiming));
			}
		}
		unsigned long long d0 = 0;
		for (int blk = 0; d0 < total; blk++) {
			const int b = blk & 1;
			const unsigned count = (unsigned)std::min<unsigned long long>(total - d0, cap);
			if (blk >= 2)                            /* the staging pair is free once block blk-2 has been copied */
				for (int g = 0; g < ns; g++)
					CK(cudaEventSynchronize(sg.done[g][b]));
			mf_rand_block(&gen, sg.host[b], count);
			for (int g = 0; g < ns; g++) {
				mf_shard *s = &pl->sh[g];
				CK(cudaSetDevice(s->dev));
				CK(cudaMemcpyAsync(sg.dev[g][b], sg.host[b], (size_t)count * sizeof(uint32_t),
				                   cudaMemcpyHostToDevice, s->st));
				k_init_factors<<<(count + 255) / 256, 256, 0, s->st>>>(sg.dev[g][b], d0, count, nl, pl->F, pl->items,
				                                                       pl->ld, s->user_lo, s->user_hi, s->L[0], s->R[0]);
				CK(cudaGetLastError());
				CK(cudaEventRecord(sg.done[g][b], s->st));
			}
			d0 += count;
		}
		for (int g = 0; g < ns; g++) {
			CK(cudaSetDevice(pl->sh[g].dev));
			CK(cudaStreamSynchronize(pl->sh[g].st));
		}
	}
	for (int g = 0; g < ns; g++) {
		mf_shard *s = &pl->sh[g];
		CK(cudaSetDevice(s->dev));
		const size_t nu = (size_t)(s->user_hi - s->user_lo);
		/* rows without a nonzero are never rewritten, so both buffers start equal */
		CK(cudaMemcpyAsync(s->L[1], s->L[0], nu * pl->ld * 8, cudaMemcpyDeviceToDevice, s->st));
		CK(cudaMemcpyAsync(s->R[1], s->R[0], (size_t)pl->items * pl->ld * 8, cudaMemcpyDeviceToDevice, s->st));
	}
	for (int g = 0; g < ns; g++)
		CK(cudaStreamSynchronize(pl->sh[g].st));
	pl->cur = 0;
	pl->have_factors = 1;
	return MFGPU_OK;
}

/* device -> host copies of the current factors; side = true: on the shard's second stream,
 * behind everything enqueued on the first so far, so that they run beside what is enqueued on
 * the first stream next (the arg-max) */
static int enqueue_get_factors(mfgpu_plan *pl, double *L, double *R, bool side)
{
	const size_t F = (size_t)pl->F, ld = (size_t)pl->ld;
	const int c = pl->cur;
	for (int g = 0; g < local_shards(pl); g++) {
		mf_shard *s = &pl->sh[g];
		CK(cudaSetDevice(s->dev));
		cudaStream_t st = s->st;
		if (side) {
			CK(cudaEventRecord(s->ev_a, s->st));
			CK(cudaStreamWaitEvent(s->st2, s->ev_a, 0));
			st = s->st2;
		}
		const size_t nu = (size_t)(s->user_hi - s->user_lo);
		if (L && F && nu)
			TRY(download_rows(L + (size_t)s->user_lo * F, s->L[c], s->L[1 - c], ld, F, nu, st));
		if (R && g == 0 && F && pl->items)
			TRY(download_rows(R, s->R[c], s->R[1 - c], ld, F, (size_t)pl->items, st));
	}
	return MFGPU_OK;
}

extern "C" int mfgpu_plan_get_factors(mfgpu_plan *pl, double *L, double *R)
{
	if (!pl)
		return mf_fail(MFGPU_EINVAL, "null plan");
	TRY(enqueue_get_factors(pl, L, R, false));
	for (int g = 0; g < local_shards(pl); g++)
		CK(cudaStreamSynchronize(pl->sh[g].st));
	return mf_check_comm_error(pl);
}

extern "C" int mfgpu_plan_get_results(mfgpu_plan *pl, double *L, double *R, int *best)
{
	if (!pl || !best)
		return mf_fail(MFGPU_EINVAL, "null argument");
	TRY(enqueue_get_factors(pl, L, R, true));        /* copy engine ... */
	TRY(mfgpu_plan_recommend(pl, best));             /* ... beside the arg-max kernels */
	for (int g = 0; g < local_shards(pl); g++) {
		CK(cudaSetDevice(pl->sh[g].dev));
		CK(cudaStreamSynchronize(pl->sh[g].st2));
	}
	return mf_check_comm_error(pl);
}

extern "C" int mfgpu_plan_sync(mfgpu_plan *pl)
{
	if (!pl)
		return mf_fail(MFGPU_EINVAL, "null plan");
	for (int g = 0; g < local_shards(pl); g++) {
		CK(cudaSetDevice(pl->sh[g].dev));
		CK(cudaStreamSynchronize(pl->sh[g].st));
	}
	return mf_check_comm_error(pl);
}

extern "C" int mfgpu_plan_user_range(mfgpu_plan *pl, int shard, int *lo, int *hi)
{
	if (!pl || shard < 0 || shard >= local_shards(pl))
		return mf_fail(MFGPU_EINVAL, "bad shard index");
	if (lo) *lo = pl->sh[shard].user_lo;
	if (hi) *hi = pl->sh[shard].user_hi;
	return MFGPU_OK;
}

extern "C" long long mfgpu_plan_launch_count(mfgpu_plan *pl) { return pl ? pl->launches : 0; }

extern "C" int mfgpu_plan_exchange_kind(mfgpu_plan *pl)
{
	if (!pl || pl->nshards < 2)
		return 0;
	if (!pl->comm)
		return pl->peer_events ? 3 : 5;
	return pl->xchg_ipc == 3 ? 5 : pl->xchg_ipc == 2 ? 4 : pl->xchg_ipc ? 2 : 1;
}

extern "C" long long mfgpu_plan_bytes_per_iter(mfgpu_plan *pl)
{
	if (!pl)
		return 0;
	/* 16*F*(U_t + I_t) + 12*nnz over the shards held by this process */
	long long b = 0;
	for (int g = 0; g < local_shards(pl); g++) {
		const mf_pattern &p = pl->sh[g].pat;
		b += 16LL * pl->F * ((long long)p.n_rows + p.n_cols) + 12LL * p.nnz;
	}
	return b;
}

extern "C" int mfgpu_plan_kernel_bytes(mfgpu_plan *pl, long long bytes[5])
{
	if (!pl || !bytes)
		return mf_fail(MFGPU_EINVAL, "null argument");
	for (int k = 0; k < 5; k++)
		bytes[k] = pl->sh[0].kbytes[k];
	return MFGPU_OK;
}

extern "C" int mfgpu_plan_kernel_gather_bytes(mfgpu_plan *pl, long long bytes[5])
{
	if (!pl || !bytes)
		return mf_fail(MFGPU_EINVAL, "null argument");
	for (int k = 0; k < 5; k++)
		bytes[k] = pl->sh[0].gbytes[k];
	return MFGPU_OK;
}

extern "C" int mfgpu_plan_set_option(mfgpu_plan *pl, const char *name, int value)
{
	if (!pl || !name)
		return mf_fail(MFGPU_EINVAL, "null argument");
	if (!strcmp(name, "graph")) pl->opt_graph = value;
	else if (!strcmp(name, "graph_iters")) pl->opt_graph_iters = std::max(2, value & ~1);
	else if (!strcmp(name, "long_row")) pl->opt_long_row = value;
	else if (!strcmp(name, "chunk")) pl->opt_ch = value;
	else if (!strcmp(name, "long_col")) pl->opt_long_col = value;
	else if (!strcmp(name, "chain_ring")) pl->opt_chain_ring = value;
	else if (!strcmp(name, "resident")) pl->opt_resident = value;
	else if (!strcmp(name, "res_rows")) pl->opt_res_rows = value;
	else if (!strcmp(name, "chain_ws")) pl->opt_chain_ws = value;
	else if (!strcmp(name, "dot_stages")) pl->opt_dot_stages = value;
	else if (!strcmp(name, "acc_rows")) pl->opt_acc_rows = value;
	else if (!strcmp(name, "cs_store")) pl->opt_cs_store = value;
	else if (!strcmp(name, "pair")) pl->opt_pair = value;
	else if (!strcmp(name, "pair_stages")) pl->opt_pair_stages = value;
	else if (!strcmp(name, "pair_warps")) pl->opt_pair_warps = value;
	else if (!strcmp(name, "acc_gw")) pl->opt_acc_gw = value;
	else if (!strcmp(name, "argmax_items")) pl->opt_argmax_items = value;
	else if (!strcmp(name, "xchg_fused")) pl->opt_xchg_fused = value;
	else if (!strcmp(name, "comm_timeout_ms")) pl->comm_timeout_ns = (unsigned long long)std::max(1, value) * 1000000ull;
	else if (!strcmp(name, "exchange")) {
		/* rank mode, collective: 0 = ncclAllReduce; peer memory when available: 1 / -1 = low
		 * latency (value, tag) exchange (k_xchg_ll), 2 = pull (k_apply_R_xchg), 3 = push with
		 * fences and flags (k_reduce_bcast) */
		pl->opt_exchange = value;
		if (pl->comm) {
			mf_xchg_teardown(pl);
			TRY(mf_xchg_setup(pl));
		} else if (pl->nshards > 1) {
			for (int g = 0; g < pl->nshards; g++) {
				CK(cudaSetDevice(pl->sh[g].dev));
				CK(cudaStreamSynchronize(pl->sh[g].st));
			}
			TRY(mf_xchg_setup_local(pl));
		}
	}
	else
		return mf_fail(MFGPU_EINVAL, "unknown option '%s'", name);
	for (int g = 0; g < local_shards(pl); g++) {
		CK(cudaSetDevice(pl->sh[g].dev));
		CK(cudaStreamSynchronize(pl->sh[g].st));
		if (pl->sh[g].d_ll_args) {           /* k_acc_ll's copy of the exchange's arguments: made again by the next iterate */
			CK(cudaFree(pl->sh[g].d_ll_args));
			pl->sh[g].d_ll_args = nullptr;
		}
		TRY(mf_configure_shard(pl, &pl->sh[g]));
	}
	return MFGPU_OK;
}

extern "C" int mfgpu_plan_save_factors(mfgpu_plan *pl)
{
	if (!pl || !pl->have_factors)
		return mf_fail(MFGPU_EINVAL, "no factors to save");
	for (int g = 0; g < local_shards(pl); g++) {
		mf_shard *s = &pl->sh[g];
		CK(cudaSetDevice(s->dev));
		const size_t lb = ((size_t)(s->user_hi - s->user_lo) * pl->ld + 2) * sizeof(double);
		const size_t rb = ((size_t)pl->items * pl->ld + 2) * sizeof(double);
		if (!s->L_keep) TRY(dev_alloc((void **)&s->L_keep, lb, s->st));
		if (!s->R_keep) TRY(dev_alloc((void **)&s->R_keep, rb, s->st));
		CK(cudaMemcpyAsync(s->L_keep, s->L[pl->cur], lb, cudaMemcpyDeviceToDevice, s->st));
		CK(cudaMemcpyAsync(s->R_keep, s->R[pl->cur], rb, cudaMemcpyDeviceToDevice, s->st));
	}
	return MFGPU_OK;
}

extern "C" int mfgpu_plan_restore_factors(mfgpu_plan *pl)
{
	if (!pl || !pl->sh[0].L_keep)
		return mf_fail(MFGPU_EINVAL, "mfgpu_plan_save_factors has not been called");
	for (int g = 0; g < local_shards(pl); g++) {
		mf_shard *s = &pl->sh[g];
		CK(cudaSetDevice(s->dev));
		const size_t lb = ((size_t)(s->user_hi - s->user_lo) * pl->ld + 2) * sizeof(double);
		const size_t rb = ((size_t)pl->items * pl->ld + 2) * sizeof(double);
		for (int b = 0; b < 2; b++) {
			CK(cudaMemcpyAsync(s->L[b], s->L_keep, lb, cudaMemcpyDeviceToDevice, s->st));
			CK(cudaMemcpyAsync(s->R[b], s->R_keep, rb, cudaMemcpyDeviceToDevice, s->st));
		}
	}
	return MFGPU_OK;
}

extern "C" int mfgpu_plan_timer_start(mfgpu_plan *pl)
{
	if (!pl)
		return mf_fail(MFGPU_EINVAL, "null plan");
	for (int g = 0; g < local_shards(pl); g++) {
		CK(cudaSetDevice(pl->sh[g].dev));
		CK(cudaEventRecord(pl->sh[g].ev_t0, pl->sh[g].st));
	}
	return MFGPU_OK;
}

extern "C" int mfgpu_plan_timer_stop(mfgpu_plan *pl, float *ms)
{
	if (!pl || !ms)
		return mf_fail(MFGPU_EINVAL, "null argument");
	for (int g = 0; g < local_shards(pl); g++) {
		CK(cudaSetDevice(pl->sh[g].dev));
		CK(cudaEventRecord(pl->sh[g].ev_t1, pl->sh[g].st));
	}
	float worst = 0.f;
	for (int g = 0; g < local_shards(pl); g++) {
		CK(cudaSetDevice(pl->sh[g].dev));
		CK(cudaEventSynchronize(pl->sh[g].ev_t1));
		float t = 0.f;
		CK(cudaEventElapsedTime(&t, pl->sh[g].ev_t0, pl->sh[g].ev_t1));
		worst = std::max(worst, t);
	}
	*ms = worst;
	return mf_check_comm_error(pl);
}
