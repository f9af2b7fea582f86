/*
 * csrc/mf_types.h -- kernel argument blocks shared by the host-side translation
 * units; the kernels themselves (mf_loop.cuh, mf_argmax.cuh) are included by
 * mf_iter.cu only.
 */
#pragma once

namespace mf {

constexpr int MAX_PEERS = 16;   /* = MF_MAX_SHARDS (mf_internal.h) */

/* Rows (users with >= 1 nonzero) are stored longest first and fall in three classes:
 *   [0, n_deep_rows)        dots by k_dot, accumulation by k_chain<deep>
 *   [n_deep_rows, n_long)   dots by k_dot, accumulation by k_chain<bulk>
 *   [n_long, n_rows)        everything by k_update_L (one warp per row)
 * Columns (items with >= 1 nonzero), longest first: [0, n_deep_cols) k_chain<deep>, rest bulk. */
struct LoopArgs {
	int F;            /* features                                         */
	int ld;           /* row stride of L and R in doubles (even, >= F)    */
	int CH;           /* nonzeros per chunk of the dot phase (<= 32)      */
	int NT;           /* tiles in the dot-phase ring of a warp (2..4)     */
	int CR, RST;      /* k_update_L_res: resident rows per chunk, their stride in doubles */
	int n_rows, n_long, n_deep_rows;
	int n_cols, n_deep_cols;
	int n_chunks;     /* chunks of the rows [0, n_long) for k_dot         */
	int user_lo;      /* first user of this shard; L holds rows user_lo.. */
	int cs_store;     /* 1: rows of L_new are stored evict-first (st.global.cs): they are not read again before the next iteration, L_old is (by the by-item pass) */
	double alpha2;    /* alpha * 2, exact                                 */
	const int *row_id, *row_ptr, *col_idx, *epos;
	const double *val;
	const int *col_id, *col_ptr, *row_idx;
	const int *chunk_row, *chunk_pos;   /* stored row and first by-user position of a chunk */
	const int *item_slot;   /* N shards: item -> row of G, else nullptr   */
	double *e_csc;    /* e_n at by-item positions                         */
	double *e_csr;    /* e_n at by-user positions (rows [0, n_long) only) */
};

/* One side of a k_chain launch: for every stored row/column r in [begin, end) and every
 * group of 32 features,  out[own(r)][k] = init[own(r)][k] + sum_n e[n] * src[idx[n]][k]
 * over n in [ptr[r], ptr[r+1]), added strictly in that order. */
struct ChainSide {
	int begin, end;
	const int *ptr, *idx, *id;
	const double *e;
	const double *init;       /* own matrix (old values); nullptr: start from 0.0      */
	const double *src;        /* the other factor (old values), gathered by idx         */
	double *out;
	int own_off, src_off;     /* first row held by init/out and by src                  */
	const int *out_slot;      /* nullptr, or out row = out_slot[id[r]] (N shards: G)   */
	/* rank mode, push exchange: the finished row goes straight into the receive buffer of the
	 * rank that owns the item (a store over NVLink peer memory): out = peer_out[out_owner[id[r]]],
	 * row = out_slot[id[r]] inside that buffer.  out_owner == nullptr: plain `out`. */
	const int *out_owner;
	double *peer_out[MAX_PEERS];
	/* low-latency exchange: the row is stored as (value, tag) pairs, tag = *ll_epoch + 1 (the
	 * iteration being computed); element (row, k) lives at doubles 2*(row*ld + k).  nullptr: off */
	const unsigned long long *ll_epoch;
	long ll_half;             /* the pair buffers are double: iteration t uses the half (t & 1) * ll_half doubles */
};

constexpr int AM_TI = 64;      /* users per CTA           */
constexpr int AM_TJ = 64;      /* items per tile          */
constexpr int AM_KC = 16;      /* features per smem panel */
constexpr int AM_LDS = AM_TI + 2;

struct ArgmaxArgs {
	int users_lo, users_hi;   /* users handled: [users_lo, users_hi); L row = user - l_row0 */
	int l_row0;
	int items, F, ld;
	int tiles_per_split, n_splits;
	int tile_begin, tile_end; /* item tiles handled by this launch: [tile_begin, tile_end) (an item block of one GPU, or all) */
	const int *mask_ptr;      /* [users+1] global user index */
	const int *mask_col;
	const int *first_free;    /* [users] */
	double *part_v;           /* [(users_hi-users_lo) * n_splits] */
	int *part_j;
	int *nan_first;           /* [users_hi-users_lo], zeroed: B[i][first_free[i]] is NaN */
};

} // namespace mf
