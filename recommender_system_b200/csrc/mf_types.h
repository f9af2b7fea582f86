/*
 * csrc/mf_types.h -- kernel argument blocks shared by the host-side translation
 * units; the kernels themselves (mf_loop.cuh, mf_argmax.cuh) are included by
 * mf_iter.cu only.
 */
#pragma once

namespace mf {

struct LoopArgs {
	int F;            /* features                                         */
	int ld;           /* row stride of L and R in doubles (even, >= F)    */
	int S;            /* row stride of the staged R rows in shared memory */
	int CH;           /* nonzeros staged per chunk (<= 32)                */
	int n_rows;       /* stored user rows (users with >= 1 nonzero)       */
	int n_long;       /* the first n_long stored rows get a whole CTA     */
	int n_cols;       /* stored item columns                              */
	int n_long_cols;  /* the first n_long_cols get the deep (32-row) pipeline */
	int user_lo;      /* first user of this shard; L holds rows user_lo.. */
	double alpha2;    /* alpha * 2, exact                                 */
	const int *row_id, *row_ptr, *col_idx, *epos;
	const double *val;
	const int *col_id, *col_ptr, *row_idx;
	const int *item_slot;   /* N-GPU: item -> row of G, else nullptr      */
	double *e_csc;    /* e_n at by-item positions                         */
	double *e_csr;    /* e_n at by-user positions (long rows only)        */
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
	const int *mask_ptr;      /* [users+1] global user index */
	const int *mask_col;
	const int *first_free;    /* [users] */
	double *part_v;           /* [(users_hi-users_lo) * n_splits] */
	int *part_j;
	int *nan_first;           /* [users_hi-users_lo], zeroed: B[i][first_free[i]] is NaN */
};

} // namespace mf
