/*
 * host/mf_prep.h -- host-side (plain C) preparation of the sparse pattern.
 *
 * The reference walks A as a COO list in file order (matFact.c:41-53).  The
 * device kernels need the same nonzeros grouped by user (for the L update) and
 * by item (for the R update) WITHOUT changing the order in which the updates
 * of one L row / one R row are accumulated, because that order fixes the bits
 * of the result.  Both groupings are therefore stable with respect to the file
 * order.  Nothing in this file touches the GPU.
 */
#ifndef MF_PREP_H
#define MF_PREP_H

#include <stddef.h>

#include "matfact_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
	int users, items, nnz;      /* nnz = nonzeros of THIS shard                      */
	int user_lo, user_hi;       /* users owned by this shard: [user_lo, user_hi)     */

	/* by-user grouping; rows are the users with >=1 nonzero, stored
	 * longest row first so that the longest serial chains start first */
	int n_rows;
	int *row_id;                /* [n_rows]   user index of stored row r             */
	int *row_ptr;               /* [n_rows+1] start of row r in col_idx/val/epos     */
	int *col_idx;               /* [nnz]      item of each nonzero (file order in row) */
	double *val;                /* [nnz]      rating a_ij                            */
	int *epos;                  /* [nnz]      by-user position -> by-item position   */

	/* by-item grouping, longest column first */
	int n_cols;
	int *col_id;                /* [n_cols]   item index of stored column c          */
	int *col_ptr;               /* [n_cols+1]                                        */
	int *row_idx;               /* [nnz]      user of each nonzero (file order in column) */

	/* rated-item mask of print_output (matFact.c:14-22), users of this shard */
	int *mask_ptr;              /* [users+1]  (entries outside [user_lo,user_hi] repeat the ends) */
	int *mask_col;              /* [mask_ptr[users]] ascending items per user        */
	int *first_free;            /* [users]    first item the cursor does not skip, -1 if none */

	int max_row_len, max_col_len;

	/* caller-provided workspace the arrays above live in (mf_pattern_build_ws), else NULL */
	char *ws;
	size_t ws_bytes, ws_used;
} mf_pattern;

/* Build the pattern of the entries whose row lies in [user_lo, user_hi).
 * `entries` is the full list in file order.  Returns 0 or an MFGPU_E* code. */
int mf_pattern_build(mf_pattern *p, int users, int items, int nnz,
                     const mfgpu_entry *entries, int user_lo, int user_hi);
void mf_pattern_free(mf_pattern *p);

/* The same with every array of the pattern carved out of the caller's workspace `ws` (e.g.
 * page-locked memory kept from one plan to the next: no first-touch page faults, and the
 * upload is a real DMA).  mf_pattern_ws_bytes bounds what a build can need; mf_pattern_free
 * then leaves the workspace to the caller. */
size_t mf_pattern_ws_bytes(int users, int items, int nnz, int user_lo, int user_hi);
int mf_pattern_build_ws(mf_pattern *p, int users, int items, int nnz,
                        const mfgpu_entry *entries, int user_lo, int user_hi,
                        void *ws, size_t ws_bytes);

/* What one pass over the whole entry list learns (mf_scan_entries): enough for a shard to build
 * its pattern from ITS entries only.  by_row: rows never decrease, so the entries of the users
 * [lo, hi) are the contiguous slice [first_of_row(lo), first_of_row(hi)) (row_first has users+1
 * ints; valid only when by_row).  matched: mf_mask_prefix of the whole list. */
typedef struct {
	int by_row;
	int matched;
	const int *row_first;       /* [users+1] first entry of each user's rows, when by_row */
} mf_scan;

/* The same build from the knowledge of a scan: only the entries of the slice are read. */
int mf_pattern_build_scan(mf_pattern *p, int users, int items, int nnz,
                          const mfgpu_entry *entries, int user_lo, int user_hi,
                          void *ws, size_t ws_bytes, const mf_scan *scan);

/* One threaded pass over the entries for everything plan creation needs from the WHOLE list
 * (each used to be a pass of its own): validation, nonzeros per user -> cut[nshards+1]
 * (mf_split_users), items with a nonzero -> item_slot[items] and their number (mf_touched_items),
 * whether the rows never decrease and where each user's entries start (row_first[users+1],
 * caller-provided), the matched prefix of print_output (mf_mask_prefix); and a second pass, only
 * when contrib != NULL (nshards > 1), for the contributors of every touched item
 * (mf_exchange_layout; contrib has room for `items` ints, *slots_per_rank is returned).
 * Returns the number of touched items or an MFGPU_E* code (< 0). */
int mf_scan_entries(int users, int items, int nnz, const mfgpu_entry *entries, int nshards,
                    int *cut, int *item_slot, int *row_first, mf_scan *scan,
                    int *contrib, int *slots_per_rank);

/* Number of leading entries the cursor of print_output (matFact.c:14-22) ever
 * matches: the longest prefix that is strictly increasing in (row, col).  For
 * the (row, col)-sorted, duplicate-free files of the reference this is nnz. */
int mf_mask_prefix(int users, int items, int nnz, const mfgpu_entry *entries);

/* Contiguous user ranges with balanced nonzero counts (SURVEY.md section 8e;
 * the reference's equal-count split is mpiutil.h:8-13 BLOCK_LOW).
 * cut[] has nshards+1 elements, cut[0]=0, cut[nshards]=users. */
int mf_split_users(int users, int nnz, const mfgpu_entry *entries, int nshards, int *cut);

/* Items with at least one nonzero anywhere, ascending; returns the count.
 * item_slot[j] = position of item j in that list or -1. */
int mf_touched_items(int items, int nnz, const mfgpu_entry *entries, int *item_slot);

/* Rank-mode exchange layout: slots per reducing rank (returned) and, per touched item, the set
 * of ranks that hold a nonzero of it (bit r of contrib[slot]); see mf_prep.c. */
int mf_exchange_layout(int users, int nnz, const mfgpu_entry *entries, int nranks, const int *cut,
                       const int *item_slot, int n_touched, int *contrib);

#ifdef __cplusplus
}
#endif
#endif
