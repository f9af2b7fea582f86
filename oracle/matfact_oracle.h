/*
 * oracle/matfact_oracle.h -- TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement of the reference's serial matFact algorithm, used as the
 * checker for the CUDA path.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may call into this.  The
 * product (libmatfact_b200.so, the matFact CLI) never links or loads it.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks this code against every
 * usable samples/X.out golden of the reference (11 files), the three
 * samples/inst{0,1,2}.mats traces, and (when /root/reference is present) the
 * stdout of the unmodified reference serial program compiled into oracle/_ref/.
 */
#ifndef MATFACT_ORACLE_H
#define MATFACT_ORACLE_H

#ifdef __cplusplus
extern "C" {
#endif

/* Same 16-byte layout as the reference's non_zero_entry (datatypes.h:10-15). */
typedef struct {
	int row;
	int col;
	double value;
} oracle_entry;

/* mat2d_random_fill_LR + mat2d_transpose (mat2d.c:61-72,115-124; call sites
 * matFact.c:113-120).  L is users x features, R is items x features. */
void oracle_init_LR(int users, int items, int features, double *L, double *R);

/* matrix_factorization without the trailing mat2d_prod (matFact.c:29-53).
 * L, R updated in place.  Returns 0, or -1 if scratch allocation fails. */
int oracle_factorize(int users, int items, int features, int nnz,
                     const oracle_entry *entries, int iters, double alpha,
                     double *L, double *R);

/* The N-shard ordering the multi-GPU path uses: users are
 * cut at shard_lo[0..nshards] (contiguous ranges), every shard accumulates its
 * R gradient from 0.0 in entry order, gradients are summed in shard order and
 * added to the old R.  L is updated exactly as in oracle_factorize. */
int oracle_factorize_sharded(int users, int items, int features, int nnz,
                             const oracle_entry *entries, int iters, double alpha,
                             double *L, double *R,
                             int nshards, const int *shard_lo);

/* mat2d_prod + print_output (mat2d.c:100-113, matFact.c:10-27) fused row by
 * row so that B is never stored.  best[i] = recommended item or -1 when the
 * reference prints nothing for user i. */
int oracle_recommend(int users, int items, int features, int nnz,
                     const oracle_entry *entries,
                     const double *L, const double *R, int *best);

/* One row of B = L * R^T with the reference's summation order (debug aid). */
void oracle_predict_row(int items, int features, const double *Li,
                        const double *R, double *out);

#ifdef __cplusplus
}
#endif
#endif
