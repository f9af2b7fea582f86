/*
 * oracle/matfact_oracle.c -- TEST INFRASTRUCTURE ONLY (see matfact_oracle.h).
 *
 * Plain-C restatement of the reference's serial factorization.  Every function
 * names the reference lines whose arithmetic ORDER it reproduces; the order is
 * what makes the doubles bit-identical, so nothing here may be re-associated,
 * vectorised with a different summation tree, or contracted into FMA.
 * Build with:  gcc -O2 -ffp-contract=off   (oracle/Makefile).
 */
#include "matfact_oracle.h"

#include <stdlib.h>
#include <string.h>

/* mat2d.h:4  RAND01 ; mat2d.c:61-72 fills L (users x F) and then R_init
 * (F x items) row-major from one srandom(0) stream; matFact.c:118-120 then
 * transposes R_init into R (items x F).  Here R is written straight into its
 * transposed position while the draw order stays that of the reference. */
void oracle_init_LR(int users, int items, int features, double *L, double *R)
{
	const double norm = (double)features;
	srandom(0);
	for (long p = 0; p < (long)users * features; p++)
		L[p] = ((double)random() / (double)RAND_MAX) / norm;
	for (int k = 0; k < features; k++)
		for (int j = 0; j < items; j++)
			R[(long)j * features + k] = ((double)random() / (double)RAND_MAX) / norm;
}

/* mat2d.c:126-139: result starts at 0 and takes row[k]*col[k] for ascending k,
 * multiply and add rounded separately. */
static double seq_dot(const double *a, const double *b, int n)
{
	double s = 0;
	for (int k = 0; k < n; k++)
		s += a[k] * b[k];
	return s;
}

/* matFact.c:36-53.  Per iteration the reference snapshots L and R
 * (mat2d_copy, :38-39), then walks the entries in file order; for entry
 * (i, j, a):  value = alpha*2*(a - dot(Ls_i, Rs_j))            (:45)
 *             L[i][k] = L[i][k] - value * (-Rs[j][k])          (:48-49)
 *             R[j][k] = R[j][k] - value * (-Ls[i][k])          (:50-51)
 * x - v*(-y) is evaluated as x - (-(v*y)); negation is exact, so it equals
 * x + rn(v*y) bit for bit -- written that way below. */
int oracle_factorize(int users, int items, int features, int nnz,
                     const oracle_entry *entries, int iters, double alpha,
                     double *L, double *R)
{
	const size_t nl = (size_t)users * features, nr = (size_t)items * features;
	double *Ls = malloc((nl ? nl : 1) * sizeof *Ls);
	double *Rs = malloc((nr ? nr : 1) * sizeof *Rs);
	if (!Ls || !Rs) { free(Ls); free(Rs); return -1; }

	for (int it = 0; it < iters; it++) {
		memcpy(Ls, L, nl * sizeof *Ls);
		memcpy(Rs, R, nr * sizeof *Rs);
		for (int n = 0; n < nnz; n++) {
			const size_t li = (size_t)entries[n].row * features;
			const size_t rj = (size_t)entries[n].col * features;
			const double step = alpha * 2 * (entries[n].value - seq_dot(Ls + li, Rs + rj, features));
			for (int k = 0; k < features; k++) {
				L[li + k] = L[li + k] + step * Rs[rj + k];
				R[rj + k] = R[rj + k] + step * Ls[li + k];
			}
		}
	}
	free(Ls);
	free(Rs);
	return 0;
}

/* SURVEY.md section 8(e) ordering, i.e. what N GPUs compute: shard s owns the users in
 * [shard_lo[s], shard_lo[s+1]); its R gradient G_s starts from 0.0 and takes
 * the shard's entries in file order; then R = R_old + ((G_0 + G_1) + ...).
 * The reference analogue is matFact-mpi.c:187-209 with a p x 1 grid. */
int oracle_factorize_sharded(int users, int items, int features, int nnz,
                             const oracle_entry *entries, int iters, double alpha,
                             double *L, double *R,
                             int nshards, const int *shard_lo)
{
	const size_t nl = (size_t)users * features, nr = (size_t)items * features;
	double *Ls = malloc((nl ? nl : 1) * sizeof *Ls);
	double *G = malloc((nr ? nr : 1) * (size_t)nshards * sizeof *G);
	if (!Ls || !G) { free(Ls); free(G); return -1; }

	for (int it = 0; it < iters; it++) {
		memcpy(Ls, L, nl * sizeof *Ls);
		memset(G, 0, nr * (size_t)nshards * sizeof *G);
		for (int n = 0; n < nnz; n++) {
			const int i = entries[n].row;
			int s = 0;
			while (s + 1 < nshards && i >= shard_lo[s + 1])
				s++;
			double *Gs = G + (size_t)s * nr;
			const size_t li = (size_t)i * features;
			const size_t rj = (size_t)entries[n].col * features;
			/* R is only rewritten after the loop, so it is its own snapshot */
			const double step = alpha * 2 * (entries[n].value - seq_dot(Ls + li, R + rj, features));
			for (int k = 0; k < features; k++) {
				L[li + k] = L[li + k] + step * R[rj + k];
				Gs[rj + k] = Gs[rj + k] + step * Ls[li + k];
			}
		}
		for (size_t p = 0; p < nr; p++) {
			double g = G[p];
			for (int s = 1; s < nshards; s++)
				g = g + G[(size_t)s * nr + p];
			R[p] = R[p] + g;
		}
	}
	free(Ls);
	free(G);
	return 0;
}

/* mat2d.c:100-113 for one row: dest starts at 0, then dest = dest + l*r for
 * ascending k. */
void oracle_predict_row(int items, int features, const double *Li,
                        const double *R, double *out)
{
	for (int j = 0; j < items; j++)
		out[j] = seq_dot(Li, R + (size_t)j * features, features);
}

/* matFact.c:10-27.  `cursor` is the reference's aix: it only ever advances
 * when entries[cursor] is exactly (i, j), for any entry order.  The reference
 * reads entries[nnz] once the list is exhausted; here that read is replaced by
 * "no match", which is what an arbitrary trailing element yields unless it
 * happens to equal (i, j). */
int oracle_recommend(int users, int items, int features, int nnz,
                     const oracle_entry *entries,
                     const double *L, const double *R, int *best)
{
	double *row = malloc((items ? (size_t)items : 1) * sizeof *row);
	if (!row) return -1;
	int cursor = 0;
	for (int i = 0; i < users; i++) {
		oracle_predict_row(items, features, L + (size_t)i * features, R, row);
		int top = -1;
		for (int j = 0; j < items; j++) {
			if (cursor < nnz && entries[cursor].row == i && entries[cursor].col == j) {
				cursor++;
			} else if (top == -1 || row[j] > row[top]) {
				top = j;
			}
		}
		best[i] = top;
	}
	free(row);
	return 0;
}
