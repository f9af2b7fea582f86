/*
 * oracle/ref_driver.c -- TEST/BENCH INFRASTRUCTURE ONLY.
 *
 * In-process entry to the UNMODIFIED reference implementation.  The recipe in
 * oracle/Makefile compiles the reference's own translation units where they
 * lie (/root/reference/matFact.c or matFact-omp.c with -Dmain=..., mat2d.c,
 * util.c, datatypes.c) and links them with this file into
 * oracle/_ref/libref_serial.so / libref_omp.so; no reference source is copied.
 * This file only wraps caller buffers in the reference's mat2d structs
 * (mat2d.h:6-11) and calls its matrix_factorization (matFact.c:29 /
 * matFact-omp.c:35), timing the call.  The call includes the trailing
 * mat2d_prod (matFact.c:55); callers subtract a run with iters = 0.
 */
#include "mat2d.h"
#include "datatypes.h"

#include <stdlib.h>
#include <string.h>
#include <time.h>

void matrix_factorization(mat2d *B, mat2d *L, mat2d *R, non_zero_entry *entries, int nz_size,
                          int iters, double alpha);

static double now_s(void)
{
	struct timespec ts;
	clock_gettime(CLOCK_MONOTONIC, &ts);
	return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

/* L: users x features, R: items x features (transposed factor), updated in place.
 * Returns the seconds spent inside matrix_factorization, or -1 on allocation failure. */
double ref_factorize(int users, int items, int features, int nnz, const void *entries,
                     int iters, double alpha, double *L, double *R)
{
	mat2d Lm = {users, features, L}, Rm = {items, features, R};
	mat2d *B = mat2d_new(users, items);
	non_zero_entry *e = malloc(sizeof *e * (nnz > 0 ? (size_t)nnz : 1));
	if (!B || !B->data || !e)
		return -1.0;
	memcpy(e, entries, sizeof *e * (size_t)nnz);   /* the OpenMP variant re-sorts its argument */
	const double t0 = now_s();
	matrix_factorization(B, &Lm, &Rm, e, nnz, iters, alpha);
	const double t1 = now_s();
	free(e);
	mat2d_free(B);
	return t1 - t0;
}
