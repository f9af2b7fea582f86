/*
 * oracle/oracle_main.c -- TEST INFRASTRUCTURE ONLY.
 *
 * Command-line front end of the oracle:  matfact_oracle <file.in> [dump.bin [norec]]
 * Reads the reference's input format (matFact.c:72-105: iters, alpha, features,
 * "users items nnz", then nnz lines "row col value", whitespace separated),
 * runs oracle_factorize + oracle_recommend and prints one item per line like
 * matFact.c:24-25.  With a second argument it also writes the final L then R as
 * raw little-endian doubles (the reference's stdout cannot show them), and it
 * prints "loop_s <seconds>" to stderr for the cpu_baseline leg of bench.py.
 */
#include "matfact_oracle.h"

#include <stdio.h>
#include <stdlib.h>
#include <time.h>

static void fail(const char *msg)
{
	fprintf(stderr, "Error: %s\n", msg);
	exit(-1);
}

static double now_s(void)
{
	struct timespec ts;
	clock_gettime(CLOCK_MONOTONIC, &ts);
	return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

int main(int argc, char **argv)
{
	if (argc < 2 || argc > 4)
		fail("usage: matfact_oracle file.in [dump.bin [norec]]");
	FILE *fp = fopen(argv[1], "r");
	if (!fp)
		fail("Unable to open input file.");

	int iters, features, users, items, nnz;
	double alpha;
	if (fscanf(fp, "%d", &iters) != 1) fail("Error in int argument.");
	if (fscanf(fp, "%lf", &alpha) != 1) fail("Error in double argument.");
	if (fscanf(fp, "%d", &features) != 1) fail("Error in int argument.");
	if (fscanf(fp, "%d %d %d", &users, &items, &nnz) != 3) fail("Error in multiple int argument.");

	oracle_entry *entries = malloc(sizeof *entries * (nnz > 0 ? nnz : 1));
	for (int n = 0; n < nnz; n++)
		if (fscanf(fp, "%d %d %lf", &entries[n].row, &entries[n].col, &entries[n].value) != 3)
			fail("Error in non-zero entry.");
	fclose(fp);

	double *L = malloc(sizeof *L * (size_t)users * features);
	double *R = malloc(sizeof *R * (size_t)items * features);
	int *best = malloc(sizeof *best * (users > 0 ? users : 1));
	if (!entries || !L || !R || !best)
		fail("out of memory");

	oracle_init_LR(users, items, features, L, R);
	double t0 = now_s();
	if (oracle_factorize(users, items, features, nnz, entries, iters, alpha, L, R))
		fail("out of memory");
	double t1 = now_s();
	fprintf(stderr, "loop_s %.6f\n", t1 - t0);

	if (argc >= 3) {
		FILE *out = fopen(argv[2], "wb");
		if (!out) fail("Unable to open dump file.");
		fwrite(L, sizeof *L, (size_t)users * features, out);
		fwrite(R, sizeof *R, (size_t)items * features, out);
		fclose(out);
	}

	if (argc == 4) /* "norec": factors only (the arg-max of the 1e6-item sample takes ~25 min) */
		return 0;
	if (oracle_recommend(users, items, features, nnz, entries, L, R, best))
		fail("out of memory");
	for (int i = 0; i < users; i++)
		if (best[i] != -1)
			printf("%d\n", best[i]);

	free(best); free(R); free(L); free(entries);
	return 0;
}
