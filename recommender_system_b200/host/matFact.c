/*
 * host/matFact.c -- the drop-in program:  matFact <file.in>
 *
 * Same command line, input format, error messages and initialisation sequence
 * as the reference's serial program (matFact.c:61-137, util.c:7-34); the two
 * calls matrix_factorization() + print_output() (matFact.c:124,127) are
 * replaced by one call into the C ABI of libmatfact_b200.so, which runs them
 * on the GPU(s).  B is never allocated.  Plain C: this file knows nothing
 * about CUDA.
 * The stdout contract is samples/X.out, i.e. the program the course was handed
 * (deliverables/serial/matFact.c): the recommendations and nothing else.  The
 * top-level matFact.c of the reference is built with benchmark.h's
 * ENABLE_BENCH and ends with a "time : ..." line on stdout; that line is left
 * out on purpose (MATFACT_TIMING puts phase times on stderr instead).  One
 * deliberate difference in error behaviour: negative sizes in the header are
 * refused ("Error in multiple int argument."), where the reference carries on
 * into mat2d_new with them.
 *
 * Environment (additions; the reference has none):
 *   MATFACT_NGPUS   number of B200s to shard the users over (default 1: the
 *                   bit-exact ordering; 0 = all visible)
 *   MATFACT_DUMP    file that receives the final L then R as raw doubles
 *   MATFACT_TIMING  if set, phase times go to stderr (stdout stays the
 *                   recommendations only)
 */
#include "matfact_b200.h"

#include <stdio.h>
#include <stdlib.h>
#include <time.h>

static void die(const char *error)
{
	fprintf(stderr, "Error: %s\n", error);   /* util.c:7-10 */
	exit(-1);
}

static void die_gpu(const char *what)
{
	fprintf(stderr, "Error: %s: %s\n", what, mfgpu_last_error());
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
	if (argc != 2) {
		fprintf(stderr, "Run ./matFact.out file");
		die("Missing input file name.");
	}
	const double t_start = now_s();
	/* matFact.c:72-110 / util.c:12-34: same tokens accepted, same message for the first error */
	int iters, features, users, items, non_zero;
	double alpha;
	mfgpu_entry *entries = NULL;
	const int prc = mfgpu_read_instance(argv[1], &iters, &alpha, &features, &users, &items, &non_zero, &entries);
	if (prc != 0)
		die(mfgpu_read_error(prc));
	const double t_parsed = now_s();

	const size_t nl = (size_t)users * features, nr = (size_t)items * features;
	int *best = malloc(sizeof *best * (users > 0 ? (size_t)users : 1));
	if (!best)
		die("Out of memory.");          /* the reference does not check (mat2d.c:14-19) */

	const char *env = getenv("MATFACT_NGPUS");
	const int ngpus = env ? atoi(env) : 1;
	const char *dump = getenv("MATFACT_DUMP");

	mfgpu_plan *plan = NULL;
	float loop_ms = 0.f, rec_ms = 0.f;
	if (mfgpu_plan_create(&plan, users, items, features, non_zero, entries, alpha, ngpus))
		die_gpu("cannot set up the GPU factorization");
	const double t_plan = now_s();
	/* matFact.c:113-120 (mat2d_random_fill_LR + transpose): same draws, generated into the device */
	if (mfgpu_plan_init_factors(plan))
		die_gpu("cannot initialise the factors");
	const double t_init = now_s();
	if (mfgpu_plan_iterate(plan, iters < 0 ? 0 : iters) || mfgpu_plan_last_loop_ms(plan, &loop_ms))
		die_gpu("factorization failed");
	if (dump) {
		double *L = malloc(sizeof *L * (nl ? nl : 1)), *R = malloc(sizeof *R * (nr ? nr : 1));
		if (!L || !R)
			die("Out of memory.");
		if (mfgpu_plan_get_factors(plan, L, R))
			die_gpu("cannot read the factors back");
		FILE *out = fopen(dump, "wb");
		if (!out || fwrite(L, sizeof *L, nl, out) != nl || fwrite(R, sizeof *R, nr, out) != nr ||
		    fclose(out) == EOF)
			die("Unable to write the factor dump.");
		free(R);
		free(L);
	}
	if (mfgpu_plan_recommend(plan, best) || mfgpu_plan_last_recommend_ms(plan, &rec_ms))
		die_gpu("recommendation failed");
	mfgpu_plan_destroy(plan);
	const double t_gpu = now_s();

	/* matFact.c:24-25: one line per user that has an unrated item */
	for (int i = 0; i < users; i++)
		if (best[i] != -1)
			printf("%d\n", best[i]);

	if (getenv("MATFACT_TIMING"))
		fprintf(stderr,
		        "parse_s %.6f\nplan_s %.6f\ninit_s %.6f\ngpu_total_s %.6f\nloop_ms %.3f\nargmax_ms %.3f\ntotal_s %.6f\n",
		        t_parsed - t_start, t_plan - t_parsed, t_init - t_plan, t_gpu - t_init, loop_ms, rec_ms,
		        now_s() - t_start);
	free(best);
	mfgpu_free(entries);
	return 0;
}
