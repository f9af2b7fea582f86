/*
 * tests/mpi_shim_check.c -- checks of oracle/mpi_shim (the single-node MPI stand-in the
 * reference's matFact-mpi.c / matFact-hybrid.c are built against) beyond what those two programs
 * exercise: messages longer than a ring, collectives longer than a staging chunk, a derived
 * datatype with a user operation, Gatherv with uneven counts, sub-communicators, the process
 * grid, concurrent collectives from two threads.  Run as
 *     mpirun_shim -np N mpi_shim_check
 * Rank 0 prints "OK <N>" when every rank agrees.
 */
#include <mpi.h>

#include <pthread.h>
#include <stddef.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static int rank, np, failures;

#define CHECK(cond)                                                                        \
	do {                                                                               \
		if (!(cond)) {                                                             \
			fprintf(stderr, "rank %d: line %d: %s\n", rank, __LINE__, #cond);  \
			failures++;                                                        \
		}                                                                          \
	} while (0)

typedef struct pair {
	int index;
	double value;
} pair;

static void pair_max(void *in_, void *inout_, int *len, MPI_Datatype *type)
{
	(void)type;
	pair *in = in_, *inout = inout_;
	for (int i = 0; i < *len; i++)
		if (in[i].value > inout[i].value)
			inout[i] = in[i];
}

static void check_point_to_point(void)
{
	/* 3 MB messages through 256 KB rings, around a ring of ranks, two tags in order */
	const int n = 3 << 18;
	double *a = malloc(sizeof(double) * n), *b = malloc(sizeof(double) * n);
	const int next = (rank + 1) % np, prev = (rank + np - 1) % np;
	for (int i = 0; i < n; i++)
		a[i] = rank * 1e6 + i;
	MPI_Status st;
	if (np > 1) {
		int small = 100 + rank, got = -1;
		if (rank % 2 == 0) {
			MPI_Send(&small, 1, MPI_INT, next, 7, MPI_COMM_WORLD);
			MPI_Send(a, n, MPI_DOUBLE, next, 8, MPI_COMM_WORLD);
			MPI_Recv(&got, 1, MPI_INT, prev, 7, MPI_COMM_WORLD, &st);
			MPI_Recv(b, n, MPI_DOUBLE, prev, MPI_ANY_TAG, MPI_COMM_WORLD, &st);
		} else {
			MPI_Recv(&got, 1, MPI_INT, prev, 7, MPI_COMM_WORLD, &st);
			MPI_Recv(b, n, MPI_DOUBLE, prev, MPI_ANY_TAG, MPI_COMM_WORLD, &st);
			MPI_Send(&small, 1, MPI_INT, next, 7, MPI_COMM_WORLD);
			MPI_Send(a, n, MPI_DOUBLE, next, 8, MPI_COMM_WORLD);
		}
		CHECK(got == 100 + prev);
		CHECK(st.MPI_TAG == 8 && st.MPI_SOURCE == prev);
		int good = 1;
		for (int i = 0; i < n; i++)
			good &= b[i] == prev * 1e6 + i;
		CHECK(good);
	}
	free(a);
	free(b);
}

static void check_collectives(MPI_Comm comm)
{
	int r, n;
	MPI_Comm_rank(comm, &r);
	MPI_Comm_size(comm, &n);
	/* 2.5 chunks of doubles; the sum over ranks in rank order is exact for these values */
	const int count = 5 << 16;
	double *x = malloc(sizeof(double) * count), *y = malloc(sizeof(double) * count);
	for (int i = 0; i < count; i++)
		x[i] = (double)(r + 1) * (double)(i % 1000);
	MPI_Allreduce(x, y, count, MPI_DOUBLE, MPI_SUM, comm);
	int good = 1;
	for (int i = 0; i < count; i++)
		good &= y[i] == (double)(n * (n + 1) / 2) * (double)(i % 1000);
	CHECK(good);
	MPI_Request req;
	MPI_Status st;
	memset(y, 0, sizeof(double) * count);
	MPI_Iallreduce(x, y, count, MPI_DOUBLE, MPI_MAX, comm, &req);
	MPI_Waitall(1, &req, &st);
	good = 1;
	for (int i = 0; i < count; i++)
		good &= y[i] == (double)n * (double)(i % 1000);
	CHECK(good);
	/* rounding order: ((x0 + x1) + x2) + ..., rank 0 first */
	double tiny = r == 0 ? 1.0 : 0x1p-53, sum = 0, want = 1.0;
	MPI_Allreduce(&tiny, &sum, 1, MPI_DOUBLE, MPI_SUM, comm);
	for (int q = 1; q < n; q++)
		want = want + 0x1p-53;
	CHECK(sum == want);
	/* broadcast from the last rank, 2.5 chunks */
	for (int i = 0; i < count; i++)
		y[i] = r == n - 1 ? 3.0 * i : -1.0;
	MPI_Bcast(y, count, MPI_DOUBLE, n - 1, comm);
	good = 1;
	for (int i = 0; i < count; i++)
		good &= y[i] == 3.0 * i;
	CHECK(good);
	/* derived type + user operation, reduced onto rank 0 (matFact-mpi.c:23-28,98) */
	MPI_Datatype pair_type;
	MPI_Datatype types[2] = {MPI_INT, MPI_DOUBLE};
	MPI_Aint offsets[2] = {offsetof(pair, index), offsetof(pair, value)};
	int lens[2] = {1, 1};
	MPI_Type_create_struct(2, lens, offsets, types, &pair_type);
	MPI_Type_commit(&pair_type);
	MPI_Op op;
	MPI_Op_create(pair_max, 1, &op);
	const int np_ = 100000; /* 1.6 MB of pairs */
	pair *p = malloc(sizeof(pair) * np_), *q = malloc(sizeof(pair) * np_);
	for (int i = 0; i < np_; i++)
		p[i] = (pair){r * np_ + i, (double)((i + r) % n)};
	MPI_Reduce(p, q, np_, pair_type, op, 0, comm);
	if (r == 0) {
		good = 1;
		for (int i = 0; i < np_; i++) {
			/* the largest value n-1 is held by the rank with (i + rank) % n == n - 1 */
			const int holder = ((n - 1 - i) % n + n) % n;
			good &= q[i].value == (double)(n - 1) && q[i].index == holder * np_ + i;
		}
		CHECK(good);
	}
	/* ties keep the higher rank's entry (inout = in op inout folds from the top) */
	pair t = {r, 1.0}, tr = {-1, -1.0};
	MPI_Reduce(&t, &tr, 1, pair_type, op, 0, comm);
	if (r == 0)
		CHECK(tr.index == n - 1);
	MPI_Op_free(&op);
	/* gatherv: rank q sends q + 1 pairs */
	int *counts = malloc(sizeof(int) * n), *displs = malloc(sizeof(int) * n), total = 0;
	for (int k = 0; k < n; k++) {
		counts[k] = k + 1;
		displs[k] = total;
		total += k + 1;
	}
	pair *all = malloc(sizeof(pair) * total);
	for (int i = 0; i <= r; i++)
		p[i] = (pair){r, (double)i};
	MPI_Gatherv(p, r + 1, pair_type, all, counts, displs, pair_type, 0, comm);
	if (r == 0) {
		good = 1;
		for (int k = 0; k < n; k++)
			for (int i = 0; i <= k; i++)
				good &= all[displs[k] + i].index == k && all[displs[k] + i].value == (double)i;
		CHECK(good);
	}
	MPI_Type_free(&pair_type);
	free(x); free(y); free(p); free(q); free(all); free(counts); free(displs);
}

static void check_grid(void)
{
	int dims[2] = {0, 0};
	MPI_Dims_create(np, 2, dims);
	CHECK(dims[0] * dims[1] == np && dims[0] >= dims[1]);
	static const int known[][3] = {{1, 1, 1}, {2, 2, 1}, {4, 2, 2}, {6, 3, 2}, {8, 4, 2}, {12, 4, 3},
	                               {16, 4, 4}, {32, 8, 4}, {64, 8, 8}, {7, 7, 1}};
	for (unsigned q = 0; q < sizeof known / sizeof *known; q++) {
		int d[2] = {0, 0};
		MPI_Dims_create(known[q][0], 2, d);
		CHECK(d[0] == known[q][1] && d[1] == known[q][2]);
	}
	int periods[2] = {0, 0};
	MPI_Comm cart, row, col;
	MPI_Cart_create(MPI_COMM_WORLD, 2, dims, periods, 1, &cart);
	int cr, coords[2], back, gd[2], gp[2], gc[2];
	MPI_Comm_rank(cart, &cr);
	MPI_Cart_coords(cart, cr, 2, coords);
	MPI_Cart_rank(cart, coords, &back);
	MPI_Cart_get(cart, 2, gd, gp, gc);
	CHECK(back == cr && coords[0] == cr / dims[1] && coords[1] == cr % dims[1]);
	CHECK(gd[0] == dims[0] && gd[1] == dims[1] && gc[0] == coords[0] && gc[1] == coords[1]);
	/* as mpiutil.c:97-104 */
	MPI_Comm_split(cart, coords[0], coords[1], &row);
	MPI_Comm_split(cart, coords[1], coords[0], &col);
	int rr, rs, cc, cs;
	MPI_Comm_rank(row, &rr);
	MPI_Comm_size(row, &rs);
	MPI_Comm_rank(col, &cc);
	MPI_Comm_size(col, &cs);
	CHECK(rs == dims[1] && rr == coords[1] && cs == dims[0] && cc == coords[0]);
	check_collectives(row);
	check_collectives(col);
	/* a message inside the sub-communicator */
	if (rs > 1) {
		int v = 1000 + cr, got = -1;
		MPI_Status st;
		if (rr == 0)
			MPI_Send(&v, 1, MPI_INT, rs - 1, 3, row);
		else if (rr == rs - 1) {
			MPI_Recv(&got, 1, MPI_INT, 0, 3, row, &st);
			CHECK(got == 1000 + cr - (rs - 1));
		}
	}
	/* reversed keys reverse the ranks; MPI_UNDEFINED leaves a rank out */
	MPI_Comm rev, some;
	MPI_Comm_split(MPI_COMM_WORLD, 0, -rank, &rev);
	int vr;
	MPI_Comm_rank(rev, &vr);
	CHECK(vr == np - 1 - rank);
	MPI_Comm_split(MPI_COMM_WORLD, rank % 2 ? MPI_UNDEFINED : 5, rank, &some);
	CHECK((some == MPI_COMM_NULL) == (rank % 2 == 1));
	if (some != MPI_COMM_NULL) {
		int ss;
		MPI_Comm_size(some, &ss);
		CHECK(ss == (np + 1) / 2);
		MPI_Comm_free(&some);
	}
	MPI_Comm_free(&rev);

	/* two threads of every rank in collectives of different communicators at the same time
	 * (matFact-hybrid.c:300-305) */
	struct job { MPI_Comm comm; int ok; } jobs[2] = {{row, 0}, {col, 0}};
	pthread_t th;
	extern void *threaded_sum(void *);
	pthread_create(&th, NULL, threaded_sum, &jobs[1]);
	threaded_sum(&jobs[0]);
	pthread_join(th, NULL);
	CHECK(jobs[0].ok && jobs[1].ok);
	MPI_Comm_free(&row);
	MPI_Comm_free(&col);
	MPI_Comm_free(&cart);
}

void *threaded_sum(void *arg)
{
	struct job { MPI_Comm comm; int ok; } *j = arg;
	int n, good = 1;
	MPI_Comm_size(j->comm, &n);
	const int count = 300000;
	double *x = malloc(sizeof(double) * count), *y = malloc(sizeof(double) * count);
	for (int rep = 0; rep < 20; rep++) {
		for (int i = 0; i < count; i++)
			x[i] = (double)(i % 97 + rep);
		MPI_Allreduce(x, y, count, MPI_DOUBLE, MPI_SUM, j->comm);
		for (int i = 0; i < count; i++)
			good &= y[i] == (double)n * (double)(i % 97 + rep);
	}
	free(x);
	free(y);
	j->ok = good;
	return NULL;
}

int main(int argc, char **argv)
{
	int provided = -1;
	MPI_Init_thread(&argc, &argv, MPI_THREAD_MULTIPLE, &provided);
	CHECK(provided == MPI_THREAD_MULTIPLE);
	MPI_Comm_rank(MPI_COMM_WORLD, &rank);
	MPI_Comm_size(MPI_COMM_WORLD, &np);
	const double t0 = MPI_Wtime();
	check_point_to_point();
	check_collectives(MPI_COMM_WORLD);
	check_grid();
	MPI_Barrier(MPI_COMM_WORLD);
	CHECK(MPI_Wtime() >= t0);
	int all = 0;
	MPI_Allreduce(&failures, &all, 1, MPI_INT, MPI_SUM, MPI_COMM_WORLD);
	if (rank == 0 && all == 0)
		printf("OK %d\n", np);
	MPI_Finalize();
	return all != 0;
}
