/*
 * oracle/mpi_shim/mpi_shim.c -- TEST/BENCH INFRASTRUCTURE ONLY (see mpi.h).
 *
 * Shared-memory implementation of the MPI subset of matFact-mpi.c / matFact-hybrid.c /
 * mpiutil.c.  One process per rank (mpirun_shim.c), all mapping the same memory file:
 *   - point to point: one byte ring per ordered pair of ranks; a message is a 16-byte header
 *     (communicator, tag, length) followed by its bytes, streamed through the ring, so a
 *     message may be longer than the ring and short messages are buffered (eager);
 *   - collectives: every communicator owns 2 x size staging slots of SHIM_CHUNK bytes.
 *     Allreduce / Reduce work on one chunk at a time: every rank copies its chunk into its input
 *     slot, barrier, every rank reduces one slice of the chunk over the ranks in rank order
 *     (rank 0 first) into its output slot, barrier, the ranks copy the slices out.  The order
 *     of the sum is fixed, so runs repeat bit for bit;
 *   - barrier: sense-reversing counter per communicator; waits spin, then yield.
 * Nonblocking collectives complete inside the call (MPI_Waitall has nothing left to do);
 * that is a legal schedule for programs whose ranks issue them in one order, which the
 * reference's do (matFact-mpi.c:207-209).  Two threads of one rank may be inside collectives of
 * DIFFERENT communicators at once (matFact-hybrid.c:300-305): the state is per communicator.
 */
#define _GNU_SOURCE
#include "mpi.h"
#include "shim_layout.h"

#include <errno.h>
#include <sched.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <time.h>
#include <unistd.h>

#if defined(__x86_64__) || defined(__i386__)
#include <immintrin.h>
#define cpu_relax() _mm_pause()
#else
#define cpu_relax() ((void)0)
#endif

/* ---- process state ---- */

typedef struct local_comm {
	int used;
	int ctx;    /* index into the shared communicator table */
	int size;
	int rank;   /* of this process */
	int *world; /* world rank of every member */
	int ndims;  /* > 0: a Cartesian grid, row-major */
	int dims[8], periods[8];
} local_comm;

typedef struct local_type {
	int used;
	size_t extent;
	int base; /* predefined element type when the type is one, else -1 */
} local_type;

#define MAX_LOCAL_COMMS 1024
#define MAX_LOCAL_TYPES 256
#define MAX_LOCAL_OPS 64
#define FIRST_USER_OP 16

static shim_header *H;
static unsigned char *BASE;
static int g_rank = -1, g_np;
static local_comm g_comm[MAX_LOCAL_COMMS];
static local_type g_type[MAX_LOCAL_TYPES];
static MPI_User_function *g_op[MAX_LOCAL_OPS];

static void shim_die(const char *what)
{
	fprintf(stderr, "mpi_shim (rank %d): %s\n", g_rank, what);
	if (H)
		atomic_store(&H->aborted, 1);
	_exit(86);
}

/* one step of a wait loop: spin briefly, then give the core away; leave if the job died */
static inline void wait_step(unsigned *spins)
{
	if (++*spins < 2000) {
		cpu_relax();
		return;
	}
	if ((*spins & 1023) == 0 && atomic_load_explicit(&H->aborted, memory_order_relaxed))
		_exit(87);
	sched_yield();
}

static local_comm *comm_of(MPI_Comm c)
{
	if (c < 0 || c >= MAX_LOCAL_COMMS || !g_comm[c].used)
		shim_die("invalid communicator");
	return &g_comm[c];
}

static size_t extent_of(MPI_Datatype t)
{
	if (t < 0 || t >= MAX_LOCAL_TYPES || !g_type[t].used)
		shim_die("invalid datatype");
	return g_type[t].extent;
}

static unsigned char *slot_in(const local_comm *c, int r)
{
	return BASE + H->comm[c->ctx].arena + (uint64_t)r * SHIM_CHUNK;
}

static unsigned char *slot_out(const local_comm *c, int r)
{
	return BASE + H->comm[c->ctx].arena + (uint64_t)(c->size + r) * SHIM_CHUNK;
}

/* ---- environment ---- */

static void init_tables(void)
{
	static const size_t sizes[] = {1, 1, sizeof(int), sizeof(long), sizeof(float), sizeof(double),
	                               sizeof(unsigned), sizeof(long long)};
	for (int t = 0; t < 8; t++)
		g_type[t] = (local_type){1, sizes[t], t};
	local_comm *w = &g_comm[0];
	w->used = 1;
	w->ctx = 0;
	w->size = g_np;
	w->rank = g_rank;
	w->world = malloc(sizeof(int) * (size_t)g_np);
	for (int r = 0; r < g_np; r++)
		w->world[r] = r;
}

int MPI_Init_thread(int *argc, char ***argv, int required, int *provided)
{
	(void)argc;
	(void)argv;
	if (H)
		shim_die("MPI_Init called twice");
	const char *fd_s = getenv("MPISHIM_FD"), *rank_s = getenv("MPISHIM_RANK"), *np_s = getenv("MPISHIM_NP");
	const int launched = fd_s && rank_s && np_s;
	int fd;
	if (launched) {
		fd = atoi(fd_s);
		g_rank = atoi(rank_s);
		g_np = atoi(np_s);
	} else {
		/* started without the launcher: a job of one rank */
		g_rank = 0;
		g_np = 1;
		fd = memfd_create("mpi_shim", 0);
		if (fd < 0 || ftruncate(fd, (off_t)shim_total_bytes(1)) != 0)
			shim_die("cannot create the shared memory file");
	}
	if (g_np < 1 || g_np > SHIM_MAX_RANKS || g_rank < 0 || g_rank >= g_np)
		shim_die("bad MPISHIM_RANK / MPISHIM_NP");
	const uint64_t bytes = shim_total_bytes(g_np);
	void *p = mmap(NULL, bytes, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_NORESERVE, fd, 0);
	if (p == MAP_FAILED)
		shim_die("cannot map the shared memory file");
	H = p;
	BASE = p;
	if (!launched)
		shim_format(H, 1);
	if (H->magic != SHIM_MAGIC || H->np != g_np)
		shim_die("shared memory file not formatted for this job");
	init_tables();
	if (provided)
		*provided = required;
	return MPI_SUCCESS;
}

int MPI_Init(int *argc, char ***argv) { return MPI_Init_thread(argc, argv, MPI_THREAD_SINGLE, NULL); }

int MPI_Finalize(void)
{
	MPI_Barrier(MPI_COMM_WORLD);
	return MPI_SUCCESS;
}

int MPI_Abort(MPI_Comm comm, int errorcode)
{
	(void)comm;
	if (H)
		atomic_store(&H->aborted, 1);
	_exit(errorcode ? errorcode : 1);
}

double MPI_Wtime(void)
{
	struct timespec ts;
	clock_gettime(CLOCK_MONOTONIC, &ts);
	return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

/* ---- datatypes and operations ---- */

int MPI_Type_create_struct(int count, const int blocklengths[], const MPI_Aint displacements[],
                           const MPI_Datatype types[], MPI_Datatype *newtype)
{
	size_t ub = 0, align = 1;
	for (int b = 0; b < count; b++) {
		const size_t e = extent_of(types[b]);
		if (displacements[b] < 0)
			shim_die("negative displacement in MPI_Type_create_struct");
		const size_t end = (size_t)displacements[b] + (size_t)blocklengths[b] * e;
		ub = end > ub ? end : ub;
		/* alignment of a predefined member is its size; of a struct, that of its widest member */
		const size_t a = g_type[types[b]].base >= 0 ? e : 8;
		align = a > align ? a : align;
	}
	int t = 8;
	while (t < MAX_LOCAL_TYPES && g_type[t].used)
		t++;
	if (t == MAX_LOCAL_TYPES)
		shim_die("too many datatypes");
	g_type[t] = (local_type){1, (ub + align - 1) / align * align, -1};
	*newtype = t;
	return MPI_SUCCESS;
}

int MPI_Type_commit(MPI_Datatype *type)
{
	extent_of(*type);
	return MPI_SUCCESS;
}

int MPI_Type_free(MPI_Datatype *type)
{
	if (*type >= 8 && *type < MAX_LOCAL_TYPES)
		g_type[*type].used = 0;
	*type = MPI_DATATYPE_NULL;
	return MPI_SUCCESS;
}

int MPI_Op_create(MPI_User_function *fn, int commute, MPI_Op *op)
{
	(void)commute;
	for (int o = FIRST_USER_OP; o < MAX_LOCAL_OPS; o++)
		if (!g_op[o]) {
			g_op[o] = fn;
			*op = o;
			return MPI_SUCCESS;
		}
	shim_die("too many operations");
	return MPI_ERR_OTHER;
}

int MPI_Op_free(MPI_Op *op)
{
	if (*op >= FIRST_USER_OP && *op < MAX_LOCAL_OPS)
		g_op[*op] = NULL;
	*op = MPI_OP_NULL;
	return MPI_SUCCESS;
}

#define COMBINE_LOOP(T)                                                                               \
	do {                                                                                          \
		T *restrict a = acc;                                                                  \
		const T *restrict b = in;                                                             \
		switch (op) {                                                                         \
		case MPI_SUM: for (size_t i = 0; i < n; i++) a[i] = a[i] + b[i]; break;               \
		case MPI_PROD: for (size_t i = 0; i < n; i++) a[i] = a[i] * b[i]; break;              \
		case MPI_MAX: for (size_t i = 0; i < n; i++) a[i] = b[i] > a[i] ? b[i] : a[i]; break; \
		case MPI_MIN: for (size_t i = 0; i < n; i++) a[i] = b[i] < a[i] ? b[i] : a[i]; break; \
		default: shim_die("unknown predefined operation");                                    \
		}                                                                                     \
	} while (0)

/* acc = acc op in, element by element (predefined operations on predefined types) */
static void combine(void *acc, const void *in, size_t n, MPI_Datatype type, MPI_Op op)
{
	switch (g_type[type].base) {
	case MPI_INT: COMBINE_LOOP(int); break;
	case MPI_LONG: COMBINE_LOOP(long); break;
	case MPI_FLOAT: COMBINE_LOOP(float); break;
	case MPI_DOUBLE: COMBINE_LOOP(double); break;
	case MPI_UNSIGNED: COMBINE_LOOP(unsigned); break;
	case MPI_LONG_LONG: COMBINE_LOOP(long long); break;
	case MPI_CHAR:
	case MPI_BYTE: COMBINE_LOOP(unsigned char); break;
	default: shim_die("predefined operation on a derived datatype");
	}
}

/* ---- barrier ---- */

static void barrier(const local_comm *c)
{
	if (c->size == 1)
		return;
	shim_comm_shared *s = &H->comm[c->ctx];
	const int gen = atomic_load_explicit(&s->generation, memory_order_acquire);
	if (atomic_fetch_add_explicit(&s->arrived, 1, memory_order_acq_rel) == c->size - 1) {
		atomic_store_explicit(&s->arrived, 0, memory_order_relaxed);
		atomic_store_explicit(&s->generation, gen + 1, memory_order_release);
		return;
	}
	unsigned spins = 0;
	while (atomic_load_explicit(&s->generation, memory_order_acquire) == gen)
		wait_step(&spins);
}

int MPI_Barrier(MPI_Comm comm)
{
	barrier(comm_of(comm));
	return MPI_SUCCESS;
}

/* ---- point to point ---- */

typedef struct msg_header {
	int ctx;
	int tag;
	uint64_t bytes;
} msg_header;

static shim_ring *ring_of(int src, int dst)
{
	return (shim_ring *)(BASE + H->rings + ((uint64_t)src * g_np + dst) * sizeof(shim_ring));
}

static void ring_write(shim_ring *r, const void *data, size_t n)
{
	const unsigned char *p = data;
	uint64_t head = atomic_load_explicit(&r->head, memory_order_relaxed);
	unsigned spins = 0;
	while (n) {
		const uint64_t tail = atomic_load_explicit(&r->tail, memory_order_acquire);
		const size_t room = SHIM_RING_BYTES - (size_t)(head - tail);
		if (!room) {
			wait_step(&spins);
			continue;
		}
		spins = 0;
		const size_t at = (size_t)(head % SHIM_RING_BYTES);
		size_t k = n < room ? n : room;
		k = k < SHIM_RING_BYTES - at ? k : SHIM_RING_BYTES - at;
		memcpy(r->data + at, p, k);
		head += k;
		p += k;
		n -= k;
		atomic_store_explicit(&r->head, head, memory_order_release);
	}
}

static void ring_read(shim_ring *r, void *data, size_t n)
{
	unsigned char *p = data;
	uint64_t tail = atomic_load_explicit(&r->tail, memory_order_relaxed);
	unsigned spins = 0;
	while (n) {
		const uint64_t head = atomic_load_explicit(&r->head, memory_order_acquire);
		const size_t have = (size_t)(head - tail);
		if (!have) {
			wait_step(&spins);
			continue;
		}
		spins = 0;
		const size_t at = (size_t)(tail % SHIM_RING_BYTES);
		size_t k = n < have ? n : have;
		k = k < SHIM_RING_BYTES - at ? k : SHIM_RING_BYTES - at;
		memcpy(p, r->data + at, k);
		p += k;
		tail += k;
		n -= k;
		atomic_store_explicit(&r->tail, tail, memory_order_release);
	}
}

static void send_bytes(const local_comm *c, int dest, int tag, const void *buf, size_t bytes)
{
	if (dest < 0 || dest >= c->size)
		shim_die("MPI_Send: destination out of range");
	shim_ring *r = ring_of(g_rank, c->world[dest]);
	const msg_header h = {c->ctx, tag, bytes};
	ring_write(r, &h, sizeof h);
	ring_write(r, buf, bytes);
}

static size_t recv_bytes(const local_comm *c, int source, int tag, void *buf, size_t capacity, int *got_tag)
{
	if (source < 0 || source >= c->size)
		shim_die("MPI_Recv: source out of range (MPI_ANY_SOURCE is not supported)");
	shim_ring *r = ring_of(c->world[source], g_rank);
	msg_header h;
	ring_read(r, &h, sizeof h);
	if (h.ctx != c->ctx || (tag != MPI_ANY_TAG && h.tag != tag)) {
		char msg[200];
		snprintf(msg, sizeof msg,
		         "MPI_Recv(source %d, tag %d, comm %d) met a message with tag %d on comm %d: matching out of "
		         "order is not supported",
		         source, tag, c->ctx, h.tag, h.ctx);
		shim_die(msg);
	}
	if (h.bytes > capacity)
		shim_die("MPI_Recv: message longer than the receive buffer");
	ring_read(r, buf, (size_t)h.bytes);
	if (got_tag)
		*got_tag = h.tag;
	return (size_t)h.bytes;
}

int MPI_Send(const void *buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm comm)
{
	if (tag < 0)
		shim_die("MPI_Send: negative tag");
	send_bytes(comm_of(comm), dest, tag, buf, (size_t)count * extent_of(type));
	return MPI_SUCCESS;
}

int MPI_Recv(void *buf, int count, MPI_Datatype type, int source, int tag, MPI_Comm comm, MPI_Status *status)
{
	int got_tag;
	const size_t n = recv_bytes(comm_of(comm), source, tag, buf, (size_t)count * extent_of(type), &got_tag);
	if (status) {
		status->MPI_SOURCE = source;
		status->MPI_TAG = got_tag;
		status->MPI_ERROR = MPI_SUCCESS;
		status->shim_bytes = (int)n;
	}
	return MPI_SUCCESS;
}

/* ---- collectives ---- */

int MPI_Bcast(void *buf, int count, MPI_Datatype type, int root, MPI_Comm comm)
{
	const local_comm *c = comm_of(comm);
	if (root < 0 || root >= c->size)
		shim_die("MPI_Bcast: root out of range");
	if (c->size == 1)
		return MPI_SUCCESS;
	size_t left = (size_t)count * extent_of(type);
	unsigned char *p = buf;
	do {
		const size_t k = left < SHIM_CHUNK ? left : SHIM_CHUNK;
		if (c->rank == root)
			memcpy(slot_in(c, root), p, k);
		barrier(c);
		if (c->rank != root)
			memcpy(p, slot_in(c, root), k);
		barrier(c);
		p += k;
		left -= k;
	} while (left);
	return MPI_SUCCESS;
}

/* recvbuf = sendbuf[0] op sendbuf[1] op ... op sendbuf[size-1] on `root` (root < 0: on everybody).
 * Slot reuse needs no third barrier: an input slot is rewritten only after the second barrier
 * of the chunk before (all slices reduced), an output slot only after the first barrier of
 * the next chunk (everybody has copied the last one out). */
static void reduce_impl(const void *sendbuf, void *recvbuf, size_t count, MPI_Datatype type, MPI_Op op, int root,
                        const local_comm *c)
{
	const size_t ext = extent_of(type);
	const int n = c->size, me = c->rank;
	const int receives = root < 0 || root == me;
	if (sendbuf == MPI_IN_PLACE)
		sendbuf = recvbuf;
	MPI_User_function *fn = NULL;
	if (op >= FIRST_USER_OP) {
		if (op >= MAX_LOCAL_OPS || !g_op[op])
			shim_die("invalid operation");
		fn = g_op[op];
	} else if (op < 0 || g_type[type].base < 0)
		shim_die("predefined operation on a derived datatype");
	if (n == 1) {
		if (recvbuf != sendbuf)
			memcpy(recvbuf, sendbuf, count * ext);
		return;
	}
	const size_t per_chunk = SHIM_CHUNK / ext;
	const unsigned char *src = sendbuf;
	unsigned char *dst = recvbuf;
	size_t done = 0;
	do {
		const size_t ne = count - done < per_chunk ? count - done : per_chunk;
		memcpy(slot_in(c, me), src + done * ext, ne * ext);
		barrier(c);
		/* this rank's slice of the chunk */
		const size_t lo = ne * (size_t)me / (size_t)n, hi = ne * ((size_t)me + 1) / (size_t)n;
		if (hi > lo) {
			unsigned char *acc = slot_out(c, me) + lo * ext;
			if (fn) {
				/* fn(in, inout): inout = in op inout, so fold from the highest rank down */
				memcpy(acc, slot_in(c, n - 1) + lo * ext, (hi - lo) * ext);
				for (int r = n - 2; r >= 0; r--) {
					int len = (int)(hi - lo);
					MPI_Datatype t = type;
					fn(slot_in(c, r) + lo * ext, acc, &len, &t);
				}
			} else {
				memcpy(acc, slot_in(c, 0) + lo * ext, (hi - lo) * ext);
				for (int r = 1; r < n; r++)
					combine(acc, slot_in(c, r) + lo * ext, hi - lo, type, op);
			}
		}
		barrier(c);
		if (receives)
			for (int r = 0; r < n; r++) {
				const size_t rlo = ne * (size_t)r / (size_t)n, rhi = ne * ((size_t)r + 1) / (size_t)n;
				memcpy(dst + (done + rlo) * ext, slot_out(c, r) + rlo * ext, (rhi - rlo) * ext);
			}
		done += ne;
	} while (done < count);
}

int MPI_Reduce(const void *sendbuf, void *recvbuf, int count, MPI_Datatype type, MPI_Op op, int root, MPI_Comm comm)
{
	const local_comm *c = comm_of(comm);
	if (root < 0 || root >= c->size)
		shim_die("MPI_Reduce: root out of range");
	reduce_impl(sendbuf, recvbuf, (size_t)count, type, op, root, c);
	return MPI_SUCCESS;
}

int MPI_Allreduce(const void *sendbuf, void *recvbuf, int count, MPI_Datatype type, MPI_Op op, MPI_Comm comm)
{
	reduce_impl(sendbuf, recvbuf, (size_t)count, type, op, -1, comm_of(comm));
	return MPI_SUCCESS;
}

int MPI_Iallreduce(const void *sendbuf, void *recvbuf, int count, MPI_Datatype type, MPI_Op op, MPI_Comm comm,
                   MPI_Request *request)
{
	reduce_impl(sendbuf, recvbuf, (size_t)count, type, op, -1, comm_of(comm));
	*request = MPI_REQUEST_NULL; /* already complete */
	return MPI_SUCCESS;
}

int MPI_Wait(MPI_Request *request, MPI_Status *status)
{
	*request = MPI_REQUEST_NULL;
	if (status)
		memset(status, 0, sizeof *status);
	return MPI_SUCCESS;
}

int MPI_Waitall(int count, MPI_Request requests[], MPI_Status statuses[])
{
	for (int q = 0; q < count; q++)
		MPI_Wait(&requests[q], statuses ? &statuses[q] : MPI_STATUS_IGNORE);
	return MPI_SUCCESS;
}

#define TAG_GATHERV (-1001) /* internal: user tags are >= 0 */

int MPI_Gatherv(const void *sendbuf, int sendcount, MPI_Datatype sendtype, void *recvbuf, const int recvcounts[],
                const int displs[], MPI_Datatype recvtype, int root, MPI_Comm comm)
{
	const local_comm *c = comm_of(comm);
	if (root < 0 || root >= c->size)
		shim_die("MPI_Gatherv: root out of range");
	if (c->rank != root) {
		send_bytes(c, root, TAG_GATHERV, sendbuf, (size_t)sendcount * extent_of(sendtype));
		return MPI_SUCCESS;
	}
	const size_t ext = extent_of(recvtype);
	for (int r = 0; r < c->size; r++) {
		unsigned char *at = (unsigned char *)recvbuf + (size_t)displs[r] * ext;
		if (r == root) {
			if (sendbuf != MPI_IN_PLACE)
				memmove(at, sendbuf, (size_t)sendcount * extent_of(sendtype));
		} else
			recv_bytes(c, r, TAG_GATHERV, at, (size_t)recvcounts[r] * ext, NULL);
	}
	return MPI_SUCCESS;
}

/* ---- communicators ---- */

int MPI_Comm_size(MPI_Comm comm, int *size)
{
	*size = comm_of(comm)->size;
	return MPI_SUCCESS;
}

int MPI_Comm_rank(MPI_Comm comm, int *rank)
{
	*rank = comm_of(comm)->rank;
	return MPI_SUCCESS;
}

typedef struct split_note {
	int color, key, prank; /* prank: rank in the parent */
} split_note;

static int by_key(const void *a, const void *b)
{
	const split_note *x = a, *y = b;
	/* equal keys keep the order of the parent */
	return x->key != y->key ? (x->key > y->key) - (x->key < y->key) : (x->prank > y->prank) - (x->prank < y->prank);
}

/* Collective over `parent`: the members that gave `color` form a new communicator, ordered by
 * key.  The first member allocates the shared state and tells the others its index. */
static MPI_Comm split_impl(const local_comm *parent, int color, int key)
{
	const int n = parent->size, me = parent->rank;
	const split_note mine = {color, key, me};
	split_note *all = malloc(sizeof *all * (size_t)n);
	if (n > 1) {
		memcpy(slot_in(parent, me), &mine, sizeof mine);
		barrier(parent);
		for (int r = 0; r < n; r++)
			memcpy(&all[r], slot_in(parent, r), sizeof mine);
	} else
		all[0] = mine;
	int m = 0;
	for (int r = 0; r < n; r++)
		if (all[r].color == color)
			all[m++] = all[r];
	qsort(all, (size_t)m, sizeof *all, by_key);
	int my_new = -1;
	for (int q = 0; q < m; q++)
		if (all[q].prank == me)
			my_new = q;
	int ctx = -1;
	if (color != MPI_UNDEFINED && my_new == 0) {
		ctx = atomic_fetch_add(&H->next_comm, 1);
		if (ctx >= SHIM_MAX_COMMS)
			shim_die("too many communicators");
		const uint64_t need = (uint64_t)m * 2 * SHIM_CHUNK;
		const uint64_t at = atomic_fetch_add(&H->bump, need);
		if (at + need > H->total_bytes)
			shim_die("staging memory exhausted (communicators are never recycled)");
		H->comm[ctx].arena = at;
		H->comm[ctx].size = m;
		atomic_store(&H->comm[ctx].arrived, 0);
		atomic_store(&H->comm[ctx].generation, 0);
		if (n > 1)
			memcpy(slot_out(parent, me), &ctx, sizeof ctx);
	}
	if (n > 1) {
		barrier(parent);
		if (color != MPI_UNDEFINED && my_new != 0)
			memcpy(&ctx, slot_out(parent, all[0].prank), sizeof ctx);
		barrier(parent);
	}
	if (color == MPI_UNDEFINED) {
		free(all);
		return MPI_COMM_NULL;
	}
	int h = 1;
	while (h < MAX_LOCAL_COMMS && g_comm[h].used)
		h++;
	if (h == MAX_LOCAL_COMMS)
		shim_die("too many communicator handles");
	local_comm *c = &g_comm[h];
	memset(c, 0, sizeof *c);
	c->used = 1;
	c->ctx = ctx;
	c->size = m;
	c->rank = my_new;
	c->world = malloc(sizeof(int) * (size_t)m);
	for (int q = 0; q < m; q++)
		c->world[q] = parent->world[all[q].prank];
	free(all);
	return h;
}

int MPI_Comm_split(MPI_Comm comm, int color, int key, MPI_Comm *newcomm)
{
	if (color < 0 && color != MPI_UNDEFINED)
		shim_die("MPI_Comm_split: negative color");
	*newcomm = split_impl(comm_of(comm), color, key);
	return MPI_SUCCESS;
}

int MPI_Comm_free(MPI_Comm *comm)
{
	if (*comm > 0 && *comm < MAX_LOCAL_COMMS && g_comm[*comm].used) {
		free(g_comm[*comm].world);
		g_comm[*comm].used = 0;
	}
	*comm = MPI_COMM_NULL;
	return MPI_SUCCESS;
}

/* Balanced factorisation: the prime factors of nnodes, largest first, each go to the dimension
 * that is smallest so far; free dimensions come out in non-increasing order (8 -> 4 x 2,
 * 12 -> 4 x 3, 16 -> 4 x 4, 64 -> 8 x 8).  Entries of dims that are not 0 are kept. */
int MPI_Dims_create(int nnodes, int ndims, int dims[])
{
	int fixed = 1, nfree = 0;
	for (int d = 0; d < ndims; d++) {
		if (dims[d] < 0)
			shim_die("MPI_Dims_create: negative dimension");
		if (dims[d] > 0)
			fixed *= dims[d];
		else
			nfree++;
	}
	if (nnodes <= 0 || nnodes % fixed)
		shim_die("MPI_Dims_create: nnodes is not a multiple of the fixed dimensions");
	if (!nfree)
		return MPI_SUCCESS;
	if (nfree > 16)
		shim_die("MPI_Dims_create: too many dimensions");
	int rest = nnodes / fixed, nprimes = 0, primes[32], out[16];
	for (int p = 2; p * p <= rest; p++)
		while (rest % p == 0) {
			primes[nprimes++] = p;
			rest /= p;
		}
	if (rest > 1)
		primes[nprimes++] = rest;
	for (int q = 0; q < nfree; q++)
		out[q] = 1;
	for (int q = nprimes - 1; q >= 0; q--) {
		int smallest = 0;
		for (int d = 1; d < nfree; d++)
			if (out[d] < out[smallest])
				smallest = d;
		out[smallest] *= primes[q];
	}
	for (int a = 0; a < nfree; a++)
		for (int b = a + 1; b < nfree; b++)
			if (out[b] > out[a]) {
				const int t = out[a];
				out[a] = out[b];
				out[b] = t;
			}
	for (int d = 0, q = 0; d < ndims; d++)
		if (dims[d] == 0)
			dims[d] = out[q++];
	return MPI_SUCCESS;
}

int MPI_Cart_create(MPI_Comm comm, int ndims, const int dims[], const int periods[], int reorder, MPI_Comm *newcomm)
{
	(void)reorder; /* ranks keep their order: the grid is row-major over the parent's ranks */
	const local_comm *parent = comm_of(comm);
	if (ndims < 1 || ndims > 8)
		shim_die("MPI_Cart_create: unsupported number of dimensions");
	long cells = 1;
	for (int d = 0; d < ndims; d++)
		cells *= dims[d];
	if (cells > parent->size)
		shim_die("MPI_Cart_create: grid larger than the communicator");
	const MPI_Comm h = split_impl(parent, parent->rank < cells ? 0 : MPI_UNDEFINED, parent->rank);
	if (h != MPI_COMM_NULL) {
		local_comm *c = &g_comm[h];
		c->ndims = ndims;
		for (int d = 0; d < ndims; d++) {
			c->dims[d] = dims[d];
			c->periods[d] = periods[d];
		}
	}
	*newcomm = h;
	return MPI_SUCCESS;
}

static const local_comm *grid_of(MPI_Comm comm)
{
	const local_comm *c = comm_of(comm);
	if (!c->ndims)
		shim_die("communicator has no Cartesian grid");
	return c;
}

int MPI_Cart_coords(MPI_Comm comm, int rank, int maxdims, int coords[])
{
	const local_comm *c = grid_of(comm);
	if (rank < 0 || rank >= c->size)
		shim_die("MPI_Cart_coords: rank out of range");
	for (int d = c->ndims - 1; d >= 0; d--) {
		if (d < maxdims)
			coords[d] = rank % c->dims[d];
		rank /= c->dims[d];
	}
	return MPI_SUCCESS;
}

int MPI_Cart_rank(MPI_Comm comm, const int coords[], int *rank)
{
	const local_comm *c = grid_of(comm);
	int r = 0;
	for (int d = 0; d < c->ndims; d++) {
		int x = coords[d];
		if (c->periods[d])
			x = ((x % c->dims[d]) + c->dims[d]) % c->dims[d];
		else if (x < 0 || x >= c->dims[d])
			shim_die("MPI_Cart_rank: coordinate out of range");
		r = r * c->dims[d] + x;
	}
	*rank = r;
	return MPI_SUCCESS;
}

int MPI_Cart_get(MPI_Comm comm, int maxdims, int dims[], int periods[], int coords[])
{
	const local_comm *c = grid_of(comm);
	for (int d = 0; d < c->ndims && d < maxdims; d++) {
		dims[d] = c->dims[d];
		periods[d] = c->periods[d];
	}
	return MPI_Cart_coords(comm, c->rank, maxdims, coords);
}
