/*
 * host/mf_ingest.c -- see mf_ingest.h.  Plain C, no GPU, no libc random state.
 */
#include "mf_ingest.h"

#include <ctype.h>
#include <errno.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ---- glibc random(), TYPE_3 ----------------------------------------------------------------
 *
 * srandom(seed) fills a table of 31 words with the Lehmer sequence x -> 16807 x mod (2^31 - 1)
 * (seed 0 counts as 1), sets the front index to 3 and the rear index to 0 and discards 310
 * draws; a draw adds table[rear] to table[front] (mod 2^32), returns that sum >> 1 and steps
 * both indices round the table.  Written as one sequence x: x[i] = table[(i + 3) mod 31] for
 * i < 31, then x[n] = x[n-3] + x[n-31] (mod 2^32), and draw k is x[341 + k] >> 1.  What
 * mat2d_random_fill_LR relies on (mat2d.c:61-72: srandom(0), then users*features +
 * features*items calls of random()) is exactly this sequence; the tests compare it draw for draw
 * with the C library's own random() (tests/test_host.py).
 *
 * The recurrence has a dependent chain three draws long, ~1 ns per draw however it is
 * scheduled.  With S the shift by one draw, x = (S^3 + S^31) x, hence x = (S^3 + S^31)^4 x:
 *     x[n] = x[n-12] + 4 x[n-40] + 6 x[n-68] + 4 x[n-96] + x[n-124]   (mod 2^32)
 * wherever the plain recurrence held for the 93 values before, i.e. from x[124] on (the first
 * draw is x[341]).  Twelve consecutive values are independent of each other in this form, so
 * blocks are generated eight at a time in vector registers.
 */
static inline uint32_t raw_next(mf_rand *g)
{
	const uint64_t n = g->n;
	const uint32_t v = g->h[(n - 3) & 127] + g->h[(n - 31) & 127];
	g->h[n & 127] = v;
	g->n = n + 1;
	return v;
}

void mf_rand_seed(mf_rand *g, unsigned seed)
{
	uint32_t table[31];
	int32_t word = seed ? (int32_t)seed : 1;
	table[0] = (uint32_t)word;
	for (int i = 1; i < 31; i++) {
		/* 16807 * word mod 2147483647 without overflow (Schrage) */
		const long hi = word / 127773, lo = word % 127773;
		long w = 16807 * lo - 2836 * hi;
		if (w < 0)
			w += 2147483647;
		word = (int32_t)w;
		table[i] = (uint32_t)word;
	}
	memset(g->h, 0, sizeof g->h);
	for (int i = 0; i < 31; i++)
		g->h[i] = table[(i + 3) % 31];
	g->n = 31;
	for (int i = 0; i < 310; i++)
		(void)raw_next(g);
}

uint32_t mf_rand_next(mf_rand *g) { return raw_next(g) >> 1; }

typedef uint32_t v8u __attribute__((vector_size(32), aligned(4), may_alias));
typedef int32_t v8i __attribute__((vector_size(32)));

#define MF_RAND_CHUNK 2048   /* raw values kept unshifted behind the generating front (>= 124) */

#if defined(__x86_64__) && defined(__GNUC__) && !defined(__clang__)
__attribute__((target_clones("avx2", "default")))
#endif
void mf_rand_block(mf_rand *g, uint32_t *out, size_t n)
{
	if (n < 2 * MF_RAND_CHUNK) {
		for (size_t i = 0; i < n; i++)
			out[i] = raw_next(g) >> 1;
		return;
	}
	/* raw values first: 124 from the ring, the rest from the block itself; a value is shifted
	 * to its draw once the front is a chunk past it */
	for (size_t i = 0; i < 124; i++)
		out[i] = raw_next(g);
	size_t shifted = 0;
	for (size_t c0 = 124; c0 < n; c0 += MF_RAND_CHUNK) {
		const size_t c1 = n - c0 < MF_RAND_CHUNK ? n : c0 + MF_RAND_CHUNK;
		size_t i = c0;
		/* the lag-12 operand straddles the two vectors stored last: taken from registers (a
		 * load would wait for both stores to retire) */
		v8u p2 = *(const v8u *)(out + i - 16), p1 = *(const v8u *)(out + i - 8);
		const v8i mid = {4, 5, 6, 7, 8, 9, 10, 11};
		for (; i + 8 <= c1; i += 8) {
			const v8u a = __builtin_shuffle(p2, p1, mid), b = *(const v8u *)(out + i - 40),
			          c = *(const v8u *)(out + i - 68), d = *(const v8u *)(out + i - 96),
			          e = *(const v8u *)(out + i - 124);
			const v8u rest = e + ((b + d) << 2) + (c << 2) + (c << 1);   /* off the dependent chain */
			const v8u x = a + rest;
			*(v8u *)(out + i) = x;
			p2 = p1;
			p1 = x;
		}
		for (; i < c1; i++)
			out[i] = out[i - 3] + out[i - 31];
		for (const size_t upto = c0 - 124; shifted < upto; shifted++)
			out[shifted] >>= 1;
	}
	g->n += n - 124;
	for (size_t k = 1; k <= 124; k++)
		g->h[(g->n - k) & 127] = out[n - k];
	for (; shifted < n; shifted++)
		out[shifted] >>= 1;
}

/* ---- initial factors --------------------------------------------------------------------
 *
 * The draws are one sequential stream (each needs the table entries written 3 and 31 draws
 * before), ~1 ns apiece; turning a draw into its double costs two divisions and a store, and
 * for R the store goes to a transposed position.  So: the calling thread generates the stream
 * block by block, and while it generates block b+1 a few worker threads convert block b, each
 * for its own range of outputs.  R is drawn features x items row-major (mat2d.c:67-70) and
 * used transposed (mat2d.c:115-124): a block is KB whole feature rows of draws, and a worker
 * writes, for each of its items, the KB consecutive doubles of that item's row -- whole cache
 * lines, and few enough passes over the rows of R that the TLB misses of the stride-F walk
 * (one page per item) stay small.
 */
#include <pthread.h>
#include <unistd.h>

#define MF_RAND_MAX 2147483647.0     /* RAND_MAX of glibc, mat2d.h:4 */
#define MF_FILL_THREADS_MAX 16
#define MF_BLOCK_DRAWS ((size_t)1 << 22)   /* draws per block: 16 MB of uint32 */

typedef struct {
	const uint32_t *buf;      /* the block's draws                                       */
	double *dst;              /* L: first output of the block; R: the matrix            */
	double norm;
	size_t n;                 /* L: draws in the block                                   */
	size_t I, F, k0, kb;      /* R: items, features, first feature and features of the block */
	int transposed;
	int t, nt;                /* this worker, workers                                    */
} fill_job;

static void *fill_worker(void *arg)
{
	const fill_job *j = (const fill_job *)arg;
	const double norm = j->norm;
	if (!j->transposed) {
		const size_t lo = j->n * (size_t)j->t / (size_t)j->nt, hi = j->n * (size_t)(j->t + 1) / (size_t)j->nt;
		for (size_t q = lo; q < hi; q++)
			j->dst[q] = ((double)j->buf[q] / MF_RAND_MAX) / norm;
		return NULL;
	}
	const size_t lo = j->I * (size_t)j->t / (size_t)j->nt, hi = j->I * (size_t)(j->t + 1) / (size_t)j->nt;
	const size_t I = j->I, F = j->F, kb = j->kb;
	for (size_t it = lo; it < hi; it++) {
		double *row = j->dst + it * F + j->k0;
		const uint32_t *src = j->buf + it;
		for (size_t kk = 0; kk < kb; kk++)
			row[kk] = ((double)src[kk * I] / MF_RAND_MAX) / norm;
	}
	return NULL;
}

static int fill_threads(size_t outputs)
{
	long n = sysconf(_SC_NPROCESSORS_ONLN);
	if (n < 1)
		n = 1;
	if (n > MF_FILL_THREADS_MAX)
		n = MF_FILL_THREADS_MAX;
	/* small jobs are not worth a thread each */
	while (n > 1 && outputs / (size_t)n < 65536)
		n--;
	return (int)n;
}

/* convert one block with nt workers while `next` (if any) is generated by the caller */
static void run_block(fill_job proto, int nt, mf_rand *g, uint32_t *next, size_t next_n)
{
	pthread_t th[MF_FILL_THREADS_MAX];
	fill_job jobs[MF_FILL_THREADS_MAX];
	int started[MF_FILL_THREADS_MAX];
	for (int t = 0; t < nt; t++) {
		jobs[t] = proto;
		jobs[t].t = t;
		jobs[t].nt = nt;
		started[t] = nt > 1 && pthread_create(&th[t], NULL, fill_worker, &jobs[t]) == 0;
	}
	if (next_n)
		mf_rand_block(g, next, next_n);
	for (int t = 0; t < nt; t++) {
		if (started[t])
			pthread_join(th[t], NULL);
		else
			fill_worker(&jobs[t]);       /* one worker, or a thread that could not be created */
	}
}

void mf_random_fill_LR(int users, int items, int features, double *L, double *R)
{
	const double norm = (double)features;
	mf_rand g;
	mf_rand_seed(&g, 0);                            /* mat2d.c:62 */
	const size_t nl = (size_t)(users > 0 ? users : 0) * (size_t)(features > 0 ? features : 0);
	const size_t I = (size_t)(items > 0 ? items : 0), F = (size_t)(features > 0 ? features : 0);

	/* features of R per block: whole cache lines of a row (multiples of 8), up to 64, within
	 * MF_BLOCK_DRAWS draws when the matrix has that many items */
	size_t KB = 8;
	while (KB < 64 && 2 * KB * I <= MF_BLOCK_DRAWS)
		KB *= 2;
	size_t cap = nl < MF_BLOCK_DRAWS ? nl : MF_BLOCK_DRAWS;
	if (I * KB > cap)
		cap = I * KB;
	if (cap == 0)
		return;
	uint32_t *buf[2];
	buf[0] = (uint32_t *)malloc(sizeof(uint32_t) * cap);
	buf[1] = (uint32_t *)malloc(sizeof(uint32_t) * cap);
	if (!buf[0] || !buf[1]) {
		/* no room for the blocks: one draw at a time (strided stores for R) */
		free(buf[0]);
		free(buf[1]);
		for (size_t n = 0; n < nl; n++)
			L[n] = ((double)mf_rand_next(&g) / MF_RAND_MAX) / norm;
		for (size_t k = 0; k < F; k++)
			for (size_t j = 0; j < I; j++)
				R[j * F + k] = ((double)mf_rand_next(&g) / MF_RAND_MAX) / norm;
		return;
	}

	/* the blocks in draw order: those of L (mat2d.c:63-66), then those of R */
	const size_t l_blocks = (nl + MF_BLOCK_DRAWS - 1) / MF_BLOCK_DRAWS;
	const size_t r_blocks = I ? (F + KB - 1) / KB : 0;
	const size_t blocks = l_blocks + r_blocks;
#define BLOCK_DRAWS(b) ((b) < l_blocks ? (nl - (b) * MF_BLOCK_DRAWS < MF_BLOCK_DRAWS ? nl - (b) * MF_BLOCK_DRAWS : MF_BLOCK_DRAWS) \
                                       : I * (F - ((b) - l_blocks) * KB < KB ? F - ((b) - l_blocks) * KB : KB))
	if (blocks)
		mf_rand_block(&g, buf[0], BLOCK_DRAWS(0));
	for (size_t b = 0; b < blocks; b++) {
		fill_job j;
		memset(&j, 0, sizeof j);
		j.buf = buf[b & 1];
		j.norm = norm;
		size_t outputs;
		if (b < l_blocks) {
			j.dst = L + b * MF_BLOCK_DRAWS;
			j.n = outputs = BLOCK_DRAWS(b);
		} else {
			j.transposed = 1;
			j.dst = R;
			j.I = I;
			j.F = F;
			j.k0 = (b - l_blocks) * KB;
			j.kb = F - j.k0 < KB ? F - j.k0 : KB;
			outputs = I * j.kb;
		}
		const size_t next_n = b + 1 < blocks ? BLOCK_DRAWS(b + 1) : 0;
		run_block(j, fill_threads(outputs), &g, buf[(b + 1) & 1], next_n);
	}
#undef BLOCK_DRAWS
	free(buf[0]);
	free(buf[1]);
}

/* ---- parser -------------------------------------------------------------------------------
 *
 * fscanf's conversions, restated over a buffer.  fscanf first delimits an input item -- the
 * longest run of characters that can still become a number -- and then converts it; the item
 * can be longer than what strtod alone would take ("0x" is read and refused, "1e+" is read
 * whole and converts to 1).  scan_int / scan_double delimit the item the way glibc's vfscanf
 * does (its corner cases are pinned by tests/ingest_check.c against fscanf itself) and convert
 * it with strtol / strtod, as vfscanf does; short plain decimals take an exact shortcut.
 */
typedef struct {
	const char *p, *end;
} cursor;

static void skip_space(cursor *c)
{
	while (c->p < c->end && isspace((unsigned char)*c->p))
		c->p++;
}

static int at(const cursor *c, size_t k) { return c->p + k < c->end ? (unsigned char)c->p[k] : -1; }

/* %d: 1 converted, 0 matching failure, -1 input failure (end of input before any character) */
static int scan_int(cursor *c, int *out)
{
	skip_space(c);
	if (c->p >= c->end)
		return -1;
	size_t n = 0;
	if (at(c, 0) == '+' || at(c, 0) == '-')
		n = 1;
	size_t digits = 0;
	while (isdigit(at(c, n + digits)) && at(c, n + digits) >= 0)
		digits++;
	if (digits == 0) {
		c->p += n;                       /* the sign has been read */
		return 0;
	}
	char tmp[64];
	const size_t len = n + digits;
	long v;
	if (digits <= 18) {
		/* fits a long: what strtol returns, without the copy */
		v = 0;
		for (size_t q = n; q < len; q++)
			v = v * 10 + (c->p[q] - '0');
		if (c->p[0] == '-')
			v = -v;
	} else if (len < sizeof tmp) {
		memcpy(tmp, c->p, len);
		tmp[len] = 0;
		v = strtol(tmp, NULL, 10);
	} else {
		/* absurdly long digit string: strtol saturates, as it does inside fscanf */
		char *big = (char *)malloc(len + 1);
		if (!big)
			return 0;
		memcpy(big, c->p, len);
		big[len] = 0;
		v = strtol(big, NULL, 10);
		free(big);
	}
	*out = (int)v;                       /* fscanf stores the long through an int * */
	c->p += len;
	return 1;
}

static int ieq(int ch, char lower) { return ch >= 0 && tolower(ch) == lower; }

/* %lf */
static int scan_double(cursor *c, double *out)
{
	skip_space(c);
	if (c->p >= c->end)
		return -1;
	size_t n = 0;
	int ok = 0;
	if (at(c, 0) == '+' || at(c, 0) == '-')
		n = 1;
	if (ieq(at(c, n), 'i')) {
		/* inf | infinity */
		static const char word[] = "infinity";
		size_t m = 0;
		while (m < 8 && ieq(at(c, n + m), word[m]))
			m++;
		ok = (m == 3 || m == 8);
		n += m;
	} else if (ieq(at(c, n), 'n')) {
		/* nan (glibc's fscanf stops there: "nan(chars)" leaves the parenthesis unread) */
		static const char word[] = "nan";
		size_t m = 0;
		while (m < 3 && ieq(at(c, n + m), word[m]))
			m++;
		ok = m == 3;
		n += m;
	} else {
		int hex = 0;
		if (at(c, n) == '0' && ieq(at(c, n + 1), 'x')) {
			hex = 1;
			n += 2;
		}
		const size_t body = n;
		size_t digits = 0, frac = 0;
		unsigned long long mant = 0;       /* decimal digits read so far, while they fit */
		for (;; n++) {
			const int ch = at(c, n);
			if (ch >= 0 && (hex ? isxdigit(ch) : isdigit(ch))) {
				digits++;
				if (digits <= 19)
					mant = mant * 10 + (unsigned)(ch - '0');
			} else
				break;
		}
		if (at(c, n) == '.') {
			n++;
			for (;; n++) {
				const int ch = at(c, n);
				if (ch >= 0 && (hex ? isxdigit(ch) : isdigit(ch))) {
					digits++;
					frac++;
					if (digits <= 19)
						mant = mant * 10 + (unsigned)(ch - '0');
				} else
					break;
			}
		}
		/* plain decimal with at most 15 digits: the integer and the power of ten are both
		 * exact doubles, so one division is the correctly rounded value -- what strtod
		 * returns (Clinger's fast path) */
		if (!hex && digits > 0 && digits <= 15 && !ieq(at(c, n), 'e')) {
			static const double p10[] = {1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9, 1e10, 1e11,
			                             1e12, 1e13, 1e14, 1e15};
			const double v = (double)mant / p10[frac];
			*out = c->p[0] == '-' ? -v : v;
			c->p += n;
			return 1;
		}
		/* glibc: a hexadecimal item is refused only when it is the bare prefix "0x"; "0x." is
		 * read and converts to 0 */
		ok = digits > 0 || (hex && n > body);
		if (digits > 0 && ieq(at(c, n), hex ? 'p' : 'e')) {
			n++;
			if (at(c, n) == '+' || at(c, n) == '-')
				n++;
			while (at(c, n) >= 0 && isdigit(at(c, n)))
				n++;
			/* glibc keeps an exponent marker without digits ("1e", "1e+") in the item and
			 * converts the mantissa: the item is read, the conversion succeeds */
		}
	}
	if (!ok) {
		c->p += n;
		return 0;
	}
	char tmp[128], *big = NULL, *s = tmp;
	if (n >= sizeof tmp) {
		big = (char *)malloc(n + 1);
		if (!big)
			return 0;
		s = big;
	}
	memcpy(s, c->p, n);
	s[n] = 0;
	*out = strtod(s, NULL);
	free(big);
	c->p += n;
	return 1;
}

int mf_parse_buffer(const char *text, size_t len, mf_instance *out)
{
	cursor c = {text, text + len};
	memset(out, 0, sizeof *out);
	if (scan_int(&c, &out->iters) != 1)                  /* matFact.c:79 */
		return MF_PARSE_INT;
	if (scan_double(&c, &out->alpha) != 1)               /* matFact.c:83 */
		return MF_PARSE_DOUBLE;
	if (scan_int(&c, &out->features) != 1)               /* matFact.c:87 */
		return MF_PARSE_INT;
	if (scan_int(&c, &out->users) != 1 || scan_int(&c, &out->items) != 1 ||
	    scan_int(&c, &out->non_zero) != 1)               /* matFact.c:93 */
		return MF_PARSE_MULTI_INT;
	/* the reference goes on with negative sizes (malloc of a negative count); here they are
	 * an error of the line that holds them */
	if (out->users < 0 || out->items < 0 || out->non_zero < 0 || out->features < 0)
		return MF_PARSE_MULTI_INT;
	const size_t nz = (size_t)out->non_zero;
	out->entries = (mfgpu_entry *)malloc(sizeof(mfgpu_entry) * (nz ? nz : 1));
	if (!out->entries)
		return MF_PARSE_NOMEM;
	for (size_t n = 0; n < nz; n++) {                    /* matFact.c:97-105 */
		mfgpu_entry *e = &out->entries[n];
		if (scan_int(&c, &e->row) != 1 || scan_int(&c, &e->col) != 1 || scan_double(&c, &e->value) != 1) {
			free(out->entries);
			out->entries = NULL;
			return MF_PARSE_ENTRY;
		}
	}
	return MF_PARSE_OK;
}

int mf_parse_instance(const char *path, mf_instance *out)
{
	memset(out, 0, sizeof *out);
	FILE *fp = fopen(path, "rb");
	if (!fp)
		return MF_PARSE_OPEN;
	size_t cap = 1 << 20, len = 0;
	char *text = (char *)malloc(cap);
	if (!text) {
		fclose(fp);
		return MF_PARSE_NOMEM;
	}
	for (;;) {
		const size_t got = fread(text + len, 1, cap - len, fp);
		len += got;
		if (len < cap)
			break;
		char *bigger = (char *)realloc(text, cap * 2);
		if (!bigger) {
			free(text);
			fclose(fp);
			return MF_PARSE_NOMEM;
		}
		text = bigger;
		cap *= 2;
	}
	const int rc = mf_parse_buffer(text, len, out);
	free(text);
	if (fclose(fp) == EOF && rc == MF_PARSE_OK) {
		free(out->entries);
		out->entries = NULL;
		return MF_PARSE_CLOSE;
	}
	return rc;
}

const char *mf_parse_message(int rc)
{
	switch (rc) {
	case MF_PARSE_OK: return "";
	case MF_PARSE_OPEN: return "Unable to open input file.";
	case MF_PARSE_INT: return "Error in int argument.";
	case MF_PARSE_DOUBLE: return "Error in double argument.";
	case MF_PARSE_MULTI_INT: return "Error in multiple int argument.";
	case MF_PARSE_ENTRY: return "Error in non-zero entry.";
	case MF_PARSE_CLOSE: return "Unable to close input file.";
	default: return "Out of memory.";
	}
}
