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
 * both indices round the table.  The front index is always rear + 3 (mod 31), so `pos` = rear
 * is the whole position.  What mat2d_random_fill_LR relies on (mat2d.c:61-72: srandom(0), then
 * users*features + features*items calls of random()) is exactly this sequence; the tests compare
 * it draw for draw with the C library's own random() (tests/test_host.py).
 */
void mf_rand_seed(mf_rand *g, unsigned seed)
{
	int32_t word = seed ? (int32_t)seed : 1;
	g->s[0] = (uint32_t)word;
	for (int i = 1; i < 31; i++) {
		/* 16807 * word mod 2147483647 without overflow (Schrage) */
		const long hi = word / 127773, lo = word % 127773;
		long w = 16807 * lo - 2836 * hi;
		if (w < 0)
			w += 2147483647;
		word = (int32_t)w;
		g->s[i] = (uint32_t)word;
	}
	g->pos = 0;
	for (int i = 0; i < 310; i++)
		(void)mf_rand_next(g);
}

uint32_t mf_rand_next(mf_rand *g)
{
	const int r = g->pos, f = r + 3 >= 31 ? r + 3 - 31 : r + 3;
	const uint32_t v = g->s[f] += g->s[r];
	g->pos = r + 1 == 31 ? 0 : r + 1;
	return v >> 1;
}

void mf_rand_block(mf_rand *g, uint32_t *out, size_t n)
{
	size_t i = 0;
	while (i < n && g->pos != 0)
		out[i++] = mf_rand_next(g);
	/* whole turns of the table: rear 0..27 feeds front 3..30, rear 28..30 feeds front 0..2 */
	uint32_t *s = g->s;
	for (; i + 31 <= n; i += 31) {
		for (int r = 0; r < 28; r++)
			out[i + r] = (s[r + 3] += s[r]) >> 1;
		for (int r = 28; r < 31; r++)
			out[i + r] = (s[r - 28] += s[r]) >> 1;
	}
	while (i < n)
		out[i++] = mf_rand_next(g);
}

/* ---- initial factors -------------------------------------------------------------------- */

#define MF_RAND_MAX 2147483647.0     /* RAND_MAX of glibc, mat2d.h:4 */

/* features of R drawn per pass: 8 doubles = one cache line of a row of the transposed R */
#define MF_KB 8

void mf_random_fill_LR(int users, int items, int features, double *L, double *R)
{
	const double norm = (double)features;
	mf_rand g;
	mf_rand_seed(&g, 0);                            /* mat2d.c:62 */
	enum { CHUNK = 1 << 14 };
	uint32_t small[CHUNK];
	/* L, row-major, in draw order (mat2d.c:63-66) */
	const size_t nl = (size_t)users * (size_t)features;
	for (size_t n = 0; n < nl; n += CHUNK) {
		const size_t c = nl - n < CHUNK ? nl - n : CHUNK;
		mf_rand_block(&g, small, c);
		for (size_t q = 0; q < c; q++)
			L[n + q] = ((double)small[q] / MF_RAND_MAX) / norm;
	}
	/* R is drawn features x items row-major (mat2d.c:67-70) and used transposed
	 * (mat2d.c:115-124): draw MF_KB feature rows at a time, then write them as MF_KB
	 * consecutive doubles of every item's row */
	if (items <= 0 || features <= 0)
		return;
	const size_t I = (size_t)items, F = (size_t)features;
	uint32_t *buf = (uint32_t *)malloc(sizeof(uint32_t) * I * MF_KB);
	if (!buf) {
		/* no room for the block: one draw at a time, strided stores */
		for (size_t k = 0; k < F; k++)
			for (size_t j = 0; j < I; j++)
				R[j * F + k] = ((double)mf_rand_next(&g) / MF_RAND_MAX) / norm;
		return;
	}
	for (size_t k0 = 0; k0 < F; k0 += MF_KB) {
		const size_t kb = F - k0 < MF_KB ? F - k0 : MF_KB;
		mf_rand_block(&g, buf, kb * I);
		if (kb == MF_KB) {
			for (size_t j = 0; j < I; j++) {
				double *row = R + j * F + k0;
				for (int kk = 0; kk < MF_KB; kk++)
					row[kk] = ((double)buf[(size_t)kk * I + j] / MF_RAND_MAX) / norm;
			}
		} else {
			for (size_t j = 0; j < I; j++)
				for (size_t kk = 0; kk < kb; kk++)
					R[j * F + k0 + kk] = ((double)buf[kk * I + j] / MF_RAND_MAX) / norm;
		}
	}
	free(buf);
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
