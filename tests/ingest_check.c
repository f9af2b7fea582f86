/*
 * tests/ingest_check.c -- differential checks of host/mf_ingest.c against the C library:
 *   1. mf_rand_* against srandom()/random(), draw for draw;
 *   2. mf_random_fill_LR against the reference's sequence written with random()
 *      (mat2d.c:61-72 + transpose :115-124), bit for bit;
 *   3. mf_parse_buffer against the reference's fscanf sequence (matFact.c:72-110, util.c:12-34)
 *      on well-formed files and on fuzzed ones (truncated numbers, stray characters, inf/nan,
 *      hex floats, overflowing integers): same error class, same values.
 * Prints "OK" and exits 0 when everything agrees.
 */
#define _GNU_SOURCE
#include "mf_ingest.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static int fail(const char *what, long where)
{
	printf("MISMATCH: %s at %ld\n", what, where);
	return 1;
}

static int check_rand(void)
{
	const unsigned seeds[] = {0, 1, 12345, 4294967295u};
	for (int q = 0; q < 4; q++) {
		mf_rand g;
		mf_rand_seed(&g, seeds[q]);
		srandom(seeds[q]);
		for (long n = 0; n < 100000; n++)
			if ((long)mf_rand_next(&g) != random())
				return fail("mf_rand_next", n);
		/* blocks of awkward sizes, continuing the same stream */
		static uint32_t buf[70000];
		const size_t sizes[] = {1, 30, 31, 32, 61, 62, 63, 1000, 69999, 7};
		for (int t = 0; t < 10; t++) {
			mf_rand_block(&g, buf, sizes[t]);
			for (size_t n = 0; n < sizes[t]; n++)
				if ((long)buf[n] != random())
					return fail("mf_rand_block", (long)n);
		}
	}
	return 0;
}

static int check_fill(void)
{
	const int shapes[][3] = {{3, 5, 2}, {30, 40, 10}, {7, 100, 8}, {5, 33, 17}, {0, 9, 4},
	                         {9, 0, 4}, {4, 4, 0}, {100, 1000, 23}, {1, 1, 1},
	                         /* several blocks of L and of R, several worker threads, block widths
	                          * of 64 / 36 and of 8 / 1 features */
	                         {170000, 30000, 100}, {100, 3000000, 9}, {3, 70000, 150}};
	for (int q = 0; q < 12; q++) {
		const int U = shapes[q][0], I = shapes[q][1], F = shapes[q][2];
		const size_t nl = (size_t)U * F, nr = (size_t)I * F;
		double *L = malloc(8 * (nl + 1)), *R = malloc(8 * (nr + 1));
		double *L0 = malloc(8 * (nl + 1)), *R0 = malloc(8 * (nr + 1));
		mf_random_fill_LR(U, I, F, L, R);
		srandom(0);
		for (size_t n = 0; n < nl; n++)
			L0[n] = ((double)random() / (double)RAND_MAX) / (double)F;
		for (int k = 0; k < F; k++)
			for (int j = 0; j < I; j++)
				R0[(size_t)j * F + k] = ((double)random() / (double)RAND_MAX) / (double)F;
		if (memcmp(L, L0, 8 * nl) || memcmp(R, R0, 8 * nr))
			return fail("mf_random_fill_LR", q);
		free(L); free(R); free(L0); free(R0);
	}
	return 0;
}

/* the reference's parser, returning the class of the first error instead of exiting */
static int parse_with_fscanf(const char *text, size_t len, mf_instance *out)
{
	memset(out, 0, sizeof *out);
	FILE *fp = fmemopen((void *)text, len ? len : 1, "r");
	if (!fp)
		return MF_PARSE_OPEN;
	if (len == 0)
		fseek(fp, 0, SEEK_END);
	int rc = MF_PARSE_OK;
	if (fscanf(fp, "%d", &out->iters) != 1) rc = MF_PARSE_INT;
	else if (fscanf(fp, "%lf", &out->alpha) != 1) rc = MF_PARSE_DOUBLE;
	else if (fscanf(fp, "%d", &out->features) != 1) rc = MF_PARSE_INT;
	else if (fscanf(fp, "%d %d %d", &out->users, &out->items, &out->non_zero) != 3) rc = MF_PARSE_MULTI_INT;
	else if (out->users < 0 || out->items < 0 || out->non_zero < 0 || out->features < 0) rc = MF_PARSE_MULTI_INT;
	if (rc == MF_PARSE_OK) {
		const size_t nz = (size_t)out->non_zero;
		out->entries = malloc(sizeof(mfgpu_entry) * (nz ? nz : 1));
		for (size_t n = 0; n < nz && rc == MF_PARSE_OK; n++)
			if (fscanf(fp, "%d %d %lf", &out->entries[n].row, &out->entries[n].col,
			           &out->entries[n].value) != 3)
				rc = MF_PARSE_ENTRY;
		if (rc != MF_PARSE_OK) {
			free(out->entries);
			out->entries = NULL;
		}
	}
	fclose(fp);
	return rc;
}

static int same(const mf_instance *a, const mf_instance *b)
{
	if (a->iters != b->iters || a->features != b->features || a->users != b->users ||
	    a->items != b->items || a->non_zero != b->non_zero || memcmp(&a->alpha, &b->alpha, 8))
		return 0;
	for (int n = 0; n < a->non_zero; n++)
		if (a->entries[n].row != b->entries[n].row || a->entries[n].col != b->entries[n].col ||
		    memcmp(&a->entries[n].value, &b->entries[n].value, 8))
			return 0;
	return 1;
}

static int compare_on(const char *text, size_t len, long id)
{
	mf_instance a, b;
	const int ra = parse_with_fscanf(text, len, &a), rb = mf_parse_buffer(text, len, &b);
	int bad = ra != rb;
	/* after an error the values read so far are not observable in the reference (it exits) */
	if (!bad && ra == MF_PARSE_OK)
		bad = !same(&a, &b);
	free(a.entries);
	free(b.entries);
	if (bad) {
		printf("input %ld (%zu bytes): fscanf class %d, mf_parse_buffer class %d\n---\n%.*s\n---\n", id, len, ra,
		       rb, (int)len, text);
		return fail("parser", id);
	}
	return 0;
}

static unsigned long long rng = 88172645463325252ull;
static unsigned rnd(void)
{
	rng ^= rng << 13;
	rng ^= rng >> 7;
	rng ^= rng << 17;
	return (unsigned)(rng >> 32);
}

static int check_parser(void)
{
	/* hand-written cases */
	const char *cases[] = {
	    "", " ", "10", "10 0.5", "10 0.5 3", "10 0.5 3 2 2", "10 0.5 3 2 2 0", "10 0.5 3 2 2 1 0 0 4.0",
	    "10\n0.001\n2\n2 3 2\n0 1 4.0\n1 2 3.5\n", "10 1e-3 2 2 3 1 0 1 5", "10 1e 2 2 3 1 0 1 5",
	    "10 1e+ 2 2 3 1 0 1 5", "10 0x1p-3 2 2 3 1 0 1 0x", "10 0x 2 2 3 0", "10 inf 2 2 3 0",
	    "10 -infinity 2 2 3 0", "10 infin 2 2 3 0", "10 nan 2 2 3 0", "10 nan(abc) 2 2 3 0",
	    "10 nan(ab 2 2 3 0", "10 .5 2 2 3 0", "10 . 2 2 3 0", "10 5. 2 2 3 0", "+10 +.5e+1 -0 2 3 0",
	    "10 0.5 2 2 3 1 0 1 4.0extra", "10 0.5 2 2 3 1 0 1", "10 0.5 2 2 3 2 0 1 4.0", "1x 0.5 2 2 3 0",
	    "10 0.5 2x 2 3 0", "10 0.5 2 2 x 0", "99999999999 0.5 2 2 3 0", "10 0.5 2 -2 3 0",
	    "10 0.5 2 2 3 1 0 99999999999999999999 1.0", "- 10 0.5", "10 0.5 2 2 3 1 0 1 1e999",
	    "10 0.5 2 2 3 1 0 1 1e-999", "10\t0.5\v2\f2\r3 0", "010 00.5 02 2 3 0", "10 0.5 2 2 3 1 0 1 +",
	    "10 0.5 2 2 3 1 0 1 -.e5", "10 1E5 2 2 3 0", "10 0X1.8P1 2 2 3 0", "10 INF 2 2 3 0", "10 NaN 2 2 3 0",
	    "10 0.5e 2 2 3 0", "10 1.5e-x 2 2 3 0", "10 0x.p1 2 2 3 0", "10 0x1p 2 2 3 0", "10 infinityx 2 2 3 0",
	};
	for (size_t q = 0; q < sizeof cases / sizeof *cases; q++)
		if (compare_on(cases[q], strlen(cases[q]), (long)q))
			return 1;
	/* fuzz: a valid file with random edits */
	static const char *frag[] = {"0", "1", "7", "12", "-3", "+4", "0.5", "1e-3", "2.", ".25", "1e", "1e+", "0x",
	                             "0x1p2", "inf", "nan", "infinity", "infi", "nan(", "nan(x)", "x", "-", "+", ".",
	                             "e5", "1.5.2", "3e4e5", "99999999999", " ", "\n", "\t", "", "1,2", "0x1.8", "1e5x",
	                             "123456789012345", "1234567890123456", "0.000001", "-0.0", "00012.5000",
	                             "4.35", "0.1", "2.675", "999999999999999.9", "-2147483648", "2147483648",
	                             "123456789012345678", "1234567890123456789", "12345678901234567890",
	                             "0.3e0", "1234.5678e-2", "5e-324", "1.7976931348623157e308"};
	const int nfrag = (int)(sizeof frag / sizeof *frag);
	char text[4096];
	for (long t = 0; t < 200000; t++) {
		const int nz = (int)(rnd() % 4);
		int len = snprintf(text, sizeof text, "%u\n0.00%u\n%u\n%u %u %d\n", rnd() % 50, rnd() % 9, rnd() % 6,
		                   rnd() % 7, rnd() % 7, nz);
		for (int n = 0; n < nz; n++)
			len += snprintf(text + len, sizeof text - (size_t)len, "%u %u %u.%u\n", rnd() % 7, rnd() % 7,
			                1 + rnd() % 5, rnd() % 10);
		/* replace a few whitespace-delimited tokens by fragments, or cut the file short */
		const int edits = (int)(rnd() % 3);
		for (int e = 0; e < edits; e++) {
			int pos = (int)(rnd() % (unsigned)(len + 1));
			while (pos < len && text[pos] != ' ' && text[pos] != '\n')
				pos++;
			const char *f = frag[rnd() % (unsigned)nfrag];
			const int fl = (int)strlen(f);
			int stop = pos + 1;
			while (stop < len && text[stop] != ' ' && text[stop] != '\n')
				stop++;
			if (pos >= len || len - (stop - pos - 1) + fl + 2 >= (int)sizeof text)
				continue;
			memmove(text + pos + 1 + fl, text + stop, (size_t)(len - stop));
			memcpy(text + pos + 1, f, (size_t)fl);
			len = len - (stop - pos - 1) + fl;
		}
		if (rnd() % 5 == 0)
			len = (int)(rnd() % (unsigned)(len + 1));
		if (compare_on(text, (size_t)len, 1000 + t))
			return 1;
	}
	return 0;
}

int main(void)
{
	if (check_rand() || check_fill() || check_parser())
		return 1;
	printf("OK\n");
	return 0;
}
