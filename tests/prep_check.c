/*
 * tests/prep_check.c -- the threaded pattern build (host/mf_prep.c) must give, for any number of
 * threads and with or without a workspace, exactly the arrays of the one-thread build.
 * Compiled and run by tests/test_host.py::test_threaded_pattern_build_is_deterministic.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "mf_prep.h"

static unsigned rng = 12345;
static unsigned rnd(void) { rng = rng * 1664525u + 1013904223u; return rng >> 8; }

static int same(const void *a, const void *b, size_t bytes, const char *what)
{
	if (bytes && memcmp(a, b, bytes)) {
		printf("MISMATCH in %s\n", what);
		return 0;
	}
	return 1;
}

static int compare(const mf_pattern *a, const mf_pattern *b, int users)
{
	int ok = a->nnz == b->nnz && a->n_rows == b->n_rows && a->n_cols == b->n_cols &&
	         a->max_row_len == b->max_row_len && a->max_col_len == b->max_col_len;
	if (!ok) {
		printf("MISMATCH in sizes\n");
		return 0;
	}
	ok &= same(a->row_id, b->row_id, sizeof(int) * a->n_rows, "row_id");
	ok &= same(a->row_ptr, b->row_ptr, sizeof(int) * (a->n_rows + 1), "row_ptr");
	ok &= same(a->col_idx, b->col_idx, sizeof(int) * a->nnz, "col_idx");
	ok &= same(a->val, b->val, sizeof(double) * a->nnz, "val");
	ok &= same(a->epos, b->epos, sizeof(int) * a->nnz, "epos");
	ok &= same(a->col_id, b->col_id, sizeof(int) * a->n_cols, "col_id");
	ok &= same(a->col_ptr, b->col_ptr, sizeof(int) * (a->n_cols + 1), "col_ptr");
	ok &= same(a->row_idx, b->row_idx, sizeof(int) * a->nnz, "row_idx");
	ok &= same(a->mask_ptr, b->mask_ptr, sizeof(int) * (users + 1), "mask_ptr");
	ok &= same(a->mask_col, b->mask_col, sizeof(int) * a->mask_ptr[users], "mask_col");
	ok &= same(a->first_free, b->first_free, sizeof(int) * users, "first_free");
	return ok;
}

int main(void)
{
	int fails = 0;
	for (int trial = 0; trial < 12; trial++) {
		const int users = 50 + rnd() % 3000, items = 20 + rnd() % 400;
		int nnz = 0, cap = users * 30;
		mfgpu_entry *e = malloc(sizeof *e * cap);
		/* sorted file with a few rows empty ... */
		for (int u = 0; u < users && nnz < cap - items; u++) {
			int c = -1, k = rnd() % 25;
			for (int q = 0; q < k; q++) {
				c += 1 + rnd() % (items / 12 + 1);
				if (c >= items)
					break;
				e[nnz].row = u; e[nnz].col = c; e[nnz].value = 1 + rnd() % 5; nnz++;
			}
		}
		/* ... and, in some trials, a disorder somewhere (duplicates, swapped entries) */
		if (trial % 3 == 1 && nnz > 10) {
			int a = rnd() % nnz, b = rnd() % nnz;
			mfgpu_entry t = e[a]; e[a] = e[b]; e[b] = t;
		}
		if (trial % 3 == 2 && nnz > 10)
			e[nnz / 2] = e[nnz / 2 - 1];
		/* ... or columns out of order inside rows that still never decrease (the two-step build) */
		if (trial % 4 == 0 && nnz > 10)
			for (int a = 1; a + 1 < nnz; a += 7)
				if (e[a].row == e[a + 1].row) {
					mfgpu_entry t = e[a]; e[a] = e[a + 1]; e[a + 1] = t;
				}
		const int lo = trial % 2 ? users / 3 : 0, hi = trial % 2 ? users - users / 4 : users;
		mf_pattern ref, got;
		setenv("MFGPU_PREP_THREADS", "1", 1);
		if (mf_pattern_build(&ref, users, items, nnz, e, lo, hi)) { printf("build failed\n"); return 2; }
		for (int T = 2; T <= 8; T += 3) {
			char buf[8];
			snprintf(buf, sizeof buf, "%d", T);
			setenv("MFGPU_PREP_THREADS", buf, 1);
			size_t wsb = mf_pattern_ws_bytes(users, items, nnz, lo, hi);
			void *ws = malloc(wsb);
			memset(ws, 0xAB, wsb);        /* the build must not rely on a zeroed workspace */
			if (mf_pattern_build_ws(&got, users, items, nnz, e, lo, hi, ws, wsb)) { printf("ws build failed\n"); return 2; }
			if (!compare(&ref, &got, users)) { printf("  (trial %d, %d threads, workspace)\n", trial, T); fails++; }
			mf_pattern_free(&got);
			free(ws);
			if (mf_pattern_build(&got, users, items, nnz, e, lo, hi)) { printf("build failed\n"); return 2; }
			if (!compare(&ref, &got, users)) { printf("  (trial %d, %d threads)\n", trial, T); fails++; }
			mf_pattern_free(&got);
		}
		if (mf_mask_prefix(users, items, nnz, e) < nnz && trial % 3 == 0 && trial % 4 != 0) { printf("unexpected disorder\n"); fails++; }
		/* the one-pass scan (mf_scan_entries) against the passes it replaces, and the build from
		 * its knowledge (only the slice of the shard's users is read) against the full build */
		for (int T = 1; T <= 8; T += 7) {
			char buf[8];
			snprintf(buf, sizeof buf, "%d", T);
			setenv("MFGPU_PREP_THREADS", buf, 1);
			for (int ns = 1; ns <= 8; ns += 7) {
				int cut[9], cut2[9], per = 0;
				int *slot = malloc(sizeof(int) * items), *slot2 = malloc(sizeof(int) * items);
				int *ctr = malloc(sizeof(int) * items), *ctr2 = malloc(sizeof(int) * items);
				int *first = malloc(sizeof(int) * (users + 1));
				mf_scan sc;
				const int nt = mf_scan_entries(users, items, nnz, e, ns, cut, slot, first, &sc, ctr, &per);
				const int nt2 = mf_touched_items(items, nnz, e, slot2);
				mf_split_users(users, nnz, e, ns, cut2);
				const int per2 = mf_exchange_layout(users, nnz, e, ns, cut2, slot2, nt2, ctr2);
				int ok = nt == nt2 && per == per2 && !memcmp(cut, cut2, sizeof(int) * (ns + 1)) &&
				         !memcmp(slot, slot2, sizeof(int) * items) && !memcmp(ctr, ctr2, sizeof(int) * (nt2 > 0 ? nt2 : 0)) &&
				         sc.matched == mf_mask_prefix(users, items, nnz, e);
				if (!ok) { printf("MISMATCH in mf_scan_entries (trial %d, %d threads, %d shards)\n", trial, T, ns); fails++; }
				size_t wsb = mf_pattern_ws_bytes(users, items, nnz, lo, hi);
				void *ws = malloc(wsb);
				if (mf_pattern_build_scan(&got, users, items, nnz, e, lo, hi, ws, wsb, &sc)) { printf("scan build failed\n"); return 2; }
				if (!compare(&ref, &got, users)) { printf("  (trial %d, %d threads, from the scan, by_row %d)\n", trial, T, sc.by_row); fails++; }
				mf_pattern_free(&got);
				free(ws); free(slot); free(slot2); free(ctr); free(ctr2); free(first);
			}
		}
		mf_pattern_free(&ref);
		free(e);
	}
	printf(fails ? "FAILED\n" : "OK\n");
	return fails != 0;
}
