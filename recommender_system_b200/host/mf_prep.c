/*
 * host/mf_prep.c -- see mf_prep.h.  Plain C, no GPU.
 */
#include "mf_prep.h"

#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#define MF_PREP_MAX_THREADS 8

/* MFGPU_TRACE=1: wall clock of the phases of a pattern build on stderr */
#include <stdio.h>
#include <time.h>
static double prep_now(void)
{
	struct timespec t;
	clock_gettime(CLOCK_MONOTONIC, &t);
	return t.tv_sec * 1e3 + t.tv_nsec * 1e-6;
}
#define PREP_LAP(what)                                                              \
	do {                                                                            \
		if (trace) {                                                                \
			const double n_ = prep_now();                                           \
			fprintf(stderr, "[mfgpu]   prep: %-22s %8.3f ms\n", what, n_ - t_last); \
			t_last = n_;                                                            \
		}                                                                           \
	} while (0)

static void *xcalloc(size_t n, size_t sz)
{
	return calloc(n ? n : 1, sz);
}

/* an array of the pattern: from the workspace when there is one; `zero` = start from zeros
 * like calloc (the big per-nonzero arrays are written in full and skip that) */
static void *palloc(mf_pattern *p, size_t n, size_t sz, int zero)
{
	if (!p->ws)
		return xcalloc(n, sz);
	size_t bytes = ((n ? n : 1) * sz + 63) & ~(size_t)63;
	if (p->ws_used + bytes > p->ws_bytes)
		return NULL;   /* reported as MFGPU_ENOMEM by the caller */
	void *q = p->ws + p->ws_used;
	p->ws_used += bytes;
	if (zero)
		memset(q, 0, bytes);
	return q;
}

size_t mf_pattern_ws_bytes(int users, int items, int nnz, int user_lo, int user_hi)
{
	size_t nu = (size_t)(user_hi - user_lo), n = (size_t)nnz, u = (size_t)users, it = (size_t)items;
	size_t ints = nu + it + (nu + 1) + (it + 1) + 4 * n + (u + 1) + u;   /* ids, ptrs, col/row idx, epos, mask */
	ints += n + 1;                                                       /* scratch of the two-step second pass */
	return ints * sizeof(int) + n * sizeof(double) + 17 * 64;
}

/* Order `n_ids` ids by key[id] descending, ties by ascending id (counting sort). */
static int order_by_length_desc(const int *count, int lo, int hi, int max_len,
                                int *out_ids)
{
	int *bucket = xcalloc((size_t)max_len + 2, sizeof *bucket);
	if (!bucket)
		return -1;
	for (int i = lo; i < hi; i++)
		if (count[i] > 0)
			bucket[max_len - count[i] + 1]++;
	for (int b = 0; b <= max_len; b++)
		bucket[b + 1] += bucket[b];
	int n = 0;
	for (int i = lo; i < hi; i++)
		if (count[i] > 0) {
			out_ids[bucket[max_len - count[i]]++] = i;
			n++;
		}
	free(bucket);
	return n;
}

int mf_mask_prefix(int users, int items, int nnz, const mfgpu_entry *entries)
{
	/* matFact.c:14-22: aix advances only when entries[aix] == (i, j) while the
	 * scan visits (i, j) in row-major order, so entry n can be matched only if
	 * every earlier entry was matched and (row, col) grew strictly. */
	int n = 0;
	long prev_r = -1, prev_c = -1;
	for (; n < nnz; n++) {
		long r = entries[n].row, c = entries[n].col;
		if (r < 0 || r >= users || c < 0 || c >= items)
			break;
		if (!(r > prev_r || (r == prev_r && c > prev_c)))
			break;
		prev_r = r;
		prev_c = c;
	}
	return n;
}

/* ---- the two passes over the entries, split over a few threads ------------------------------
 * Thread t owns the contiguous slice [n0, n1) of the file.  Pass 1 counts the slice's nonzeros
 * per row and per column; a prefix over the threads turns the counts into each thread's
 * first write position in every row and column, so that pass 2 can scatter the slices
 * concurrently and still leave every row and column in file order. */
typedef struct {
	const mfgpu_entry *entries;
	int n0, n1, users, items, user_lo, user_hi;
	int *row_cur, *col_cur;     /* [users], [items]: counts after pass 1, cursors in pass 2 */
	int local, bad;
	int first_unsorted;         /* pass 1: first n in the slice that does not grow strictly in (row, col) */
	int matched, mask_at;       /* pass 2: entries n < matched go to mask_col from position mask_at */
	mf_pattern *p;
	/* pass 2 in two steps (rows known never to decrease): */
	int *pos;                   /* [n_hi - n_lo] by-item position of entry n_base + x */
	int n_base;
	int r0, r1;                 /* stored rows this thread fills in prep_fill_rows */
	int first_free_here;        /* ... which also sets first_free of its users */
	const int *row_first;
} prep_slice;

static void *prep_count(void *arg)
{
	prep_slice *s = arg;
	int local = 0;
	long prev_r = -1, prev_c = -1;
	if (s->n0 > 0) {
		prev_r = s->entries[s->n0 - 1].row;
		prev_c = s->entries[s->n0 - 1].col;
	}
	s->first_unsorted = s->n1;
	for (int n = s->n0; n < s->n1; n++) {
		int r = s->entries[n].row, c = s->entries[n].col;
		if (r < 0 || r >= s->users || c < 0 || c >= s->items) {
			s->bad = 1;
			return NULL;
		}
		if (!(r > prev_r || (r == prev_r && c > prev_c)) && s->first_unsorted == s->n1)
			s->first_unsorted = n;
		prev_r = r;
		prev_c = c;
		if (r >= s->user_lo && r < s->user_hi) {
			s->row_cur[r]++;
			s->col_cur[c]++;
			local++;
		}
	}
	s->local = local;
	return NULL;
}

/* pass 1 for a list that mf_scan_entries has been over (valid, rows never decrease, so every
 * entry of the slice is the shard's and the rows' counts are row_first's differences): columns only */
static void *prep_count_cols(void *arg)
{
	prep_slice *s = arg;
	const mfgpu_entry *en = s->entries;
	int *col_cur = s->col_cur;
	for (int n = s->n0; n < s->n1; n++)
		col_cur[en[n].col]++;
	s->local = s->n1 - s->n0;
	s->first_unsorted = s->n1;
	return NULL;
}

static void *prep_scatter(void *arg)
{
	prep_slice *s = arg;
	mf_pattern *p = s->p;
	for (int n = s->n0; n < s->n1; n++) {
		int r = s->entries[n].row, c = s->entries[n].col;
		if (r < s->user_lo || r >= s->user_hi)
			continue;
		int pr = s->row_cur[r]++, pc = s->col_cur[c]++;
		p->col_idx[pr] = c;
		p->val[pr] = s->entries[n].value;
		p->epos[pr] = pc;
		p->row_idx[pc] = r;
		if (n < s->matched)            /* the matched prefix is (row, col)-sorted: a linear fill */
			p->mask_col[s->mask_at++] = c;
	}
	return NULL;
}

/* Pass 2 in two steps, for entry lists whose rows never decrease (a row's entries are then one
 * piece of the file, row_first).  prep_scatter above writes the by-user arrays row block by
 * row block in FILE order, and the stored rows are ordered by length: every ~15 entries three
 * arrays jump to a cold line (~20 ns per entry, 3.1 of the 4.7 ms of a build at the inst60000
 * shape).  Here the by-item side is done in file order (the only order that leaves a column's
 * rows ascending) with each entry's by-item position kept in a scratch array, and the by-user
 * arrays are then written front to back in STORED order, reading each row's piece of the file. */
static void *prep_scatter_cols(void *arg)
{
	prep_slice *s = arg;
	const mfgpu_entry *en = s->entries;
	int *row_idx = s->p->row_idx, *mask_col = s->p->mask_col;
	int *col_cur = s->col_cur, *pos = s->pos - s->n_base;
	int mask_at = s->mask_at;
	const int matched = s->matched;
	for (int n = s->n0; n < s->n1; n++) {
		const int c = en[n].col;
		const int pc = col_cur[c]++;
		row_idx[pc] = en[n].row;
		pos[n] = pc;
		if (n < matched)
			mask_col[mask_at++] = c;
	}
	return NULL;
}

static void *prep_fill_rows(void *arg)
{
	prep_slice *s = arg;
	const mf_pattern *p = s->p;
	const mfgpu_entry *en = s->entries;
	const int *pos = s->pos - s->n_base;
	for (int r = s->r0; r < s->r1; r++) {
		if (r + 8 < s->r1) {                 /* the piece of the file of a row soon to come */
			const mfgpu_entry *nx = en + s->row_first[p->row_id[r + 8]];
			__builtin_prefetch(nx);
			__builtin_prefetch(nx + 4);
			__builtin_prefetch(nx + 8);
			__builtin_prefetch(nx + 12);
		}
		const int i = p->row_id[r];
		int n = s->row_first[i];
		if (s->first_free_here) {
			/* every entry matched: the user's mask is its piece of the file, ascending and
			 * distinct, and the first item the cursor does not skip is the first a with
			 * col[a] != a (see the serial loop in mf_pattern_build_scan) */
			const int cnt = p->row_ptr[r + 1] - p->row_ptr[r];
			int a = 0;
			while (a < cnt && en[n + a].col == a)
				a++;
			p->first_free[i] = a < p->items ? a : -1;
		}
		for (int k = p->row_ptr[r]; k < p->row_ptr[r + 1]; k++, n++) {
			p->col_idx[k] = en[n].col;
			p->val[k] = en[n].value;
			p->epos[k] = pos[n];
		}
	}
	return NULL;
}

static void run_slices(prep_slice *sl, int T, void *(*fn)(void *))
{
	pthread_t th[MF_PREP_MAX_THREADS];
	int started[MF_PREP_MAX_THREADS] = {0};
	for (int t = 1; t < T; t++)
		started[t] = pthread_create(&th[t], NULL, fn, &sl[t]) == 0;
	fn(&sl[0]);
	for (int t = 1; t < T; t++) {
		if (started[t])
			pthread_join(th[t], NULL);
		else
			fn(&sl[t]);       /* no thread to be had: do the slice here */
	}
}

static int prep_threads(int nnz)
{
	long cpus = sysconf(_SC_NPROCESSORS_ONLN);
	int T = nnz / 131072;                 /* a slice is worth a thread from ~10^5 entries */
	if (T > MF_PREP_MAX_THREADS) T = MF_PREP_MAX_THREADS;
	if (cpus > 0 && T > cpus) T = (int)cpus;
	const char *env = getenv("MFGPU_PREP_THREADS");
	if (env && atoi(env) > 0)
		T = atoi(env) > MF_PREP_MAX_THREADS ? MF_PREP_MAX_THREADS : atoi(env);
	return T < 1 ? 1 : T;
}

int mf_pattern_build(mf_pattern *p, int users, int items, int nnz,
                     const mfgpu_entry *entries, int user_lo, int user_hi)
{
	return mf_pattern_build_ws(p, users, items, nnz, entries, user_lo, user_hi, NULL, 0);
}

int mf_pattern_build_ws(mf_pattern *p, int users, int items, int nnz,
                        const mfgpu_entry *entries, int user_lo, int user_hi,
                        void *ws, size_t ws_bytes)
{
	return mf_pattern_build_scan(p, users, items, nnz, entries, user_lo, user_hi, ws, ws_bytes, NULL);
}

int mf_pattern_build_scan(mf_pattern *p, int users, int items, int nnz,
                          const mfgpu_entry *entries, int user_lo, int user_hi,
                          void *ws, size_t ws_bytes, const mf_scan *scan)
{
	memset(p, 0, sizeof *p);
	p->ws = (char *)ws;
	p->ws_bytes = ws ? ws_bytes : 0;
	if (users < 0 || items < 0 || nnz < 0 || user_lo < 0 || user_hi > users || user_lo > user_hi)
		return MFGPU_EINVAL;
	if (nnz > 0 && !entries)
		return MFGPU_EINVAL;
	p->users = users;
	p->items = items;
	p->user_lo = user_lo;
	p->user_hi = user_hi;

	/* the entries that can belong to this shard: all of them, or -- rows known never to decrease
	 * (mf_scan_entries) -- the slice of its users */
	int n_lo = 0, n_hi = nnz;
	if (scan && scan->by_row && scan->row_first) {
		n_lo = scan->row_first[user_lo];
		n_hi = scan->row_first[user_hi];
	}
	/* rows known never to decrease: the lighter passes (prep_count_cols, prep_scatter_cols + prep_fill_rows) */
	const int two_step = scan && scan->by_row && scan->row_first && !getenv("MFGPU_PREP_ONE_STEP");
	const int trace = getenv("MFGPU_TRACE") != NULL;
	double t_last = trace ? prep_now() : 0.0;
	const int T = prep_threads(n_hi - n_lo);
	prep_slice sl[MF_PREP_MAX_THREADS];
	int *cur = NULL;            /* threads 1..T-1: their per-row / per-column arrays */
	int *pos = NULL;            /* pass 2 in two steps: by-item position of every entry of the slice */
	int pos_heap = 0;
	int *row_count = xcalloc((size_t)users, sizeof *row_count);
	int *col_count = xcalloc((size_t)items, sizeof *col_count);
	int *row_slot = xcalloc((size_t)users, sizeof *row_slot);
	int *col_slot = xcalloc((size_t)items, sizeof *col_slot);
	int rc = MFGPU_ENOMEM;
	if (!row_count || !col_count || !row_slot || !col_slot)
		goto done;
	if (T > 1) {
		cur = xcalloc((size_t)(T - 1) * ((size_t)users + items), sizeof *cur);
		if (!cur)
			goto done;
	}
	for (int t = 0; t < T; t++) {
		sl[t].entries = entries;
		sl[t].n0 = n_lo + (int)((long)(n_hi - n_lo) * t / T);
		sl[t].n1 = n_lo + (int)((long)(n_hi - n_lo) * (t + 1) / T);
		sl[t].users = users;
		sl[t].items = items;
		sl[t].user_lo = user_lo;
		sl[t].user_hi = user_hi;
		/* thread 0 counts straight into row_slot / col_slot */
		sl[t].row_cur = t == 0 ? row_slot : cur + (size_t)(t - 1) * ((size_t)users + items);
		sl[t].col_cur = t == 0 ? col_slot : sl[t].row_cur + users;
		sl[t].local = 0;
		sl[t].bad = 0;
		sl[t].p = p;
	}
	PREP_LAP("alloc");
	run_slices(sl, T, two_step ? prep_count_cols : prep_count);
	PREP_LAP("count (threads)");
	int local = 0;
	for (int t = 0; t < T; t++) {
		if (sl[t].bad) {
			rc = MFGPU_EINVAL;
			goto done;
		}
		local += sl[t].local;
	}
	/* matFact.c:14-22: the entries the cursor of print_output ever matches (mf_mask_prefix) */
	int matched = nnz;
	for (int t = T - 1; t >= 0; t--)
		if (sl[t].first_unsorted < sl[t].n1)
			matched = sl[t].first_unsorted;
	if (scan)
		matched = scan->matched;         /* of the whole list, not of the slice */
	for (int i = user_lo; i < user_hi; i++) {
		int c = 0;
		if (two_step)
			c = scan->row_first[i + 1] - scan->row_first[i];
		else
			for (int t = 0; t < T; t++)
				c += sl[t].row_cur[i];
		row_count[i] = c;
	}
	for (int j = 0; j < items; j++) {
		int c = 0;
		for (int t = 0; t < T; t++)
			c += sl[t].col_cur[j];
		col_count[j] = c;
	}
	p->nnz = local;
	for (int i = user_lo; i < user_hi; i++)
		if (row_count[i] > p->max_row_len)
			p->max_row_len = row_count[i];
	for (int j = 0; j < items; j++)
		if (col_count[j] > p->max_col_len)
			p->max_col_len = col_count[j];

	PREP_LAP("sum counts");
	p->row_id = palloc(p, (size_t)(user_hi - user_lo), sizeof(int), 0);
	p->col_id = palloc(p, (size_t)items, sizeof(int), 0);
	if (!p->row_id || !p->col_id)
		goto done;
	p->n_rows = order_by_length_desc(row_count, user_lo, user_hi, p->max_row_len, p->row_id);
	p->n_cols = order_by_length_desc(col_count, 0, items, p->max_col_len, p->col_id);
	if (p->n_rows < 0 || p->n_cols < 0)
		goto done;

	PREP_LAP("order by length");
	p->row_ptr = palloc(p, (size_t)p->n_rows + 1, sizeof(int), 1);
	p->col_ptr = palloc(p, (size_t)p->n_cols + 1, sizeof(int), 1);
	p->col_idx = palloc(p, (size_t)local, sizeof(int), 0);
	p->row_idx = palloc(p, (size_t)local, sizeof(int), 0);
	p->epos = palloc(p, (size_t)local, sizeof(int), 0);
	p->val = palloc(p, (size_t)local, sizeof(double), 0);
	if (!p->row_ptr || !p->col_ptr || !p->col_idx || !p->row_idx || !p->epos || !p->val)
		goto done;
	/* first write position of thread t in row / column x = start of x + what the threads
	 * before t put there; the scatter then keeps the file order inside every row and column */
	for (int r = 0; r < p->n_rows; r++) {
		const int i = p->row_id[r];
		p->row_ptr[r + 1] = p->row_ptr[r] + row_count[i];
		int at = p->row_ptr[r];
		for (int t = 0; t < T && !two_step; t++) {
			const int k = sl[t].row_cur[i];
			sl[t].row_cur[i] = at;
			at += k;
		}
	}
	for (int c = 0; c < p->n_cols; c++) {
		const int j = p->col_id[c];
		p->col_ptr[c + 1] = p->col_ptr[c] + col_count[j];
		int at = p->col_ptr[c];
		for (int t = 0; t < T; t++) {
			const int k = sl[t].col_cur[j];
			sl[t].col_cur[j] = at;
			at += k;
		}
	}
	PREP_LAP("pointers");
	/* rated-item mask, matFact.c:14-22: per user the items of its matched entries */
	p->mask_ptr = palloc(p, (size_t)users + 1, sizeof(int), 1);
	p->first_free = palloc(p, (size_t)users, sizeof(int), 0);
	if (!p->mask_ptr || !p->first_free)
		goto done;
	if (matched == nnz) {
		for (int i = user_lo; i < user_hi; i++)
			p->mask_ptr[i + 1] = row_count[i];
	} else {
		for (int n = n_lo; n < matched && n < n_hi; n++) {
			int r = entries[n].row;
			if (r >= user_lo && r < user_hi)
				p->mask_ptr[r + 1]++;
		}
	}
	for (int i = 0; i < users; i++)
		p->mask_ptr[i + 1] += p->mask_ptr[i];
	p->mask_col = palloc(p, (size_t)p->mask_ptr[users], sizeof(int), 0);
	if (!p->mask_col)
		goto done;
	for (int t = 0, at = 0; t < T; t++) {
		sl[t].matched = matched;
		sl[t].mask_at = at;             /* local entries before the slice */
		at += sl[t].local;
	}
	PREP_LAP("mask pointers");
	/* every entry matched (the sorted files of the reference): first_free by the threads of the
	 * by-user step; users without entries (not among the stored rows) are set here */
	const int fused_ff = two_step && matched == nnz;
	if (fused_ff)
		for (int i = 0; i < users; i++)
			p->first_free[i] = (i >= user_lo && i < user_hi && items > 0) ? 0 : -1;
	if (two_step) {
		/* from the workspace when there is room (page-locked and recycled between plans: no
		 * page faults on 4 bytes per nonzero), else from the heap */
		if (p->ws)
			pos = palloc(p, (size_t)(n_hi - n_lo + 1), sizeof *pos, 0);
		if (!pos) {
			pos = malloc(sizeof *pos * (size_t)(n_hi - n_lo + 1));
			pos_heap = 1;
		}
		if (!pos) {
			rc = MFGPU_ENOMEM;
			goto done;
		}
		for (int t = 0; t < T; t++) {
			sl[t].pos = pos;
			sl[t].first_free_here = fused_ff;
			sl[t].n_base = n_lo;
			sl[t].row_first = scan->row_first;
			/* stored rows by equal shares of the nonzeros */
			const int want = (int)((long)local * t / T);
			int a = 0, b = p->n_rows;
			while (a < b) {                      /* first stored row that starts at or after `want` */
				const int m = (a + b) / 2;
				if (p->row_ptr[m] < want)
					a = m + 1;
				else
					b = m;
			}
			sl[t].r0 = a;
			if (t > 0)
				sl[t - 1].r1 = a;
		}
		sl[0].r0 = 0;
		sl[T - 1].r1 = p->n_rows;
		run_slices(sl, T, prep_scatter_cols);
		PREP_LAP("by item (threads)");
		run_slices(sl, T, prep_fill_rows);
		PREP_LAP("by user (threads)");
	} else {
		run_slices(sl, T, prep_scatter);
		PREP_LAP("scatter (threads)");
	}

	for (int i = 0; i < users && !fused_ff; i++) {
		int f = -1;
		if (i >= user_lo && i < user_hi) {
			const int *m = p->mask_col + p->mask_ptr[i];
			int cnt = p->mask_ptr[i + 1] - p->mask_ptr[i];
			/* first j with m[j] != j (m ascending, distinct): almost always j = 0 or 1, so a
			 * forward scan beats a bisection (one predictable branch, one cache line) */
			int a = 0;
			while (a < cnt && m[a] == a)
				a++;
			f = (a < items) ? a : -1;
		}
		p->first_free[i] = f;
	}
	PREP_LAP("first_free");
	rc = MFGPU_OK;
done:
	if (pos_heap)
		free(pos);
	free(cur);
	free(row_count);
	free(col_count);
	free(row_slot);
	free(col_slot);
	if (rc != MFGPU_OK)
		mf_pattern_free(p);
	return rc;
}

void mf_pattern_free(mf_pattern *p)
{
	if (p->ws) {                 /* the arrays belong to the caller's workspace */
		memset(p, 0, sizeof *p);
		return;
	}
	free(p->row_id); free(p->row_ptr); free(p->col_idx); free(p->val); free(p->epos);
	free(p->col_id); free(p->col_ptr); free(p->row_idx);
	free(p->mask_ptr); free(p->mask_col); free(p->first_free);
	memset(p, 0, sizeof *p);
}

/* ---- mf_scan_entries: one pass for the whole list ------------------------------------------ */
typedef struct {
	const mfgpu_entry *entries;
	int n0, n1, users, items;
	int *row_cnt;               /* [users] of this thread   */
	unsigned char *col_seen;    /* [items] of this thread   */
	int bad, row_desc, first_unsorted;
	/* second pass */
	const int *cut, *item_slot;
	int nshards;
	int *contrib;               /* [n_touched] of this thread */
} scan_slice;

static void *scan_count(void *arg)
{
	scan_slice *s = arg;
	long prev_r = -1, prev_c = -1;
	if (s->n0 > 0) {
		prev_r = s->entries[s->n0 - 1].row;
		prev_c = s->entries[s->n0 - 1].col;
	}
	s->first_unsorted = s->n1;
	for (int n = s->n0; n < s->n1; n++) {
		const int r = s->entries[n].row, c = s->entries[n].col;
		if (r < 0 || r >= s->users || c < 0 || c >= s->items) {
			s->bad = 1;
			return NULL;
		}
		if (r < prev_r)
			s->row_desc = 1;
		if (!(r > prev_r || (r == prev_r && c > prev_c)) && s->first_unsorted == s->n1)
			s->first_unsorted = n;
		prev_r = r;
		prev_c = c;
		s->row_cnt[r]++;
		s->col_seen[c] = 1;
	}
	return NULL;
}

static void *scan_contrib(void *arg)
{
	scan_slice *s = arg;
	int r = 0;
	for (int n = s->n0; n < s->n1; n++) {
		const int u = s->entries[n].row;
		while (r + 1 < s->nshards && u >= s->cut[r + 1])
			r++;
		while (r > 0 && u < s->cut[r])
			r--;
		s->contrib[s->item_slot[s->entries[n].col]] |= 1 << r;
	}
	return NULL;
}

static void run_scan(scan_slice *sl, int T, void *(*fn)(void *))
{
	pthread_t th[MF_PREP_MAX_THREADS];
	int started[MF_PREP_MAX_THREADS] = {0};
	for (int t = 1; t < T; t++)
		started[t] = pthread_create(&th[t], NULL, fn, &sl[t]) == 0;
	fn(&sl[0]);
	for (int t = 1; t < T; t++) {
		if (started[t])
			pthread_join(th[t], NULL);
		else
			fn(&sl[t]);
	}
}

int mf_scan_entries(int users, int items, int nnz, const mfgpu_entry *entries, int nshards,
                    int *cut, int *item_slot, int *row_first, mf_scan *scan,
                    int *contrib, int *slots_per_rank)
{
	if (users < 0 || items < 0 || nnz < 0 || nshards < 1 || nshards > 31 || !cut || !item_slot || !row_first || !scan)
		return MFGPU_EINVAL;
	if (nnz > 0 && !entries)
		return MFGPU_EINVAL;
	const int T = prep_threads(nnz);
	scan_slice sl[MF_PREP_MAX_THREADS];
	int *cnt = xcalloc((size_t)T * (size_t)users, sizeof *cnt);
	unsigned char *seen = xcalloc((size_t)T * (size_t)items, 1);
	int *ctr = NULL;
	int rc = MFGPU_ENOMEM, n_touched = 0;
	if (!cnt || !seen)
		goto done;
	memset(sl, 0, sizeof sl);
	for (int t = 0; t < T; t++) {
		sl[t].entries = entries;
		sl[t].n0 = (int)((long)nnz * t / T);
		sl[t].n1 = (int)((long)nnz * (t + 1) / T);
		sl[t].users = users;
		sl[t].items = items;
		sl[t].row_cnt = cnt + (size_t)t * users;
		sl[t].col_seen = seen + (size_t)t * items;
	}
	run_scan(sl, T, scan_count);
	int by_row = 1, matched = nnz;
	for (int t = T - 1; t >= 0; t--) {
		if (sl[t].bad) {
			rc = MFGPU_EINVAL;
			goto done;
		}
		if (sl[t].row_desc)
			by_row = 0;
		if (sl[t].first_unsorted < sl[t].n1)
			matched = sl[t].first_unsorted;
	}
	/* nonzeros per user -> where each user's entries start (by_row) and the balanced cut */
	long run = 0;
	int s = 1;
	cut[0] = 0;
	for (int i = 0; i < users; i++) {
		row_first[i] = (int)run;
		/* smallest i with (nonzeros before i) >= s*nnz/nshards, as mf_split_users */
		while (s < nshards && run >= (long)((double)nnz * s / nshards + 0.5))
		{
			cut[s] = nnz ? i : (int)((long)s * users / nshards);
			s++;
		}
		for (int t = 0; t < T; t++)
			run += sl[t].row_cnt[i];
	}
	row_first[users] = (int)run;
	while (s < nshards) {
		cut[s] = nnz ? users : (int)((long)s * users / nshards);
		s++;
	}
	cut[nshards] = users;
	for (int j = 0; j < items; j++) {
		int any = 0;
		for (int t = 0; t < T; t++)
			any |= sl[t].col_seen[j];
		item_slot[j] = any ? n_touched++ : -1;
	}
	scan->by_row = by_row;
	scan->matched = matched;
	scan->row_first = row_first;
	if (contrib) {
		ctr = xcalloc((size_t)T * (size_t)(n_touched ? n_touched : 1), sizeof *ctr);
		if (!ctr)
			goto done;
		for (int t = 0; t < T; t++) {
			sl[t].cut = cut;
			sl[t].item_slot = item_slot;
			sl[t].nshards = nshards;
			sl[t].contrib = ctr + (size_t)t * n_touched;
		}
		run_scan(sl, T, scan_contrib);
		for (int k = 0; k < n_touched; k++) {
			int m = 0;
			for (int t = 0; t < T; t++)
				m |= sl[t].contrib[k];
			contrib[k] = m;
		}
		if (slots_per_rank) {
			const int per = (n_touched + nshards - 1) / nshards;
			*slots_per_rank = per > 0 ? per : 1;
		}
	}
	rc = n_touched;
done:
	free(cnt);
	free(seen);
	free(ctr);
	return rc;
}

int mf_split_users(int users, int nnz, const mfgpu_entry *entries, int nshards, int *cut)
{
	if (users < 0 || nnz < 0 || nshards < 1 || !cut)
		return MFGPU_EINVAL;
	long *pre = xcalloc((size_t)users + 1, sizeof *pre);
	if (!pre)
		return MFGPU_ENOMEM;
	for (int n = 0; n < nnz; n++) {
		int r = entries[n].row;
		if (r < 0 || r >= users) {
			free(pre);
			return MFGPU_EINVAL;
		}
		pre[r + 1]++;
	}
	for (int i = 0; i < users; i++)
		pre[i + 1] += pre[i];
	cut[0] = 0;
	int i = 0;
	for (int s = 1; s < nshards; s++) {
		/* smallest i with pre[i] >= s*nnz/nshards, never moving backwards */
		long target = (long)((double)nnz * s / nshards + 0.5);
		while (i < users && pre[i] < target)
			i++;
		/* with no nonzeros at all fall back to the reference's equal-count split */
		cut[s] = nnz ? i : (int)((long)s * users / nshards);
	}
	cut[nshards] = users;
	free(pre);
	return MFGPU_OK;
}

int mf_touched_items(int items, int nnz, const mfgpu_entry *entries, int *item_slot)
{
	for (int j = 0; j < items; j++)
		item_slot[j] = -1;
	for (int n = 0; n < nnz; n++) {
		int c = entries[n].col;
		if (c < 0 || c >= items)
			return MFGPU_EINVAL;
		item_slot[c] = 0;
	}
	int cnt = 0;
	for (int j = 0; j < items; j++)
		if (item_slot[j] == 0)
			item_slot[j] = cnt++;
	return cnt;
}

/* Layout of the rank-mode exchange (csrc/mf_iter.cu, k_xchg_ll / k_reduce_bcast): the touched
 * items (item_slot >= 0, mf_touched_items) are dealt to the ranks in blocks of
 * per = ceil(n_touched / nranks) consecutive slots -- rank slot / per adds up the
 * contributions to the item and stores its new row into every replica -- and contrib[slot]
 * has bit r set when rank r (users [cut[r], cut[r+1])) holds a nonzero of the item: the only
 * contributions the reducing rank waits for.  Returns per, or an MFGPU_E* code (< 0). */
int mf_exchange_layout(int users, int nnz, const mfgpu_entry *entries, int nranks, const int *cut,
                       const int *item_slot, int n_touched, int *contrib)
{
	if (nranks < 1 || nranks > 31 || n_touched < 0 || !cut || !item_slot || !contrib)
		return MFGPU_EINVAL;
	for (int t = 0; t < n_touched; t++)
		contrib[t] = 0;
	int r = 0;
	for (int n = 0; n < nnz; n++) {
		const int u = entries[n].row;
		if (u < 0 || u >= users)
			return MFGPU_EINVAL;
		/* entries usually come sorted by user: the rank of the previous one is a good guess */
		while (r + 1 < nranks && u >= cut[r + 1])      /* cut[nranks] = users > u */
			r++;
		while (r > 0 && u < cut[r])                    /* cut[0] = 0 <= u */
			r--;
		const int t = item_slot[entries[n].col];
		if (t < 0 || t >= n_touched)
			return MFGPU_EINVAL;
		contrib[t] |= 1 << r;
	}
	const int per = (n_touched + nranks - 1) / nranks;
	return per > 0 ? per : 1;
}

