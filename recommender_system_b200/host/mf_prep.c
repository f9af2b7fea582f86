/*
 * host/mf_prep.c -- see mf_prep.h.  Plain C, no GPU.
 */
#include "mf_prep.h"

#include <stdlib.h>
#include <string.h>

static void *xcalloc(size_t n, size_t sz)
{
	return calloc(n ? n : 1, sz);
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

int mf_pattern_build(mf_pattern *p, int users, int items, int nnz,
                     const mfgpu_entry *entries, int user_lo, int user_hi)
{
	memset(p, 0, sizeof *p);
	if (users < 0 || items < 0 || nnz < 0 || user_lo < 0 || user_hi > users || user_lo > user_hi)
		return MFGPU_EINVAL;
	if (nnz > 0 && !entries)
		return MFGPU_EINVAL;
	p->users = users;
	p->items = items;
	p->user_lo = user_lo;
	p->user_hi = user_hi;

	int *row_count = xcalloc((size_t)users, sizeof *row_count);
	int *col_count = xcalloc((size_t)items, sizeof *col_count);
	int *row_slot = xcalloc((size_t)users, sizeof *row_slot);
	int *col_slot = xcalloc((size_t)items, sizeof *col_slot);
	int rc = MFGPU_ENOMEM;
	if (!row_count || !col_count || !row_slot || !col_slot)
		goto done;

	int local = 0;
	for (int n = 0; n < nnz; n++) {
		int r = entries[n].row, c = entries[n].col;
		if (r < 0 || r >= users || c < 0 || c >= items) {
			rc = MFGPU_EINVAL;
			goto done;
		}
		if (r >= user_lo && r < user_hi) {
			row_count[r]++;
			col_count[c]++;
			local++;
		}
	}
	p->nnz = local;
	for (int i = user_lo; i < user_hi; i++)
		if (row_count[i] > p->max_row_len)
			p->max_row_len = row_count[i];
	for (int j = 0; j < items; j++)
		if (col_count[j] > p->max_col_len)
			p->max_col_len = col_count[j];

	p->row_id = xcalloc((size_t)(user_hi - user_lo), sizeof(int));
	p->col_id = xcalloc((size_t)items, sizeof(int));
	if (!p->row_id || !p->col_id)
		goto done;
	p->n_rows = order_by_length_desc(row_count, user_lo, user_hi, p->max_row_len, p->row_id);
	p->n_cols = order_by_length_desc(col_count, 0, items, p->max_col_len, p->col_id);
	if (p->n_rows < 0 || p->n_cols < 0)
		goto done;

	p->row_ptr = xcalloc((size_t)p->n_rows + 1, sizeof(int));
	p->col_ptr = xcalloc((size_t)p->n_cols + 1, sizeof(int));
	p->col_idx = xcalloc((size_t)local, sizeof(int));
	p->row_idx = xcalloc((size_t)local, sizeof(int));
	p->epos = xcalloc((size_t)local, sizeof(int));
	p->val = xcalloc((size_t)local, sizeof(double));
	if (!p->row_ptr || !p->col_ptr || !p->col_idx || !p->row_idx || !p->epos || !p->val)
		goto done;
	for (int r = 0; r < p->n_rows; r++) {
		p->row_ptr[r + 1] = p->row_ptr[r] + row_count[p->row_id[r]];
		row_slot[p->row_id[r]] = p->row_ptr[r];       /* running write cursor */
	}
	for (int c = 0; c < p->n_cols; c++) {
		p->col_ptr[c + 1] = p->col_ptr[c] + col_count[p->col_id[c]];
		col_slot[p->col_id[c]] = p->col_ptr[c];
	}
	/* one pass in file order keeps the file order inside every row and column */
	for (int n = 0; n < nnz; n++) {
		int r = entries[n].row, c = entries[n].col;
		if (r < user_lo || r >= user_hi)
			continue;
		int pr = row_slot[r]++, pc = col_slot[c]++;
		p->col_idx[pr] = c;
		p->val[pr] = entries[n].value;
		p->epos[pr] = pc;
		p->row_idx[pc] = r;
	}

	/* rated-item mask, matFact.c:14-22 */
	p->mask_ptr = xcalloc((size_t)users + 1, sizeof(int));
	p->first_free = xcalloc((size_t)users, sizeof(int));
	if (!p->mask_ptr || !p->first_free)
		goto done;
	int matched = mf_mask_prefix(users, items, nnz, entries);
	for (int n = 0; n < matched; n++) {
		int r = entries[n].row;
		if (r >= user_lo && r < user_hi)
			p->mask_ptr[r + 1]++;
	}
	for (int i = 0; i < users; i++)
		p->mask_ptr[i + 1] += p->mask_ptr[i];
	p->mask_col = xcalloc((size_t)p->mask_ptr[users], sizeof(int));
	if (!p->mask_col)
		goto done;
	{
		int w = 0; /* the matched prefix is (row, col)-sorted, so a linear fill is in order */
		for (int n = 0; n < matched; n++) {
			int r = entries[n].row;
			if (r >= user_lo && r < user_hi)
				p->mask_col[w++] = entries[n].col;
		}
	}
	for (int i = 0; i < users; i++) {
		int f = -1;
		if (i >= user_lo && i < user_hi) {
			const int *m = p->mask_col + p->mask_ptr[i];
			int cnt = p->mask_ptr[i + 1] - p->mask_ptr[i];
			/* first j with m[j] != j (m ascending, distinct) */
			int a = 0, b = cnt;
			while (a < b) {
				int mid = a + (b - a) / 2;
				if (m[mid] == mid) a = mid + 1; else b = mid;
			}
			f = (a < items) ? a : -1;
		}
		p->first_free[i] = f;
	}
	rc = MFGPU_OK;
done:
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
	free(p->row_id); free(p->row_ptr); free(p->col_idx); free(p->val); free(p->epos);
	free(p->col_id); free(p->col_ptr); free(p->row_idx);
	free(p->mask_ptr); free(p->mask_col); free(p->first_free);
	memset(p, 0, sizeof *p);
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
