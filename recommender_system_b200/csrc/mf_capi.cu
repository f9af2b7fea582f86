/*
 * csrc/mf_capi.cu -- implementation of include/matfact_b200.h (libmatfact_b200.so).
 *
 * Host side of the CUDA path: owns the device buffers, builds the two stable
 * groupings of the nonzeros (host/mf_prep.c), launches the kernels of
 * mf_loop.cuh / mf_argmax.cuh, captures the iteration loop in CUDA graphs, and
 * for N shards exchanges the partial R gradient (peer memory inside one
 * process, NCCL between processes).  There is no CPU fallback anywhere in this
 * file: without an sm_100 device every compute entry returns MFGPU_ENODEV.
 */
#include "mf_internal.h"
#include "../host/mf_ingest.h"

#include <dlfcn.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <new>

/* ---- errors ------------------------------------------------------------------ */

static thread_local char g_err[512] = "";

int mf_fail(int code, const char *fmt, ...)
{
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(g_err, sizeof g_err, fmt, ap);
	va_end(ap);
	return code;
}
#define fail mf_fail

extern "C" const char *mfgpu_last_error(void) { return g_err; }
extern "C" const char *mfgpu_version(void) { return "matfact_b200 0.1 (sm_100a)"; }

/* Devices of compute capability 10.x, in CUDA order. */
int mf_usable_devices(int *ids, int cap)
{
	int n = 0, found = 0;
	if (cudaGetDeviceCount(&n) != cudaSuccess) {
		cudaGetLastError();
		return 0;
	}
	for (int d = 0; d < n; d++) {
		int major = 0;
		if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, d) != cudaSuccess)
			continue;
		if (major == 10) {
			if (ids && found < cap)
				ids[found] = d;
			found++;
		}
	}
	return found;
}

extern "C" int mfgpu_device_count(void) { return mf_usable_devices(nullptr, 0); }

/* ---- NCCL binding (dlopen; the single-process paths never need it) ------------ */

mf_nccl_api g_nccl;

int mf_load_nccl(void)
{
	if (g_nccl.handle)
		return MFGPU_OK;
	void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
	if (!h)
		h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
	if (!h)
		return fail(MFGPU_ECOMM, "cannot load libnccl.so.2: %s", dlerror());
#define SYM(field, name)                                                    \
	*(void **)(&g_nccl.field) = dlsym(h, name);                             \
	if (!g_nccl.field)                                                      \
		return fail(MFGPU_ECOMM, "libnccl.so.2 lacks %s", name)
	SYM(GetUniqueId, "ncclGetUniqueId");
	SYM(CommInitRank, "ncclCommInitRank");
	SYM(CommDestroy, "ncclCommDestroy");
	SYM(AllReduce, "ncclAllReduce");
	SYM(AllGather, "ncclAllGather");
	SYM(GetErrorString, "ncclGetErrorString");
#undef SYM
	g_nccl.handle = h;
	return MFGPU_OK;
}

extern "C" int mfgpu_comm_unique_id(unsigned char uid[MFGPU_UID_BYTES])
{
	TRY(mf_load_nccl());
	mf_ncclUniqueId id;
	NCK(g_nccl.GetUniqueId(&id));
	memcpy(uid, id.internal, MFGPU_UID_BYTES);
	return MFGPU_OK;
}

extern "C" int mfgpu_comm_create(mfgpu_comm **out, int nranks, int rank,
                                 const unsigned char uid[MFGPU_UID_BYTES], int device)
{
	if (!out || nranks < 1 || rank < 0 || rank >= nranks || nranks > MF_MAX_SHARDS)
		return fail(MFGPU_EINVAL, "mfgpu_comm_create: bad rank %d of %d", rank, nranks);
	int major = 0;
	if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device) != cudaSuccess ||
	    major != 10) {
		cudaGetLastError();
		return fail(MFGPU_ENODEV, "device %d is not an sm_100 GPU", device);
	}
	TRY(mf_load_nccl());
	CK(cudaSetDevice(device));
	mfgpu_comm *c = new (std::nothrow) mfgpu_comm();
	if (!c)
		return fail(MFGPU_ENOMEM, "out of host memory");
	c->nranks = nranks;
	c->rank = rank;
	c->device = device;
	mf_ncclUniqueId id;
	memcpy(id.internal, uid, MFGPU_UID_BYTES);
	int r = g_nccl.CommInitRank(&c->nccl, nranks, id, rank);
	if (r != 0) {
		delete c;
		return fail(MFGPU_ECOMM, "ncclCommInitRank failed: %s", g_nccl.GetErrorString(r));
	}
	*out = c;
	return MFGPU_OK;
}

extern "C" void mfgpu_comm_destroy(mfgpu_comm *c)
{
	if (!c)
		return;
	mf_arena_reset(c);               /* collective, like the destruction of the communicator */
	if (c->nccl && g_nccl.handle)
		g_nccl.CommDestroy(c->nccl);
	delete c;
}

/* ---- host-side helpers exported for tests and tools ---------------------------- */

extern "C" int mfgpu_split_users(int users, int nnz, const mfgpu_entry *entries, int nshards, int *cut)
{
	int rc = mf_split_users(users, nnz, entries, nshards, cut);
	return rc == MFGPU_OK ? rc : fail(rc, "mfgpu_split_users: bad argument");
}

extern "C" int mfgpu_exchange_layout(int users, int items, int nnz, const mfgpu_entry *entries, int nranks,
                                     int *item_slot, int *contrib, int *slots_per_rank)
{
	if (!item_slot || !contrib || !slots_per_rank || nranks < 1 || nranks > MF_MAX_SHARDS || items < 0)
		return fail(MFGPU_EINVAL, "mfgpu_exchange_layout: bad argument");
	int cut[MF_MAX_SHARDS + 1];
	int rc = mf_split_users(users, nnz, entries, nranks, cut);
	if (rc != MFGPU_OK)
		return fail(rc, "mfgpu_exchange_layout: bad argument");
	const int cnt = mf_touched_items(items, nnz, entries, item_slot);
	if (cnt < 0)
		return fail(cnt, "mfgpu_exchange_layout: an entry lies outside the users x items matrix");
	const int per = mf_exchange_layout(users, nnz, entries, nranks, cut, item_slot, cnt, contrib);
	if (per < 0)
		return fail(per, "mfgpu_exchange_layout: bad argument");
	*slots_per_rank = per;
	return cnt;
}

/* mat2d.c:61-72 (draw order: all of L row-major, then R_init = features x items
 * row-major) with R written straight into the transposed position that
 * mat2d_transpose (mat2d.c:115-124, matFact.c:118-120) would give it. */
extern "C" void mfgpu_random_fill_LR(int users, int items, int features, double *L, double *R)
{
	mf_random_fill_LR(users, items, features, L, R);      /* host/mf_ingest.c */
}

extern "C" int mfgpu_read_instance(const char *path, int *iters, double *alpha, int *features,
                                   int *users, int *items, int *non_zero, mfgpu_entry **entries)
{
	if (!path || !iters || !alpha || !features || !users || !items || !non_zero || !entries)
		return MF_PARSE_OPEN;
	mf_instance in;
	const int rc = mf_parse_instance(path, &in);
	*entries = nullptr;
	if (rc != MF_PARSE_OK)
		return rc;
	*iters = in.iters;
	*alpha = in.alpha;
	*features = in.features;
	*users = in.users;
	*items = in.items;
	*non_zero = in.non_zero;
	*entries = in.entries;
	return 0;
}

extern "C" const char *mfgpu_read_error(int rc) { return mf_parse_message(rc); }

extern "C" void mfgpu_free(void *p) { free(p); }

extern "C" long mfgpu_selftest_random(unsigned seed, long n)
{
	mf_rand g;
	mf_rand_seed(&g, seed);
	srandom(seed);
	for (long i = 0; i < n; i++)
		if ((long)mf_rand_next(&g) != random())
			return i;
	return -1;
}

extern "C" int mfgpu_mask_stats(int users, int items, int nnz, const mfgpu_entry *entries,
                                int *mask_count, int *first_free)
{
	mf_pattern p;
	int rc = mf_pattern_build(&p, users, items, nnz, entries, 0, users);
	if (rc != MFGPU_OK)
		return fail(rc, "mfgpu_mask_stats: bad argument or out of memory");
	for (int i = 0; i < users; i++) {
		if (mask_count) mask_count[i] = p.mask_ptr[i + 1] - p.mask_ptr[i];
		if (first_free) first_free[i] = p.first_free[i];
	}
	int total = p.mask_ptr[users];
	mf_pattern_free(&p);
	return total;
}
