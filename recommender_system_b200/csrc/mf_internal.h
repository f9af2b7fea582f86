/*
 * csrc/mf_internal.h -- private state behind the opaque handles of
 * include/matfact_b200.h.  Not part of the ABI.
 */
#pragma once
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>

#include "matfact_b200.h"
#include "../host/mf_prep.h"
#include "mf_types.h"

#define MF_MAX_SHARDS 16

/* sync block of a rank (unsigned long long words; peers write into it over NVLink):
 *   pull:  [0..1] done[b]  [2..3] consumed[b]  [4] CTAs finished   [7] barrier scratch
 *   push:  [16+h] done_from[h]   rank h has stored all its contributions of the iteration here
 *          [32+h] ready_from[h]  rank h has stored its rows of R_new here
 *          [48]   epoch          iterations completed by this rank
 *          [49]   CTAs of k_reduce_bcast finished
 *          [50]   error word: a wait for a peer ran out of time (host: MFGPU_ECOMM) */
#define MF_SYNC_WORDS 64
#define MF_SYNC_DONE 16
#define MF_SYNC_READY 32
#define MF_SYNC_EPOCH 48
#define MF_SYNC_CTAS 49
#define MF_SYNC_ERR 50      /* set by a device-side wait that ran out of time (value: 1 + the rank waited for, or 99) */

/* ---- NCCL, bound at run time (only rank mode needs it) ---------------------- */
typedef struct ncclComm *mf_ncclComm_t;
typedef struct { char internal[128]; } mf_ncclUniqueId;
struct mf_nccl_api {
	void *handle;
	int (*GetUniqueId)(mf_ncclUniqueId *);
	int (*CommInitRank)(mf_ncclComm_t *, int, mf_ncclUniqueId, int);
	int (*CommDestroy)(mf_ncclComm_t);
	int (*AllReduce)(const void *, void *, size_t, int, int, mf_ncclComm_t, cudaStream_t);
	int (*AllGather)(const void *, void *, size_t, int, mf_ncclComm_t, cudaStream_t);
	const char *(*GetErrorString)(int);
};

/* Exchange buffers of a rank, kept by the communicator from one plan to the next: the two G
 * buffers, the two replicas of R, the flag block, and the peer mappings of all of them (CUDA
 * IPC).  cudaMalloc / cudaIpcOpenMemHandle / cudaFree of these cost tens of milliseconds per
 * plan on eight ranks; a plan whose sizes fit borrows them instead (mf_plan.cu).  Identical on
 * every rank by construction: all ranks create the same plans in the same order. */
struct mf_xchg_arena {
	size_t g_bytes, r_bytes;                /* capacity of each G / R buffer          */
	double *G[2], *R[2];
	unsigned long long *sync;
	int mapped;                             /* 0 not tried, 1 peers mapped, -1 IPC unavailable */
	double *peer_G[2][MF_MAX_SHARDS], *peer_R[2][MF_MAX_SHARDS];
	unsigned long long *peer_sync[MF_MAX_SHARDS];
	void *ipc_mapped[5 * MF_MAX_SHARDS];
	int n_ipc_mapped;
	int in_use;                             /* a live plan has borrowed it            */
};

struct mfgpu_comm {
	int nranks, rank, device;
	mf_ncclComm_t nccl;
	mf_xchg_arena arena;
};

void mf_arena_reset(mfgpu_comm *c);        /* mf_plan.cu; collective over the ranks */

struct mf_shard {
	int dev;
	int me;                                  /* index of this shard among the nshards of the problem (rank mode: the rank) */
	unsigned long long *d_sync;              /* this shard's sync block (MF_SYNC_WORDS words), mapped by every peer */
	cudaStream_t st, st2;                    /* st2: the long-row launch runs beside the short rows */
	void *d_ll_args;                         /* mf::LLArgs[2] for k_acc_ll (by factor parity), or null */
	bool ll_fused;                           /* the iteration being enqueued has its exchange inside k_acc_ll */
	mf_pattern pat;
	int user_lo, user_hi;
	/* sparse pattern on the device */
	int *d_row_id, *d_row_ptr, *d_col_idx, *d_epos, *d_col_id, *d_col_ptr, *d_row_idx;
	double *d_val, *d_e_csc, *d_e_csr;
	int *d_mask_ptr, *d_mask_col, *d_first_free;
	int *d_item_slot, *d_touched_item;       /* N-shard only */
	int *d_push_owner, *d_push_row;          /* rank mode, push exchange: item -> reducing rank, row in its receive buffer */
	int *d_contrib;                          /* rank mode: per touched item, bit h = rank h has a nonzero in it */
	double *G_scratch;                       /* profiling launches of the chain kernels write here */
	int r_plain;                             /* 1: R[] from cudaMalloc (exportable over CUDA IPC), 0: from the pool */
	int borrowed;                            /* 1: G[], R[] belong to the communicator's arena */
	int *d_chunk_row, *d_chunk_pos;          /* chunks of the long rows (k_dot) */
	int *d_units;                            /* k_dot_units: 8 ints per unit (mf_loop.cuh) */
	int n_units, n_mid;                      /* units of the paired launch; rows [n_long, n_mid) stay with k_update_L */
	size_t pair_smem;
	int pair_warps;                          /* warps per CTA of k_dot_units (4, 5, 6 or 8) */
	int pair_nt;                             /* tiles in the ring of a k_dot_pair warp (2..6) */
	int split;                               /* 1: dots (k_dot_pair, k_dot) and accumulations (k_acc) are separate launches */
	unsigned long long *d_ticket;            /* k_acc: unit counter, zeroed before every launch */
	/* factors, double buffered; L holds the shard's users only */
	double *L[2], *R[2];
	double *G[2];                            /* N-shard only: partial R gradient */
	double *L_keep, *R_keep;                 /* mfgpu_plan_save_factors           */
	mf::LoopArgs la;
	int warps;                               /* 8 or 1: k_update_L template      */
	int acc_rows;                            /* 1 or 3: rows in flight ahead in its accumulation phase */
	int k1_blocks, dot_blocks;
	int res_ng;                              /* > 0: k_update_L_res<res_ng> does the short rows */
	size_t res_smem;
	void *host_ws;                           /* page-locked workspace of `pat`, from the arena cache */
	size_t host_ws_bytes;
	long long kbytes[5];                     /* algorithmic bytes per launch of each kernel */
	long long gbytes[5];                     /* bytes each launch requests from L2 (a gathered row per nonzero and pass) */
	size_t k1_smem;
	cudaGraphExec_t gexec[2];                /* by starting parity                */
	cudaEvent_t ev0, ev1, ev_g[2], ev_t0, ev_t1, ev_fork, ev_join, ev_a, ev_b;
	/* arg-max scratch */
	double *part_v;
	int *part_j, *nan_first, *d_best;
	int am_splits;
	/* item-sharded arg-max (one process, N GPUs, users << items): this GPU's item block against ALL users */
	double *am_L;                            /* users x ld: the L rows of every shard */
	int *am_mask_ptr, *am_mask_col, *am_first_free;   /* rated-item mask of all users */
	double *am_part_v, *am_out_v;
	int *am_part_j, *am_nan, *am_out_j;
	int am2_splits, am_tile0, am_tile1;
};

struct mfgpu_plan {
	int users, items, F, ld, nnz;
	double alpha;
	int nshards;
	int cut[MF_MAX_SHARDS + 1];
	mf_shard sh[MF_MAX_SHARDS];
	mfgpu_comm *comm;            /* rank mode, else NULL   */
	const mf_scan *scan;         /* during plan creation: what the pass over the entries found */
	int n_touched;               /* items with >=1 nonzero */
	int cur;                     /* buffer holding the current factors */
	int have_factors;
	int graph_len;               /* iterations per captured graph once one exists (0: none yet) */
	long long launches;
	float loop_ms, rec_ms;
	/* Rank mode over peer memory (mf_xchg_setup): every rank maps the G and R buffers and the
	 * flag block of every other rank (CUDA IPC).  xchg_ipc:
	 *   0  ncclAllReduce + k_apply_R
	 *   1  pull: one kernel reads every rank's G over NVLink, adds in rank order, applies
	 *   2  push: the chain kernels store each finished gradient row into the receive buffer of
	 *      the rank that reduces that item; that rank adds the N contributions in rank order,
	 *      applies, and stores its rows of R_new into every replica (k_reduce_bcast); fences and
	 *      flag words order the two phases
	 *   3  low latency (default): as push, but every element travels as a (value, tag) pair in
	 *      one 16-byte store and the readers poll for the tag of the iteration: no fences, no
	 *      flag words; one launch (k_xchg_ll) reduces the own block and unpacks the others' */
	int xchg_ipc;
	int peer_events;                               /* one process, N GPUs: 1 = the event-ordered pull exchange (k_apply_R_peer) instead of the low-latency one */
	long ll_half;                                  /* doubles per parity half of a (value, tag) pair buffer (G[0], G[1]) */
	unsigned long long comm_timeout_ns;            /* device-side waits for a peer give up after this long */
	mf_xchg_arena *arena;                          /* borrowed exchange buffers, or NULL (own) */
	int slots_per_rank;                            /* push: touched items reduced by one rank (the last may own fewer) */
	double *peer_G[2][MF_MAX_SHARDS];              /* G[b] of rank h (own rank: the local buffer) */
	double *peer_R[2][MF_MAX_SHARDS];              /* R[b] of rank h */
	unsigned long long *peer_sync[MF_MAX_SHARDS];  /* sync block of rank h, MF_SYNC_WORDS words   */
	unsigned long long *d_sync;                    /* own sync block                              */
	void *ipc_mapped[5 * MF_MAX_SHARDS];
	int n_ipc_mapped;
	/* options */
	int opt_graph, opt_graph_iters, opt_long_row, opt_long_col, opt_ch, opt_chain_ring, opt_resident, opt_res_rows, opt_chain_ws, opt_exchange, opt_dot_stages, opt_acc_rows, opt_cs_store, opt_pair, opt_pair_stages, opt_pair_warps, opt_acc_gw, opt_comm_timeout_ms, opt_argmax_items, opt_xchg_fused;
};

/* ---- shared helpers (defined in mf_capi.cu) ----------------------------------- */
int mf_fail(int code, const char *fmt, ...);
int mf_load_nccl(void);
int mf_usable_devices(int *ids, int cap);
int mf_configure_shard(mfgpu_plan *pl, mf_shard *s);      /* mf_plan.cu */
int mf_xchg_setup(mfgpu_plan *pl);                          /* mf_plan.cu: collective over the ranks */
void mf_xchg_teardown(mfgpu_plan *pl);
int mf_check_comm_error(mfgpu_plan *pl);                    /* mf_iter.cu: after a sync; MFGPU_ECOMM when a device-side wait timed out */
/* MFGPU_TRACE=1: wall-clock of the host-side phases (plan creation / destruction, graph
 * builds) on stderr */
#include <chrono>
#include <cstdio>
#include <cstdlib>
static inline bool mf_trace_on()
{
	static const bool on = getenv("MFGPU_TRACE") != nullptr;
	return on;
}
struct TraceClock {
	std::chrono::steady_clock::time_point t = std::chrono::steady_clock::now();
	void lap(const char *what)
	{
		if (!mf_trace_on())
			return;
		const auto n = std::chrono::steady_clock::now();
		fprintf(stderr, "[mfgpu] %-28s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(n - t).count());
		t = n;
	}
};

int mf_kernels_prepare(int warps, size_t k1_smem, int res_ng, size_t res_smem, size_t pair_smem, int pair_warps, int pair_nt);   /* mf_iter.cu: kernels live in that TU only */
int mf_kernels_preload(int F, int split, int nshards);      /* mf_iter.cu */
extern mf_nccl_api g_nccl;

#define CK(call)                                                                        \
	do {                                                                                \
		cudaError_t e_ = (call);                                                        \
		if (e_ != cudaSuccess)                                                          \
			return mf_fail(e_ == cudaErrorMemoryAllocation ? MFGPU_ENOMEM : MFGPU_ECUDA, \
			               "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_),      \
			               __FILE__, __LINE__);                                         \
	} while (0)

#define TRY(call)                 \
	do {                          \
		int rc_ = (call);         \
		if (rc_ != MFGPU_OK)      \
			return rc_;           \
	} while (0)

#define NCK(call)                                                                          \
	do {                                                                                   \
		int r_ = (call);                                                                   \
		if (r_ != 0)                                                                       \
			return mf_fail(MFGPU_ECOMM, "%s failed: %s", #call, g_nccl.GetErrorString(r_)); \
	} while (0)
