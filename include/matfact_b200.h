/*
 * matfact_b200.h -- C ABI of the B200 factorization path (libmatfact_b200.so).
 *
 * The reference (vladstojna/recommender-system) has no plugin or FFI layer; the
 * seam this library replaces is two consecutive calls in its main():
 *
 *     matrix_factorization(B, L, R, entries, non_zero, iters, alpha);   matFact.c:124 (def. :29-59)
 *     print_output(B, entries);                                         matFact.c:127 (def. :10-27)
 *
 * together with the three mat2d routines those call (mat2d_copy mat2d.c:34-36,
 * mat2d_dot_product :126-139, mat2d_prod :100-113).  Everything here is plain
 * C: pointers, ints, doubles.  All matrices are row-major doubles exactly as
 * the reference's mat2d (mat2d.h:6-11): L is users x features, R is
 * items x features (the reference keeps R transposed, matFact.c:117-120).
 * B (users x items) is never materialised.
 *
 * Every function returns 0 on success or a negative MFGPU_E* code;
 * mfgpu_last_error() gives the message (thread local).  There is no CPU
 * fallback: without a usable sm_100 device every compute entry fails with
 * MFGPU_ENODEV.  The reference reports errors by die() -> exit(-1)
 * (util.c:7-10); the host program (host/matFact.c) maps non-zero returns to the
 * same behaviour.
 */
#ifndef MATFACT_B200_H
#define MATFACT_B200_H

#ifdef __cplusplus
extern "C" {
#endif

#define MFGPU_OK        0
#define MFGPU_EINVAL   -1   /* bad argument (negative size, entry out of range, ...) */
#define MFGPU_ENOMEM   -2   /* host or device allocation failed                      */
#define MFGPU_ENODEV   -3   /* no CUDA device / not an sm_100 device                 */
#define MFGPU_ECUDA    -4   /* a CUDA runtime call failed                            */
#define MFGPU_ECOMM    -5   /* a collective (NCCL) call failed, or a peer did not deliver in time */

/* Same layout as the reference's non_zero_entry (datatypes.h:10-15;
 * deliverables/serial/matFact.c:148-153): 16 bytes, no padding. */
typedef struct {
	int row;      /* user, 0-based  */
	int col;      /* item, 0-based  */
	double value; /* rating a_ij    */
} mfgpu_entry;

typedef struct mfgpu_plan mfgpu_plan;   /* device-resident problem           */
typedef struct mfgpu_comm mfgpu_comm;   /* one rank of a multi-process group */

/* ---- library state --------------------------------------------------------- */

/* Number of usable devices (compute capability 10.x); 0 if none. */
int mfgpu_device_count(void);
/* Message of the last failure on the calling thread ("" if none). */
const char *mfgpu_last_error(void);
/* Library version string. */
const char *mfgpu_version(void);

/* ---- one-shot entries with HOST buffers (what the reference's main() calls) -- */

/* Replaces matrix_factorization (matFact.c:29-53, without the trailing
 * mat2d_prod): runs `iters` iterations over `entries` (file order; the
 * accumulation order of the reference is preserved for any order) and
 * overwrites L and R in place.  ngpus <= 0 means "all visible devices";
 * ngpus == 1 reproduces the reference's doubles bit for bit. */
int mfgpu_factorize(int users, int items, int features, int nnz,
                    const mfgpu_entry *entries, int iters, double alpha,
                    double *L, double *R, int ngpus);

/* Replaces mat2d_prod + print_output (matFact.c:55,127; mat2d.c:100-113;
 * matFact.c:10-27): best[i] = the item the reference would print for user i,
 * or -1 where it prints nothing (every item rated). */
int mfgpu_recommend(int users, int items, int features, int nnz,
                    const mfgpu_entry *entries,
                    const double *L, const double *R, int *best, int ngpus);

/* Both of the above with L and R staying on the device in between.
 * L, R are in/out; best is out. */
int mfgpu_run(int users, int items, int features, int nnz,
              const mfgpu_entry *entries, int iters, double alpha,
              double *L, double *R, int *best, int ngpus);

/* ---- allocation-free masked arg-max on DEVICE buffers ----------------------- */

/* No counterpart in the reference: added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  The masked arg-max of mfgpu_recommend / mfgpu_plan_recommend on the CURRENT
 * device, enqueued on `stream` (a cudaStream_t; NULL = the default stream), with every pointer a
 * DEVICE pointer supplied by the caller:
 *   L (users x ld), R (items x ld) row-major doubles, rows 16-byte aligned (ld even, >= features);
 *   mask_ptr[users+1], mask_col[mask_ptr[users]]: per user the ascending items print_output's
 *   cursor skips (matFact.c:14-22) -- for a (row, col)-sorted duplicate-free entry list the
 *   row pointers and columns of the entries themselves;
 *   first_free[users]: first item not in the user's list, -1 if none (mfgpu_mask_stats gives it);
 *   best[users]: out, as mfgpu_recommend.
 * The call allocates nothing: the kernels' scratch lives in one built-in device array of
 * MFGPU_NOALLOC_SCRATCH_BYTES, held by every GPU that loads the library until the process
 * exits.  Limit: users * 16 + 64 <= MFGPU_NOALLOC_SCRATCH_BYTES (262 140 users); larger
 * inputs return MFGPU_EINVAL.  All calls share the array: calls that overlap on one GPU (two
 * streams, two threads) race; calls on one stream are ordered and safe.  Returns once the
 * work is enqueued; synchronise `stream` before reading best. */
#define MFGPU_NOALLOC_SCRATCH_BYTES (4u << 20)
int mfgpu_recommend_noalloc(int users, int items, int features, int ld,
                            const double *L, const double *R, const int *mask_ptr,
                            const int *mask_col, const int *first_free, int *best,
                            void *stream);

/* ---- plan API: keep the problem resident in HBM across calls ---------------- */

/* Builds the by-user / by-item groupings of `entries`, uploads them, and
 * allocates the factor buffers on `ngpus` devices of this process (users are
 * cut into contiguous, nonzero-balanced ranges, one per device). */
int mfgpu_plan_create(mfgpu_plan **out, int users, int items, int features, int nnz,
                      const mfgpu_entry *entries, double alpha, int ngpus);
/* Same for ONE rank of a multi-process group (one process per GPU): this
 * process owns the users of rank comm->rank and a replica of R. */
int mfgpu_plan_create_rank(mfgpu_plan **out, int users, int items, int features, int nnz,
                           const mfgpu_entry *entries, double alpha, mfgpu_comm *comm);
void mfgpu_plan_destroy(mfgpu_plan *plan);

/* Host -> device copy of the full L (users x features) and R (items x features). */
int mfgpu_plan_set_factors(mfgpu_plan *plan, const double *L, const double *R);
/* The reference's initial factors (mat2d_random_fill_LR + mat2d_transpose, mat2d.c:61-72,115-124,
 * called matFact.c:113-120) generated straight into the device buffers: the same values, bit
 * for bit, as mfgpu_random_fill_LR followed by mfgpu_plan_set_factors, without the host copies
 * of L and R (the random() stream is produced on the calling thread and crosses as 4-byte
 * draws; the divisions run on the device, overlapped with the generation of the next block).
 * In rank mode every rank calls it; each keeps its own rows of L and all of R. */
int mfgpu_plan_init_factors(mfgpu_plan *plan);
/* Device -> host.  In rank mode only the rows of L owned by this rank are
 * written (the others are left untouched); R is complete on every rank. */
int mfgpu_plan_get_factors(mfgpu_plan *plan, double *L, double *R);
/* mfgpu_plan_get_factors and mfgpu_plan_recommend in one call: the copies of the factors run on
 * the copy engine beside the arg-max kernels.  L and R may be NULL (not copied). */
int mfgpu_plan_get_results(mfgpu_plan *plan, double *L, double *R, int *best);
/* Enqueue `iters` iterations; returns once they are enqueued. */
int mfgpu_plan_iterate(mfgpu_plan *plan, int iters);
/* Wait for everything enqueued on the plan. */
int mfgpu_plan_sync(mfgpu_plan *plan);
/* Device time (CUDA events on the plan's stream) of the last mfgpu_plan_iterate
 * call, max over the plan's devices.  Implies a sync. */
int mfgpu_plan_last_loop_ms(mfgpu_plan *plan, float *ms);
/* Masked arg-max on the resident factors; best has `users` ints.  In rank mode
 * only the entries of the rank's users are written. */
int mfgpu_plan_recommend(mfgpu_plan *plan, int *best);
/* Device time of the last mfgpu_plan_recommend call. */
int mfgpu_plan_last_recommend_ms(mfgpu_plan *plan, float *ms);
/* [lo, hi) of the users owned by local shard `shard` (0 in rank mode). */
int mfgpu_plan_user_range(mfgpu_plan *plan, int shard, int *lo, int *hi);
/* Kernel launches issued by this plan so far (for bench.py's gpu_launches). */
long long mfgpu_plan_launch_count(mfgpu_plan *plan);
/* How the partial R gradients of the shards meet: 0 = one shard, nothing to exchange;
 * 1 = ncclAllReduce + apply (one process per GPU); 2 = one kernel that reads every rank's
 * gradient over NVLink peer memory (CUDA IPC) and applies the sum (one process per GPU);
 * 3 = the event-ordered pull inside one process driving N GPUs (option exchange = 2 there);
 * 4 = chain kernels store the gradient rows into the rank that owns the item, the owner
 * reduces, applies and stores its rows of R_new into every replica, ordered by fences and flag
 * words; 5 = the same flow with (value, tag) pairs polled by the readers, one launch (the
 * default of both N-GPU modes). */
int mfgpu_plan_exchange_kind(mfgpu_plan *plan);
/* Algorithmic bytes of ONE iteration, SURVEY.md section 8(d):
 * 16*F*(U_t + I_t) + 12*nnz, with U_t / I_t the users / items having >=1 nonzero. */
long long mfgpu_plan_bytes_per_iter(mfgpu_plan *plan);
/* Keep a device-side copy of the current factors / put that copy back (both
 * buffers).  Lets a benchmark repeat the same factorization without a host
 * round trip; the copy is allocated on the first save. */
int mfgpu_plan_save_factors(mfgpu_plan *plan);
int mfgpu_plan_restore_factors(mfgpu_plan *plan);
/* Average device time (CUDA events on the plan's stream, shard 0) of `reps`
 * back-to-back launches of each loop kernel on the current factors; the factors
 * are left unchanged.  ms[0] = by-user pass of the short rows (k_update_L),
 * ms[1] = dots of the long rows (k_dot), ms[2] / ms[3] = sequential accumulations
 * (k_chain for the bulk of the chains / k_chain_ws for the longest ones), ms[4] =
 * exchange + apply launched on its own (rank mode, else 0; collective: every rank
 * must call it; for the default exchange the figure includes a re-send of the rank's
 * contributions, which an iteration does inside its chain kernels). */
int mfgpu_plan_profile_kernels(mfgpu_plan *plan, int reps, float ms[5]);
/* Algorithmic (compulsory) bytes of one launch of each of those kernels on shard 0. */
int mfgpu_plan_kernel_bytes(mfgpu_plan *plan, long long bytes[5]);
/* Bytes each of those launches requests from L2: as above, but one gathered row of the other
 * factor for every nonzero and pass instead of once per distinct row (the diagnostic traffic
 * model of the loop: the gathers, not the compulsory HBM bytes, are what bounds it). */
int mfgpu_plan_kernel_gather_bytes(mfgpu_plan *plan, long long bytes[5]);
/* Device stopwatch on the plan's stream(s): start records an event, stop records
 * another, waits for it and returns the elapsed device time (max over the
 * plan's devices).  Everything enqueued on the plan in between is covered. */
int mfgpu_plan_timer_start(mfgpu_plan *plan);
int mfgpu_plan_timer_stop(mfgpu_plan *plan, float *ms);
/* Launch-configuration knobs (tests and experiments; none of them changes a bit
 * of the result).  value 0 restores the default where one is derived from the input.
 *   "graph"        1/0   iterations replayed from CUDA graphs (default 1)
 *   "graph_iters"  n     iterations captured per graph (default 16)
 *   "long_row"     n     rows longer than n nonzeros have their dots done by k_dot
 *   "long_col"     n     chains longer than n nonzeros go to the long-chain kernel
 *   "chunk"        n     nonzeros per chunk of the tiled dot phase (1..32)
 *   "resident"     0/1/2 short rows with their gathered R rows resident in shared
 *                        memory: never / when eight warps fit an SM (default) / always
 *   "res_rows"     n     resident rows per chunk (1..32)
 *   "chain_ws"     1/0   warp-specialised kernel for the longest chains (default 1)
 *   "chain_ring"   1/0   k_chain gathers through a shared-memory ring / registers
 *   "dot_stages"   2/4   tiles in the dot-phase ring of a k_update_L / k_dot warp (default 2)
 *   "acc_rows"     1/3   rows in flight ahead of the addition in k_update_L (default 1)
 *   "pair"         1/0/-1 split mode: the dots by k_dot_units (two short rows to a warp), the
 *                        accumulations of both sides by k_acc / the fused kernels / by shape
 *                        (default: F >= 128, or F >= 64 on a shard below 6e5 nonzeros, and
 *                        rows of at least 8 nonzeros on average)
 *   "pair_warps"   4/5/6/8 warps per CTA of k_dot_units (default 4)
 *   "pair_stages"  2..6  tiles in the ring of a k_dot_units warp (default 2)
 *   "acc_gw"       1..8  feature groups per unit of k_acc (default: all, at most 8)
 *   "cs_store"     1/0   rows of L_new stored evict-first by the fused kernels (default 0)
 *   "argmax_items" 1/0/-1 one process, N GPUs: the arg-max sharded by items with a
 *                        (value, index) merge / by users / by shape (default: items >= 8 users)
 *   "xchg_fused"   1/0   N shards, split mode, (value, tag) exchange: the exchange inside the
 *                        accumulation kernel (k_acc_ll: columns | reduce | rows | unpack)
 *                        instead of a launch of its own (default 0, or MFGPU_XCHG_FUSED: measured
 *                        slower; same bits either way, a rank may choose for itself)
 *   "comm_timeout_ms" n  device-side waits for a peer give up after n ms (default 30000, or
 *                        MFGPU_COMM_TIMEOUT_MS); the next sync then returns MFGPU_ECOMM
 *   "exchange"     N shards.  Rank mode (collective: every rank sets it at the same point):
 *                        0 ncclAllReduce + apply; otherwise over peer memory (CUDA IPC) when
 *                        every rank can map the others: 1 or -1 (default) (value, tag) pairs
 *                        polled by the readers, one launch; 2 pull (one kernel reads every
 *                        rank's gradient); 3 push with fences and flag words.  One process,
 *                        N GPUs: 2 = the event-ordered pull, anything else the (value, tag)
 *                        exchange.  All peer-memory transports add in shard order
 *                        (bit-identical results).
 * May be set between iterate calls (the plan is reconfigured); unknown names return
 * MFGPU_EINVAL. */
int mfgpu_plan_set_option(mfgpu_plan *plan, const char *name, int value);

/* ---- multi-process groups (one process per GPU, e.g. under torchrun) -------- */

#define MFGPU_UID_BYTES 128
/* Rank 0 creates the id and ships it to the other ranks by any side channel
 * (bench.py broadcasts it with torch.distributed). */
int mfgpu_comm_unique_id(unsigned char uid[MFGPU_UID_BYTES]);
int mfgpu_comm_create(mfgpu_comm **out, int nranks, int rank,
                      const unsigned char uid[MFGPU_UID_BYTES], int device);
void mfgpu_comm_destroy(mfgpu_comm *comm);

/* ---- host-side helpers exported for tests and tools (no GPU needed) --------- */

/* Contiguous user ranges with balanced nonzero counts; cut has nshards+1 ints. */
int mfgpu_split_users(int users, int nnz, const mfgpu_entry *entries, int nshards, int *cut);
/* Layout of the rank-mode exchange for `nranks` ranks (mfgpu_split_users cut): item_slot[j] =
 * position of item j among the items with a nonzero (-1: none), contrib[slot] = bit r set when
 * rank r holds a nonzero of that item (contrib needs room for `items` ints).  Slot t is reduced
 * by rank t / slots_per_rank.  Returns the number of touched items or < 0. */
int mfgpu_exchange_layout(int users, int items, int nnz, const mfgpu_entry *entries, int nranks,
                          int *item_slot, int *contrib, int *slots_per_rank);
/* The reference's initialisation, kept on the host: srandom(0), L then R_init
 * drawn as (double)random()/RAND_MAX/features, R stored transposed
 * (mat2d.c:61-72,115-124; matFact.c:113-120).  The draws come from a lock-free
 * restatement of glibc's generator (host/mf_ingest.c), identical draw for draw, and R is
 * written straight into its transposed layout a cache line of features at a time. */
void mfgpu_random_fill_LR(int users, int items, int features, double *L, double *R);
/* The input parser of the reference's main (matFact.c:72-110: fscanf "%d", "%lf", "%d",
 * "%d %d %d", then non_zero x "%d %d %lf"; util.c:12-34) over the file read in one piece,
 * with the acceptance of glibc's fscanf for every token (host/mf_ingest.c, pinned by a
 * differential test against fscanf).  Returns 0 and a malloc'ed entry list (release it with
 * mfgpu_free), or the class of the first error: mfgpu_read_error(rc) is the reference's
 * die() message for it ("Unable to open input file.", "Error in int argument.", ...). */
int mfgpu_read_instance(const char *path, int *iters, double *alpha, int *features,
                        int *users, int *items, int *non_zero, mfgpu_entry **entries);
const char *mfgpu_read_error(int rc);
void mfgpu_free(void *p);
/* n draws of the library's restatement of glibc's random() after srandom(seed) compared with
 * the C library's own: index of the first difference, -1 if none (self-test of the
 * initialisation sequence on the machine at hand; uses the process-wide srandom state). */
long mfgpu_selftest_random(unsigned seed, long n);
/* Rated-item mask of print_output for user range [0, users): fills
 * mask_count[i] and first_free[i]; returns total matched entries or <0. */
int mfgpu_mask_stats(int users, int items, int nnz, const mfgpu_entry *entries,
                     int *mask_count, int *first_free);

#ifdef __cplusplus
}
#endif
#endif
