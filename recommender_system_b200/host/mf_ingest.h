/*
 * host/mf_ingest.h -- host-side ingest of `matFact <file.in>` (plain C, no GPU):
 * the reference's initialisation sequence and its input parser, restated so that
 * they are no longer what a run waits for.
 *
 *   mf_rand_*            glibc's random() (TYPE_3 additive feedback generator) as a
 *                        lock-free block generator, draw for draw what
 *                        srandom(seed); random(); random(); ... returns
 *   mf_random_fill_LR    mat2d_random_fill_LR + mat2d_transpose (mat2d.c:61-72,115-124,
 *                        called matFact.c:113-120): same draws in the same order, R written
 *                        directly in its transposed (items x features) layout, a block of
 *                        features at a time so that every store fills whole cache lines
 *   mf_parse_instance    the fscanf sequence of matFact.c:72-110 / util.c:12-34 over a buffer
 *                        read in one piece, with the acceptance of fscanf's %d and %lf
 *                        (longest-prefix rule included) and the same error classes
 */
#ifndef MF_INGEST_H
#define MF_INGEST_H

#include <stddef.h>
#include <stdint.h>

#include "matfact_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
	uint32_t h[128];    /* the last 128 values of the generator's sequence x: h[m & 127] = x[m] */
	uint64_t n;         /* index of the next value                                              */
} mf_rand;

void mf_rand_seed(mf_rand *g, unsigned seed);                 /* srandom(seed) */
uint32_t mf_rand_next(mf_rand *g);                            /* random()      */
void mf_rand_block(mf_rand *g, uint32_t *out, size_t n);      /* n x random()  */

/* L: users x features, R: items x features (the reference's transposed R), both row-major */
void mf_random_fill_LR(int users, int items, int features, double *L, double *R);

/* Result classes of mf_parse_instance: the reference's die() messages, in its order of tests */
enum {
	MF_PARSE_OK = 0,
	MF_PARSE_OPEN,          /* "Unable to open input file."        matFact.c:69-70  */
	MF_PARSE_INT,           /* "Error in int argument."            util.c:12-15     */
	MF_PARSE_DOUBLE,        /* "Error in double argument."         util.c:17-20     */
	MF_PARSE_MULTI_INT,     /* "Error in multiple int argument."   util.c:22-27     */
	MF_PARSE_ENTRY,         /* "Error in non-zero entry."          util.c:29-34     */
	MF_PARSE_CLOSE,         /* "Unable to close input file."       matFact.c:108-109 */
	MF_PARSE_NOMEM          /* out of memory (the reference does not check)          */
};

typedef struct {
	int iters, features, users, items, non_zero;
	double alpha;
	mfgpu_entry *entries;   /* malloc'ed, non_zero of them (at least one slot); caller frees */
} mf_instance;

int mf_parse_instance(const char *path, mf_instance *out);
/* the same over a buffer (text[len] need not be NUL) -- what the tests feed */
int mf_parse_buffer(const char *text, size_t len, mf_instance *out);
const char *mf_parse_message(int rc);

#ifdef __cplusplus
}
#endif
#endif
