/*
 * oracle/mpi_shim/shim_layout.h -- the shared memory file of one mpirun_shim job
 * (TEST/BENCH INFRASTRUCTURE ONLY, see mpi.h).
 *
 *   [ header | comm table | np x np message rings | staging arenas (bump-allocated) ]
 *
 * The file is sparse (memfd + ftruncate): only the pages a job touches are backed.
 */
#ifndef MPI_SHIM_LAYOUT_H
#define MPI_SHIM_LAYOUT_H

#include <stdatomic.h>
#include <stdint.h>

#define SHIM_MAGIC 0x4d50495348494d31ull /* "MPISHIM1" */
#define SHIM_MAX_RANKS 256
#define SHIM_MAX_COMMS 4096
#define SHIM_RING_BYTES (256u << 10)  /* bytes in flight from one rank to another */
#define SHIM_CHUNK (1u << 20)         /* staging bytes per rank and direction of a communicator */
#define SHIM_ARENAS_PER_RANK 64       /* room for this many communicators of full size */

typedef struct shim_ring {
	_Alignas(64) _Atomic uint64_t head; /* bytes written, producer only */
	_Alignas(64) _Atomic uint64_t tail; /* bytes read, consumer only */
	_Alignas(64) unsigned char data[SHIM_RING_BYTES];
} shim_ring;

typedef struct shim_comm_shared {
	_Alignas(64) _Atomic int arrived;
	_Alignas(64) _Atomic int generation;
	uint64_t arena; /* offset of 2 x size slots of SHIM_CHUNK bytes: inputs, then outputs */
	int size;
} shim_comm_shared;

typedef struct shim_header {
	uint64_t magic;
	uint64_t total_bytes;
	uint64_t rings; /* offset of ring (src, dst) = rings + (src * np + dst) * sizeof(shim_ring) */
	int np;
	_Alignas(64) _Atomic uint64_t bump; /* next free arena offset */
	_Atomic int next_comm;
	_Atomic int aborted;
	shim_comm_shared comm[SHIM_MAX_COMMS];
} shim_header;

static inline uint64_t shim_total_bytes(int np)
{
	uint64_t n = sizeof(shim_header);
	n = (n + 4095) & ~4095ull;
	n += (uint64_t)np * np * sizeof(shim_ring);
	n = (n + 4095) & ~4095ull;
	n += (uint64_t)SHIM_ARENAS_PER_RANK * np * 2 * SHIM_CHUNK;
	return n;
}

/* called by whoever creates the file, before any rank starts */
static inline void shim_format(shim_header *h, int np)
{
	h->magic = SHIM_MAGIC;
	h->np = np;
	h->total_bytes = shim_total_bytes(np);
	h->rings = (sizeof(shim_header) + 4095) & ~4095ull;
	uint64_t arenas = (h->rings + (uint64_t)np * np * sizeof(shim_ring) + 4095) & ~4095ull;
	/* communicator 0 is MPI_COMM_WORLD */
	h->comm[0].arena = arenas;
	h->comm[0].size = np;
	atomic_store(&h->bump, arenas + (uint64_t)np * 2 * SHIM_CHUNK);
	atomic_store(&h->next_comm, 1);
	atomic_store(&h->aborted, 0);
}

#endif
