/*
 * oracle/mpi_shim/mpirun_shim.c -- TEST/BENCH INFRASTRUCTURE ONLY (see mpi.h).
 *
 *   mpirun_shim -np N program [arguments...]
 *
 * Creates the job's shared memory file (anonymous, sparse), starts N copies of `program` with
 * MPISHIM_FD / MPISHIM_RANK / MPISHIM_NP in their environment and waits for them.  If a rank
 * exits with an error or dies, the rest of the job is told (abort flag) and terminated, and
 * the launcher returns that rank's status -- what a job relies on when one rank calls the
 * reference's die() (util.c:7-10).  Ranks also die with the launcher (PDEATHSIG), so a caller's
 * timeout that kills the launcher leaves nothing spinning.
 */
#define _GNU_SOURCE
#include "shim_layout.h"

#include <signal.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/prctl.h>
#include <sys/wait.h>
#include <unistd.h>

static int usage(void)
{
	fprintf(stderr, "usage: mpirun_shim -np N program [arguments...]\n");
	return 2;
}

int main(int argc, char **argv)
{
	int np = 0, at = 1;
	while (at < argc && argv[at][0] == '-') {
		if ((!strcmp(argv[at], "-np") || !strcmp(argv[at], "-n")) && at + 1 < argc) {
			np = atoi(argv[at + 1]);
			at += 2;
		} else if (!strcmp(argv[at], "--")) {
			at++;
			break;
		} else
			return usage();
	}
	if (np < 1 || np > SHIM_MAX_RANKS || at >= argc)
		return usage();

	const uint64_t bytes = shim_total_bytes(np);
	const int fd = memfd_create("mpi_shim", 0); /* no MFD_CLOEXEC: the ranks inherit it */
	if (fd < 0 || ftruncate(fd, (off_t)bytes) != 0) {
		perror("mpirun_shim: shared memory file");
		return 1;
	}
	shim_header *h = mmap(NULL, bytes, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_NORESERVE, fd, 0);
	if (h == MAP_FAILED) {
		perror("mpirun_shim: mmap");
		return 1;
	}
	shim_format(h, np);

	pid_t pid[SHIM_MAX_RANKS];
	char fd_s[16], np_s[16], rank_s[16];
	snprintf(fd_s, sizeof fd_s, "%d", fd);
	snprintf(np_s, sizeof np_s, "%d", np);
	fflush(NULL);
	for (int r = 0; r < np; r++) {
		pid[r] = fork();
		if (pid[r] < 0) {
			perror("mpirun_shim: fork");
			for (int q = 0; q < r; q++)
				kill(pid[q], SIGTERM);
			return 1;
		}
		if (pid[r] == 0) {
			prctl(PR_SET_PDEATHSIG, SIGKILL);   /* a killed launcher takes its ranks with it */
			snprintf(rank_s, sizeof rank_s, "%d", r);
			setenv("MPISHIM_FD", fd_s, 1);
			setenv("MPISHIM_NP", np_s, 1);
			setenv("MPISHIM_RANK", rank_s, 1);
			execvp(argv[at], argv + at);
			perror("mpirun_shim: exec");
			_exit(127);
		}
	}

	int left = np, result = 0;
	while (left > 0) {
		int status;
		const pid_t done = wait(&status);
		if (done < 0)
			break;
		int r = -1;
		for (int q = 0; q < np; q++)
			if (pid[q] == done)
				r = q;
		if (r < 0)
			continue;
		pid[r] = 0;
		left--;
		const int bad = WIFSIGNALED(status) ? 128 + WTERMSIG(status) : WEXITSTATUS(status);
		if (bad && !result) {
			/* first failure: take the job down */
			result = bad;
			atomic_store(&h->aborted, 1);
			for (int q = 0; q < np; q++)
				if (pid[q] > 0)
					kill(pid[q], SIGTERM);
		}
	}
	return result;
}
