/*
 * ccmpi_host -- CPU shared-memory transport (libccmpi_host.so).
 *
 * Stands in for "mpi4py + a system MPI" (reference requirements.txt:4, README.md:11-24) when
 * the arrays live in host memory or the machine has no GPU: one POSIX shared-memory segment
 * per job, a staging area per rank for collectives (copy in -> barrier -> combine -> barrier,
 * the way MPI's own shared-memory collectives work), and a FIFO ring per ordered rank pair for
 * Send / Recv / Isend / Irecv / Sendrecv (mpi_wrapper/comm.py:86-193).
 *
 * Communicators are described by the caller: `members` lists the world ranks of the group in
 * group-rank order, `me` is the caller's index in it, `slot` is the group's barrier slot
 * (any id < CCMPI_HOST_MAX_COMMS that all members agree on and no overlapping group shares).
 * All calls return 0 or a negative error; ccmpi_host_last_error() explains.
 */
#ifndef CCMPI_HOST_H_
#define CCMPI_HOST_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CCMPI_HOST_MAX_RANKS 64
#define CCMPI_HOST_MAX_COMMS 256

const char* ccmpi_host_last_error(void);

/* Attach to (creating if necessary) the job's segment. Returns a handle > 0. */
int64_t ccmpi_host_attach(const char* job, int rank, int nranks, size_t stage_bytes, size_t ring_bytes,
                          double timeout_s);
int ccmpi_host_detach(int64_t h, int unlink_segment);

int ccmpi_host_barrier(int64_t h, int slot, const int* members, int n, int me);

/* dtype / op use the ccmpi_dtype / ccmpi_op numbering of ccmpi_b200.h */
int ccmpi_host_allreduce(int64_t h, int slot, const int* members, int n, int me, const void* src, void* dst,
                         size_t count, int dtype, int op);
int ccmpi_host_allgather(int64_t h, int slot, const int* members, int n, int me, const void* src, void* dst,
                         size_t bytes_per_rank);
int ccmpi_host_reduce_scatter(int64_t h, int slot, const int* members, int n, int me, const void* src, void* dst,
                              size_t count_per_rank, int dtype, int op);
int ccmpi_host_alltoall(int64_t h, int slot, const int* members, int n, int me, const void* src, void* dst,
                        size_t seg_bytes);

/* point-to-point, world ranks; (comm_id, tag) travel in the message header and are checked */
int ccmpi_host_send(int64_t h, int dst, int comm_id, int tag, const void* buf, size_t bytes);
int ccmpi_host_recv(int64_t h, int src, int comm_id, int tag, void* buf, size_t bytes);
int64_t ccmpi_host_isend(int64_t h, int dst, int comm_id, int tag, const void* buf, size_t bytes);
int64_t ccmpi_host_irecv(int64_t h, int src, int comm_id, int tag, void* buf, size_t bytes);
int ccmpi_host_test(int64_t h, int64_t req);              /* 1 done, 0 pending, <0 error */
int ccmpi_host_waitall(int64_t h, const int64_t* reqs, int n);
/* size of the next message queued from `src` (blocks until a header is there) */
int64_t ccmpi_host_probe(int64_t h, int src);
/* non-blocking: 1 and the header fields of the next message queued from `src`, 0 when none is there
 * (or a posted receive is ahead of it), <0 on error */
int ccmpi_host_iprobe(int64_t h, int src, int* comm_id, int* tag, int64_t* bytes);

/* Reserve `count` consecutive communicator ids from the job-wide counter in the segment; returns the
 * first one (>= 1; 0 is the world) or <0. One member of the parent calls this and tells the others. */
int ccmpi_host_reserve_comm_ids(int64_t h, int count);

/* Out-of-band completion counters, one per ordered pair of ranks (device point-to-point path: a
 * descriptor travels through the ring, the payload stays in the sender's symmetric heap until the
 * receiver has copied it over NVLink). ack_post: "I am done with one more message from `src`";
 * ack_count: how many of this rank's messages to `dst` the receiver is done with. */
int ccmpi_host_ack_post(int64_t h, int src);
int64_t ccmpi_host_ack_count(int64_t h, int dst);

#ifdef __cplusplus
}
#endif
#endif /* CCMPI_HOST_H_ */
