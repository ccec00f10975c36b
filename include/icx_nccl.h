/*
 * icx_nccl.h -- the exchange steps of the multi-GPU paths of intercession-b200 as C entry points taking the caller's
 * ncclComm_t, so that a C++ server can run BASELINE.json configs 4 and 5 without Python or torch.
 *
 * NCCL is bound at run time (libnccl.so.2 via dlopen): libicx.so does not link it.  One process per GPU (ncclCommInitRank)
 * or several devices in one process (ncclCommInitAll + ncclGroupStart / End) both work.
 */
#ifndef ICX_NCCL_H
#define ICX_NCCL_H

#include "icx.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ncclComm* ncclComm_t;        /* as in nccl.h */

/* 1 if an NCCL library could be bound in this process. */
ICX_API int icx_nccl_available(void);

/*
 * Config 4: world-line candidates are sharded over the ranks with no data-path collective; this is the ONE
 * gather at the end.  Replaces nothing in the reference (it has a single parallel cosmos, README.md:162 "(TBD)"); generalises
 * ParallelNetworkDynamo's congruence verdict (parallel_network_dynamo.cpp:513-581) to a batch.
 * Takes the results of the scans enqueued on `tl` (icx_divergence_scan_enqueue) for this rank's n_local world-lines, all-gathers
 * n_per_rank (>= n_local, same on every rank) entries of 16 bytes per rank on the scan's stream, and returns, rank-major
 * ([rank * n_per_rank + k]): "flagged by some scan since the last reset", the last scan's incongruent count and score.
 * Synchronises the stream.
 */
ICX_API int icx_scores_allgather(IcxTimeline* tl, uint32_t n_local, uint32_t n_per_rank, int n_ranks, ncclComm_t comm,
                                 uint8_t* flags_ever, uint32_t* incongruent, double* score);

/*
 * Config 5: TimelineApi::push_past_timestream between timeslices on different GPUs
 * (server_network_dynamo.cpp:642-679).  send = icx_timeline_slot_export + ncclSend, recv = ncclRecv + icx_timeline_slot_import,
 * all enqueued on `stream`; device_scratch is a caller-owned device buffer of scratch_bytes = ICX_SNAPSHOT_CAPACITY(n, rows)
 * bytes, the same on both sides.  A rank that both sends and receives in a tick brackets the two calls with
 * icx_nccl_group_start / _end so that neighbours cannot deadlock; NCCL launches a group's operations at its end, so the import
 * of a receive posted inside a group is enqueued by icx_nccl_group_end (use ONE stream per communicator inside a group).
 */
ICX_API int icx_timeline_slot_send(IcxTimeline* tl, uint16_t coherency, ncclComm_t comm, int peer,
                                   void* device_scratch, size_t scratch_bytes, void* stream);
ICX_API int icx_timeline_slot_recv(IcxTimeline* tl, uint16_t coherency, uint32_t n_entities, ncclComm_t comm, int peer,
                                   void* device_scratch, size_t scratch_bytes, void* stream);
ICX_API int icx_nccl_group_start(void);
ICX_API int icx_nccl_group_end(void);

#ifdef __cplusplus
}
#endif
#endif /* ICX_NCCL_H */
