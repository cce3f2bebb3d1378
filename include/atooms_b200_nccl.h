/*
 * atooms_b200_nccl.h — NCCL fast path of the slab decomposition (CUDA library only).
 *
 * No counterpart in the reference (it has no domain decomposition and no data
 * collectives).  The generic path in atooms_b200/dd.py moves
 * the halo with torch.distributed from Python, one round trip per collective;
 * these entry points enqueue the same per-step sequence
 *     fused step kernel -> ncclAllReduce(max, 1 float) -> commit
 *     -> pack -> ncclSend/ncclRecv with the two ring neighbours -> unpack
 * from C on the context's stream.  NCCL is resolved with dlopen at run time.
 */
#ifndef ATOOMS_B200_NCCL_H
#define ATOOMS_B200_NCCL_H

#include "atooms_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* 128-byte ncclUniqueId, to be created on one rank and broadcast by the caller. */
int b2_nccl_unique_id(char out[128]);
/* Create this context's communicator (collective over all ranks).               */
int b2_nccl_init(b2_ctx *ctx, const char id[128], int rank, int nranks);
/* Describe the halo of the resident domain: device arrays (caller-owned, kept
 * alive until the next call) of the LOCAL ids sent to the left / right ring
 * neighbour each step, and how many ghosts arrive from the left / right; ghosts
 * are the last n_ghost_l + n_ghost_r local ids, left block first.               */
int b2_nccl_set_halo(b2_ctx *ctx, const int32_t *send_l_dev, int64_t n_send_l,
                     const int32_t *send_r_dev, int64_t n_send_r,
                     int64_t n_ghost_l, int64_t n_ghost_r);
/* Enqueue `count` guarded steps with their communication (after b2_dd_begin);
 * returns without waiting.  Every rank must call it with the same count.       */
int b2_nccl_run(b2_ctx *ctx, int64_t count);

/* ---- peer-memory halo: the same per-step sequence without a communication library ----
 *
 * Each rank owns a mailbox block in its own HBM (b2_peer_export returns its 64-byte CUDA
 * IPC handle); after the caller has gathered every rank's handle, b2_peer_connect maps
 * the others' blocks.  b2_peer_run then enqueues, per step,
 *     fused step kernel, which stores this rank's max |v|^2 into every rank's mailbox
 *     -> a one-warp kernel that waits for all ranks' values (the all-reduce) and commits
 *     -> a kernel that gathers the boundary positions straight INTO the two neighbours'
 *        inboxes over NVLink and raises their arrival flags
 *     -> a kernel that waits for both neighbours' flags and scatters its inbox into the ghosts
 * i.e. 4 launches and no library call per step (NCCL path: 8 launches, two of them
 * collectives).  One rank per GPU is required (kernels of different ranks wait on each
 * other).  `halo_capacity` (records per inbox, the same on every rank) must cover
 * b2_nccl_set_halo's counts, which describes the halo for this path too.  Waits time out
 * after ~10 s and surface through b2_dd_poll.                                            */
int b2_peer_export(b2_ctx *ctx, int64_t halo_capacity, int rank, int nranks, char handle_out[64]);
int b2_peer_connect(b2_ctx *ctx, const char *handles /* nranks x 64 bytes, by rank */);
int b2_peer_run(b2_ctx *ctx, int64_t count);

/* ---- slab refresh on the device (peer-memory path only) --------------------------------
 *
 * The reference has no counterpart (its engines are single-domain); this replaces the
 * host-driven refresh of atooms_b200/dd.py — download, classify, exchange through
 * torch.distributed, re-upload: 1.7 ms for 1.2M particles per GPU — by five kernels and one
 * host synchronisation.  The context then also owns what travels with a particle besides
 * position, velocity, mass and species: its global id and its count of box crossings.
 *
 * b2_peer_set_ids     hands over the global ids (and the crossings so far) of the resident
 *                     particles, by local id; needed again after every b2_upload.
 * b2_peer_refresh     owned particles whose x / L_x + 1/2 left [x_lo, x_hi) move to the ring
 *                     neighbour they now belong to, those within `margin` (same units) of a
 *                     face are sent as ghosts, and the context re-ingests
 *                     [staying | from left | from right | ghosts left | ghosts right]
 *                     with the per-step halo (b2_nccl_set_halo) set up to match; the list is
 *                     rebuilt by the next b2_dd_begin.  Every rank must call it at the same
 *                     point.  counts_out = {owned, ghosts left, ghosts right, sent left per
 *                     step, sent right per step, migrants out, migrants in}.
 *                     Limits: migrants per face <= halo_capacity / 4 + 1024, ghosts per face
 *                     <= halo_capacity, and the new domain must fit the arrays allocated by
 *                     the last b2_upload (on a connected context it leaves 12 % headroom);
 *                     beyond them the call fails with B2_ESTATE and the domain must be
 *                     re-uploaded.
 * b2_peer_update      b2_update for the first n_first local ids only (the owned particles of a slab;
 *                     the ghosts keep their coordinates until the next b2_peer_refresh, which the
 *                     caller must run before stepping on).  pos / vel: n_first x 3 doubles in
 *                     pinned host or device memory; either may be NULL.
 * b2_peer_state       global id, total crossings, species (with B2_GHOST) and mass of the
 *                     resident particles, by local id, into device arrays (any may be NULL). */
int b2_peer_set_ids(b2_ctx *ctx, const int64_t *gid_dev, const int32_t *image_dev /* or NULL */);
int b2_peer_refresh(b2_ctx *ctx, double x_lo, double x_hi, double margin, int64_t counts_out[7]);
int b2_peer_update(b2_ctx *ctx, int64_t n_first, const double *pos, const double *vel);
int b2_peer_state(b2_ctx *ctx, int64_t *gid_dev, int32_t *image_dev, int32_t *species_dev, double *mass_dev);

#ifdef __cplusplus
}
#endif
#endif
