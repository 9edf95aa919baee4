/*
 * zk_internal.h — shared between the translation units of libzkbatch.so.
 * Not part of the public ABI (that is include/zkbatch.h).
 */
#ifndef ZK_INTERNAL_H
#define ZK_INTERNAL_H

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/zkbatch.h"

/* ---- geometry ---------------------------------------------------------------------- */
#define ZK_GRAN_MIN_LOG2     11   /* splitter granule: 2 KiB .. 8 KiB (= one decode chunk) */
#define ZK_GRAN_MAX_LOG2     13
#define ZK_MAX_BATCH_BYTES ((1ull << 34) - 1)  /* keeps every row index below 2^32 */

#define ZK_NONE     0xFFFFFFFFFFFFFFFFull  /* no live chain */
#define ZK_INHERIT  0xFFFFFFFFFFFFFFFEull  /* granule did not touch the chain */

/* device error bits (d_ctrl->error) */
#define ZK_DERR_SESSIONS  1u  /* session offsets not monotone / not covering the batch */
#define ZK_DERR_COUNT     4u  /* frame count disagreed between split and emit */
#define ZK_DERR_MISPREDICT 8u /* a chunk's side rows differ from the splitter's prediction: place the rest by the exact route */

struct zk_ctrl {
	unsigned int       error;
	unsigned int       n_dirty;     /* granules whose state changed in the last scan (dirty or re-selected) */
	unsigned int       first_bad;   /* k_decode<true>: lowest chunk whose rows differ from the prediction (cleared with n_dirty during the split) */
	unsigned int       n_rewalk;    /* of those, granules that need a new walk */
	unsigned long long n_frames;
	unsigned long long n_stat, n_buf, n_acl;
	unsigned long long enc_bytes;
	unsigned long long n_long;   /* children vectors longer than a tile that the decode pass left for k_long_children (cleared together with n_slow and prof) */
	unsigned long long n_slow;   /* frames that needed the generic (exact) parser */
	unsigned long long prof[8];  /* ZK_PROFILE_PHASES builds: cycles per decode phase, chunk count */
	unsigned int       pred_rows[3]; /* the splitter's predicted STAT / BUF / ACL row totals of the batch */
	unsigned int       dir_mask;     /* 1 = some session is client->server, 2 = some is server->client (k_check_sessions) */
	unsigned int       first_dirty[2]; /* n_dirty / n_rewalk after the first scan of a split (k_split_round) */
	unsigned long long xid_stored, xid_filled; /* xid table build: requests that stored an opcode; slots that hold one */
};

struct zkb_ctx {
	int           device;
	cudaStream_t  stream;
	char          err[512];
	uint64_t      launches;
	uint64_t      split_fixups;   /* FIX iterations the splitter has needed so far */
	uint64_t      slow_frames;    /* frames the last decode sent through the generic parser */
	int           no_predict;     /* ZKBATCH_NO_PREDICT=1: every batch takes the exact route (count, scan, place) */
	int           dir;            /* batch split last: 1 = only client->server streams, 2 = only server->client, 0 = both */
	uint64_t      pred_rows[3];   /* predicted side-table rows of the batch split last (exact for well-formed traffic) */
	uint64_t      mispredicts;    /* batches whose tail was placed again by the exact route because a row prediction failed */
	unsigned long long prof[8];
	unsigned long long split_prof[8];
	unsigned int split_iters[32][2];  /* last split: (n_dirty, n_rewalk) after each scan */
	/* device scratch (grown on demand, never shrunk) */
	void         *d_scratch;   size_t d_scratch_bytes;
	void         *d_span;      size_t d_span_bytes;    /* encoder: first packet of every output span */
	zk_ctrl      *d_ctrl;
	zk_ctrl      *h_ctrl;      /* pinned */
	/* host-variant staging */
	void         *d_in;        size_t d_in_bytes;      /* batch bytes + session arrays + xid table */
	void         *d_out;       size_t d_out_bytes;     /* every output column */
	void         *h_out;       size_t h_out_bytes;     /* pinned mirror of d_out */
	void         *h_in;        size_t h_in_bytes;      /* pinned bounce buffer for small arrays */
	int           sm_count;
	/* zkb_replay_host: upload / download streams, double-buffered slices */
	cudaStream_t  s_in, s_out;
	cudaEvent_t   ev_in[2], ev_done[2], ev_out[2];
	void         *rp_in[2];    size_t rp_in_bytes[2];    /* device: slice inputs */
	void         *rp_out[2];   size_t rp_out_bytes[2];   /* device: slice outputs */
	void         *rp_host[2];  size_t rp_host_bytes[2];  /* pinned: slice results */
	void         *rp_stage[2]; size_t rp_stage_bytes[2]; /* pinned: rebased session arrays */
	/* optional per-phase CUDA-event timing (zkb_ctx_set_timing) */
	int           replay_compact; /* zkb_replay_host leaves derivable columns out of the download */
	int           timing;
	cudaEvent_t   ev[6];          /* split begin/end, emit begin/end, decode begin/end */
	float         phase_ms[3];    /* split, emit, decode of the last decode call */
};

#define ZK_EV(ctx, i)                                                                    \
	do { if ((ctx)->timing) cudaEventRecord((ctx)->ev[i], (ctx)->stream); } while (0)

#define ZK_CUDA(ctx, call)                                                              \
	do {                                                                                \
		cudaError_t e__ = (call);                                                       \
		if (e__ != cudaSuccess) {                                                       \
			snprintf((ctx)->err, sizeof ((ctx)->err), "%s:%d: %s: %s", __FILE__, __LINE__, \
			    #call, cudaGetErrorString(e__));                                        \
			return ZKB_ECUDA;                                                           \
		}                                                                               \
	} while (0)

static inline size_t zk_align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

int zk_ensure(zkb_ctx *ctx, void **p, size_t *have, size_t need, int pinned_host);

/* decode pipeline on device pointers; defined in zk_decode.cu */
int zk_decode_device(zkb_ctx *ctx, const uint8_t *d_bytes, uint64_t nbytes,
    const zkb_sessions *d_sess, const zkb_xidtab *d_xid, zkb_decoded *d_out);
int zk_split_device(zkb_ctx *ctx, const uint8_t *d_bytes, uint64_t nbytes,
    const zkb_sessions *d_sess, const zkb_xidtab *d_xid, uint64_t *n_frames);
int zk_fields_device(zkb_ctx *ctx, const uint8_t *d_bytes, uint64_t nbytes,
    const zkb_sessions *d_sess, const zkb_xidtab *d_xid, uint64_t n_frames, zkb_decoded *d_out);
int zk_xidtab_build_device(zkb_ctx *ctx, const zkb_decoded *d_req, uint32_t nsessions,
    const zkb_xidtab *d_layout, int8_t *d_ops_out);
/* encode pipeline on device pointers; defined in zk_encode.cu */
int zk_encode_device(zkb_ctx *ctx, const zkb_requests *d_req, uint8_t *d_out, uint64_t out_cap,
    uint64_t *d_frame_off, uint8_t *d_status, uint64_t *out_len);
size_t zk_decode_scratch_need(uint64_t nbytes, uint32_t nsessions);
size_t zk_encode_scratch_need(uint64_t n);

#endif
