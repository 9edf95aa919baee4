/*
 * zk_stats.cu — per-session reductions over decoded response columns, sm_100a.
 *
 * The batched form of what ZKSession does with every 'packet' event of an attached
 * connection (reference lib/zk-session.js:227-236, :389-419): track the greatest zxid and
 * count notifications by event.  FRAMES rows are in (session, stream position) order, so a
 * session is a contiguous run of rows: one thread per row, a segmented scan by session id
 * inside each warp, and one set of atomics per (warp, session) run.  Streaming, HBM bound:
 * 16 bytes read per row (meta, session, zxid), 8 more for NOTIFICATION rows (a0, a1).
 */
#include <stdio.h>
#include <string.h>
#include <limits.h>
#include "zk_internal.h"

#define ZKB_EXPORT extern "C" __attribute__((visibility("default")))

#define LAUNCH_CHECK(ctx)                                                                  \
	do { (ctx)->launches++; ZK_CUDA(ctx, cudaGetLastError()); } while (0)

namespace {

__global__ void k_stats_init(int64_t *max_zxid, uint32_t *n_replies, uint32_t *notifications,
    uint32_t *first_err, uint32_t nsess)
{
	const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
	if (s >= nsess) return;
	max_zxid[s] = LLONG_MIN;
	n_replies[s] = 0;
#pragma unroll
	for (int k = 0; k < ZKB_NOTIFY_SLOTS; ++k) notifications[(size_t)s * ZKB_NOTIFY_SLOTS + k] = 0;
	if (first_err) first_err[s] = 0xFFFFFFFFu;
}

/* first BAD_DECODE row of each session (ZKB_STATS_UNTIL_ERROR) */
__global__ void k_stats_first_error(const uint32_t *meta, const uint32_t *session, uint64_t n, uint32_t nsess,
    uint32_t *first_err)
{
	const uint64_t f = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (f >= n) return;
	if (ZKB_META_STATUS(__ldg(meta + f)) == ZKB_ST_OK) return;
	const uint32_t s = __ldg(session + f);
	if (s < nsess) atomicMin(first_err + s, (uint32_t)f);
}

__global__ void __launch_bounds__(256) k_stats_reduce(const uint32_t *meta, const uint32_t *session,
    const int64_t *zxid, const int32_t *a0, const int32_t *a1, uint64_t n, uint32_t nsess,
    const uint32_t *first_err, zkb_session_stats out)
{
	const uint64_t f = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	const uint32_t lane = threadIdx.x & 31;
	uint32_t s = 0xFFFFFFFFu;
	long long z = LLONG_MIN;
	/* seven counters of at most 32 each, one byte apiece: byte 0 = replies, byte 1 + k = slot k */
	unsigned long long cnt = 0;
	if (f < n) {
		const uint32_t m = __ldg(meta + f);
		const uint32_t sf = __ldg(session + f);
		if (sf < nsess) {
			s = sf;
			bool counts = ZKB_META_STATUS(m) == ZKB_ST_OK && ZKB_META_KIND(m) == ZKB_KIND_RESPONSE;
			if (counts && first_err && f >= (uint64_t)__ldg(first_err + sf)) counts = false;
			if (counts) {
				if (ZKB_META_OPCODE(m) != ZKB_OP_NOTIFICATION) {
					/* zk-session.js:229-235 */
					z = __ldg(zxid + f);
					cnt = 1ull;
				} else {
					/* processNotification, zk-session.js:389-419.  An error reply's body is never
					 * read (zk-buffer.js:296): its a0/a1 are 0 and state 0 is not SYNC_CONNECTED. */
					const int32_t type = __ldg(a0 + f), state = __ldg(a1 + f);
					int slot;
					if (state != 3) slot = ZKB_NOTIFY_IGNORED;
					else if (type >= 1 && type <= 4) slot = type;
					else slot = ZKB_NOTIFY_BAD_TYPE;
					cnt = 1ull << (8 * (1 + slot));
				}
			}
		}
	}
	/* inclusive segmented scan: rows are sorted by session, so equal ids are adjacent */
#pragma unroll
	for (int d = 1; d < 32; d <<= 1) {
		const uint32_t so = __shfl_up_sync(0xffffffffu, s, d);
		const long long zo = __shfl_up_sync(0xffffffffu, z, d);
		const unsigned long long co = __shfl_up_sync(0xffffffffu, cnt, d);
		if (lane >= (uint32_t)d && so == s) {
			z = zo > z ? zo : z;
			cnt += co;
		}
	}
	const uint32_t sn = __shfl_down_sync(0xffffffffu, s, 1);
	const bool tail = (lane == 31) || (sn != s);
	if (!tail || s == 0xFFFFFFFFu || cnt == 0) return;
	const uint32_t nr = (uint32_t)(cnt & 0xff);
	if (nr) {
		atomicMax(reinterpret_cast<long long *>(out.max_zxid) + s, z);
		atomicAdd(out.n_replies + s, nr);
	}
#pragma unroll
	for (int k = 0; k < ZKB_NOTIFY_SLOTS; ++k) {
		const uint32_t c = (uint32_t)((cnt >> (8 * (1 + k))) & 0xff);
		if (c) atomicAdd(out.notifications + (size_t)s * ZKB_NOTIFY_SLOTS + k, c);
	}
}

}  /* namespace */

static int zk_session_stats_device(zkb_ctx *ctx, const zkb_decoded *dec, uint32_t nsessions, uint32_t flags,
    const zkb_session_stats *out)
{
	int rc;
	if (nsessions == 0) return ZKB_OK;
	const uint64_t n = dec->n_frames;
	uint32_t *first_err = NULL;
	if (flags & ZKB_STATS_UNTIL_ERROR) {
		if ((rc = zk_ensure(ctx, &ctx->d_scratch, &ctx->d_scratch_bytes, (size_t)nsessions * 4 + 256, 0))) return rc;
		first_err = (uint32_t *)ctx->d_scratch;
	}
	k_stats_init<<<(nsessions + 255) / 256, 256, 0, ctx->stream>>>(out->max_zxid, out->n_replies,
	    out->notifications, first_err, nsessions);
	LAUNCH_CHECK(ctx);
	if (n) {
		const unsigned blocks = (unsigned)((n + 255) / 256);
		if (first_err) {
			k_stats_first_error<<<blocks, 256, 0, ctx->stream>>>(dec->meta, dec->session, n, nsessions, first_err);
			LAUNCH_CHECK(ctx);
		}
		k_stats_reduce<<<blocks, 256, 0, ctx->stream>>>(dec->meta, dec->session, dec->zxid, dec->a0, dec->a1, n,
		    nsessions, first_err, *out);
		LAUNCH_CHECK(ctx);
	}
	return ZKB_OK;
}

static int stats_args_ok(zkb_ctx *ctx, const zkb_decoded *dec, uint32_t nsessions, uint32_t flags,
    const zkb_session_stats *out)
{
	if (!ctx) return 0;
	if (!dec || !out || (flags & ~ZKB_STATS_UNTIL_ERROR) ||
	    (nsessions && (!out->max_zxid || !out->n_replies || !out->notifications)) ||
	    (dec->n_frames && (!dec->meta || !dec->session || !dec->zxid || !dec->a0 || !dec->a1))) {
		snprintf(ctx->err, sizeof (ctx->err), "zkb_session_stats: NULL argument or unknown flag");
		return 0;
	}
	if (dec->n_frames > 0xFFFFFFFFull) {
		snprintf(ctx->err, sizeof (ctx->err), "zkb_session_stats: more than 2^32 frame rows");
		return 0;
	}
	return 1;
}

ZKB_EXPORT int zkb_session_stats_dev(zkb_ctx *ctx, const zkb_decoded *dec, uint32_t nsessions, uint32_t flags,
    zkb_session_stats *out)
{
	if (!stats_args_ok(ctx, dec, nsessions, flags, out)) return ZKB_EINVAL;
	ZK_CUDA(ctx, cudaSetDevice(ctx->device));
	int rc = zk_session_stats_device(ctx, dec, nsessions, flags, out);
	if (rc) return rc;
	ZK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return ZKB_OK;
}

ZKB_EXPORT int zkb_session_stats_host(zkb_ctx *ctx, const zkb_decoded *dec, uint32_t nsessions, uint32_t flags,
    zkb_session_stats *out)
{
	int rc;
	if (!stats_args_ok(ctx, dec, nsessions, flags, out)) return ZKB_EINVAL;
	if (nsessions == 0) return ZKB_OK;
	ZK_CUDA(ctx, cudaSetDevice(ctx->device));
	const uint64_t n = dec->n_frames;
	/* staging in the context's input arena: five columns in, three arrays out */
	size_t off = 0;
	auto take = [&off](size_t bytes) { size_t o = off; off += zk_align_up(bytes + 16, 256); return o; };
	const size_t o_meta = take(n * 4), o_sess = take(n * 4), o_zxid = take(n * 8), o_a0 = take(n * 4), o_a1 = take(n * 4);
	const size_t o_mx = take((size_t)nsessions * 8), o_nr = take((size_t)nsessions * 4);
	const size_t o_nt = take((size_t)nsessions * 4 * ZKB_NOTIFY_SLOTS);
	if ((rc = zk_ensure(ctx, &ctx->d_in, &ctx->d_in_bytes, off, 0))) return rc;
	uint8_t *base = (uint8_t *)ctx->d_in;
	zkb_decoded d;
	memset(&d, 0, sizeof (d));
	d.n_frames = n;
	d.meta = (uint32_t *)(base + o_meta); d.session = (uint32_t *)(base + o_sess);
	d.zxid = (int64_t *)(base + o_zxid); d.a0 = (int32_t *)(base + o_a0); d.a1 = (int32_t *)(base + o_a1);
	if (n) {
		ZK_CUDA(ctx, cudaMemcpyAsync(d.meta, dec->meta, n * 4, cudaMemcpyHostToDevice, ctx->stream));
		ZK_CUDA(ctx, cudaMemcpyAsync(d.session, dec->session, n * 4, cudaMemcpyHostToDevice, ctx->stream));
		ZK_CUDA(ctx, cudaMemcpyAsync(d.zxid, dec->zxid, n * 8, cudaMemcpyHostToDevice, ctx->stream));
		ZK_CUDA(ctx, cudaMemcpyAsync(d.a0, dec->a0, n * 4, cudaMemcpyHostToDevice, ctx->stream));
		ZK_CUDA(ctx, cudaMemcpyAsync(d.a1, dec->a1, n * 4, cudaMemcpyHostToDevice, ctx->stream));
	}
	zkb_session_stats ds;
	ds.max_zxid = (int64_t *)(base + o_mx); ds.n_replies = (uint32_t *)(base + o_nr);
	ds.notifications = (uint32_t *)(base + o_nt);
	if ((rc = zk_session_stats_device(ctx, &d, nsessions, flags, &ds))) return rc;
	ZK_CUDA(ctx, cudaMemcpyAsync(out->max_zxid, ds.max_zxid, (size_t)nsessions * 8, cudaMemcpyDeviceToHost, ctx->stream));
	ZK_CUDA(ctx, cudaMemcpyAsync(out->n_replies, ds.n_replies, (size_t)nsessions * 4, cudaMemcpyDeviceToHost, ctx->stream));
	ZK_CUDA(ctx, cudaMemcpyAsync(out->notifications, ds.notifications, (size_t)nsessions * 4 * ZKB_NOTIFY_SLOTS,
	    cudaMemcpyDeviceToHost, ctx->stream));
	ZK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return ZKB_OK;
}
