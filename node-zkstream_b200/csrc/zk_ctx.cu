/*
 * zk_ctx.cu — context management and the extern "C" entry points of libzkbatch.so
 * (declared in include/zkbatch.h).  The *_dev entry points take device pointers; the
 * *_host entry points stage host buffers through context-owned device memory and hand
 * back context-owned pinned host columns.  There is no CPU codec anywhere in here:
 * without a CUDA device zkb_ctx_create fails and nothing else is callable.
 */
#include <new>
#include <stdlib.h>

#include "zk_internal.h"

/* the library is built with -fvisibility=hidden: only the ABI of include/zkbatch.h is exported */
#define ZKB_EXPORT extern "C" __attribute__((visibility("default")))

int zk_ensure(zkb_ctx *ctx, void **p, size_t *have, size_t need, int pinned_host)
{
	if (*have >= need && *p) return ZKB_OK;
	/* grow geometrically so a stream of slightly larger batches does not reallocate each time */
	size_t want = need + need / 4 + 4096;
	if (*p) {
		if (pinned_host) cudaFreeHost(*p); else cudaFree(*p);
		*p = NULL; *have = 0;
	}
	cudaError_t e = pinned_host ? cudaHostAlloc(p, want, cudaHostAllocDefault) : cudaMalloc(p, want);
	if (e != cudaSuccess) {
		/* retry with the exact size before giving up */
		(void)cudaGetLastError();
		want = need;
		e = pinned_host ? cudaHostAlloc(p, want, cudaHostAllocDefault) : cudaMalloc(p, want);
	}
	if (e != cudaSuccess) {
		(void)cudaGetLastError();
		*p = NULL;
		snprintf(ctx->err, sizeof (ctx->err), "allocation of %zu %s bytes failed: %s", want,
		    pinned_host ? "pinned host" : "device", cudaGetErrorString(e));
		return ZKB_ENOMEM;
	}
	*have = want;
	return ZKB_OK;
}

ZKB_EXPORT int zkb_ctx_create(int device_id, zkb_ctx **out)
{
	if (!out) return ZKB_EINVAL;
	*out = NULL;
	int ndev = 0;
	if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0 || device_id < 0 || device_id >= ndev) {
		(void)cudaGetLastError();
		return ZKB_ENODEV;
	}
	if (cudaSetDevice(device_id) != cudaSuccess) { (void)cudaGetLastError(); return ZKB_ENODEV; }
	zkb_ctx *ctx = new (std::nothrow) zkb_ctx();
	if (!ctx) return ZKB_ENOMEM;
	memset(ctx, 0, sizeof (*ctx));
	{ const char *np = getenv("ZKBATCH_NO_PREDICT"); ctx->no_predict = np && np[0] == '1'; }
	ctx->device = device_id;
	strcpy(ctx->err, "");
	cudaDeviceProp prop;
	if (cudaGetDeviceProperties(&prop, device_id) != cudaSuccess) { delete ctx; return ZKB_ENODEV; }
	ctx->sm_count = prop.multiProcessorCount;
	if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess ||
	    cudaMalloc((void **)&ctx->d_ctrl, sizeof (zk_ctrl)) != cudaSuccess ||
	    cudaHostAlloc((void **)&ctx->h_ctrl, sizeof (zk_ctrl), cudaHostAllocDefault) != cudaSuccess) {
		(void)cudaGetLastError();
		zkb_ctx_destroy(ctx);
		return ZKB_ENOMEM;
	}
	*out = ctx;
	return ZKB_OK;
}

ZKB_EXPORT void zkb_ctx_destroy(zkb_ctx *ctx)
{
	if (!ctx) return;
	cudaSetDevice(ctx->device);
	if (ctx->stream) { cudaStreamSynchronize(ctx->stream); cudaStreamDestroy(ctx->stream); }
	if (ctx->d_scratch) cudaFree(ctx->d_scratch);
	if (ctx->d_span) cudaFree(ctx->d_span);
	if (ctx->d_ctrl) cudaFree(ctx->d_ctrl);
	if (ctx->h_ctrl) cudaFreeHost(ctx->h_ctrl);
	if (ctx->d_in) cudaFree(ctx->d_in);
	if (ctx->d_out) cudaFree(ctx->d_out);
	if (ctx->h_out) cudaFreeHost(ctx->h_out);
	if (ctx->h_in) cudaFreeHost(ctx->h_in);
	for (int i = 0; i < 6; i++) if (ctx->ev[i]) cudaEventDestroy(ctx->ev[i]);
	for (int i = 0; i < 2; i++) {
		if (ctx->ev_in[i]) cudaEventDestroy(ctx->ev_in[i]);
		if (ctx->ev_done[i]) cudaEventDestroy(ctx->ev_done[i]);
		if (ctx->ev_out[i]) cudaEventDestroy(ctx->ev_out[i]);
		if (ctx->rp_in[i]) cudaFree(ctx->rp_in[i]);
		if (ctx->rp_out[i]) cudaFree(ctx->rp_out[i]);
		if (ctx->rp_host[i]) cudaFreeHost(ctx->rp_host[i]);
		if (ctx->rp_stage[i]) cudaFreeHost(ctx->rp_stage[i]);
	}
	if (ctx->s_in) cudaStreamDestroy(ctx->s_in);
	if (ctx->s_out) cudaStreamDestroy(ctx->s_out);
	delete ctx;
}

ZKB_EXPORT const char *zkb_last_error(const zkb_ctx *ctx) { return ctx ? ctx->err : "no context"; }
ZKB_EXPORT uint64_t zkb_kernel_launches(const zkb_ctx *ctx) { return ctx ? ctx->launches : 0; }
ZKB_EXPORT uint64_t zkb_row_mispredicts(const zkb_ctx *ctx) { return ctx ? ctx->mispredicts : 0; }
ZKB_EXPORT uint64_t zkb_split_fixups(const zkb_ctx *ctx) { return ctx ? ctx->split_fixups : 0; }
ZKB_EXPORT uint64_t zkb_last_slow_frames(const zkb_ctx *ctx) { return ctx ? ctx->slow_frames : 0; }
#ifdef ZK_PROFILE_PHASES
/* profiling builds only (not part of include/zkbatch.h) */
extern "C" __attribute__((visibility("default"))) void zkb_debug_phase_cycles(const zkb_ctx *ctx, unsigned long long out[8])
{
	memcpy(out, ctx->prof, sizeof (ctx->prof));
}
extern "C" __attribute__((visibility("default"))) void zkb_debug_split_cycles(const zkb_ctx *ctx, unsigned long long out[8])
{
	memcpy(out, ctx->split_prof, sizeof (ctx->split_prof));
}
extern "C" __attribute__((visibility("default"))) void zkb_debug_split_iters(const zkb_ctx *ctx, unsigned int out[64])
{
	memcpy(out, ctx->split_iters, sizeof (ctx->split_iters));
}
#endif
ZKB_EXPORT int zkb_ctx_set_timing(zkb_ctx *ctx, int on)
{
	if (!ctx) return ZKB_EINVAL;
	ZK_CUDA(ctx, cudaSetDevice(ctx->device));
	if (on && !ctx->ev[0])
		for (int i = 0; i < 6; i++) ZK_CUDA(ctx, cudaEventCreate(&ctx->ev[i]));
	ctx->timing = on ? 1 : 0;
	return ZKB_OK;
}
ZKB_EXPORT int zkb_ctx_set_replay_compact(zkb_ctx *ctx, int on)
{
	if (!ctx) return ZKB_EINVAL;
	ctx->replay_compact = on ? 1 : 0;
	return ZKB_OK;
}
ZKB_EXPORT int zkb_last_phase_ms(const zkb_ctx *ctx, float out[3])
{
	if (!ctx || !out) return ZKB_EINVAL;
	out[0] = ctx->phase_ms[0]; out[1] = ctx->phase_ms[1]; out[2] = ctx->phase_ms[2];
	return ZKB_OK;
}
ZKB_EXPORT void *zkb_ctx_stream(zkb_ctx *ctx) { return ctx ? (void *)ctx->stream : NULL; }
ZKB_EXPORT int zkb_ctx_sync(zkb_ctx *ctx)
{
	if (!ctx) return ZKB_EINVAL;
	ZK_CUDA(ctx, cudaSetDevice(ctx->device));
	ZK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return ZKB_OK;
}
ZKB_EXPORT uint64_t zkb_decode_scratch_bytes(uint64_t nbytes, uint32_t nsessions)
{
	return zk_decode_scratch_need(nbytes, nsessions);
}

ZKB_EXPORT int zkb_decode_dev(zkb_ctx *ctx, const uint8_t *bytes, uint64_t nbytes,
    const zkb_sessions *sess, const zkb_xidtab *xid, zkb_decoded *out)
{
	if (!ctx) return ZKB_EINVAL;
	ZK_CUDA(ctx, cudaSetDevice(ctx->device));
	return zk_decode_device(ctx, bytes, nbytes, sess, xid, out);
}

/* ---- carving columns out of one allocation ------------------------------------------ */

struct Carver {
	uint8_t *base;
	size_t   off;
	template <typename T> T *take(uint64_t n)
	{
		T *p = base ? reinterpret_cast<T *>(base + off) : NULL;
		off += zk_align_up((size_t)n * sizeof (T) + 16, 256);
		return p;
	}
};

static size_t carve_decoded(zkb_decoded *d, uint8_t *base, uint64_t cf, uint64_t cs, uint64_t cb,
    uint64_t ca, uint32_t nsess)
{
	Carver c = { base, 0 };
	d->cap_frames = cf; d->cap_stat = cs; d->cap_buf = cb; d->cap_acl = ca;
	d->frame_off = c.take<uint64_t>(cf); d->frame_len = c.take<int32_t>(cf);
	d->session = c.take<uint32_t>(cf); d->meta = c.take<uint32_t>(cf);
	d->xid = c.take<int32_t>(cf); d->zxid = c.take<int64_t>(cf); d->err = c.take<int32_t>(cf);
	d->a0 = c.take<int32_t>(cf); d->a1 = c.take<int32_t>(cf);
	d->stat_frame = c.take<uint32_t>(cs);
	d->czxid = c.take<int64_t>(cs); d->mzxid = c.take<int64_t>(cs);
	d->ctime = c.take<int64_t>(cs); d->mtime = c.take<int64_t>(cs);
	d->version = c.take<int32_t>(cs); d->cversion = c.take<int32_t>(cs);
	d->aversion = c.take<int32_t>(cs); d->ephemeralOwner = c.take<int64_t>(cs);
	d->dataLength = c.take<int32_t>(cs); d->numChildren = c.take<int32_t>(cs);
	d->pzxid = c.take<int64_t>(cs);
	d->buf_frame = c.take<uint32_t>(cb); d->buf_off = c.take<uint64_t>(cb);
	d->buf_len = c.take<int32_t>(cb);
	d->acl_frame = c.take<uint32_t>(ca); d->acl_perms = c.take<int32_t>(ca);
	d->sess_consumed = c.take<uint64_t>(nsess); d->sess_bad_length_at = c.take<uint64_t>(nsess);
	d->sess_first_frame = c.take<uint32_t>((uint64_t)nsess + 1);
	return c.off;
}

/* copy `n` rows of every column from a device-side layout to a host-side layout */
static int copy_decoded_d2h(zkb_ctx *ctx, const zkb_decoded *dv, zkb_decoded *hv, uint32_t nsess)
{
#define CP(col, n)                                                                            \
	if ((n) > 0)                                                                              \
		ZK_CUDA(ctx, cudaMemcpyAsync(hv->col, dv->col, (size_t)(n) * sizeof (*hv->col),        \
		    cudaMemcpyDeviceToHost, ctx->stream))
	const uint64_t nf = dv->n_frames, ns = dv->n_stat, nb = dv->n_buf, na = dv->n_acl;
	CP(frame_off, nf); CP(frame_len, nf); CP(session, nf); CP(meta, nf); CP(xid, nf); CP(zxid, nf);
	CP(err, nf); CP(a0, nf); CP(a1, nf);
	CP(stat_frame, ns); CP(czxid, ns); CP(mzxid, ns); CP(ctime, ns); CP(mtime, ns); CP(version, ns);
	CP(cversion, ns); CP(aversion, ns); CP(ephemeralOwner, ns); CP(dataLength, ns);
	CP(numChildren, ns); CP(pzxid, ns);
	CP(buf_frame, nb); CP(buf_off, nb); CP(buf_len, nb);
	CP(acl_frame, na); CP(acl_perms, na);
	CP(sess_consumed, nsess); CP(sess_bad_length_at, nsess); CP(sess_first_frame, (uint64_t)nsess + 1);
#undef CP
	hv->n_frames = nf; hv->n_stat = ns; hv->n_buf = nb; hv->n_acl = na;
	return ZKB_OK;
}

/* compact replay results: the columns that cannot be derived, buf_rel in place of buf_off */
static size_t carve_compact(zkb_decoded *d, uint8_t *base, uint64_t cf, uint64_t cs, uint64_t cb,
    uint64_t ca, uint32_t nsess)
{
	Carver c = { base, 0 };
	memset(d, 0, sizeof (*d));
	d->cap_frames = cf; d->cap_stat = cs; d->cap_buf = cb; d->cap_acl = ca;
	d->frame_len = c.take<int32_t>(cf); d->meta = c.take<uint32_t>(cf);
	d->xid = c.take<int32_t>(cf); d->zxid = c.take<int64_t>(cf); d->err = c.take<int32_t>(cf);
	d->a0 = c.take<int32_t>(cf); d->a1 = c.take<int32_t>(cf);
	d->czxid = c.take<int64_t>(cs); d->mzxid = c.take<int64_t>(cs);
	d->ctime = c.take<int64_t>(cs); d->mtime = c.take<int64_t>(cs);
	d->version = c.take<int32_t>(cs); d->cversion = c.take<int32_t>(cs);
	d->aversion = c.take<int32_t>(cs); d->ephemeralOwner = c.take<int64_t>(cs);
	d->dataLength = c.take<int32_t>(cs); d->numChildren = c.take<int32_t>(cs);
	d->pzxid = c.take<int64_t>(cs);
	d->buf_rel = c.take<uint32_t>(cb); d->buf_len = c.take<int32_t>(cb);
	d->acl_perms = c.take<int32_t>(ca);
	d->sess_consumed = c.take<uint64_t>(nsess); d->sess_bad_length_at = c.take<uint64_t>(nsess);
	d->sess_first_frame = c.take<uint32_t>((uint64_t)nsess + 1);
	return c.off;
}

/* buf_rel, written over the device copy of buf_frame (same index, same thread: no hazard) */
__global__ void k_buf_rel(zkb_decoded d)
{
	const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= d.n_buf) return;
	const uint32_t owner = d.buf_frame[i];
	d.buf_frame[i] = (uint32_t)(d.buf_off[i] - d.frame_off[owner]);
}

static int copy_compact_d2h(zkb_ctx *ctx, const zkb_decoded *dv, zkb_decoded *hv, uint32_t nsess)
{
#define CP(dst, src, n)                                                                       \
	if ((n) > 0)                                                                              \
		ZK_CUDA(ctx, cudaMemcpyAsync(hv->dst, dv->src, (size_t)(n) * sizeof (*hv->dst),           \
		    cudaMemcpyDeviceToHost, ctx->stream))
	const uint64_t nf = dv->n_frames, ns = dv->n_stat, nb = dv->n_buf, na = dv->n_acl;
	CP(frame_len, frame_len, nf); CP(meta, meta, nf); CP(xid, xid, nf); CP(zxid, zxid, nf);
	CP(err, err, nf); CP(a0, a0, nf); CP(a1, a1, nf);
	CP(czxid, czxid, ns); CP(mzxid, mzxid, ns); CP(ctime, ctime, ns); CP(mtime, mtime, ns); CP(version, version, ns);
	CP(cversion, cversion, ns); CP(aversion, aversion, ns); CP(ephemeralOwner, ephemeralOwner, ns);
	CP(dataLength, dataLength, ns); CP(numChildren, numChildren, ns); CP(pzxid, pzxid, ns);
	CP(buf_rel, buf_frame, nb); CP(buf_len, buf_len, nb);
	CP(acl_perms, acl_perms, na);
	CP(sess_consumed, sess_consumed, nsess); CP(sess_bad_length_at, sess_bad_length_at, nsess);
	CP(sess_first_frame, sess_first_frame, (uint64_t)nsess + 1);
#undef CP
	hv->n_frames = nf; hv->n_stat = ns; hv->n_buf = nb; hv->n_acl = na;
	return ZKB_OK;
}

ZKB_EXPORT int zkb_decode_host(zkb_ctx *ctx, const uint8_t *bytes, uint64_t nbytes,
    const zkb_sessions *sess, const zkb_xidtab *xid, zkb_decoded *out)
{
	int rc;
	if (!ctx) return ZKB_EINVAL;
	if (!sess || !sess->off || !out || (nbytes && !bytes)) {
		snprintf(ctx->err, sizeof (ctx->err), "zkb_decode_host: NULL argument");
		return ZKB_EINVAL;
	}
	ZK_CUDA(ctx, cudaSetDevice(ctx->device));
	const uint32_t nsess = sess->nsessions;
	const bool have_xid = xid && xid->ops && xid->base && xid->count && xid->off;

	/* ---- inputs: one device allocation, carved ---- */
	Carver ci = { NULL, 0 };
	ci.take<uint8_t>(nbytes + 32);
	ci.take<uint64_t>((uint64_t)nsess + 1);
	ci.take<uint8_t>(nsess); ci.take<uint32_t>(nsess);
	ci.take<int32_t>(nsess); ci.take<uint32_t>(nsess); ci.take<uint64_t>(nsess);
	ci.take<int8_t>(have_xid ? xid->nops : 0);
	if ((rc = zk_ensure(ctx, &ctx->d_in, &ctx->d_in_bytes, ci.off, 0))) return rc;
	ci.base = (uint8_t *)ctx->d_in; ci.off = 0;
	uint8_t *d_bytes = ci.take<uint8_t>(nbytes + 32);
	uint64_t *d_off = ci.take<uint64_t>((uint64_t)nsess + 1);
	uint8_t *d_flags = ci.take<uint8_t>(nsess);
	uint32_t *d_hs = ci.take<uint32_t>(nsess);
	int32_t *d_xbase = ci.take<int32_t>(nsess);
	uint32_t *d_xcount = ci.take<uint32_t>(nsess);
	uint64_t *d_xoff = ci.take<uint64_t>(nsess);
	int8_t *d_xops = ci.take<int8_t>(have_xid ? xid->nops : 0);

#define H2D(dst, src, n)                                                                      \
	if ((n) > 0)                                                                              \
		ZK_CUDA(ctx, cudaMemcpyAsync(dst, src, (size_t)(n), cudaMemcpyHostToDevice, ctx->stream))
	H2D(d_bytes, bytes, nbytes);
	H2D(d_off, sess->off, ((size_t)nsess + 1) * 8);
	if (sess->flags) H2D(d_flags, sess->flags, nsess);
	if (sess->handshake_frames) H2D(d_hs, sess->handshake_frames, (size_t)nsess * 4);
	if (have_xid) {
		H2D(d_xbase, xid->base, (size_t)nsess * 4);
		H2D(d_xcount, xid->count, (size_t)nsess * 4);
		H2D(d_xoff, xid->off, (size_t)nsess * 8);
		H2D(d_xops, xid->ops, xid->nops);
	}
#undef H2D
	zkb_sessions dsess;
	dsess.nsessions = nsess; dsess.off = d_off;
	dsess.flags = sess->flags ? d_flags : NULL;
	dsess.handshake_frames = sess->handshake_frames ? d_hs : NULL;
	zkb_xidtab dxid;
	memset(&dxid, 0, sizeof (dxid));
	if (have_xid) {
		dxid.base = d_xbase; dxid.count = d_xcount; dxid.off = d_xoff; dxid.ops = d_xops;
		dxid.nops = xid->nops;
	}

	/* ---- phase A gives the exact frame count, so FRAMES columns are sized exactly ---- */
	uint64_t n_frames = 0;
	if ((rc = zk_split_device(ctx, d_bytes, nbytes, &dsess, have_xid ? &dxid : NULL, &n_frames))) return rc;

	/* the splitter's row prediction is exact for well-formed traffic; a batch it got wrong
	 * comes back with ZKB_EOVERFLOW and the true counts, and is decoded once more */
	uint64_t cs = ctx->pred_rows[0] + 16, cb = ctx->pred_rows[1] + 16, ca = ctx->pred_rows[2] + 16;
	if (nbytes == 0) cs = cb = ca = 16;
	/* a damaged count field can promise anything: no more rows than the bytes could hold */
	if (cs > nbytes / 68 + 16) cs = nbytes / 68 + 16;
	if (cb > nbytes / 4 + 16) cb = nbytes / 4 + 16;
	if (ca > nbytes / 12 + 16) ca = nbytes / 12 + 16;
	zkb_decoded dv;
	for (int attempt = 0;; ++attempt) {
		memset(&dv, 0, sizeof (dv));
		size_t need = carve_decoded(&dv, NULL, n_frames, cs, cb, ca, nsess);
		if ((rc = zk_ensure(ctx, &ctx->d_out, &ctx->d_out_bytes, need, 0))) return rc;
		carve_decoded(&dv, (uint8_t *)ctx->d_out, n_frames, cs, cb, ca, nsess);
		rc = zk_fields_device(ctx, d_bytes, nbytes, &dsess, have_xid ? &dxid : NULL, n_frames, &dv);
		if (rc == ZKB_EOVERFLOW && attempt == 0) {
			cs = dv.n_stat; cb = dv.n_buf; ca = dv.n_acl;
			continue;
		}
		if (rc) return rc;
		break;
	}

	/* ---- results: exact-size pinned host columns ---- */
	zkb_decoded hv;
	memset(&hv, 0, sizeof (hv));
	size_t hneed = carve_decoded(&hv, NULL, dv.n_frames, dv.n_stat, dv.n_buf, dv.n_acl, nsess);
	if ((rc = zk_ensure(ctx, &ctx->h_out, &ctx->h_out_bytes, hneed, 1))) return rc;
	carve_decoded(&hv, (uint8_t *)ctx->h_out, dv.n_frames, dv.n_stat, dv.n_buf, dv.n_acl, nsess);
	if ((rc = copy_decoded_d2h(ctx, &dv, &hv, nsess))) return rc;
	ZK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	*out = hv;
	return ZKB_OK;
}

/* ---- xid table ------------------------------------------------------------------------ */

ZKB_EXPORT int zkb_xidtab_build_dev(zkb_ctx *ctx, const zkb_decoded *req, uint32_t nsessions,
    const zkb_xidtab *layout, int8_t *ops_out)
{
	if (!ctx) return ZKB_EINVAL;
	ZK_CUDA(ctx, cudaSetDevice(ctx->device));
	return zk_xidtab_build_device(ctx, req, nsessions, layout, ops_out);
}

ZKB_EXPORT int zkb_xidtab_build_host(zkb_ctx *ctx, const zkb_decoded *req, uint32_t nsessions,
    const zkb_xidtab *layout, int8_t *ops_out)
{
	int rc;
	if (!ctx) return ZKB_EINVAL;
	if (!req || !layout || !ops_out) return ZKB_EINVAL;
	ZK_CUDA(ctx, cudaSetDevice(ctx->device));
	const uint64_t nf = req->n_frames, nops = layout->nops;
	Carver ci = { NULL, 0 };
	ci.take<uint32_t>(nf); ci.take<int32_t>(nf); ci.take<uint32_t>(nf); ci.take<int32_t>(nf);
	ci.take<int32_t>(nsessions); ci.take<uint32_t>(nsessions); ci.take<uint64_t>(nsessions);
	ci.take<int8_t>(nops);
	if ((rc = zk_ensure(ctx, &ctx->d_in, &ctx->d_in_bytes, ci.off, 0))) return rc;
	ci.base = (uint8_t *)ctx->d_in; ci.off = 0;
	zkb_decoded dreq;
	memset(&dreq, 0, sizeof (dreq));
	dreq.n_frames = nf;
	dreq.meta = ci.take<uint32_t>(nf); dreq.frame_len = ci.take<int32_t>(nf);
	dreq.session = ci.take<uint32_t>(nf); dreq.xid = ci.take<int32_t>(nf);
	int32_t *d_base = ci.take<int32_t>(nsessions);
	uint32_t *d_count = ci.take<uint32_t>(nsessions);
	uint64_t *d_off = ci.take<uint64_t>(nsessions);
	int8_t *d_ops = ci.take<int8_t>(nops);
#define H2D(dst, src, n)                                                                      \
	if ((n) > 0)                                                                              \
		ZK_CUDA(ctx, cudaMemcpyAsync(dst, src, (size_t)(n), cudaMemcpyHostToDevice, ctx->stream))
	H2D(dreq.meta, req->meta, nf * 4); H2D(dreq.frame_len, req->frame_len, nf * 4);
	H2D(dreq.session, req->session, nf * 4); H2D(dreq.xid, req->xid, nf * 4);
	H2D(d_base, layout->base, (size_t)nsessions * 4);
	H2D(d_count, layout->count, (size_t)nsessions * 4);
	H2D(d_off, layout->off, (size_t)nsessions * 8);
#undef H2D
	zkb_xidtab dl;
	memset(&dl, 0, sizeof (dl));
	dl.base = d_base; dl.count = d_count; dl.off = d_off; dl.nops = nops;
	if ((rc = zk_xidtab_build_device(ctx, &dreq, nsessions, &dl, d_ops))) return rc;
	if (nops)
		ZK_CUDA(ctx, cudaMemcpyAsync(ops_out, d_ops, nops, cudaMemcpyDeviceToHost, ctx->stream));
	ZK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return ZKB_OK;
}

/* ---- encode ----------------------------------------------------------------------------- */

ZKB_EXPORT int zkb_encode_dev(zkb_ctx *ctx, const zkb_requests *req, uint8_t *out, uint64_t out_cap,
    uint64_t *frame_off, uint8_t *status, uint64_t *out_len)
{
	if (!ctx) return ZKB_EINVAL;
	ZK_CUDA(ctx, cudaSetDevice(ctx->device));
	return zk_encode_device(ctx, req, out, out_cap, frame_off, status, out_len);
}

ZKB_EXPORT int zkb_encode_host(zkb_ctx *ctx, const zkb_requests *req, const uint8_t **out,
    const uint64_t **frame_off, const uint8_t **status, uint64_t *out_len)
{
	int rc;
	if (!ctx) return ZKB_EINVAL;
	if (!req || !out || !frame_off || !status || !out_len) return ZKB_EINVAL;
	ZK_CUDA(ctx, cudaSetDevice(ctx->device));
	const uint64_t n = req->n, ne = req->n_elem, na = req->n_acl, al = req->arena_len;
	Carver ci = { NULL, 0 };
	zkb_requests d;
	memset(&d, 0, sizeof (d));
	/* size pass then carve pass */
	for (int pass = 0; pass < 2; ++pass) {
		ci.off = 0;
		d.xid = ci.take<int32_t>(n); d.op = ci.take<int8_t>(n);
		d.arena = ci.take<uint8_t>(al);
		d.path_off = ci.take<uint64_t>(n); d.path_len = ci.take<int32_t>(n);
		d.data_off = ci.take<uint64_t>(n); d.data_len = ci.take<int32_t>(n);
		d.a0 = ci.take<int32_t>(n); d.a1 = ci.take<int32_t>(n);
		d.zxid = ci.take<int64_t>(n); d.zxid2 = ci.take<int64_t>(n);
		d.vec_first = ci.take<uint32_t>(n); d.vec_count = ci.take<int32_t>(n);
		d.n0 = ci.take<int32_t>(n); d.n1 = ci.take<int32_t>(n);
		d.elem_off = ci.take<uint64_t>(ne); d.elem_len = ci.take<int32_t>(ne);
		d.acl_perms = ci.take<int32_t>(na);
		d.acl_scheme_off = ci.take<uint64_t>(na); d.acl_scheme_len = ci.take<int32_t>(na);
		d.acl_id_off = ci.take<uint64_t>(na); d.acl_id_len = ci.take<int32_t>(na);
		if (pass == 0) {
			if ((rc = zk_ensure(ctx, &ctx->d_in, &ctx->d_in_bytes, ci.off, 0))) return rc;
			ci.base = (uint8_t *)ctx->d_in;
		}
	}
	d.n = n; d.arena_len = al; d.n_elem = ne; d.n_acl = na;
#define H2D(field, cnt)                                                                       \
	if ((cnt) > 0 && req->field)                                                              \
		ZK_CUDA(ctx, cudaMemcpyAsync((void *)d.field, req->field, (size_t)(cnt) * sizeof (*req->field), \
		    cudaMemcpyHostToDevice, ctx->stream))
	H2D(xid, n); H2D(op, n); H2D(arena, al); H2D(path_off, n); H2D(path_len, n); H2D(data_off, n);
	H2D(data_len, n); H2D(a0, n); H2D(a1, n); H2D(zxid, n); H2D(zxid2, n); H2D(vec_first, n);
	H2D(vec_count, n); H2D(n0, n); H2D(n1, n); H2D(elem_off, ne); H2D(elem_len, ne);
	H2D(acl_perms, na); H2D(acl_scheme_off, na); H2D(acl_scheme_len, na); H2D(acl_id_off, na);
	H2D(acl_id_len, na);
#undef H2D
	/* outputs: frame_off[n+1], status[n], then the bytes once their total is known */
	Carver co = { NULL, 0 };
	co.take<uint64_t>(n + 1); co.take<uint8_t>(n);
	size_t fixed = co.off;
	if ((rc = zk_ensure(ctx, &ctx->d_out, &ctx->d_out_bytes, fixed + 256, 0))) return rc;
	co.base = (uint8_t *)ctx->d_out; co.off = 0;
	uint64_t *d_fo = co.take<uint64_t>(n + 1);
	uint8_t *d_st = co.take<uint8_t>(n);
	uint64_t total = 0;
	uint64_t cap = ctx->d_out_bytes - fixed;
	uint8_t *d_bytes = (uint8_t *)ctx->d_out + fixed;
	rc = zk_encode_device(ctx, &d, d_bytes, cap, d_fo, d_st, &total);
	if (rc == ZKB_EOVERFLOW) {
		/* growing d_out moves it: redo the whole encode into the larger buffer */
		if ((rc = zk_ensure(ctx, &ctx->d_out, &ctx->d_out_bytes, fixed + total + 256, 0))) return rc;
		co.base = (uint8_t *)ctx->d_out; co.off = 0;
		d_fo = co.take<uint64_t>(n + 1);
		d_st = co.take<uint8_t>(n);
		d_bytes = (uint8_t *)ctx->d_out + fixed;
		cap = ctx->d_out_bytes - fixed;
		rc = zk_encode_device(ctx, &d, d_bytes, cap, d_fo, d_st, &total);
	}
	if (rc) return rc;
	Carver ch = { NULL, 0 };
	ch.take<uint64_t>(n + 1); ch.take<uint8_t>(n); ch.take<uint8_t>(total);
	if ((rc = zk_ensure(ctx, &ctx->h_out, &ctx->h_out_bytes, ch.off, 1))) return rc;
	ch.base = (uint8_t *)ctx->h_out; ch.off = 0;
	uint64_t *h_fo = ch.take<uint64_t>(n + 1);
	uint8_t *h_st = ch.take<uint8_t>(n);
	uint8_t *h_bytes = ch.take<uint8_t>(total);
	ZK_CUDA(ctx, cudaMemcpyAsync(h_fo, d_fo, (size_t)(n + 1) * 8, cudaMemcpyDeviceToHost, ctx->stream));
	if (n) ZK_CUDA(ctx, cudaMemcpyAsync(h_st, d_st, (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
	if (total) ZK_CUDA(ctx, cudaMemcpyAsync(h_bytes, d_bytes, (size_t)total, cudaMemcpyDeviceToHost, ctx->stream));
	ZK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	*out = h_bytes; *frame_off = h_fo; *status = h_st; *out_len = total;
	return ZKB_OK;
}

/* ---- pipelined paired replay -------------------------------------------------------------- */

/* slice-local -> global: byte offsets get the slice's byte base, sessions the session base */
__global__ void k_rebase(zkb_decoded d, uint64_t byte_base, uint32_t sess_base, uint32_t nsess)
{
	const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i < d.n_frames) { d.frame_off[i] += byte_base; d.session[i] += sess_base; }
	if (i < d.n_buf) d.buf_off[i] += byte_base;
	if (i < nsess) {
		d.sess_consumed[i] += byte_base;
		if (d.sess_bad_length_at[i] != ZKB_NO_OFFSET) d.sess_bad_length_at[i] += byte_base;
	}
}

struct ReplaySlice {
	uint32_t s0, s1;
	uint64_t req_lo, req_hi, rsp_lo, rsp_hi;
	uint64_t nops;
};

struct SliceDev {      /* device pointers of one slice's inputs */
	uint8_t *req, *rsp;
	uint64_t *req_off, *rsp_off, *x_off;
	uint8_t *req_flags, *rsp_flags;
	uint32_t *req_hs, *rsp_hs, *x_count;
	int32_t *x_base;
	int8_t *x_ops;
};

static size_t carve_slice_in(SliceDev *d, uint8_t *base, const ReplaySlice &sl)
{
	Carver c = { base, 0 };
	const uint32_t n = sl.s1 - sl.s0;
	d->req = c.take<uint8_t>(sl.req_hi - sl.req_lo + 64);
	d->rsp = c.take<uint8_t>(sl.rsp_hi - sl.rsp_lo + 64);
	d->req_off = c.take<uint64_t>((uint64_t)n + 1); d->rsp_off = c.take<uint64_t>((uint64_t)n + 1);
	d->x_off = c.take<uint64_t>(n);
	d->req_flags = c.take<uint8_t>(n); d->rsp_flags = c.take<uint8_t>(n);
	d->req_hs = c.take<uint32_t>(n); d->rsp_hs = c.take<uint32_t>(n); d->x_count = c.take<uint32_t>(n);
	d->x_base = c.take<int32_t>(n);
	d->x_ops = c.take<int8_t>(sl.nops);
	return c.off;
}

/* decode one direction of a slice into ctx->rp_out[...] region starting at *cursor */
static int replay_decode_dir(zkb_ctx *ctx, const uint8_t *d_bytes, uint64_t nbytes, const zkb_sessions *ds,
    const zkb_xidtab *dx, uint8_t *out_base, size_t out_cap, size_t *used, zkb_decoded *dv)
{
	int rc;
	uint64_t n_frames = 0;
	const uint32_t nsess = ds->nsessions;
	if ((rc = zk_split_device(ctx, d_bytes, nbytes, ds, dx, &n_frames))) return rc;
	/* the splitter's row prediction is exact for well-formed traffic; a batch it got wrong
	 * comes back with ZKB_EOVERFLOW and the true counts, and is decoded once more */
	uint64_t cs = ctx->pred_rows[0] + 16, cb = ctx->pred_rows[1] + 16, ca = ctx->pred_rows[2] + 16;
	if (nbytes == 0) cs = cb = ca = 16;
	/* a damaged count field can promise anything: no more rows than the bytes could hold */
	if (cs > nbytes / 68 + 16) cs = nbytes / 68 + 16;
	if (cb > nbytes / 4 + 16) cb = nbytes / 4 + 16;
	if (ca > nbytes / 12 + 16) ca = nbytes / 12 + 16;
	for (int attempt = 0;; ++attempt) {
		memset(dv, 0, sizeof (*dv));
		size_t need = carve_decoded(dv, NULL, n_frames, cs, cb, ca, nsess);
		if (need > out_cap) { *used = need; return ZKB_EOVERFLOW; }
		carve_decoded(dv, out_base, n_frames, cs, cb, ca, nsess);
		rc = zk_fields_device(ctx, d_bytes, nbytes, ds, dx, n_frames, dv);
		if (rc == ZKB_EOVERFLOW && attempt == 0) { cs = dv->n_stat; cb = dv->n_buf; ca = dv->n_acl; continue; }
		/* still too small: tell the caller what the true row counts need, so it grows enough */
		if (rc == ZKB_EOVERFLOW) *used = carve_decoded(dv, NULL, n_frames, dv->n_stat + 16, dv->n_buf + 16, dv->n_acl + 16, nsess);
		if (rc) return rc;
		*used = need;
		return ZKB_OK;
	}
}

ZKB_EXPORT int zkb_replay_host(zkb_ctx *ctx,
    const uint8_t *req_bytes, uint64_t req_nbytes, const zkb_sessions *req_sess,
    const uint8_t *rsp_bytes, uint64_t rsp_nbytes, const zkb_sessions *rsp_sess,
    const int32_t *xid_base, const uint32_t *xid_count, uint64_t slice_bytes,
    zkb_slice_cb cb, void *user)
{
	int rc;
	if (!ctx) return ZKB_EINVAL;
	if (!req_sess || !rsp_sess || !req_sess->off || !rsp_sess->off || !xid_base || !xid_count || !cb ||
	    req_sess->nsessions != rsp_sess->nsessions) {
		snprintf(ctx->err, sizeof (ctx->err), "zkb_replay_host: bad arguments");
		return ZKB_EINVAL;
	}
	const uint32_t nsess = req_sess->nsessions;
	if (req_sess->off[0] != 0 || rsp_sess->off[0] != 0 || req_sess->off[nsess] != req_nbytes ||
	    rsp_sess->off[nsess] != rsp_nbytes) {
		snprintf(ctx->err, sizeof (ctx->err), "zkb_replay_host: session offsets do not cover the buffers");
		return ZKB_EINVAL;
	}
	for (uint32_t s = 0; s < nsess; s++) {
		if (req_sess->off[s] > req_sess->off[s + 1] || rsp_sess->off[s] > rsp_sess->off[s + 1]) {
			snprintf(ctx->err, sizeof (ctx->err), "zkb_replay_host: session offsets not monotone");
			return ZKB_EINVAL;
		}
	}
	ZK_CUDA(ctx, cudaSetDevice(ctx->device));
	if (!ctx->s_in) {
		ZK_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->s_in, cudaStreamNonBlocking));
		ZK_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->s_out, cudaStreamNonBlocking));
		for (int i = 0; i < 2; i++) {
			ZK_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_in[i], cudaEventDisableTiming));
			ZK_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_done[i], cudaEventDisableTiming));
			ZK_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_out[i], cudaEventDisableTiming));
		}
	}
	if (slice_bytes == 0) slice_bytes = 512ull << 20;

	/* ---- cut the sessions into slices ---- */
	uint32_t nslices = 0, cap_slices = 16;
	ReplaySlice *sl = (ReplaySlice *)malloc(cap_slices * sizeof (ReplaySlice));
	if (!sl) return ZKB_ENOMEM;
	size_t max_in = 0;
	uint32_t max_n = 0;
	for (uint32_t s = 0; s < nsess || nslices == 0;) {
		ReplaySlice x;
		x.s0 = s;
		uint64_t bytes = 0;
		x.nops = 0;
		while (s < nsess) {
			uint64_t b = (req_sess->off[s + 1] - req_sess->off[s]) + (rsp_sess->off[s + 1] - rsp_sess->off[s]);
			if (s > x.s0 && bytes + b > slice_bytes) break;
			bytes += b; x.nops += xid_count[s]; s++;
		}
		x.s1 = s;
		x.req_lo = req_sess->off[x.s0]; x.req_hi = req_sess->off[x.s1];
		x.rsp_lo = rsp_sess->off[x.s0]; x.rsp_hi = rsp_sess->off[x.s1];
		if (nslices == cap_slices) {
			cap_slices *= 2;
			ReplaySlice *n2 = (ReplaySlice *)realloc(sl, cap_slices * sizeof (ReplaySlice));
			if (!n2) { free(sl); return ZKB_ENOMEM; }
			sl = n2;
		}
		sl[nslices++] = x;
		SliceDev tmp;
		size_t need = carve_slice_in(&tmp, NULL, x);
		if (need > max_in) max_in = need;
		if (x.s1 - x.s0 > max_n) max_n = x.s1 - x.s0;
		if (nsess == 0) break;
	}

#define RP_FAIL(code) do { rc = (code); goto out; } while (0)
#define RP_CUDA(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) { \
	snprintf(ctx->err, sizeof (ctx->err), "%s:%d: %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
	RP_FAIL(ZKB_ECUDA); } } while (0)

	rc = ZKB_OK;
	{
		/* buffers: two slice inputs, two slice outputs (device + pinned mirror), staging */
		const size_t stage_need = zk_align_up(((size_t)max_n + 1) * 8, 256) * 3 + 4096;
		size_t out_guess = (size_t)(slice_bytes + slice_bytes / 4) + (64u << 20);
		for (int i = 0; i < 2; i++) {
			if ((rc = zk_ensure(ctx, &ctx->rp_in[i], &ctx->rp_in_bytes[i], max_in + 4096, 0))) goto out;
			if ((rc = zk_ensure(ctx, &ctx->rp_stage[i], &ctx->rp_stage_bytes[i], stage_need, 1))) goto out;
			if (ctx->rp_out_bytes[i] < out_guess) {
				if ((rc = zk_ensure(ctx, &ctx->rp_out[i], &ctx->rp_out_bytes[i], out_guess, 0))) goto out;
				if ((rc = zk_ensure(ctx, &ctx->rp_host[i], &ctx->rp_host_bytes[i], out_guess, 1))) goto out;
			}
		}
		SliceDev sd[2];
		zkb_decoded hreq[2], hrsp[2];
		uint32_t pending = UINT32_MAX;     /* slice whose download is in flight */

		/* upload of slice i into buffer i & 1 (asynchronous, on s_in) */
		auto upload = [&](uint32_t i) -> int {
			const int b = i & 1;
			const ReplaySlice &x = sl[i];
			const uint32_t n = x.s1 - x.s0;
			carve_slice_in(&sd[b], (uint8_t *)ctx->rp_in[b], x);
			/* rebased session offsets and the table layout are staged in pinned memory */
			uint8_t *st = (uint8_t *)ctx->rp_stage[b];
			uint64_t *h_req_off = (uint64_t *)st;
			uint64_t *h_rsp_off = (uint64_t *)(st + zk_align_up(((size_t)n + 1) * 8, 256));
			uint64_t *h_x_off = (uint64_t *)(st + 2 * zk_align_up(((size_t)n + 1) * 8, 256));
			uint64_t xo = 0;
			for (uint32_t k = 0; k <= n; k++) {
				h_req_off[k] = req_sess->off[x.s0 + k] - x.req_lo;
				h_rsp_off[k] = rsp_sess->off[x.s0 + k] - x.rsp_lo;
				if (k < n) { h_x_off[k] = xo; xo += xid_count[x.s0 + k]; }
			}
			cudaError_t e;
#define UP(dst, src, bytes) if ((bytes) > 0 && (e = cudaMemcpyAsync(dst, src, (size_t)(bytes), cudaMemcpyHostToDevice, ctx->s_in)) != cudaSuccess) { \
	snprintf(ctx->err, sizeof (ctx->err), "zkb_replay_host upload: %s", cudaGetErrorString(e)); return ZKB_ECUDA; }
			UP(sd[b].req, req_bytes + x.req_lo, x.req_hi - x.req_lo);
			UP(sd[b].rsp, rsp_bytes + x.rsp_lo, x.rsp_hi - x.rsp_lo);
			UP(sd[b].req_off, h_req_off, ((size_t)n + 1) * 8);
			UP(sd[b].rsp_off, h_rsp_off, ((size_t)n + 1) * 8);
			UP(sd[b].x_off, h_x_off, (size_t)n * 8);
			if (req_sess->flags) UP(sd[b].req_flags, req_sess->flags + x.s0, n);
			if (rsp_sess->flags) UP(sd[b].rsp_flags, rsp_sess->flags + x.s0, n);
			if (req_sess->handshake_frames) UP(sd[b].req_hs, req_sess->handshake_frames + x.s0, (size_t)n * 4);
			if (rsp_sess->handshake_frames) UP(sd[b].rsp_hs, rsp_sess->handshake_frames + x.s0, (size_t)n * 4);
			UP(sd[b].x_base, xid_base + x.s0, (size_t)n * 4);
			UP(sd[b].x_count, xid_count + x.s0, (size_t)n * 4);
#undef UP
			if ((e = cudaEventRecord(ctx->ev_in[b], ctx->s_in)) != cudaSuccess) return ZKB_ECUDA;
			return ZKB_OK;
		};
		auto deliver = [&](uint32_t i) -> int {
			const int b = i & 1;
			cudaError_t e = cudaEventSynchronize(ctx->ev_out[b]);
			if (e != cudaSuccess) { snprintf(ctx->err, sizeof (ctx->err), "zkb_replay_host download: %s", cudaGetErrorString(e)); return ZKB_ECUDA; }
			return cb(user, sl[i].s0, sl[i].s1 - sl[i].s0, &hreq[b], &hrsp[b]);
		};

		if (nsess) {
			/* the input buffer may still be read by nothing: first slice goes straight up */
			if ((rc = upload(0))) goto out;
		}
		for (uint32_t i = 0; i < nslices && nsess; i++) {
			const int b = i & 1;
			const ReplaySlice &x = sl[i];
			const uint32_t n = x.s1 - x.s0;
			/* next slice's upload runs under this slice's kernels; its buffer was last read by
			 * slice i-1's kernels, which have finished (the decode calls below are synchronous) */
			if (i + 1 < nslices && (rc = upload(i + 1))) goto out;
			RP_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->ev_in[b], 0));
			/* this slice's device outputs were last downloaded for slice i-2 */
			if (i >= 2) RP_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->ev_out[b], 0));

			zkb_sessions ds_req = { n, sd[b].req_off, req_sess->flags ? sd[b].req_flags : NULL,
			    req_sess->handshake_frames ? sd[b].req_hs : NULL };
			zkb_sessions ds_rsp = { n, sd[b].rsp_off, rsp_sess->flags ? sd[b].rsp_flags : NULL,
			    rsp_sess->handshake_frames ? sd[b].rsp_hs : NULL };
			zkb_xidtab dx;
			dx.base = sd[b].x_base; dx.count = sd[b].x_count; dx.off = sd[b].x_off; dx.ops = sd[b].x_ops; dx.nops = x.nops;
			zkb_decoded dreq, drsp;
			size_t used_req = 0, used_rsp = 0;
			for (int attempt = 0;; ++attempt) {
				rc = replay_decode_dir(ctx, sd[b].req, x.req_hi - x.req_lo, &ds_req, NULL,
				    (uint8_t *)ctx->rp_out[b], ctx->rp_out_bytes[b], &used_req, &dreq);
				if (rc == ZKB_OK) {
					used_req = zk_align_up(used_req, 256);
					if ((rc = zk_xidtab_build_device(ctx, &dreq, n, &dx, sd[b].x_ops))) goto out;
					rc = replay_decode_dir(ctx, sd[b].rsp, x.rsp_hi - x.rsp_lo, &ds_rsp, x.nops ? &dx : NULL,
					    (uint8_t *)ctx->rp_out[b] + used_req, ctx->rp_out_bytes[b] - used_req, &used_rsp, &drsp);
				}
				if (rc == ZKB_EOVERFLOW && attempt < 3) {
					/* the slice needs more room than estimated: wait for the other buffer's
					 * download, then grow this pair (device + pinned mirror) */
					if (pending != UINT32_MAX) { int r2 = deliver(pending); pending = UINT32_MAX; if (r2) RP_FAIL(r2); }
					RP_CUDA(cudaStreamSynchronize(ctx->s_out));
					const size_t want = (used_req + used_rsp) * 2 + (64u << 20);
					if ((rc = zk_ensure(ctx, &ctx->rp_out[b], &ctx->rp_out_bytes[b], want, 0))) goto out;
					if ((rc = zk_ensure(ctx, &ctx->rp_host[b], &ctx->rp_host_bytes[b], want, 1))) goto out;
					continue;
				}
				if (rc) goto out;
				break;
			}
			const uint64_t m1 = dreq.n_frames > dreq.n_buf ? dreq.n_frames : dreq.n_buf;
			const uint64_t m2 = drsp.n_frames > drsp.n_buf ? drsp.n_frames : drsp.n_buf;
			k_rebase<<<(unsigned)(((m1 > n ? m1 : n) + 255) / 256), 256, 0, ctx->stream>>>(dreq, x.req_lo, x.s0, n);
			k_rebase<<<(unsigned)(((m2 > n ? m2 : n) + 255) / 256), 256, 0, ctx->stream>>>(drsp, x.rsp_lo, x.s0, n);
			ctx->launches += 2;
			if (ctx->replay_compact) {
				if (dreq.n_buf) { k_buf_rel<<<(unsigned)((dreq.n_buf + 255) / 256), 256, 0, ctx->stream>>>(dreq); ctx->launches++; }
				if (drsp.n_buf) { k_buf_rel<<<(unsigned)((drsp.n_buf + 255) / 256), 256, 0, ctx->stream>>>(drsp); ctx->launches++; }
			}
			RP_CUDA(cudaGetLastError());
			RP_CUDA(cudaEventRecord(ctx->ev_done[b], ctx->stream));

			/* download on s_out: exact row counts, into this buffer's pinned mirror */
			RP_CUDA(cudaStreamWaitEvent(ctx->s_out, ctx->ev_done[b], 0));
			{
				cudaStream_t keep = ctx->stream;
				ctx->stream = ctx->s_out;      /* copy_decoded_d2h queues on ctx->stream */
				memset(&hreq[b], 0, sizeof (zkb_decoded)); memset(&hrsp[b], 0, sizeof (zkb_decoded));
				if (ctx->replay_compact) {
					size_t h1 = carve_compact(&hreq[b], (uint8_t *)ctx->rp_host[b], dreq.n_frames, dreq.n_stat, dreq.n_buf, dreq.n_acl, n);
					carve_compact(&hrsp[b], (uint8_t *)ctx->rp_host[b] + zk_align_up(h1, 256), drsp.n_frames, drsp.n_stat, drsp.n_buf, drsp.n_acl, n);
					rc = copy_compact_d2h(ctx, &dreq, &hreq[b], n);
					if (!rc) rc = copy_compact_d2h(ctx, &drsp, &hrsp[b], n);
				} else {
					size_t h1 = carve_decoded(&hreq[b], (uint8_t *)ctx->rp_host[b], dreq.n_frames, dreq.n_stat, dreq.n_buf, dreq.n_acl, n);
					carve_decoded(&hrsp[b], (uint8_t *)ctx->rp_host[b] + zk_align_up(h1, 256), drsp.n_frames, drsp.n_stat, drsp.n_buf, drsp.n_acl, n);
					rc = copy_decoded_d2h(ctx, &dreq, &hreq[b], n);
					if (!rc) rc = copy_decoded_d2h(ctx, &drsp, &hrsp[b], n);
				}
				ctx->stream = keep;
				if (rc) goto out;
			}
			RP_CUDA(cudaEventRecord(ctx->ev_out[b], ctx->s_out));

			/* hand over the previous slice while this one downloads */
			if (pending != UINT32_MAX) { int r2 = deliver(pending); pending = UINT32_MAX; if (r2) RP_FAIL(r2); }
			pending = i;
			/* the kernels of this slice are done (decode calls are synchronous): its input buffer
			 * is free for slice i+2, which upload() will queue in the next iteration */
		}
		if (pending != UINT32_MAX) { int r2 = deliver(pending); if (r2) RP_FAIL(r2); }
	}
out:
	cudaStreamSynchronize(ctx->s_in);
	cudaStreamSynchronize(ctx->s_out);
	cudaStreamSynchronize(ctx->stream);
	free(sl);
	return rc;
#undef RP_FAIL
#undef RP_CUDA
}
