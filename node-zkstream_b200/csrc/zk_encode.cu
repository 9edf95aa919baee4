/*
 * zk_encode.cu — batched ZKEncodeStream for sm_100a.
 *
 * Replaces, for a whole batch of packets at once,
 *   ZKEncodeStream._transform             lib/zk-streams.js:121-148
 *   ZKBuffer.writeRequest + sub-writers   lib/zk-buffer.js:97-136,143-273,384-426
 *   ZKBuffer.writeConnectRequest/Response lib/zk-buffer.js:32-39,50-56
 *   JuteBuffer.write*, writeLengthPrefixed lib/jute-buffer.js:111-165,181-189
 *
 *   k_enc_size   one thread per packet: frame size from the descriptors (and whether the
 *                reference would have thrown 'Unsupported opcode')
 *   k_esc_*      exclusive prefix sum of the sizes -> frame_off (64-bit)
 *   k_enc_write  one warp per packet: header fields and payload bytes are scattered to
 *                frame_off[i]; payload copies are lane-strided so a warp's stores land in
 *                consecutive bytes.
 */
#include "zk_internal.h"

namespace {

struct EncArgs {
	zkb_requests r;
	uint32_t    *sizes;
	uint64_t    *frame_off;
	uint8_t     *status;
	uint8_t     *out;
	uint64_t     out_cap;
	zk_ctrl     *ctrl;
};

/* writeBuffer (jute-buffer.js:123-134): 4-byte length (-1 when empty) + bytes */
__device__ __forceinline__ uint32_t sz_buf(int32_t len) { return 4u + (len > 0 ? (uint32_t)len : 0u); }

__device__ uint32_t body_size(const zkb_requests &r, uint64_t i, int op, bool &bad)
{
	bad = false;
	if (op == ZKB_ENC_CONNECT_REQUEST) return 4 + 8 + 4 + 8 + sz_buf(r.data_len[i]);
	if (op == ZKB_ENC_CONNECT_RESPONSE) return 4 + 4 + 8 + sz_buf(r.data_len[i]);
	uint32_t n = 8;   /* xid + opcode */
	switch (op) {
	case ZKB_OP_GET_CHILDREN: case ZKB_OP_GET_CHILDREN2: case ZKB_OP_GET_DATA: case ZKB_OP_EXISTS:
		n += sz_buf(r.path_len[i]) + 1; break;
	case ZKB_OP_CREATE: {
		n += sz_buf(r.path_len[i]) + sz_buf(r.data_len[i]) + 4;
		const int32_t cnt = r.vec_count[i];
		for (int32_t k = 0; k < cnt; ++k) {
			uint64_t a = (uint64_t)r.vec_first[i] + (uint64_t)k;
			n += 4 + sz_buf(r.acl_scheme_len[a]) + sz_buf(r.acl_id_len[a]);
		}
		n += 4; break;
	}
	case ZKB_OP_DELETE: n += sz_buf(r.path_len[i]) + 4; break;
	case ZKB_OP_GET_ACL: case ZKB_OP_SYNC: n += sz_buf(r.path_len[i]); break;
	case ZKB_OP_SET_DATA: n += sz_buf(r.path_len[i]) + sz_buf(r.data_len[i]) + 4; break;
	case ZKB_OP_SET_WATCHES: {
		n += 8 + 12;
		const int32_t cnt = r.vec_count[i];
		for (int32_t k = 0; k < cnt; ++k) n += sz_buf(r.elem_len[(uint64_t)r.vec_first[i] + (uint64_t)k]);
		break;
	}
	case ZKB_OP_CLOSE_SESSION: case ZKB_OP_PING: break;
	default: bad = true; return 0;   /* zk-buffer.js:133-134 throws: nothing is pushed */
	}
	return n;
}

__global__ void k_enc_size(EncArgs a)
{
	uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= a.r.n) return;
	bool bad;
	uint32_t body = body_size(a.r, i, a.r.op[i], bad);
	a.sizes[i] = bad ? 0u : 4u + body;
	a.status[i] = bad ? ZKB_ENC_UNSUPPORTED : ZKB_ENC_OK;
}

/* ---- 64-bit exclusive prefix sum of 32-bit sizes --------------------------------- */
#define ESC_THREADS 1024

__device__ uint64_t esc_block_incl(uint64_t x, uint64_t *total, uint64_t *s_w)
{
	const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
	for (int d = 1; d < 32; d <<= 1) {
		uint64_t y = __shfl_up_sync(0xffffffffu, x, d);
		if (lane >= d) x += y;
	}
	if (lane == 31) s_w[wid] = x;
	__syncthreads();
	uint64_t pre = 0, tot = 0;
	for (int w = 0; w < ESC_THREADS / 32; ++w) {
		if (w == wid) pre = tot;
		tot += s_w[w];
	}
	__syncthreads();
	*total = tot;
	return pre + x;
}

__global__ void __launch_bounds__(ESC_THREADS) k_esc_reduce(const uint32_t *sizes, uint64_t n, uint64_t *blk)
{
	__shared__ uint64_t s_w[32];
	uint64_t i = (uint64_t)blockIdx.x * ESC_THREADS + threadIdx.x;
	uint64_t tot;
	esc_block_incl(i < n ? sizes[i] : 0, &tot, s_w);
	if (threadIdx.x == 0) blk[blockIdx.x] = tot;
}

__global__ void __launch_bounds__(ESC_THREADS) k_esc_blocks(uint64_t *blk, uint32_t nblk)
{
	__shared__ uint64_t s_w[32];
	uint64_t carry = 0;
	for (uint32_t b0 = 0; b0 < nblk; b0 += ESC_THREADS) {
		uint32_t i = b0 + threadIdx.x;
		uint64_t x = i < nblk ? blk[i] : 0;
		uint64_t tot;
		uint64_t inc = esc_block_incl(x, &tot, s_w);
		if (i < nblk) blk[i] = carry + inc - x;
		carry += tot;
	}
}

__global__ void __launch_bounds__(ESC_THREADS) k_esc_apply(const uint32_t *sizes, uint64_t n,
    const uint64_t *blk, uint64_t *frame_off, zk_ctrl *ctrl)
{
	__shared__ uint64_t s_w[32];
	uint64_t i = (uint64_t)blockIdx.x * ESC_THREADS + threadIdx.x;
	uint64_t x = i < n ? sizes[i] : 0;
	uint64_t tot;
	uint64_t inc = esc_block_incl(x, &tot, s_w);
	if (i >= n) return;
	uint64_t ex = blk[blockIdx.x] + inc - x;
	frame_off[i] = ex;
	if (i == n - 1) { frame_off[n] = ex + x; ctrl->enc_bytes = ex + x; }
}

/* ---- warp-cooperative writer: `pos` is uniform across the warp --------------------- */
struct WW {
	uint8_t *out;
	uint64_t pos;
	int      lane;
	__device__ __forceinline__ void w_int(int32_t v)       /* writeInt jute-buffer.js:141-147 */
	{
		if (lane < 4) out[pos + lane] = (uint8_t)((uint32_t)v >> (24 - 8 * lane));
		pos += 4;
	}
	__device__ __forceinline__ void w_long(int64_t v)      /* writeLong(8-byte Buffer) :149-165 */
	{
		if (lane < 8) out[pos + lane] = (uint8_t)((uint64_t)v >> (56 - 8 * lane));
		pos += 8;
	}
	__device__ __forceinline__ void w_byte(int v)          /* writeByte / writeBool :111-121 */
	{
		if (lane == 0) out[pos] = (uint8_t)v;
		pos += 1;
	}
	__device__ __forceinline__ void w_buf(const uint8_t *arena, uint64_t off, int32_t len)
	{
		if (len <= 0) { w_int(-1); return; }                /* :127-130 empty encodes as -1 */
		w_int(len);
		const uint8_t *src = arena + off;
		uint8_t *dst = out + pos;
		/* head bytes up to a 4-byte aligned destination, then 4 bytes per lane, then the tail */
		uint32_t head = (uint32_t)((4 - ((uintptr_t)dst & 3)) & 3);
		if (head > (uint32_t)len) head = (uint32_t)len;
		if ((uint32_t)lane < head) dst[lane] = src[lane];
		const uint32_t body = ((uint32_t)len - head) & ~3u;
		const uint8_t *s2 = src + head;
		uint32_t *d2 = reinterpret_cast<uint32_t *>(dst + head);
		for (uint32_t w = lane; w < body / 4; w += 32) {
			const uint8_t *q = s2 + 4 * w;
			d2[w] = (uint32_t)q[0] | ((uint32_t)q[1] << 8) | ((uint32_t)q[2] << 16) | ((uint32_t)q[3] << 24);
		}
		const uint32_t done = head + body;
		if (done + (uint32_t)lane < (uint32_t)len) dst[done + lane] = src[done + lane];
		pos += (uint32_t)len;
	}
};

__global__ void __launch_bounds__(256) k_enc_write(EncArgs a)
{
	const uint64_t i = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	if (i >= a.r.n) return;
	if (a.status[i] != ZKB_ENC_OK) return;
	const zkb_requests &r = a.r;
	const uint64_t start = a.frame_off[i];
	const uint32_t size = a.sizes[i];
	if (start + size > a.out_cap) return;
	WW w;
	w.out = a.out; w.pos = start; w.lane = threadIdx.x & 31;
	const int op = r.op[i];
	w.w_int((int32_t)(size - 4));                       /* writeLengthPrefixed jute-buffer.js:181-189 */
	if (op == ZKB_ENC_CONNECT_REQUEST) {                /* zk-buffer.js:32-39 */
		w.w_int(r.a0[i]); w.w_long(r.zxid2[i]); w.w_int(r.a1[i]); w.w_long(r.zxid[i]);
		w.w_buf(r.arena, r.data_off[i], r.data_len[i]);
		return;
	}
	if (op == ZKB_ENC_CONNECT_RESPONSE) {               /* zk-buffer.js:50-56 */
		w.w_int(r.a0[i]); w.w_int(r.a1[i]); w.w_long(r.zxid[i]);
		w.w_buf(r.arena, r.data_off[i], r.data_len[i]);
		return;
	}
	w.w_int(r.xid[i]);                                  /* writeRequest zk-buffer.js:97-99 */
	w.w_int(op);
	switch (op) {
	case ZKB_OP_GET_CHILDREN: case ZKB_OP_GET_CHILDREN2: case ZKB_OP_GET_DATA: case ZKB_OP_EXISTS:
		w.w_buf(r.arena, r.path_off[i], r.path_len[i]);
		w.w_byte(r.a0[i] ? 1 : 0);
		break;
	case ZKB_OP_CREATE: {
		w.w_buf(r.arena, r.path_off[i], r.path_len[i]);
		w.w_buf(r.arena, r.data_off[i], r.data_len[i]);
		const int32_t cnt = r.vec_count[i];
		w.w_int(cnt);                                   /* writeACL :384-393 */
		for (int32_t k = 0; k < cnt; ++k) {
			uint64_t q = (uint64_t)r.vec_first[i] + (uint64_t)k;
			w.w_int(r.acl_perms[q]);
			w.w_buf(r.arena, r.acl_scheme_off[q], r.acl_scheme_len[q]);
			w.w_buf(r.arena, r.acl_id_off[q], r.acl_id_len[q]);
		}
		w.w_int(r.a0[i]);
		break;
	}
	case ZKB_OP_DELETE:
		w.w_buf(r.arena, r.path_off[i], r.path_len[i]);
		w.w_int(r.a0[i]);
		break;
	case ZKB_OP_GET_ACL: case ZKB_OP_SYNC:
		w.w_buf(r.arena, r.path_off[i], r.path_len[i]);
		break;
	case ZKB_OP_SET_DATA:
		w.w_buf(r.arena, r.path_off[i], r.path_len[i]);
		w.w_buf(r.arena, r.data_off[i], r.data_len[i]);
		w.w_int(r.a0[i]);
		break;
	case ZKB_OP_SET_WATCHES: {                          /* :255-273 */
		const int32_t n0 = r.n0[i], n1 = r.n1[i], n2 = r.vec_count[i] - n0 - n1;
		uint64_t e = r.vec_first[i];
		w.w_long(r.zxid[i]);
		w.w_int(n0);
		for (int32_t k = 0; k < n0; ++k, ++e) w.w_buf(r.arena, r.elem_off[e], r.elem_len[e]);
		w.w_int(n1);
		for (int32_t k = 0; k < n1; ++k, ++e) w.w_buf(r.arena, r.elem_off[e], r.elem_len[e]);
		w.w_int(n2);
		for (int32_t k = 0; k < n2; ++k, ++e) w.w_buf(r.arena, r.elem_off[e], r.elem_len[e]);
		break;
	}
	default:
		break;   /* CLOSE_SESSION, PING: header only (:129-132) */
	}
}

} /* namespace */

size_t zk_encode_scratch_need(uint64_t n)
{
	uint64_t nblk = (n + ESC_THREADS - 1) / ESC_THREADS;
	return zk_align_up(n * 4, 256) + zk_align_up((nblk + 1) * 8, 256) + 256;
}

int zk_encode_device(zkb_ctx *ctx, const zkb_requests *req, uint8_t *d_out, uint64_t out_cap,
    uint64_t *d_frame_off, uint8_t *d_status, uint64_t *out_len)
{
	int rc;
	if (!req || !d_frame_off || !d_status || !out_len) {
		snprintf(ctx->err, sizeof (ctx->err), "zkb_encode: NULL argument");
		return ZKB_EINVAL;
	}
	*out_len = 0;
	const uint64_t n = req->n;
	if (n == 0) {
		ZK_CUDA(ctx, cudaMemsetAsync(d_frame_off, 0, 8, ctx->stream));
		ZK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
		return ZKB_OK;
	}
	if (n > 0xfffffff0ull) {
		snprintf(ctx->err, sizeof (ctx->err), "zkb_encode: too many packets");
		return ZKB_ERANGE;
	}
	if ((rc = zk_ensure(ctx, &ctx->d_scratch, &ctx->d_scratch_bytes, zk_encode_scratch_need(n), 0))) return rc;
	const uint32_t nblk = (uint32_t)((n + ESC_THREADS - 1) / ESC_THREADS);
	EncArgs a;
	a.r = *req;
	a.sizes = (uint32_t *)ctx->d_scratch;
	uint64_t *blk = (uint64_t *)((uint8_t *)ctx->d_scratch + zk_align_up(n * 4, 256));
	a.frame_off = d_frame_off; a.status = d_status; a.out = d_out; a.out_cap = out_cap;
	a.ctrl = ctx->d_ctrl;

	ZK_CUDA(ctx, cudaMemsetAsync(ctx->d_ctrl, 0, sizeof (zk_ctrl), ctx->stream));
	k_enc_size<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(a);
	ctx->launches++; ZK_CUDA(ctx, cudaGetLastError());
	k_esc_reduce<<<nblk, ESC_THREADS, 0, ctx->stream>>>(a.sizes, n, blk);
	ctx->launches++; ZK_CUDA(ctx, cudaGetLastError());
	k_esc_blocks<<<1, ESC_THREADS, 0, ctx->stream>>>(blk, nblk);
	ctx->launches++; ZK_CUDA(ctx, cudaGetLastError());
	k_esc_apply<<<nblk, ESC_THREADS, 0, ctx->stream>>>(a.sizes, n, blk, d_frame_off, ctx->d_ctrl);
	ctx->launches++; ZK_CUDA(ctx, cudaGetLastError());
	ZK_CUDA(ctx, cudaMemcpyAsync(ctx->h_ctrl, ctx->d_ctrl, sizeof (zk_ctrl), cudaMemcpyDeviceToHost,
	    ctx->stream));
	ZK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	*out_len = ctx->h_ctrl->enc_bytes;
	if (*out_len > out_cap || (*out_len && !d_out)) {
		snprintf(ctx->err, sizeof (ctx->err), "zkb_encode: %llu bytes needed, capacity %llu",
		    (unsigned long long)*out_len, (unsigned long long)out_cap);
		return ZKB_EOVERFLOW;
	}
	const uint64_t threads = n * 32;
	k_enc_write<<<(unsigned)((threads + 255) / 256), 256, 0, ctx->stream>>>(a);
	ctx->launches++; ZK_CUDA(ctx, cudaGetLastError());
	ZK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return ZKB_OK;
}
