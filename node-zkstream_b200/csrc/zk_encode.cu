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
 *   k_enc_spans  which packet covers the first byte of every 4 KiB span of the output
 *   k_enc_write  one warp per 4 KiB span of the output: lane l composes packet (first + l) of
 *                the span in a shared-memory window — scalar fields and short strings on its
 *                own, long buffers as jobs the whole warp copies afterwards — and the window
 *                goes to HBM with 16-byte coalesced stores.  A packet that straddles spans is
 *                composed by each of them, clipped to its own bytes.
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

/* ---- span writer ----------------------------------------------------------------- */
#define ENC_SPAN        4096u     /* output bytes per warp */
#define ENC_WARPS       4
#define ENC_INLINE_MAX  48        /* buffers up to this long are copied by the packet's own lane */

__global__ void k_enc_spans(EncArgs a, uint32_t *span_first)
{
	const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= a.r.n) return;
	const uint64_t start = a.frame_off[i], end = a.frame_off[i + 1];
	for (uint64_t s = (start + ENC_SPAN - 1) / ENC_SPAN; s * ENC_SPAN < end; ++s) span_first[s] = (uint32_t)i;
}

/* One lane's view of the span window: bytes of its packet that fall outside [s0, s1) are
 * dropped (another span's warp writes them). */
struct LaneOut {
	uint8_t *win;
	uint64_t s0, s1;
	uint64_t pos;
	__device__ __forceinline__ void put_byte(uint32_t v)      /* writeByte / writeBool jute-buffer.js:111-121 */
	{
		if (pos >= s0 && pos < s1) win[pos - s0] = (uint8_t)v;
		pos += 1;
	}
	__device__ __forceinline__ void put_int(int32_t v)        /* writeInt :141-147 */
	{
		if (pos >= s0 && pos + 4 <= s1) {
			uint8_t *q = win + (pos - s0);
			q[0] = (uint8_t)((uint32_t)v >> 24); q[1] = (uint8_t)((uint32_t)v >> 16);
			q[2] = (uint8_t)((uint32_t)v >> 8); q[3] = (uint8_t)v;
			pos += 4;
		} else {
#pragma unroll
			for (int k = 0; k < 4; ++k) put_byte((uint32_t)v >> (24 - 8 * k));
		}
	}
	__device__ __forceinline__ void put_long(int64_t v)       /* writeLong(8-byte Buffer) :149-165 */
	{
		put_int((int32_t)((uint64_t)v >> 32));
		put_int((int32_t)(uint64_t)v);
	}
	/* the part [k0, k1) of a len-byte field at `pos` that lies inside the span */
	__device__ __forceinline__ void clip(uint32_t len, uint32_t &k0, uint32_t &k1) const
	{
		k0 = pos < s0 ? (uint32_t)((s0 - pos) < (uint64_t)len ? (s0 - pos) : (uint64_t)len) : 0u;
		k1 = pos + len <= s1 ? len : (pos < s1 ? (uint32_t)(s1 - pos) : 0u);
		if (k1 < k0) k1 = k0;
	}
	/* writeBuffer (:123-134): length (-1 when empty), then the bytes — here, or as a job */
	__device__ __forceinline__ void put_buf(const uint8_t *arena, uint64_t off, int32_t len, bool may_defer,
	    uint64_t &job_src, uint32_t &job_len, uint64_t &job_dst)
	{
		if (len <= 0) { put_int(-1); return; }
		put_int(len);
		if (may_defer && len > ENC_INLINE_MAX) {
			job_src = off; job_len = (uint32_t)len; job_dst = pos;
		} else {
			uint32_t k0, k1;
			clip((uint32_t)len, k0, k1);
			const uint8_t *src = arena + off;
			uint8_t *dst = win + (pos + k0 - s0);
			for (uint32_t k = k0; k < k1; ++k) *dst++ = __ldg(src + k);
		}
		pos += (uint32_t)len;
	}
};

__global__ void __launch_bounds__(ENC_WARPS * 32) k_enc_write(EncArgs a, const uint32_t *span_first, uint64_t out_len,
    int aligned16)
{
	__shared__ __align__(16) uint8_t s_win[ENC_WARPS][ENC_SPAN];
	const uint32_t lane = threadIdx.x & 31;
	const uint64_t span = (uint64_t)blockIdx.x * ENC_WARPS + (threadIdx.x >> 5);
	const uint64_t s0 = span * ENC_SPAN;
	if (s0 >= out_len) return;
	const uint64_t s1 = s0 + ENC_SPAN < out_len ? s0 + ENC_SPAN : out_len;
	uint8_t *win = s_win[threadIdx.x >> 5];
	const zkb_requests &r = a.r;

	for (uint64_t base = span_first[span];; base += 32) {
		const uint64_t i = base + lane;
		const bool valid = i < r.n;
		const uint64_t start = valid ? a.frame_off[i] : ~0ull;
		if (!__any_sync(0xffffffffu, valid && start < s1)) break;
		uint64_t j_src[2] = { 0, 0 }, j_dst[2] = { 0, 0 };
		uint32_t j_len[2] = { 0, 0 };
		if (valid && start < s1 && a.status[i] == ZKB_ENC_OK) {
			const uint64_t end = a.frame_off[i + 1];
			if (end > s0 && end > start) {
				LaneOut w;
				w.win = win; w.s0 = s0; w.s1 = s1; w.pos = start;
				const int op = r.op[i];
				w.put_int((int32_t)(end - start - 4));             /* writeLengthPrefixed :181-189 */
				/* what follows the header, by opcode (zk-buffer.js:97-136; connect: :32-39, :50-56) */
				bool f_path = false, f_data = false, f_acl = false, f_byte = false, f_int = false, f_watch = false;
				if (op == ZKB_ENC_CONNECT_REQUEST) {
					w.put_int(r.a0[i]); w.put_long(r.zxid2[i]); w.put_int(r.a1[i]); w.put_long(r.zxid[i]);
					f_data = true;
				} else if (op == ZKB_ENC_CONNECT_RESPONSE) {
					w.put_int(r.a0[i]); w.put_int(r.a1[i]); w.put_long(r.zxid[i]);
					f_data = true;
				} else {
					w.put_int(r.xid[i]);
					w.put_int(op);
					switch (op) {
					case ZKB_OP_GET_CHILDREN: case ZKB_OP_GET_CHILDREN2: case ZKB_OP_GET_DATA: case ZKB_OP_EXISTS:
						f_path = true; f_byte = true; break;
					case ZKB_OP_CREATE: f_path = true; f_data = true; f_acl = true; f_int = true; break;
					case ZKB_OP_DELETE: f_path = true; f_int = true; break;
					case ZKB_OP_GET_ACL: case ZKB_OP_SYNC: f_path = true; break;
					case ZKB_OP_SET_DATA: f_path = true; f_data = true; f_int = true; break;
					case ZKB_OP_SET_WATCHES: f_watch = true; break;
					default: break;   /* CLOSE_SESSION, PING: header only (:129-132) */
					}
				}
				if (f_watch) {                                      /* :255-273 */
					const int32_t n0 = r.n0[i], n1 = r.n1[i], n2 = r.vec_count[i] - n0 - n1;
					uint64_t e = r.vec_first[i];
					uint64_t x0; uint32_t x1; uint64_t x2;
					w.put_long(r.zxid[i]);
					w.put_int(n0);
					for (int32_t k = 0; k < n0; ++k, ++e) w.put_buf(r.arena, r.elem_off[e], r.elem_len[e], false, x0, x1, x2);
					w.put_int(n1);
					for (int32_t k = 0; k < n1; ++k, ++e) w.put_buf(r.arena, r.elem_off[e], r.elem_len[e], false, x0, x1, x2);
					w.put_int(n2);
					for (int32_t k = 0; k < n2; ++k, ++e) w.put_buf(r.arena, r.elem_off[e], r.elem_len[e], false, x0, x1, x2);
				}
				if (f_path) w.put_buf(r.arena, r.path_off[i], r.path_len[i], true, j_src[0], j_len[0], j_dst[0]);
				if (f_data) w.put_buf(r.arena, r.data_off[i], r.data_len[i], true, j_src[1], j_len[1], j_dst[1]);
				if (f_acl) {                                        /* writeACL :384-393 */
					const int32_t cnt = r.vec_count[i];
					uint64_t x0; uint32_t x1; uint64_t x2;
					w.put_int(cnt);
					for (int32_t k = 0; k < cnt; ++k) {
						const uint64_t q = (uint64_t)r.vec_first[i] + (uint64_t)k;
						w.put_int(r.acl_perms[q]);
						w.put_buf(r.arena, r.acl_scheme_off[q], r.acl_scheme_len[q], false, x0, x1, x2);
						w.put_buf(r.arena, r.acl_id_off[q], r.acl_id_len[q], false, x0, x1, x2);
					}
				}
				if (f_byte) w.put_byte(r.a0[i] ? 1 : 0);
				if (f_int) w.put_int(r.a0[i]);
			}
		}
		/* the long buffers of these 32 packets, copied by the whole warp, 4 bytes per lane */
#pragma unroll
		for (int slot = 0; slot < 2; ++slot) {
			uint32_t m = __ballot_sync(0xffffffffu, j_len[slot] != 0);
			while (m) {
				const int src_lane = __ffs(m) - 1;
				m &= m - 1;
				const uint64_t src = __shfl_sync(0xffffffffu, j_src[slot], src_lane);
				const uint64_t dst = __shfl_sync(0xffffffffu, j_dst[slot], src_lane);
				const uint32_t len = __shfl_sync(0xffffffffu, j_len[slot], src_lane);
				const uint64_t lo = dst > s0 ? dst : s0;
				const uint64_t hi = dst + len < s1 ? dst + len : s1;
				if (hi <= lo) continue;
				const uint64_t so = src + (lo - dst);            /* first source byte, arena-relative */
				const uint8_t *sp = r.arena + so;
				uint8_t *dp = win + (lo - s0);
				const uint32_t cnt = (uint32_t)(hi - lo);
				/* head bytes up to a 4-byte aligned window address, aligned words built from two
				 * aligned source words, tail bytes; the last source word is only read whole when
				 * it lies inside the arena */
				uint32_t head = (4u - ((uint32_t)(lo - s0) & 3u)) & 3u;
				if (head > cnt) head = cnt;
				if (lane < head) dp[lane] = __ldg(sp + lane);
				uint32_t nw = (cnt - head) >> 2;
				while (nw && so + head + 4ull * nw + 3 >= r.arena_len) --nw;
				const uint32_t mis = (uint32_t)((uintptr_t)(sp + head) & 3u);
				const uint32_t *sw = reinterpret_cast<const uint32_t *>(sp + head - mis);
				uint32_t *dw = reinterpret_cast<uint32_t *>(dp + head);
				for (uint32_t q = lane; q < nw; q += 32) {
					const uint32_t x = __ldg(sw + q);
					dw[q] = mis ? __funnelshift_r(x, __ldg(sw + q + 1), mis * 8) : x;
				}
				for (uint32_t k = head + 4 * nw + lane; k < cnt; k += 32) dp[k] = __ldg(sp + k);
			}
		}
		if (!__any_sync(0xffffffffu, valid && start < s1 && lane == 31)) break;   /* the last lane was past the span */
	}
	__syncwarp();
	/* window -> HBM */
	const uint32_t L = (uint32_t)(s1 - s0);
	uint8_t *dst = a.out + s0;
	if (aligned16) {
		const uint32_t full = L & ~15u;
		for (uint32_t k = lane * 16; k < full; k += 512)
			*reinterpret_cast<uint4 *>(dst + k) = *reinterpret_cast<const uint4 *>(win + k);
		for (uint32_t k = full + lane; k < L; k += 32) dst[k] = win[k];
	} else {
		for (uint32_t k = lane; k < L; k += 32) dst[k] = win[k];
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
	if (*out_len == 0) return ZKB_OK;
	const uint64_t nspans = (*out_len + ENC_SPAN - 1) / ENC_SPAN;
	if ((rc = zk_ensure(ctx, &ctx->d_span, &ctx->d_span_bytes, (size_t)nspans * 4 + 256, 0))) return rc;
	uint32_t *span_first = (uint32_t *)ctx->d_span;
	k_enc_spans<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(a, span_first);
	ctx->launches++; ZK_CUDA(ctx, cudaGetLastError());
	k_enc_write<<<(unsigned)((nspans + ENC_WARPS - 1) / ENC_WARPS), ENC_WARPS * 32, 0, ctx->stream>>>(a, span_first,
	    *out_len, ((uintptr_t)d_out & 15) == 0);
	ctx->launches++; ZK_CUDA(ctx, cudaGetLastError());
	ZK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return ZKB_OK;
}
