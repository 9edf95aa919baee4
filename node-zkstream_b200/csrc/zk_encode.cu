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

__global__ void k_enc_spans(EncArgs a, uint32_t *span_first)
{
	const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= a.r.n) return;
	const uint64_t start = a.frame_off[i], end = a.frame_off[i + 1];
	for (uint64_t s = (start + ENC_SPAN - 1) / ENC_SPAN; s * ENC_SPAN < end; ++s) span_first[s] = (uint32_t)i;
}

/* ---- the span writer ------------------------------------------------------------------------
 * One warp per 4 KiB span of the output.  Lane l takes packet (first + l) of the span and
 *   1. writes the packet's integers (length prefix, xid, opcode, field lengths, flags ...) into the
 *      span's shared-memory window, four byte stores each, in a fixed order of stages - header,
 *      path, data, ACL, tail - so that lanes with different opcodes run the same code together;
 *   2. files every buffer / string it has to copy as a job { arena offset, window range };
 * then the warp splits into four groups of eight lanes that take jobs off the list until it is
 * empty (aligned words built from two aligned source words, bytes at the ragged ends), and the
 * window goes to HBM with 16-byte coalesced stores.  A packet that straddles spans is composed by
 * each of them, clipped to its own bytes.  ConnectRequest / ConnectResponse / SET_WATCHES packets
 * and ACL entries beyond the third take a plain per-lane path (LaneOut::put_buf_inline). */
#define ENC_JOBS 88u

struct EncJob { uint64_t src; uint32_t dlo, dhi; };     /* arena offset of the byte that lands at window offset dlo */

#define ENC_SHORT 64u      /* jobs up to this many bytes are copied by one lane each, longer ones by eight lanes */
struct __align__(16) EncSm {
	uint8_t  win[ENC_SPAN];
	EncJob   job[ENC_JOBS];  /* short jobs from the bottom, long jobs from the top */
	uint32_t nshort, nlong, next;
};

/* One lane's view of the span window: bytes of its packet that fall outside the span are dropped
 * (another span's warp writes them).  Positions are relative to the span's first byte. */
struct LaneOut {
	EncSm   *sm;
	int64_t  rel;        /* where the next byte goes, relative to the window (may be negative) */
	int64_t  L;          /* bytes of the window in use */
	__device__ __forceinline__ void put_byte(uint32_t v)      /* writeByte / writeBool jute-buffer.js:111-121 */
	{
		if (rel >= 0 && rel < L) sm->win[rel] = (uint8_t)v;
		rel += 1;
	}
	__device__ __forceinline__ void put_int(int32_t v)        /* writeInt :141-147 */
	{
		if (rel >= 0 && rel + 4 <= L) {
			uint8_t *q = sm->win + rel;
			q[0] = (uint8_t)((uint32_t)v >> 24); q[1] = (uint8_t)((uint32_t)v >> 16);
			q[2] = (uint8_t)((uint32_t)v >> 8); q[3] = (uint8_t)v;
			rel += 4;
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
	/* the bytes of a buffer, copied by this lane */
	__device__ __forceinline__ void copy_inline(const uint8_t *arena, uint64_t off, uint32_t len)
	{
		int64_t k0 = rel < 0 ? -rel : 0, k1 = rel + (int64_t)len <= L ? (int64_t)len : L - rel;
		if (k0 > (int64_t)len) k0 = len;
		for (int64_t k = k0; k < k1; ++k) sm->win[rel + k] = __ldg(arena + off + (uint64_t)k);
		rel += len;
	}
	/* writeBuffer (:123-134): length (-1 when empty), then the bytes, as a job for the warp */
	__device__ __forceinline__ void put_buf(const uint8_t *arena, uint64_t off, int32_t len)
	{
		if (len <= 0) { put_int(-1); return; }
		put_int(len);
		const int64_t lo = rel < 0 ? 0 : rel, hi = rel + len < L ? rel + len : L;
		if (hi > lo) {
			const bool small = (uint32_t)(hi - lo) <= ENC_SHORT;
			const uint32_t j = atomicAdd(small ? &sm->nshort : &sm->nlong, 1u);
			/* the two ends may not meet: either counter alone stays below half the list */
			if (j < ENC_JOBS / 2) {
				EncJob e;
				e.src = off + (uint64_t)(lo - rel); e.dlo = (uint32_t)lo; e.dhi = (uint32_t)hi;
				sm->job[small ? j : ENC_JOBS - 1 - j] = e;
				rel += len;
				return;
			}
			copy_inline(arena, off, (uint32_t)len);      /* list full: this lane copies */
			return;
		}
		rel += len;
	}
	__device__ __forceinline__ void put_buf_inline(const uint8_t *arena, uint64_t off, int32_t len)
	{
		if (len <= 0) { put_int(-1); return; }
		put_int(len);
		copy_inline(arena, off, (uint32_t)len);
	}
};

/* one lane copies one short job: every lane of the warp has its own job in flight, so the arena
 * reads of 32 strings overlap instead of queueing behind each other */
__device__ __forceinline__ void enc_copy_short(EncSm &sm, const uint8_t *arena, uint64_t arena_len, const EncJob e)
{
	const uint32_t n = e.dhi - e.dlo;
	uint8_t *dp = sm.win + e.dlo;
	const uint8_t *sp = arena + e.src;
	if (n <= 8 && e.src + 16 <= arena_len) {
		/* the ACL's "world" / "anyone" and the like: two or three aligned words, shifted into place,
		 * then at most eight byte stores - no loop, no byte loads */
		const uint32_t mis8 = (uint32_t)((uintptr_t)sp & 3u) * 8u;
		const uint32_t *sw = reinterpret_cast<const uint32_t *>(sp - ((uintptr_t)sp & 3u));
		const uint32_t x0 = __ldg(sw), x1 = __ldg(sw + 1), x2 = __ldg(sw + 2);
		const uint32_t lo = __funnelshift_r(x0, x1, mis8), hi = __funnelshift_r(x1, x2, mis8);
#pragma unroll
		for (uint32_t k = 0; k < 8; ++k)
			if (k < n) dp[k] = (uint8_t)((k < 4 ? lo : hi) >> ((k & 3u) * 8u));
		return;
	}
	if (n < 8 || e.src + n + 8 > arena_len) {
		for (uint32_t k = 0; k < n; ++k) dp[k] = __ldg(sp + k);
		return;
	}
	/* up to three bytes to the next word of the window, whole words, up to three bytes after them; the
	 * odd bytes come out of aligned words like the rest (no byte loads, no byte loops) */
	const uint32_t head = (4u - (e.dlo & 3u)) & 3u;
	{
		const uint32_t m0 = (uint32_t)((uintptr_t)sp & 3u);
		const uint32_t *s0 = reinterpret_cast<const uint32_t *>(sp - m0);
		const uint32_t h = __funnelshift_r(__ldg(s0), __ldg(s0 + 1), m0 * 8);
		if (head > 0) dp[0] = (uint8_t)h;
		if (head > 1) dp[1] = (uint8_t)(h >> 8);
		if (head > 2) dp[2] = (uint8_t)(h >> 16);
	}
	const uint32_t nw = (n - head) >> 2;
	const uint32_t mis = (uint32_t)((uintptr_t)(sp + head) & 3u);
	const uint32_t *sw = reinterpret_cast<const uint32_t *>(sp + head - mis);
	uint32_t *dw = reinterpret_cast<uint32_t *>(dp + head);
	uint32_t x0 = __ldg(sw);
#pragma unroll 4
	for (uint32_t q = 0; q < nw; ++q) {
		const uint32_t x1 = mis ? __ldg(sw + q + 1) : 0u;
		dw[q] = mis ? __funnelshift_r(x0, x1, mis * 8) : x0;
		x0 = mis ? x1 : __ldg(sw + q + 1);
	}
	/* x0 = sw[nw]: the word the tail starts in (the read of sw[nw + 1] stays inside src + n + 8) */
	const uint32_t rem = n - head - 4 * nw;
	if (rem) {
		const uint32_t t = mis ? __funnelshift_r(x0, __ldg(sw + nw + 1), mis * 8) : x0;
		uint8_t *tp = dp + head + 4 * nw;
		tp[0] = (uint8_t)t;
		if (rem > 1) tp[1] = (uint8_t)(t >> 8);
		if (rem > 2) tp[2] = (uint8_t)(t >> 16);
	}
}

/* eight lanes copy one job: bytes [e.dlo, e.dhi) of the window from the arena */
__device__ __forceinline__ void enc_copy_job(EncSm &sm, const uint8_t *arena, uint64_t arena_len, const EncJob e, const uint32_t gl)
{
	const uint32_t n = e.dhi - e.dlo;
	uint8_t *dp = sm.win + e.dlo;
	const uint8_t *sp = arena + e.src;
	if (n < 24 || e.src + n + 8 > arena_len) {
		for (uint32_t k = gl; k < n; k += 8) dp[k] = __ldg(sp + k);
		return;
	}
	const uint32_t head = (4u - (e.dlo & 3u)) & 3u;
	if (gl < head) dp[gl] = __ldg(sp + gl);
	const uint32_t nw = (n - head) >> 2;
	const uint32_t mis = (uint32_t)((uintptr_t)(sp + head) & 3u);
	const uint32_t *sw = reinterpret_cast<const uint32_t *>(sp + head - mis);
	uint32_t *dw = reinterpret_cast<uint32_t *>(dp + head);
	if (mis) {
		for (uint32_t q = gl; q < nw; q += 8) dw[q] = __funnelshift_r(__ldg(sw + q), __ldg(sw + q + 1), mis * 8);
	} else {
		for (uint32_t q = gl; q < nw; q += 8) dw[q] = __ldg(sw + q);
	}
	for (uint32_t k = head + 4 * nw + gl; k < n; k += 8) dp[k] = __ldg(sp + k);
}

__global__ void __launch_bounds__(ENC_WARPS * 32, 10) k_enc_write(EncArgs a, const uint32_t *span_first, uint64_t out_len,
    int aligned16)
{
	__shared__ EncSm s_all[ENC_WARPS];
	const uint32_t lane = threadIdx.x & 31;
	const uint64_t span = (uint64_t)blockIdx.x * ENC_WARPS + (threadIdx.x >> 5);
	const uint64_t s0 = span * ENC_SPAN;
	if (s0 >= out_len) return;
	const uint64_t s1 = s0 + ENC_SPAN < out_len ? s0 + ENC_SPAN : out_len;
	EncSm &sm = s_all[threadIdx.x >> 5];
	const zkb_requests &r = a.r;

	for (uint64_t base = span_first[span];; base += 32) {
		const uint64_t i = base + lane;
		const bool valid = i < r.n;
		const uint64_t start = valid ? a.frame_off[i] : ~0ull;
		if (!__any_sync(0xffffffffu, valid && start < s1)) break;
		if (lane == 0) { sm.nshort = 0; sm.nlong = 0; sm.next = 0; }
		__syncwarp();
		uint64_t end = 0;
		bool live = false;
		if (valid && start < s1 && a.status[i] == ZKB_ENC_OK) {
			end = a.frame_off[i + 1];
			live = end > s0 && end > start;
		}
		int op = 0;
		LaneOut w;
		w.sm = &sm; w.L = (int64_t)(s1 - s0); w.rel = 0;
		/* which stages the packet goes through (zk-buffer.js:97-136; connect: :32-39, :50-56) */
		bool f_path = false, f_data = false, f_acl = false, f_byte = false, f_int = false, f_other = false;
		/* the packet's scalar columns, fetched together (one round trip instead of one per stage);
		 * a column the caller left out reads as zero, as before */
		int32_t c_xid = 0, c_a0 = 0, c_plen = 0, c_dlen = 0, c_vcnt = 0;
		uint64_t c_poff = 0, c_doff = 0, c_vfirst = 0;
		if (live) {
			op = r.op[i];
			if (r.xid) c_xid = r.xid[i];
			if (r.a0) c_a0 = r.a0[i];
			if (r.path_len) { c_plen = r.path_len[i]; c_poff = r.path_off[i]; }
			if (r.data_len) { c_dlen = r.data_len[i]; c_doff = r.data_off[i]; }
			if (r.vec_count) { c_vcnt = r.vec_count[i]; c_vfirst = r.vec_first[i]; }
			w.rel = (int64_t)start - (int64_t)s0;
			w.put_int((int32_t)(end - start - 4));                 /* writeLengthPrefixed :181-189 */
			switch (op) {
			case ZKB_OP_GET_CHILDREN: case ZKB_OP_GET_CHILDREN2: case ZKB_OP_GET_DATA: case ZKB_OP_EXISTS:
				f_path = true; f_byte = true; break;
			case ZKB_OP_CREATE: f_path = true; f_data = true; f_acl = true; f_int = true; break;
			case ZKB_OP_DELETE: f_path = true; f_int = true; break;
			case ZKB_OP_GET_ACL: case ZKB_OP_SYNC: f_path = true; break;
			case ZKB_OP_SET_DATA: f_path = true; f_data = true; f_int = true; break;
			case ZKB_OP_CLOSE_SESSION: case ZKB_OP_PING: break;        /* header only (:129-132) */
			default: f_other = true; break;                            /* connect packets, SET_WATCHES */
			}
		}
		if (live && !f_other) { w.put_int(c_xid); w.put_int(op); }
		if (f_other) {
			if (op == ZKB_ENC_CONNECT_REQUEST) {
				w.put_int(r.a0[i]); w.put_long(r.zxid2[i]); w.put_int(r.a1[i]); w.put_long(r.zxid[i]);
				w.put_buf_inline(r.arena, r.data_off[i], r.data_len[i]);
			} else if (op == ZKB_ENC_CONNECT_RESPONSE) {
				w.put_int(r.a0[i]); w.put_int(r.a1[i]); w.put_long(r.zxid[i]);
				w.put_buf_inline(r.arena, r.data_off[i], r.data_len[i]);
			} else {                                                   /* SET_WATCHES :255-273 */
				const int32_t n0 = r.n0[i], n1 = r.n1[i], n2 = r.vec_count[i] - n0 - n1;
				uint64_t e = r.vec_first[i];
				w.put_int(r.xid[i]); w.put_int(op);
				w.put_long(r.zxid[i]);
				w.put_int(n0);
				for (int32_t k = 0; k < n0; ++k, ++e) w.put_buf_inline(r.arena, r.elem_off[e], r.elem_len[e]);
				w.put_int(n1);
				for (int32_t k = 0; k < n1; ++k, ++e) w.put_buf_inline(r.arena, r.elem_off[e], r.elem_len[e]);
				w.put_int(n2);
				for (int32_t k = 0; k < n2; ++k, ++e) w.put_buf_inline(r.arena, r.elem_off[e], r.elem_len[e]);
			}
		}
		if (f_path) w.put_buf(r.arena, c_poff, c_plen);
		if (f_data) w.put_buf(r.arena, c_doff, c_dlen);
		if (f_acl) {                                                   /* writeACL :384-393 */
			const int32_t cnt = c_vcnt;
			w.put_int(cnt);
			for (int32_t k = 0; k < cnt; ++k) {
				const uint64_t q = c_vfirst + (uint64_t)k;
				w.put_int(r.acl_perms[q]);
				if (k < 3) {
					w.put_buf(r.arena, r.acl_scheme_off[q], r.acl_scheme_len[q]);
					w.put_buf(r.arena, r.acl_id_off[q], r.acl_id_len[q]);
				} else {
					w.put_buf_inline(r.arena, r.acl_scheme_off[q], r.acl_scheme_len[q]);
					w.put_buf_inline(r.arena, r.acl_id_off[q], r.acl_id_len[q]);
				}
			}
		}
		if (f_byte) w.put_byte(c_a0 ? 1 : 0);
		if (f_int) w.put_int(c_a0);
		__syncwarp();
		/* the copies of these 32 packets: short ones one per lane, long ones taken by four groups of
		 * eight lanes until none is left */
		{
			const uint32_t ns = sm.nshort < ENC_JOBS / 2 ? sm.nshort : ENC_JOBS / 2;
			const uint32_t nl = sm.nlong < ENC_JOBS / 2 ? sm.nlong : ENC_JOBS / 2;
			for (uint32_t j = lane; j < ns; j += 32) enc_copy_short(sm, r.arena, r.arena_len, sm.job[j]);
			const uint32_t gl = lane & 7u;
			const uint32_t gmask = 0xffu << (lane & 24u);      /* the groups leave the loop at different times */
			for (;;) {
				uint32_t j = 0;
				if (gl == 0) j = atomicAdd(&sm.next, 1u);
				j = __shfl_sync(gmask, j, lane & 24u);
				if (j >= nl) break;
				enc_copy_job(sm, r.arena, r.arena_len, sm.job[ENC_JOBS - 1 - j], gl);
			}
		}
		__syncwarp();
		if (!__any_sync(0xffffffffu, valid && start < s1 && lane == 31)) break;   /* the last lane was past the span */
	}
	__syncwarp();
	/* window -> HBM */
	const uint32_t L = (uint32_t)(s1 - s0);
	uint8_t *dst = a.out + s0;
	const uint8_t *win = sm.win;
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
