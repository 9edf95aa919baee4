/*
 * zk_decode.cu — batch frame splitter + field decoder for ZooKeeper jute streams, sm_100a.
 *
 * Replaces, for a whole batch at once, the reference's per-frame loop
 *   ZKDecodeStream._transform / handlePacket   lib/zk-streams.js:39-100
 *   ZKBuffer.readResponse / readRequest / readConnect*   lib/zk-buffer.js:22-95,281-442
 *   JuteBuffer.read*                            lib/jute-buffer.js:46-109
 *
 * Pipeline (all on the context's stream):
 *   k_split<SPEC>   one thread per splitter granule walks the length-prefix chain
 *                   (zk-streams.js:47-64) from a *guessed* entry point.  The walk only
 *                   touches the 4-byte prefixes, so it reads a few % of the batch.
 *   k_scan_*        a prefix scan over granules carries every granule's exit position to
 *                   its successors and sums frame counts; a granule whose guess disagrees
 *                   with what its predecessors say is marked dirty.
 *   k_split<FIX>    dirty granules re-walk from the carried position; repeat until no
 *                   granule is dirty.  By induction from granule 0 (whose entry is a
 *                   session start) the result is then exactly the sequential split.
 *   k_split<EMIT>   writes frame_off / frame_len / session rows and the per-session
 *                   consumed / bad_length_at / first_frame results.
 *   k_decode        one CTA per 32 KiB tile: stages the tile in shared memory with 16-byte
 *                   coalesced loads, one thread per frame parses the schema chosen by
 *                   kind/opcode, writes the FRAMES columns, and places STAT/BUF/ACL rows
 *                   through a decoupled look-back over per-tile row counts, so the input
 *                   is read from HBM once and the side tables come out dense and ordered.
 */
#include "zk_internal.h"

namespace {

/* ================================================================================== */
/* small helpers                                                                      */

__device__ __forceinline__ uint32_t bswap32(uint32_t v) { return __byte_perm(v, 0, 0x0123); }

/* big-endian int32 at an arbitrary byte position of a 4-byte-aligned global buffer */
__device__ __forceinline__ uint32_t g_be32(const uint8_t *base, uint64_t pos)
{
	const uint32_t *w = reinterpret_cast<const uint32_t *>(base + (pos & ~3ull));
	uint32_t sh = (uint32_t)(pos & 3) * 8;
	uint32_t lo = __ldg(w);
	uint32_t v = lo;
	if (sh) v = __funnelshift_r(lo, __ldg(w + 1), sh);
	return bswap32(v);
}

__device__ __forceinline__ bool len_ok(int32_t len) { return len >= 0 && len <= ZKB_MAX_PACKET; }

/* ERR_LOOKUP has an entry (lib/zk-consts.js:26-47) — only used to rank split guesses */
__device__ __forceinline__ bool err_known(int32_t e)
{
	return (e <= 0 && e >= -8) || (e <= -100 && e >= -103) || e == -108 || (e <= -110 && e >= -115);
}

/* OP_CODE_LOOKUP has an entry (lib/zk-consts.js:84-109) */
__device__ __forceinline__ bool op_known(int32_t v)
{
	return (v >= 0 && v <= 14 && v != 10) || (v >= 100 && v <= 102) || v == -10 || v == -11 || v == -1;
}

__device__ __forceinline__ uint4 ld_stream16(const uint4 *p)
{
	uint4 r;
	asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
	    : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
	return r;
}

__device__ __forceinline__ uint32_t ld_acquire(const uint32_t *p)
{
	uint32_t v;
	asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
	return v;
}
__device__ __forceinline__ void st_release(uint32_t *p, uint32_t v)
{
	asm volatile("st.release.gpu.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}

/* ================================================================================== */
/* splitter                                                                           */

enum { MODE_SPEC = 0, MODE_FIX = 1, MODE_EMIT = 2 };

struct SplitArgs {
	const uint8_t  *bytes;
	uint64_t        nbytes;
	const uint64_t *sess_off;
	uint32_t        nsess;
	const uint8_t  *sess_flags;
	uint32_t        glog2;
	uint32_t        n_gran;
	uint64_t       *g_assumed;   /* entry position the last walk of the granule assumed */
	uint64_t       *g_out;       /* exit: next frame start >= granule end | NONE | INHERIT */
	uint64_t       *g_carry;     /* scan result: exit of the nearest predecessor that had one */
	uint32_t       *g_cnt;       /* frames starting in the granule */
	uint32_t       *g_base;      /* exclusive prefix of g_cnt */
	uint8_t        *g_dirty;
	/* EMIT */
	uint64_t       *frame_off;
	int32_t        *frame_len;
	uint32_t       *session;
	uint64_t        cap_frames;
	uint64_t       *sess_consumed;
	uint64_t       *sess_bad;
	uint32_t       *sess_first;
	zk_ctrl        *ctrl;
};

/* first index in off[0..n) with off[idx] >= key (n = nsess + 1, off[n-1] = nbytes > key) */
__device__ __forceinline__ uint32_t lower_bound_u64(const uint64_t *off, uint32_t n, uint64_t key)
{
	uint32_t lo = 0, hi = n;
	while (lo < hi) {
		uint32_t mid = (lo + hi) >> 1;
		if (__ldg(off + mid) < key) lo = mid + 1; else hi = mid;
	}
	return lo;
}

/* Does a frame header that makes sense for this direction start at p?  Ranking only:
 * a wrong answer costs a FIX iteration, never a wrong result. */
__device__ bool header_plausible(const uint8_t *bytes, uint64_t p, uint64_t send, bool server,
    uint64_t &next)
{
	if (p + 4 > send) return false;
	uint32_t b0 = __ldg(bytes + p);
	if (b0 > 1) return false;                       /* len <= 16 MiB => first byte 0 or 1 */
	int32_t len = (int32_t)g_be32(bytes, p);
	if (!len_ok(len)) return false;
	if (p + 4 + (uint64_t)len > send) return false;
	if (server) {
		if (len < 8) return false;
		if (!op_known((int32_t)g_be32(bytes, p + 8))) return false;
	} else {
		if (len < 16) return false;
		if (!err_known((int32_t)g_be32(bytes, p + 16))) return false;
	}
	next = p + 4 + (uint64_t)len;
	return true;
}

__device__ uint64_t guess_entry(const SplitArgs &a, uint32_t s, uint64_t gs, uint64_t ge, uint64_t send)
{
	bool server = a.sess_flags && (a.sess_flags[s] & ZKB_SESS_SERVER);
	uint64_t lim = ge < send ? ge : send;
	for (uint64_t p = gs; p < lim; ++p) {
		uint64_t next;
		if (!header_plausible(a.bytes, p, send, server, next)) continue;
		if (next == send || next + 4 > send) return p;
		if (len_ok((int32_t)g_be32(a.bytes, next))) return p;
	}
	return ZK_NONE;
}

/* The loop of ZKDecodeStream._transform (zk-streams.js:47-64) restricted to the frames
 * that start before `ge`.  Returns where the chain stands afterwards. */
template <int MODE>
__device__ uint64_t walk_chain(const SplitArgs &a, uint32_t s, uint64_t pos, uint64_t send,
    uint64_t ge, uint32_t &count, uint32_t base)
{
	for (;;) {
		uint64_t bad = ZK_NONE;
		bool stop = false;
		int32_t len = 0;
		if (pos == send) {
			stop = true;
		} else {
			if (pos >= ge) return pos;
			uint64_t rem = send - pos;
			if (rem < 4) {
				stop = true;                                /* zd_used < 4 */
			} else {
				len = (int32_t)g_be32(a.bytes, pos);
				if (!len_ok(len)) { stop = true; bad = pos; } /* :49-53 BAD_LENGTH */
				else if (rem < 4 + (uint64_t)len) stop = true;   /* frame not complete yet */
			}
		}
		if (stop) {
			if (MODE == MODE_EMIT) {
				a.sess_consumed[s] = pos;
				a.sess_bad[s] = bad;
			}
			return ZK_NONE;
		}
		if (MODE == MODE_EMIT) {
			uint64_t row = (uint64_t)base + count;
			if (row < a.cap_frames) {
				a.frame_off[row] = pos;
				a.frame_len[row] = len;
				a.session[row] = s;
			}
		}
		count++;
		pos += 4 + (uint64_t)len;
	}
}

template <int MODE>
__global__ void __launch_bounds__(128) k_split(SplitArgs a)
{
	uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
	if (g >= a.n_gran) return;
	if (MODE == MODE_FIX && !a.g_dirty[g]) return;
	const uint64_t gs = (uint64_t)g << a.glog2;
	const bool last = (g == a.n_gran - 1);
	const uint64_t ge = last ? a.nbytes : gs + (1ull << a.glog2);

	uint32_t sf = lower_bound_u64(a.sess_off, a.nsess + 1, gs);
	if (sf > a.nsess) return;   /* offsets do not cover the batch: k_check_sessions reports it */
	bool cont = (__ldg(a.sess_off + sf) != gs);   /* granule starts inside session sf-1 */
	uint64_t assumed = ZK_NONE, pos = ZK_NONE;
	uint32_t count = 0;
	bool touched = false;
	const uint32_t base = (MODE == MODE_EMIT) ? a.g_base[g] : 0;

	if (cont) {
		uint32_t s = sf - 1;
		uint64_t send = __ldg(a.sess_off + sf);
		assumed = (MODE == MODE_SPEC) ? guess_entry(a, s, gs, ge, send) : a.g_carry[g];
		if (assumed >= gs && assumed < ge) {
			touched = true;
			pos = walk_chain<MODE>(a, s, assumed, send, ge, count, base);
		}
	}
	for (uint32_t s = sf; s < a.nsess; ++s) {
		uint64_t so = __ldg(a.sess_off + s);
		if (!(so < ge || last)) break;
		touched = true;
		if (MODE == MODE_EMIT) a.sess_first[s] = base + count;
		pos = walk_chain<MODE>(a, s, so, __ldg(a.sess_off + s + 1), ge, count, base);
	}
	if (MODE == MODE_EMIT) {
		if (count != a.g_cnt[g]) atomicOr(&a.ctrl->error, ZK_DERR_COUNT);
	} else {
		a.g_assumed[g] = assumed;
		a.g_out[g] = touched ? pos : ZK_INHERIT;
		a.g_cnt[g] = count;
	}
}

/* ---- scan over granules: value = (last exit that is not INHERIT, sum of counts) ------ */

struct ScanVal { uint64_t v; uint32_t c; uint32_t pad; };

__device__ __forceinline__ ScanVal sv_comb(ScanVal l, ScanVal r)
{
	ScanVal o;
	o.v = (r.v != ZK_INHERIT) ? r.v : l.v;
	o.c = l.c + r.c;
	o.pad = 0;
	return o;
}
__device__ __forceinline__ ScanVal sv_shfl_up(ScanVal x, int d)
{
	ScanVal o;
	o.v = __shfl_up_sync(0xffffffffu, x.v, d);
	o.c = __shfl_up_sync(0xffffffffu, x.c, d);
	o.pad = 0;
	return o;
}

#define SCAN_THREADS 1024

/* inclusive scan across the block; returns this thread's inclusive value, *total = block sum */
__device__ ScanVal block_scan_incl(ScanVal x, ScanVal *total, ScanVal *s_w /* [32] */)
{
	const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
	for (int d = 1; d < 32; d <<= 1) {
		ScanVal y = sv_shfl_up(x, d);
		if (lane >= d) x = sv_comb(y, x);
	}
	if (lane == 31) s_w[wid] = x;
	__syncthreads();
	ScanVal pre; pre.v = ZK_INHERIT; pre.c = 0; pre.pad = 0;
	ScanVal tot = pre;
	const int nw = blockDim.x >> 5;
	for (int w = 0; w < nw; ++w) {
		if (w == wid) pre = tot;
		tot = sv_comb(tot, s_w[w]);
	}
	__syncthreads();
	*total = tot;
	return sv_comb(pre, x);
}

__global__ void __launch_bounds__(SCAN_THREADS) k_scan_reduce(SplitArgs a, ScanVal *blk)
{
	__shared__ ScanVal s_w[32];
	uint32_t g = blockIdx.x * SCAN_THREADS + threadIdx.x;
	ScanVal x; x.v = ZK_INHERIT; x.c = 0; x.pad = 0;
	if (g < a.n_gran) { x.v = a.g_out[g]; x.c = a.g_cnt[g]; }
	ScanVal tot;
	block_scan_incl(x, &tot, s_w);
	if (threadIdx.x == 0) blk[blockIdx.x] = tot;
}

/* exclusive scan of the block aggregates, in place, by one block */
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_blocks(ScanVal *blk, uint32_t nblk)
{
	__shared__ ScanVal s_w[32];
	ScanVal carry; carry.v = ZK_INHERIT; carry.c = 0; carry.pad = 0;
	for (uint32_t b0 = 0; b0 < nblk; b0 += SCAN_THREADS) {
		uint32_t i = b0 + threadIdx.x;
		ScanVal x; x.v = ZK_INHERIT; x.c = 0; x.pad = 0;
		if (i < nblk) x = blk[i];
		ScanVal tot;
		ScanVal inc = block_scan_incl(x, &tot, s_w);
		/* exclusive = carry (+) inclusive of the previous element */
		ScanVal prev = sv_shfl_up(inc, 1);
		__shared__ ScanVal s_last[32];
		if ((threadIdx.x & 31) == 31) s_last[threadIdx.x >> 5] = inc;
		__syncthreads();
		ScanVal ex;
		if (threadIdx.x == 0) { ex.v = ZK_INHERIT; ex.c = 0; ex.pad = 0; }
		else if ((threadIdx.x & 31) == 0) ex = s_last[(threadIdx.x >> 5) - 1];
		else ex = prev;
		if (i < nblk) blk[i] = sv_comb(carry, ex);
		carry = sv_comb(carry, tot);
		__syncthreads();
	}
}

__global__ void __launch_bounds__(SCAN_THREADS) k_scan_apply(SplitArgs a, const ScanVal *blk)
{
	__shared__ ScanVal s_w[32];
	__shared__ ScanVal s_last[32];
	uint32_t g = blockIdx.x * SCAN_THREADS + threadIdx.x;
	ScanVal x; x.v = ZK_INHERIT; x.c = 0; x.pad = 0;
	if (g < a.n_gran) { x.v = a.g_out[g]; x.c = a.g_cnt[g]; }
	ScanVal tot;
	ScanVal inc = block_scan_incl(x, &tot, s_w);
	ScanVal prev = sv_shfl_up(inc, 1);
	if ((threadIdx.x & 31) == 31) s_last[threadIdx.x >> 5] = inc;
	__syncthreads();
	ScanVal ex;
	if (threadIdx.x == 0) { ex.v = ZK_INHERIT; ex.c = 0; ex.pad = 0; }
	else if ((threadIdx.x & 31) == 0) ex = s_last[(threadIdx.x >> 5) - 1];
	else ex = prev;
	ex = sv_comb(blk[blockIdx.x], ex);
	if (g >= a.n_gran) return;
	const uint64_t gs = (uint64_t)g << a.glog2;
	const uint64_t ge = (g == a.n_gran - 1) ? a.nbytes : gs + (1ull << a.glog2);
	uint64_t carry = (ex.v == ZK_INHERIT) ? ZK_NONE : ex.v;
	a.g_carry[g] = carry;
	a.g_base[g] = ex.c;
	/* consistent iff both name the same in-granule position, or neither names one */
	uint64_t as = a.g_assumed[g];
	uint64_t na = (as >= gs && as < ge) ? as : ZK_NONE;
	uint64_t nc = (carry >= gs && carry < ge) ? carry : ZK_NONE;
	uint8_t dirty = (na != nc);
	a.g_dirty[g] = dirty;
	if (dirty) atomicAdd(&a.ctrl->n_dirty, 1u);
	if (g == a.n_gran - 1) a.ctrl->n_frames = (unsigned long long)ex.c + x.c;
}

/* checks off[0] == 0, off[n] == nbytes, monotone */
__global__ void k_check_sessions(const uint64_t *off, uint32_t nsess, uint64_t nbytes, zk_ctrl *ctrl)
{
	uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i > nsess) return;
	bool bad = false;
	if (i == 0 && off[0] != 0) bad = true;
	if (i == nsess && off[i] != nbytes) bad = true;
	if (i < nsess && off[i] > off[i + 1]) bad = true;
	if (bad) atomicOr(&ctrl->error, ZK_DERR_SESSIONS);
}

/* nbytes == 0: every session is empty */
__global__ void k_empty_sessions(uint32_t nsess, uint64_t *consumed, uint64_t *bad, uint32_t *first)
{
	uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < nsess) { consumed[i] = 0; bad[i] = ZK_NONE; }
	if (i <= nsess) first[i] = 0;
}

/* ================================================================================== */
/* field decode                                                                       */

struct DecArgs {
	const uint8_t  *bytes;
	uint64_t        nbytes;
	uint64_t        nbytes_pad;   /* nbytes rounded up to 16: readable extent */
	const uint8_t  *sess_flags;
	const uint32_t *sess_hs;
	const int32_t  *xt_base;
	const uint32_t *xt_count;
	const uint64_t *xt_off;
	const int8_t   *xt_ops;
	const uint32_t *g_base;
	uint32_t        glog2;
	uint32_t        n_gran;
	uint32_t        n_tiles;
	uint64_t        n_frames;
	zkb_decoded     o;
	uint32_t       *lb_flag;      /* [n_tiles] */
	uint32_t       *lb_agg;       /* [n_tiles][4] */
	uint32_t       *lb_pfx;       /* [n_tiles][4] */
	zk_ctrl        *ctrl;
};

/* Byte source: the staged tile when the bytes are in the window, global memory otherwise
 * (only frames longer than the overhang ever take the second branch). */
struct Rd {
	const uint8_t *sm;
	uint64_t       lo;
	uint32_t       wlen;
	const uint8_t *g;
	__device__ __forceinline__ uint32_t be32(uint64_t pos) const
	{
		uint64_t r = pos - lo;
		if (r + 4 <= (uint64_t)wlen) {
			const uint32_t *w = reinterpret_cast<const uint32_t *>(sm + ((uint32_t)r & ~3u));
			return bswap32(__funnelshift_r(w[0], w[1], ((uint32_t)r & 3u) * 8));
		}
		return g_be32(g, pos);
	}
	__device__ __forceinline__ uint32_t u8(uint64_t pos) const
	{
		uint64_t r = pos - lo;
		if (r < (uint64_t)wlen) return sm[(uint32_t)r];
		return __ldg(g + pos);
	}
	__device__ __forceinline__ int64_t be64(uint64_t pos) const
	{
		return (int64_t)(((uint64_t)be32(pos) << 32) | (uint64_t)be32(pos + 4));
	}
};

/* JuteBuffer cursor (lib/jute-buffer.js:14-21): jb_offset = cur, jb_size = size */
struct Cur {
	const Rd *rd;
	uint64_t  pay;    /* absolute position of payload byte 0 */
	uint32_t  size;
	uint32_t  cur;
	uint32_t  fail;   /* detail byte of the first exception, 0 = none */
};

#define DETAIL(cause, field) ((uint32_t)(cause) | ((uint32_t)(field) << 3))

/* readInt (jute-buffer.js:58-62): RangeError past the end */
__device__ __forceinline__ bool c_int(Cur &c, int field, int32_t &v)
{
	if (c.cur + 4 > c.size) { c.fail = DETAIL(ZKB_CAUSE_RANGE, field); return false; }
	v = (int32_t)c.rd->be32(c.pay + c.cur);
	c.cur += 4;
	return true;
}
/* readLongBuffer (jute-buffer.js:64-69): assert.ok(off + 8 <= size) */
__device__ __forceinline__ bool c_long(Cur &c, int field, int64_t &v)
{
	if (c.cur + 8 > c.size) { c.fail = DETAIL(ZKB_CAUSE_ASSERT, field); return false; }
	v = c.rd->be64(c.pay + c.cur);
	c.cur += 8;
	return true;
}
/* readBool (jute-buffer.js:51-56) */
__device__ __forceinline__ bool c_bool(Cur &c, int field, int32_t &v)
{
	if (!(c.cur < c.size)) { c.fail = DETAIL(ZKB_CAUSE_ASSERT, field); return false; }
	uint32_t b = c.rd->u8(c.pay + c.cur);
	c.cur += 1;
	if (b > 1) { c.fail = DETAIL(ZKB_CAUSE_ASSERT, field); return false; }
	v = (int32_t)b;
	return true;
}
/* readBuffer / readUString (jute-buffer.js:97-109): negative length reads as empty */
__device__ __forceinline__ bool c_buf(Cur &c, int field, uint32_t &pos, int32_t &len)
{
	int32_t l;
	if (!c_int(c, field, l)) return false;
	if (l < 0) l = 0;
	if (c.cur + (uint32_t)l > c.size) { c.fail = DETAIL(ZKB_CAUSE_ASSERT, field); return false; }
	pos = c.cur;
	len = l;
	c.cur += (uint32_t)l;
	return true;
}
/* readStat (zk-buffer.js:428-442): 4 longs, 3 ints, long, 2 ints, long = 68 bytes */
__device__ __forceinline__ bool c_stat(Cur &c, uint32_t &pos)
{
	pos = c.cur;
	if (c.cur + 68 <= c.size) { c.cur += 68; return true; }
	/* find the first field that does not fit: longs assert, ints throw RangeError */
	const uint32_t is_long = 0x48Fu;   /* bit i set: field i is a long (0-3, 7, 10) */
	uint32_t p = c.cur;
	for (int i = 0; i < 11; ++i) {
		uint32_t w = ((is_long >> i) & 1) ? 8u : 4u;
		if (p + w > c.size) {
			c.fail = DETAIL(w == 8 ? ZKB_CAUSE_ASSERT : ZKB_CAUSE_RANGE, ZKB_F_STAT);
			return false;
		}
		p += w;
	}
	return false;
}

/* what one frame contributes; small enough to live in registers across the look-back */
struct Parsed {
	uint32_t meta;
	int32_t  xid, err, a0, a1;
	int64_t  zxid;
	uint32_t n_stat, n_buf, n_acl;
	uint32_t stat_pos;
	uint32_t fb0_pos, fb1_pos;   /* up to two scalar buffer fields (path, data | passwd) */
	int32_t  fb0_len, fb1_len;
	uint32_t n_fixed;
	uint32_t vec_pos;
	uint32_t vec_kind;     /* VEC_* */
};
enum { VEC_NONE = 0, VEC_STR1 = 1, VEC_ACL = 2, VEC_STR3 = 3 };

struct SideRows {
	zkb_decoded o;
	uint32_t frame;
	uint32_t buf_row, acl_row;
};

/* vector<ustring> (zk-buffer.js:334-337,238-252).  EMIT = false: validate and count. */
template <bool EMIT>
__device__ bool walk_strvec(Cur &c, int32_t &n_out, uint32_t &n_buf, SideRows *sr)
{
	int32_t count;
	if (!c_int(c, ZKB_F_COUNT, count)) return false;
	n_out = count > 0 ? count : 0;
	for (int32_t i = 0; i < count; ++i) {
		uint32_t pos; int32_t len;
		if (!c_buf(c, ZKB_F_ELEM, pos, len)) return false;
		if (EMIT) {
			uint64_t row = sr->buf_row++;
			if (row < sr->o.cap_buf) {
				sr->o.buf_frame[row] = sr->frame;
				sr->o.buf_off[row] = c.pay + pos;
				sr->o.buf_len[row] = len;
			}
		}
		n_buf++;
	}
	return true;
}

/* readACL (zk-buffer.js:372-382): count x { perms, scheme, id } */
template <bool EMIT>
__device__ bool walk_acl(Cur &c, int32_t &n_out, uint32_t &n_buf, uint32_t &n_acl, SideRows *sr)
{
	int32_t count;
	if (!c_int(c, ZKB_F_COUNT, count)) return false;
	n_out = count > 0 ? count : 0;
	for (int32_t i = 0; i < count; ++i) {
		int32_t perms; uint32_t p0, p1; int32_t l0, l1;
		if (!c_int(c, ZKB_F_ACL_PERMS, perms)) return false;
		if (!c_buf(c, ZKB_F_ACL_SCHEME, p0, l0)) return false;
		if (!c_buf(c, ZKB_F_ACL_ID, p1, l1)) return false;
		if (EMIT) {
			uint64_t ar = sr->acl_row++;
			if (ar < sr->o.cap_acl) { sr->o.acl_frame[ar] = sr->frame; sr->o.acl_perms[ar] = perms; }
			uint64_t row = sr->buf_row; sr->buf_row += 2;
			if (row < sr->o.cap_buf) {
				sr->o.buf_frame[row] = sr->frame; sr->o.buf_off[row] = c.pay + p0; sr->o.buf_len[row] = l0;
			}
			if (row + 1 < sr->o.cap_buf) {
				sr->o.buf_frame[row + 1] = sr->frame; sr->o.buf_off[row + 1] = c.pay + p1; sr->o.buf_len[row + 1] = l1;
			}
		}
		n_buf += 2;
		n_acl++;
	}
	return true;
}

/* SPECIAL_XIDS (zk-buffer.js:275-279) */
__device__ __forceinline__ int special_xid(int32_t xid)
{
	switch (xid) {
	case ZKB_XID_NOTIFICATION: return ZKB_OP_NOTIFICATION;
	case ZKB_XID_PING: return ZKB_OP_PING;
	case ZKB_XID_AUTH: return ZKB_OP_AUTH;
	case ZKB_XID_SET_WATCHES: return ZKB_OP_SET_WATCHES;
	default: return ZKB_OP_NONE;
	}
}

__device__ __forceinline__ bool fixed_buf(Cur &c, int field, Parsed &p)
{
	uint32_t pos; int32_t len;
	if (!c_buf(c, field, pos, len)) return false;
	if (p.n_fixed == 0) { p.fb0_pos = pos; p.fb0_len = len; }
	else { p.fb1_pos = pos; p.fb1_len = len; }
	p.n_fixed++;
	p.n_buf++;
	return true;
}

/* readResponse (zk-buffer.js:281-331) */
__device__ void parse_response(Cur &c, Parsed &p, int &opcode, const DecArgs &a, uint32_t s)
{
	if (!c_int(c, ZKB_F_XID, p.xid)) return;
	if (!c_long(c, ZKB_F_ZXID, p.zxid)) return;
	if (!c_int(c, ZKB_F_ERR, p.err)) return;
	int op = special_xid(p.xid);                   /* :288 specials win */
	if (op == ZKB_OP_NONE && a.xt_ops) {           /* :289-290 xidMap[pkt.xid] */
		int64_t idx = (int64_t)p.xid - (int64_t)__ldg(a.xt_base + s);
		if (idx >= 0 && idx < (int64_t)__ldg(a.xt_count + s))
			op = __ldg(a.xt_ops + __ldg(a.xt_off + s) + (uint64_t)idx);
	}
	opcode = op;
	if (op == ZKB_OP_NONE) { c.fail = DETAIL(ZKB_CAUSE_NO_MATCH, ZKB_F_OPCODE); return; } /* :291 */
	if (p.err != 0) return;                        /* :292 only 'OK' replies have a body */
	switch (op) {
	case ZKB_OP_GET_CHILDREN:
	case ZKB_OP_GET_CHILDREN2:
		p.vec_kind = VEC_STR1; p.vec_pos = c.cur;
		if (!walk_strvec<false>(c, p.a0, p.n_buf, nullptr)) return;
		if (op == ZKB_OP_GET_CHILDREN2) { if (!c_stat(c, p.stat_pos)) return; p.n_stat = 1; }
		break;
	case ZKB_OP_CREATE:
		fixed_buf(c, ZKB_F_PATH, p);
		break;
	case ZKB_OP_GET_ACL:
		p.vec_kind = VEC_ACL; p.vec_pos = c.cur;
		if (!walk_acl<false>(c, p.a0, p.n_buf, p.n_acl, nullptr)) return;
		if (!c_stat(c, p.stat_pos)) return;
		p.n_stat = 1;
		break;
	case ZKB_OP_GET_DATA:
		if (!fixed_buf(c, ZKB_F_DATA, p)) return;
		if (!c_stat(c, p.stat_pos)) return;
		p.n_stat = 1;
		break;
	case ZKB_OP_NOTIFICATION:
		if (!c_int(c, ZKB_F_NTYPE, p.a0)) return;
		if (!c_int(c, ZKB_F_NSTATE, p.a1)) return;
		fixed_buf(c, ZKB_F_PATH, p);
		break;
	case ZKB_OP_EXISTS:
	case ZKB_OP_SET_DATA:
		if (!c_stat(c, p.stat_pos)) return;
		p.n_stat = 1;
		break;
	case ZKB_OP_SET_WATCHES:
	case ZKB_OP_PING:
	case ZKB_OP_SYNC:
	case ZKB_OP_DELETE:
	case ZKB_OP_CLOSE_SESSION:
		break;
	default:
		c.fail = DETAIL(ZKB_CAUSE_UNSUPPORTED, ZKB_F_OPCODE);   /* :326-327 */
		break;
	}
}

/* readRequest (zk-buffer.js:58-95,138-253) */
__device__ void parse_request(Cur &c, Parsed &p, int &opcode)
{
	int32_t type;
	if (!c_int(c, ZKB_F_XID, p.xid)) return;
	if (!c_int(c, ZKB_F_OPCODE, type)) return;
	int op = op_known(type) ? type : ZKB_OP_NONE;
	opcode = op;
	switch (op) {
	case ZKB_OP_GET_CHILDREN:
	case ZKB_OP_GET_CHILDREN2:
	case ZKB_OP_GET_DATA:
	case ZKB_OP_EXISTS:
		if (!fixed_buf(c, ZKB_F_PATH, p)) return;
		c_bool(c, ZKB_F_WATCH, p.a0);
		break;
	case ZKB_OP_CREATE:
		if (!fixed_buf(c, ZKB_F_PATH, p)) return;
		if (!fixed_buf(c, ZKB_F_DATA, p)) return;
		p.vec_kind = VEC_ACL; p.vec_pos = c.cur;
		if (!walk_acl<false>(c, p.a1, p.n_buf, p.n_acl, nullptr)) return;
		c_int(c, ZKB_F_FLAGS, p.a0);
		break;
	case ZKB_OP_DELETE:
		if (!fixed_buf(c, ZKB_F_PATH, p)) return;
		c_int(c, ZKB_F_VERSION, p.a0);
		break;
	case ZKB_OP_GET_ACL:
	case ZKB_OP_SYNC:
		fixed_buf(c, ZKB_F_PATH, p);
		break;
	case ZKB_OP_SET_DATA:
		if (!fixed_buf(c, ZKB_F_PATH, p)) return;
		if (!fixed_buf(c, ZKB_F_DATA, p)) return;
		c_int(c, ZKB_F_VERSION, p.a0);
		break;
	case ZKB_OP_SET_WATCHES:
		if (!c_long(c, ZKB_F_REL_ZXID, p.zxid)) return;
		p.vec_kind = VEC_STR3; p.vec_pos = c.cur;
		if (!walk_strvec<false>(c, p.a0, p.n_buf, nullptr)) return;
		if (!walk_strvec<false>(c, p.a1, p.n_buf, nullptr)) return;
		walk_strvec<false>(c, p.err, p.n_buf, nullptr);
		break;
	default:
		c.fail = DETAIL(ZKB_CAUSE_UNSUPPORTED, ZKB_F_OPCODE);   /* :91-92 */
		break;
	}
}

/* handlePacket (zk-streams.js:66-99) for one frame; no side rows are written here */
__device__ void parse_frame(const Rd &rd, const DecArgs &a, uint64_t foff, int32_t flen,
    uint32_t s, uint32_t idx_in_session, Parsed &p)
{
	Cur c;
	c.rd = &rd; c.pay = foff + 4; c.size = (uint32_t)flen; c.cur = 0; c.fail = 0;
	p.xid = p.err = p.a0 = p.a1 = 0; p.zxid = 0;
	p.n_stat = p.n_buf = p.n_acl = 0; p.n_fixed = 0; p.vec_kind = VEC_NONE;
	p.stat_pos = 0; p.vec_pos = 0;
	p.fb0_pos = p.fb1_pos = 0; p.fb0_len = p.fb1_len = 0;
	const bool server = a.sess_flags && (__ldg(a.sess_flags + s) & ZKB_SESS_SERVER);
	const uint32_t hs = a.sess_hs ? __ldg(a.sess_hs + s) : 1u;
	int opcode = ZKB_OP_NONE;
	uint32_t kind;
	if (idx_in_session < hs) {
		if (server) {                              /* readConnectRequest zk-buffer.js:22-30 */
			kind = ZKB_KIND_CONNECT_REQUEST;
			int64_t last;
			if (c_int(c, ZKB_F_PROTOCOL_VERSION, p.xid) && c_long(c, ZKB_F_LAST_ZXID, last)) {
				p.a0 = (int32_t)((uint64_t)last >> 32);
				p.a1 = (int32_t)((uint64_t)last & 0xffffffffu);
				if (c_int(c, ZKB_F_TIMEOUT, p.err) && c_long(c, ZKB_F_SESSION_ID, p.zxid))
					fixed_buf(c, ZKB_F_PASSWD, p);
			}
		} else {                                   /* readConnectResponse zk-buffer.js:41-48 */
			kind = ZKB_KIND_CONNECT_RESPONSE;
			if (c_int(c, ZKB_F_PROTOCOL_VERSION, p.xid) && c_int(c, ZKB_F_TIMEOUT, p.err) &&
			    c_long(c, ZKB_F_SESSION_ID, p.zxid))
				fixed_buf(c, ZKB_F_PASSWD, p);
		}
	} else if (server) {
		kind = ZKB_KIND_REQUEST;
		parse_request(c, p, opcode);
	} else {
		kind = ZKB_KIND_RESPONSE;
		parse_response(c, p, opcode, a, s);
	}
	uint32_t status = ZKB_ST_OK;
	if (c.fail) {
		status = ZKB_ST_BAD_DECODE;
		p.n_stat = p.n_buf = p.n_acl = 0;   /* the packet is never pushed */
	}
	p.meta = status | ((uint32_t)(uint8_t)(int8_t)opcode << 8) | (kind << 16) | (c.fail << 24);
}

/* second visit of an OK frame: write its STAT / BUF / ACL rows */
__device__ void emit_side(const Rd &rd, const DecArgs &a, uint64_t foff, int32_t flen,
    const Parsed &p, uint32_t frame, uint32_t stat_row, uint32_t buf_row, uint32_t acl_row)
{
	const uint64_t pay = foff + 4;
	const zkb_decoded &o = a.o;
	if (p.n_stat && stat_row < o.cap_stat) {
		uint64_t q = pay + p.stat_pos;
		o.stat_frame[stat_row] = frame;
		o.czxid[stat_row] = rd.be64(q);
		o.mzxid[stat_row] = rd.be64(q + 8);
		o.ctime[stat_row] = rd.be64(q + 16);
		o.mtime[stat_row] = rd.be64(q + 24);
		o.version[stat_row] = (int32_t)rd.be32(q + 32);
		o.cversion[stat_row] = (int32_t)rd.be32(q + 36);
		o.aversion[stat_row] = (int32_t)rd.be32(q + 40);
		o.ephemeralOwner[stat_row] = rd.be64(q + 44);
		o.dataLength[stat_row] = (int32_t)rd.be32(q + 52);
		o.numChildren[stat_row] = (int32_t)rd.be32(q + 56);
		o.pzxid[stat_row] = rd.be64(q + 60);
	}
	SideRows sr;
	sr.o = o; sr.frame = frame; sr.buf_row = buf_row; sr.acl_row = acl_row;
	if (p.n_fixed > 0) {
		uint64_t row = sr.buf_row++;
		if (row < o.cap_buf) {
			o.buf_frame[row] = frame; o.buf_off[row] = pay + p.fb0_pos; o.buf_len[row] = p.fb0_len;
		}
	}
	if (p.n_fixed > 1) {
		uint64_t row = sr.buf_row++;
		if (row < o.cap_buf) {
			o.buf_frame[row] = frame; o.buf_off[row] = pay + p.fb1_pos; o.buf_len[row] = p.fb1_len;
		}
	}
	if (p.vec_kind != VEC_NONE) {
		Cur c;
		c.rd = &rd; c.pay = pay; c.size = (uint32_t)flen; c.cur = p.vec_pos; c.fail = 0;
		int32_t n; uint32_t nb = 0, na = 0;
		if (p.vec_kind == VEC_STR1) {
			walk_strvec<true>(c, n, nb, &sr);
		} else if (p.vec_kind == VEC_ACL) {
			walk_acl<true>(c, n, nb, na, &sr);
		} else {
			walk_strvec<true>(c, n, nb, &sr);
			walk_strvec<true>(c, n, nb, &sr);
			walk_strvec<true>(c, n, nb, &sr);
		}
	}
}

/* exclusive scan of three counters over the block; tot[] = block totals */
__device__ void block_scan3(uint32_t v0, uint32_t v1, uint32_t v2, uint32_t &e0, uint32_t &e1,
    uint32_t &e2, uint32_t tot[3], uint32_t (*s_w)[ZK_DECODE_THREADS / 32])
{
	const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
	uint32_t i0 = v0, i1 = v1, i2 = v2;
	for (int d = 1; d < 32; d <<= 1) {
		uint32_t y0 = __shfl_up_sync(0xffffffffu, i0, d);
		uint32_t y1 = __shfl_up_sync(0xffffffffu, i1, d);
		uint32_t y2 = __shfl_up_sync(0xffffffffu, i2, d);
		if (lane >= d) { i0 += y0; i1 += y1; i2 += y2; }
	}
	if (lane == 31) { s_w[0][wid] = i0; s_w[1][wid] = i1; s_w[2][wid] = i2; }
	__syncthreads();
	uint32_t p0 = 0, p1 = 0, p2 = 0, t0 = 0, t1 = 0, t2 = 0;
	for (int w = 0; w < ZK_DECODE_THREADS / 32; ++w) {
		if (w == wid) { p0 = t0; p1 = t1; p2 = t2; }
		t0 += s_w[0][w]; t1 += s_w[1][w]; t2 += s_w[2][w];
	}
	__syncthreads();
	e0 = p0 + i0 - v0; e1 = p1 + i1 - v1; e2 = p2 + i2 - v2;
	tot[0] = t0; tot[1] = t1; tot[2] = t2;
}

#define ZK_SPIN_LIMIT (1u << 26)

__global__ void __launch_bounds__(ZK_DECODE_THREADS) k_decode(DecArgs a)
{
	__shared__ __align__(16) uint8_t s_tile[ZK_TILE_BYTES + ZK_TILE_OVERHANG + 16];
	__shared__ uint32_t s_w[3][ZK_DECODE_THREADS / 32];
	__shared__ uint32_t s_bc[4];
	const uint32_t tid = threadIdx.x;

	if (tid == 0) s_bc[0] = atomicAdd(&a.ctrl->ticket, 1u);
	__syncthreads();
	const uint32_t tile = s_bc[0];
	if (tile >= a.n_tiles) return;
	const uint64_t ts = (uint64_t)tile * ZK_TILE_BYTES;
	uint64_t avail = a.nbytes_pad - ts;
	const uint32_t wlen = avail < (uint64_t)(ZK_TILE_BYTES + ZK_TILE_OVERHANG)
	    ? (uint32_t)avail : (ZK_TILE_BYTES + ZK_TILE_OVERHANG);

	/* ---- stage the tile: 16-byte coalesced streaming loads ---- */
	{
		const uint4 *src = reinterpret_cast<const uint4 *>(a.bytes + ts);
		uint4 *dst = reinterpret_cast<uint4 *>(s_tile);
		const uint32_t nvec = wlen >> 4;
		uint32_t i = tid;
		for (; i + 3 * ZK_DECODE_THREADS < nvec; i += 4 * ZK_DECODE_THREADS) {
			uint4 v0 = ld_stream16(src + i);
			uint4 v1 = ld_stream16(src + i + ZK_DECODE_THREADS);
			uint4 v2 = ld_stream16(src + i + 2 * ZK_DECODE_THREADS);
			uint4 v3 = ld_stream16(src + i + 3 * ZK_DECODE_THREADS);
			dst[i] = v0; dst[i + ZK_DECODE_THREADS] = v1;
			dst[i + 2 * ZK_DECODE_THREADS] = v2; dst[i + 3 * ZK_DECODE_THREADS] = v3;
		}
		for (; i < nvec; i += ZK_DECODE_THREADS) dst[i] = ld_stream16(src + i);
	}
	Rd rd;
	rd.sm = s_tile; rd.lo = ts; rd.wlen = wlen; rd.g = a.bytes;

	const uint32_t g0 = (uint32_t)(ts >> a.glog2);
	const uint32_t g1 = g0 + (ZK_TILE_BYTES >> a.glog2);
	const uint32_t f0 = a.g_base[g0];
	const uint32_t f1 = (g1 < a.n_gran) ? a.g_base[g1] : (uint32_t)a.n_frames;
	const uint32_t nfr = f1 - f0;
	const bool single = nfr <= ZK_DECODE_THREADS;
	__syncthreads();

	/* ---- pass 1: parse, write FRAMES columns, count side rows ---- */
	Parsed p;
	uint64_t foff = 0; int32_t flen = 0;
	uint32_t ex0 = 0, ex1 = 0, ex2 = 0;
	uint32_t tile_tot[3] = { 0, 0, 0 };
	for (uint32_t r0 = 0; r0 < nfr || r0 == 0; r0 += ZK_DECODE_THREADS) {
		const uint32_t k = r0 + tid;
		const bool valid = k < nfr;
		p.n_stat = p.n_buf = p.n_acl = 0;
		if (valid) {
			const uint32_t f = f0 + k;
			foff = a.o.frame_off[f];
			flen = a.o.frame_len[f];
			const uint32_t s = a.o.session[f];
			parse_frame(rd, a, foff, flen, s, f - __ldg(a.o.sess_first_frame + s), p);
			a.o.meta[f] = p.meta;
			a.o.xid[f] = p.xid;
			a.o.zxid[f] = p.zxid;
			a.o.err[f] = p.err;
			a.o.a0[f] = p.a0;
			a.o.a1[f] = p.a1;
		}
		uint32_t t[3];
		block_scan3(p.n_stat, p.n_buf, p.n_acl, ex0, ex1, ex2, t, s_w);
		tile_tot[0] += t[0]; tile_tot[1] += t[1]; tile_tot[2] += t[2];
		if (nfr == 0) break;
	}

	/* ---- decoupled look-back over tiles: exclusive prefix of (stat, buf, acl) rows ---- */
	if (tid < 32) {
		uint32_t e0 = 0, e1 = 0, e2 = 0;
		if (tile > 0) {
			if (tid == 0) {
				a.lb_agg[tile * 4 + 0] = tile_tot[0];
				a.lb_agg[tile * 4 + 1] = tile_tot[1];
				a.lb_agg[tile * 4 + 2] = tile_tot[2];
				st_release(a.lb_flag + tile, ZK_LB_AGG);
			}
			int64_t j = (int64_t)tile - 1;
			uint32_t spins = 0;
			for (;;) {
				const int64_t idx = j - (int64_t)tid;
				uint32_t st = ZK_LB_PFX;
				if (idx >= 0) st = ld_acquire(a.lb_flag + idx);
				if (__any_sync(0xffffffffu, st == ZK_LB_EMPTY)) {
					if (++spins > ZK_SPIN_LIMIT) {
						if (tid == 0) atomicOr(&a.ctrl->error, ZK_DERR_LOOKBACK);
						break;
					}
					__nanosleep(64);
					continue;
				}
				const uint32_t pm = __ballot_sync(0xffffffffu, st == ZK_LB_PFX);
				const int first = __ffs(pm) - 1;   /* nearest predecessor holding a full prefix */
				uint32_t c0 = 0, c1 = 0, c2 = 0;
				if (idx >= 0 && (first < 0 || (int)tid <= first)) {
					const uint32_t *src = (first >= 0 && (int)tid == first) ? a.lb_pfx : a.lb_agg;
					c0 = __ldcg(src + idx * 4 + 0);
					c1 = __ldcg(src + idx * 4 + 1);
					c2 = __ldcg(src + idx * 4 + 2);
				}
				for (int d = 16; d > 0; d >>= 1) {
					c0 += __shfl_xor_sync(0xffffffffu, c0, d);
					c1 += __shfl_xor_sync(0xffffffffu, c1, d);
					c2 += __shfl_xor_sync(0xffffffffu, c2, d);
				}
				e0 += c0; e1 += c1; e2 += c2;
				if (first >= 0) break;
				j -= 32;
			}
		}
		if (tid == 0) {
			a.lb_pfx[tile * 4 + 0] = e0 + tile_tot[0];
			a.lb_pfx[tile * 4 + 1] = e1 + tile_tot[1];
			a.lb_pfx[tile * 4 + 2] = e2 + tile_tot[2];
			st_release(a.lb_flag + tile, ZK_LB_PFX);
			s_bc[1] = e0; s_bc[2] = e1; s_bc[3] = e2;
			if (tile == a.n_tiles - 1) {
				a.ctrl->n_stat = (unsigned long long)e0 + tile_tot[0];
				a.ctrl->n_buf = (unsigned long long)e1 + tile_tot[1];
				a.ctrl->n_acl = (unsigned long long)e2 + tile_tot[2];
			}
		}
	}
	__syncthreads();
	uint32_t run0 = s_bc[1], run1 = s_bc[2], run2 = s_bc[3];
	if (tile_tot[0] + tile_tot[1] + tile_tot[2] == 0) return;

	/* ---- pass 2: side rows ---- */
	if (single) {
		if (tid < nfr && (p.n_stat | p.n_buf | p.n_acl))
			emit_side(rd, a, foff, flen, p, f0 + tid, run0 + ex0, run1 + ex1, run2 + ex2);
		return;
	}
	for (uint32_t r0 = 0; r0 < nfr; r0 += ZK_DECODE_THREADS) {
		const uint32_t k = r0 + tid;
		const bool valid = k < nfr;
		p.n_stat = p.n_buf = p.n_acl = 0;
		if (valid) {
			const uint32_t f = f0 + k;
			foff = a.o.frame_off[f];
			flen = a.o.frame_len[f];
			const uint32_t s = a.o.session[f];
			parse_frame(rd, a, foff, flen, s, f - __ldg(a.o.sess_first_frame + s), p);
		}
		uint32_t t[3];
		block_scan3(p.n_stat, p.n_buf, p.n_acl, ex0, ex1, ex2, t, s_w);
		if (valid && (p.n_stat | p.n_buf | p.n_acl))
			emit_side(rd, a, foff, flen, p, f0 + k, run0 + ex0, run1 + ex1, run2 + ex2);
		run0 += t[0]; run1 += t[1]; run2 += t[2];
	}
}

/* ================================================================================== */
/* xid -> opcode table from a decoded request stream                                   */

__global__ void k_xid_scatter(const zkb_decoded req, uint64_t n_frames, uint32_t nsess,
    const int32_t *base, const uint32_t *count, const uint64_t *off, unsigned long long *tmp)
{
	uint64_t f = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (f >= n_frames) return;
	uint32_t m = req.meta[f];
	if (ZKB_META_KIND(m) != ZKB_KIND_REQUEST) return;
	if (req.frame_len[f] < 8) return;
	uint32_t s = req.session[f];
	if (s >= nsess) return;
	int64_t idx = (int64_t)req.xid[f] - (int64_t)base[s];
	if (idx < 0 || idx >= (int64_t)count[s]) return;
	/* zcf_xidMap[xid] = opcode, last write wins (zk-streams.js:145): keep the highest frame */
	unsigned long long key = ((f + 1) << 8) | (unsigned long long)((m >> 8) & 0xff);
	atomicMax(tmp + off[s] + (uint64_t)idx, key);
}

__global__ void k_xid_finish(const unsigned long long *tmp, uint64_t n, int8_t *ops)
{
	uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	unsigned long long k = tmp[i];
	ops[i] = k ? (int8_t)(k & 0xff) : (int8_t)ZKB_OP_NONE;
}

} /* namespace */

/* ================================================================================== */
/* host side                                                                          */

static uint32_t zk_pick_glog2(uint64_t nbytes)
{
	/* aim for ~256K granules so the splitter has enough threads, within [1 KiB, 32 KiB] */
	uint64_t want = (nbytes + 262143) / 262144;
	uint32_t l = ZK_GRAN_MIN_LOG2;
	while (l < ZK_GRAN_MAX_LOG2 && (1ull << l) < want) l++;
	return l;
}

struct ScratchLayout {
	size_t g_assumed, g_out, g_carry, g_cnt, g_base, g_dirty, blk, lb_flag, lb_agg, lb_pfx, total;
	uint32_t glog2, n_gran, n_tiles, n_blk;
};

static ScratchLayout zk_layout(uint64_t nbytes)
{
	ScratchLayout L;
	memset(&L, 0, sizeof (L));
	L.glog2 = zk_pick_glog2(nbytes);
	L.n_gran = (uint32_t)((nbytes + (1ull << L.glog2) - 1) >> L.glog2);
	L.n_tiles = (uint32_t)((nbytes + ZK_TILE_BYTES - 1) / ZK_TILE_BYTES);
	L.n_blk = (L.n_gran + SCAN_THREADS - 1) / SCAN_THREADS;
	size_t o = 0;
	L.g_assumed = o; o += zk_align_up((size_t)L.n_gran * 8, 256);
	L.g_out = o;     o += zk_align_up((size_t)L.n_gran * 8, 256);
	L.g_carry = o;   o += zk_align_up((size_t)L.n_gran * 8, 256);
	L.g_cnt = o;     o += zk_align_up((size_t)L.n_gran * 4, 256);
	L.g_base = o;    o += zk_align_up((size_t)(L.n_gran + 1) * 4, 256);
	L.g_dirty = o;   o += zk_align_up((size_t)L.n_gran, 256);
	L.blk = o;       o += zk_align_up((size_t)(L.n_blk + 1) * sizeof (ScanVal), 256);
	L.lb_flag = o;   o += zk_align_up((size_t)L.n_tiles * 4, 256);
	L.lb_agg = o;    o += zk_align_up((size_t)L.n_tiles * 16, 256);
	L.lb_pfx = o;    o += zk_align_up((size_t)L.n_tiles * 16, 256);
	L.total = o;
	return L;
}

size_t zk_decode_scratch_need(uint64_t nbytes, uint32_t nsessions)
{
	(void)nsessions;
	return zk_layout(nbytes).total + 256;
}

static int zk_read_ctrl(zkb_ctx *ctx)
{
	ZK_CUDA(ctx, cudaMemcpyAsync(ctx->h_ctrl, ctx->d_ctrl, sizeof (zk_ctrl), cudaMemcpyDeviceToHost,
	    ctx->stream));
	ZK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return ZKB_OK;
}

#define LAUNCH_CHECK(ctx)                                                         \
	do {                                                                          \
		(ctx)->launches++;                                                        \
		ZK_CUDA(ctx, cudaGetLastError());                                         \
	} while (0)

static int zk_fill_split_args(zkb_ctx *ctx, SplitArgs &sa, const ScratchLayout &L,
    const uint8_t *d_bytes, uint64_t nbytes, const zkb_sessions *sess, const zkb_decoded *out)
{
	uint8_t *S = (uint8_t *)ctx->d_scratch;
	memset(&sa, 0, sizeof (sa));
	sa.bytes = d_bytes; sa.nbytes = nbytes; sa.sess_off = sess->off; sa.nsess = sess->nsessions;
	sa.sess_flags = sess->flags; sa.glog2 = L.glog2; sa.n_gran = L.n_gran;
	sa.g_assumed = (uint64_t *)(S + L.g_assumed); sa.g_out = (uint64_t *)(S + L.g_out);
	sa.g_carry = (uint64_t *)(S + L.g_carry); sa.g_cnt = (uint32_t *)(S + L.g_cnt);
	sa.g_base = (uint32_t *)(S + L.g_base); sa.g_dirty = (uint8_t *)(S + L.g_dirty);
	if (out) {
		sa.frame_off = out->frame_off; sa.frame_len = out->frame_len; sa.session = out->session;
		sa.cap_frames = out->cap_frames;
		sa.sess_consumed = out->sess_consumed; sa.sess_bad = out->sess_bad_length_at;
		sa.sess_first = out->sess_first_frame;
	}
	sa.ctrl = ctx->d_ctrl;
	return ZKB_OK;
}

/* Phase A: find every frame boundary; on success *n_frames is exact. */
int zk_split_device(zkb_ctx *ctx, const uint8_t *d_bytes, uint64_t nbytes,
    const zkb_sessions *sess, uint64_t *n_frames)
{
	int rc;
	*n_frames = 0;
	if (!sess || !sess->off) {
		snprintf(ctx->err, sizeof (ctx->err), "zkb_decode: NULL argument");
		return ZKB_EINVAL;
	}
	if (nbytes > ZK_MAX_BATCH_BYTES) {
		snprintf(ctx->err, sizeof (ctx->err), "zkb_decode: batch of %llu bytes exceeds the %llu byte limit",
		    (unsigned long long)nbytes, (unsigned long long)ZK_MAX_BATCH_BYTES);
		return ZKB_ERANGE;
	}
	if (nbytes && ((uintptr_t)d_bytes & 15)) {
		snprintf(ctx->err, sizeof (ctx->err), "zkb_decode: bytes must be 16-byte aligned");
		return ZKB_EINVAL;
	}
	const uint32_t nsess = sess->nsessions;
	ZK_CUDA(ctx, cudaMemsetAsync(ctx->d_ctrl, 0, sizeof (zk_ctrl), ctx->stream));
	k_check_sessions<<<(nsess + 1 + 255) / 256, 256, 0, ctx->stream>>>(sess->off, nsess, nbytes,
	    ctx->d_ctrl);
	LAUNCH_CHECK(ctx);
	if (nbytes == 0) {
		if ((rc = zk_read_ctrl(ctx))) return rc;
		if (ctx->h_ctrl->error & ZK_DERR_SESSIONS) {
			snprintf(ctx->err, sizeof (ctx->err), "zkb_decode: session offsets invalid");
			return ZKB_EINVAL;
		}
		return ZKB_OK;
	}
	const ScratchLayout L = zk_layout(nbytes);
	if ((rc = zk_ensure(ctx, &ctx->d_scratch, &ctx->d_scratch_bytes, L.total + 256, 0))) return rc;
	SplitArgs sa;
	zk_fill_split_args(ctx, sa, L, d_bytes, nbytes, sess, NULL);
	ScanVal *blk = (ScanVal *)((uint8_t *)ctx->d_scratch + L.blk);

	const uint32_t split_blocks = (L.n_gran + 127) / 128;
	k_split<MODE_SPEC><<<split_blocks, 128, 0, ctx->stream>>>(sa);
	LAUNCH_CHECK(ctx);
	for (uint32_t iter = 0;; ++iter) {
		k_scan_reduce<<<L.n_blk, SCAN_THREADS, 0, ctx->stream>>>(sa, blk);
		LAUNCH_CHECK(ctx);
		k_scan_blocks<<<1, SCAN_THREADS, 0, ctx->stream>>>(blk, L.n_blk);
		LAUNCH_CHECK(ctx);
		k_scan_apply<<<L.n_blk, SCAN_THREADS, 0, ctx->stream>>>(sa, blk);
		LAUNCH_CHECK(ctx);
		if ((rc = zk_read_ctrl(ctx))) return rc;
		if (ctx->h_ctrl->error & ZK_DERR_SESSIONS) {
			snprintf(ctx->err, sizeof (ctx->err), "zkb_decode: session offsets invalid");
			return ZKB_EINVAL;
		}
		if (ctx->h_ctrl->n_dirty == 0) break;
		if (iter > L.n_gran + 2) {
			snprintf(ctx->err, sizeof (ctx->err), "zkb_decode: splitter did not converge");
			return ZKB_EINTERNAL;
		}
		ctx->split_fixups++;
		ZK_CUDA(ctx, cudaMemsetAsync(&ctx->d_ctrl->n_dirty, 0, sizeof (unsigned int), ctx->stream));
		k_split<MODE_FIX><<<split_blocks, 128, 0, ctx->stream>>>(sa);
		LAUNCH_CHECK(ctx);
	}
	*n_frames = ctx->h_ctrl->n_frames;
	return ZKB_OK;
}

/* Phase B: write the frame rows and decode every frame.  Must follow zk_split_device on
 * the same batch (it reuses the granule tables in the context's scratch). */
int zk_fields_device(zkb_ctx *ctx, const uint8_t *d_bytes, uint64_t nbytes,
    const zkb_sessions *sess, const zkb_xidtab *xid, uint64_t n_frames, zkb_decoded *out)
{
	int rc;
	const uint32_t nsess = sess->nsessions;
	out->n_frames = n_frames;
	out->n_stat = out->n_buf = out->n_acl = 0;
	if (nbytes == 0) {
		k_empty_sessions<<<(nsess + 1 + 255) / 256, 256, 0, ctx->stream>>>(nsess, out->sess_consumed,
		    out->sess_bad_length_at, out->sess_first_frame);
		LAUNCH_CHECK(ctx);
		ZK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
		return ZKB_OK;
	}
	if (n_frames > out->cap_frames) {
		snprintf(ctx->err, sizeof (ctx->err), "zkb_decode: %llu frames, capacity %llu",
		    (unsigned long long)n_frames, (unsigned long long)out->cap_frames);
		return ZKB_EOVERFLOW;
	}
	const ScratchLayout L = zk_layout(nbytes);
	uint8_t *S = (uint8_t *)ctx->d_scratch;
	SplitArgs sa;
	zk_fill_split_args(ctx, sa, L, d_bytes, nbytes, sess, out);
	const uint32_t split_blocks = (L.n_gran + 127) / 128;
	ZK_CUDA(ctx, cudaMemsetAsync(&ctx->d_ctrl->ticket, 0, sizeof (unsigned int), ctx->stream));
	k_split<MODE_EMIT><<<split_blocks, 128, 0, ctx->stream>>>(sa);
	LAUNCH_CHECK(ctx);
	/* sess_first_frame[nsess] = n_frames */
	ZK_CUDA(ctx, cudaMemcpyAsync(out->sess_first_frame + nsess, &ctx->d_ctrl->n_frames, sizeof (uint32_t),
	    cudaMemcpyDeviceToDevice, ctx->stream));

	DecArgs da;
	memset(&da, 0, sizeof (da));
	da.bytes = d_bytes; da.nbytes = nbytes; da.nbytes_pad = (nbytes + 15) & ~15ull;
	da.sess_flags = sess->flags; da.sess_hs = sess->handshake_frames;
	if (xid && xid->ops) {
		da.xt_base = xid->base; da.xt_count = xid->count; da.xt_off = xid->off; da.xt_ops = xid->ops;
	}
	da.g_base = sa.g_base; da.glog2 = L.glog2; da.n_gran = L.n_gran; da.n_tiles = L.n_tiles;
	da.n_frames = n_frames; da.o = *out;
	da.lb_flag = (uint32_t *)(S + L.lb_flag); da.lb_agg = (uint32_t *)(S + L.lb_agg);
	da.lb_pfx = (uint32_t *)(S + L.lb_pfx); da.ctrl = ctx->d_ctrl;
	ZK_CUDA(ctx, cudaMemsetAsync(da.lb_flag, 0, (size_t)L.n_tiles * 4, ctx->stream));
	k_decode<<<L.n_tiles, ZK_DECODE_THREADS, 0, ctx->stream>>>(da);
	LAUNCH_CHECK(ctx);
	if ((rc = zk_read_ctrl(ctx))) return rc;
	if (ctx->h_ctrl->error) {
		snprintf(ctx->err, sizeof (ctx->err), "zkb_decode: device error bits 0x%x", ctx->h_ctrl->error);
		return ZKB_EINTERNAL;
	}
	out->n_stat = ctx->h_ctrl->n_stat;
	out->n_buf = ctx->h_ctrl->n_buf;
	out->n_acl = ctx->h_ctrl->n_acl;
	if (out->n_stat > out->cap_stat || out->n_buf > out->cap_buf || out->n_acl > out->cap_acl) {
		snprintf(ctx->err, sizeof (ctx->err),
		    "zkb_decode: side tables need stat=%llu buf=%llu acl=%llu rows, capacity %llu/%llu/%llu",
		    (unsigned long long)out->n_stat, (unsigned long long)out->n_buf,
		    (unsigned long long)out->n_acl, (unsigned long long)out->cap_stat,
		    (unsigned long long)out->cap_buf, (unsigned long long)out->cap_acl);
		return ZKB_EOVERFLOW;
	}
	return ZKB_OK;
}

int zk_decode_device(zkb_ctx *ctx, const uint8_t *d_bytes, uint64_t nbytes,
    const zkb_sessions *sess, const zkb_xidtab *xid, zkb_decoded *out)
{
	uint64_t n_frames = 0;
	if (!out) return ZKB_EINVAL;
	int rc = zk_split_device(ctx, d_bytes, nbytes, sess, &n_frames);
	if (rc) return rc;
	return zk_fields_device(ctx, d_bytes, nbytes, sess, xid, n_frames, out);
}

int zk_xidtab_build_device(zkb_ctx *ctx, const zkb_decoded *req, uint32_t nsessions,
    const zkb_xidtab *layout, int8_t *d_ops_out)
{
	int rc;
	if (!req || !layout || !d_ops_out) return ZKB_EINVAL;
	const uint64_t n = layout->nops;
	if (n == 0) return ZKB_OK;
	if ((rc = zk_ensure(ctx, &ctx->d_scratch, &ctx->d_scratch_bytes, n * 8 + 256, 0))) return rc;
	unsigned long long *tmp = (unsigned long long *)ctx->d_scratch;
	ZK_CUDA(ctx, cudaMemsetAsync(tmp, 0, n * 8, ctx->stream));
	if (req->n_frames) {
		k_xid_scatter<<<(unsigned)((req->n_frames + 255) / 256), 256, 0, ctx->stream>>>(*req,
		    req->n_frames, nsessions, layout->base, layout->count, layout->off, tmp);
		LAUNCH_CHECK(ctx);
	}
	k_xid_finish<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(tmp, n, d_ops_out);
	LAUNCH_CHECK(ctx);
	ZK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return ZKB_OK;
}
