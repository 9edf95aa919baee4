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
 *                   (zk-streams.js:47-64) from a *guessed* entry point and records, besides
 *                   its exit and frame count, where the chain first stands inside every
 *                   256-byte sub-range of the granule.  Only frame headers are touched.
 *   k_scan_*        a prefix scan over granules carries every granule's exit position to
 *                   its successors and sums frame counts; a granule whose guess disagrees
 *                   with what its predecessors say is marked dirty.
 *   k_split<FIX>    a dirty granule with a trustworthy carried position re-walks from it and
 *                   carries the corrected chain on through the granules that follow until it
 *                   merges with what is already recorded; scan again; repeat until no granule
 *                   is dirty (one round on ordinary traffic).  By induction from each session
 *                   start the result is then exactly the sequential split.
 *   k_chunk_meta    two 16-byte records per decode chunk, gathered from the per-granule tables.
 *   k_decode        one warp per 8 KiB chunk (four per CTA): TMA bulk copy of the chunk into
 *                   shared memory; each lane walks the chain through its own 256-byte
 *                   sub-range to place the frame starts; lane k reads frames k, k+32, ... in a few
 *                   converged stages, writes all FRAMES columns and - from the first rows the
 *                   splitter predicted for the chunk plus one warp scan per round - every STAT /
 *                   BUF / ACL row where it belongs (zk_decode_direct.cuh).  The input is read from
 *                   HBM once and the side tables come out dense and ordered.
 *   k_decode_exact  the exact route for batches in which a frame broke what its header promised:
 *                   count (bitmap body), k_rows_scan, place - no chunk waits for another.
 */
#include <stdlib.h>

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


/* ---- TMA bulk copy global -> shared, completion on an mbarrier ------------------------- */
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count) : "memory");
	asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar)
{
	asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
	    :: "r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
	asm volatile(
	    "{\n"
	    ".reg .pred p;\n"
	    "ZK_WAIT_%=:\n"
	    "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
	    "@p bra ZK_DONE_%=;\n"
	    "bra ZK_WAIT_%=;\n"
	    "ZK_DONE_%=:\n"
	    "}\n" :: "r"(smem_u32(bar)), "r"(parity) : "memory");
}

/* ================================================================================== */
/* splitter                                                                           */

enum { MODE_SPEC = 0, MODE_FIX = 1 };

struct SplitArgs {
	const uint8_t  *bytes;
	uint64_t        nbytes;
	const uint64_t *sess_off;
	uint32_t        nsess;
	const uint8_t  *sess_flags;
	const uint32_t *sess_hs;
	const int32_t  *xt_base;     /* xid table of the batch (responses), NULL = none */
	const uint32_t *xt_count;
	const uint64_t *xt_off;
	const int8_t   *xt_ops;
	uint32_t        glog2;
	uint32_t        n_gran;
	/* two walk results per granule ([version * n_gran + g]): version 0 is the walk from the
	 * granule's own guess, version 1 the latest walk from a carried position; g_sel says
	 * which one the scan uses */
	uint64_t       *g_assumed;   /* entry position the walk assumed */
	uint64_t       *g_out;       /* exit: next frame start >= granule end | NONE | INHERIT */
	uint32_t       *g_cnt;       /* frames starting in the granule */
	uint4          *g_rows;      /* [version][g] { stat, buf, acl } rows those frames should produce (predict_rows) */
	uint4          *g_rbase;     /* [n_gran + 1] exclusive prefix of the selected g_rows; [n_gran] = totals */
	uint8_t        *g_sel;
	uint64_t       *g_carry;     /* scan result: exit of the nearest predecessor that had one */
	uint32_t       *g_base;      /* exclusive prefix of g_cnt */
	uint8_t        *g_dirty;
	uint32_t       *g_src;       /* 1 + index of the granule whose exit is this granule's carry; 0 = none */
	uint32_t       *g_sf;        /* first session that starts at or after the granule's first byte */
	uint32_t       *g_openlocal; /* [version][g]: frames of the granule before the first frame of the
	                                last session that starts inside it (valid if one does) */
	uint32_t       *g_ssrc;      /* 1 + index of the last granule before this one in which a session starts */
	uint16_t       *g_sub;       /* [version][g][granule/256]: where the chain first stands inside each
	                                256-byte sub-range (offset from the granule start), 0xFFFF = nowhere */
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
 * a wrong answer costs a FIX iteration, never a wrong result.  Both words are loaded
 * before either is tested so the two misses overlap. */
/* replies: is this xid one the session's table covers, or a special one (zk-buffer.js:275-279)?  A reply
 * with any other xid cannot be decoded ('reply packet must match a request'), so a candidate that has
 * one is far more likely payload that happens to look like a header */
struct XidWin { int64_t base, count; bool on; };
__device__ __forceinline__ bool xid_plausible(const XidWin &w, int32_t xid)
{
	if (!w.on) return true;
	if (xid == ZKB_XID_NOTIFICATION || xid == ZKB_XID_PING || xid == ZKB_XID_AUTH || xid == ZKB_XID_SET_WATCHES) return true;
	const int64_t idx = (int64_t)xid - w.base;
	return idx >= 0 && idx < w.count;
}

__device__ bool header_plausible(const uint8_t *bytes, uint64_t p, uint64_t send, bool server,
    uint64_t &next, const XidWin &xw)
{
	const uint32_t fld = server ? 8u : 16u;      /* type (requests) / err (responses) */
	if (p + 4 + fld + 4 > send) return false;
	const int32_t len = (int32_t)g_be32(bytes, p);
	const int32_t f = (int32_t)g_be32(bytes, p + 4 + fld - 4);
	if (len < (int32_t)fld || len > ZKB_MAX_PACKET) return false;
	if (p + 4 + (uint64_t)len > send) return false;
	if (server ? !op_known(f) : !err_known(f)) return false;
	if (!server && xw.on && !xid_plausible(xw, (int32_t)g_be32(bytes, p + 4))) return false;
	next = p + 4 + (uint64_t)len;
	return true;
}

/* Follow the chain `hops` more frames from q: a false start almost never survives this,
 * a true one always does (a legitimately incomplete or final frame ends the check). */
__device__ bool chain_plausible(const uint8_t *bytes, uint64_t q, uint64_t send, bool server, int hops, const XidWin &xw)
{
	for (int k = 0; k < hops; ++k) {
		if (q == send || q + 4 > send) return true;
		uint64_t n2;
		if (!header_plausible(bytes, q, send, server, n2, xw)) {
			/* the chain may end in a frame that is not complete yet, but only one whose
			 * header is sane: a wild length that overruns the session proves nothing */
			const uint32_t fld = server ? 8u : 16u;
			if (q + 8 + fld > send) return true;
			const int32_t l = (int32_t)g_be32(bytes, q);
			const int32_t f = (int32_t)g_be32(bytes, q + fld);
			return l >= (int32_t)fld && l <= ZKB_MAX_PACKET && q + 4 + (uint64_t)l > send &&
			    (server ? op_known(f) : err_known(f));
		}
		q = n2;
	}
	return true;
}

#ifndef ZK_GUESS_HOPS
#define ZK_GUESS_HOPS 3
#endif
#ifndef ZK_GUESS_SPAN
#define ZK_GUESS_SPAN 8192u
#endif
#ifndef ZK_GUESS_HOPS_SERVER
#define ZK_GUESS_HOPS_SERVER 1   /* requests: the local test already checks the path */
#endif
struct GuessCtx {
	const uint8_t *bytes;
	uint64_t lim, send, padend;
	bool server;
	XidWin xw;
};

/* 0x80 in every zero byte of w */
__device__ __forceinline__ uint32_t mark_zero(uint32_t w) { return ~(((w & 0x7f7f7f7fu) + 0x7f7f7f7fu) | w | 0x7f7f7f7fu); }
__device__ __forceinline__ uint32_t nib_of(uint32_t t) { return (((t >> 7) * 0x00204081u) >> 21) & 0xfu; }
/* bit i = byte i of w is zero (exact: no false positives), i = 0 is the byte at the lowest address */
__device__ __forceinline__ uint32_t nib_zero(uint32_t w)
{
	const uint32_t t = ~(((w & 0x7f7f7f7fu) + 0x7f7f7f7fu) | w | 0x7f7f7f7fu);   /* 0x80 in every zero byte */
	return (((t >> 7) * 0x00204081u) >> 21) & 0xfu;
}
__device__ __forceinline__ uint32_t mask16_zero(const uint4 &v)
{
	/* payload bytes (paths, random data) mostly hold no zero at all: one test spares the extraction */
	const uint32_t t0 = mark_zero(v.x), t1 = mark_zero(v.y), t2 = mark_zero(v.z), t3 = mark_zero(v.w);
	if (!(t0 | t1 | t2 | t3)) return 0u;
	return nib_of(t0) | (nib_of(t1) << 4) | (nib_of(t2) << 8) | (nib_of(t3) << 12);
}
__device__ __forceinline__ uint32_t mask16_ff(const uint4 &v)
{
	const uint32_t t0 = mark_zero(~v.x), t1 = mark_zero(~v.y), t2 = mark_zero(~v.z), t3 = mark_zero(~v.w);
	if (!(t0 | t1 | t2 | t3)) return 0u;
	return nib_of(t0) | (nib_of(t1) << 4) | (nib_of(t2) << 8) | (nib_of(t3) << 12);
}
__device__ __forceinline__ uint32_t mask16_le1(const uint4 &v)
{
	return nib_zero(v.x & 0xfefefefeu) | (nib_zero(v.y & 0xfefefefeu) << 4) | (nib_zero(v.z & 0xfefefefeu) << 8) |
	    (nib_zero(v.w & 0xfefefefeu) << 12);
}

/* local test of the candidate at p: sane length that fits the session, err code / opcode (and, with a
 * table, an xid the session can have) where the header would have them.  The bytes come through
 * the L1: the scan has just loaded them. */
__device__ bool guess_local(const GuessCtx &g, uint64_t p)
{
	if (p >= g.lim || p + 24 > g.padend) return false;
	const int32_t len = (int32_t)g_be32(g.bytes, p);
	if (len < (g.server ? 8 : 16) || len > ZKB_MAX_PACKET) return false;
	if (p + 4 + (uint64_t)len > g.send) return false;
	const int32_t f = (int32_t)g_be32(g.bytes, p + (g.server ? 8 : 16));
	if (g.server ? !op_known(f) : !err_known(f)) return false;
	if (!g.server && g.xw.on && !xid_plausible(g.xw, (int32_t)g_be32(g.bytes, p + 4))) return false;
#ifndef ZK_GUESS_NO_PATHCHECK
	/* a request that carries a path has its length next and '/' as its first byte; a real frame
	 * that fails this (an empty path) only costs the granule a fix-up */
	if (g.server && ((f >= 1 && f <= 6) || f == 8 || f == 9 || f == 12)) {
		const int32_t pl = (int32_t)g_be32(g.bytes, p + 12);
		if (pl < 1 || pl + 12 > len || __ldg(g.bytes + p + 16) != 0x2fu) return false;
	}
#endif
	return true;
}

/* First position in [gs, min(ge, send)) that looks like a frame start.  The granule is scanned 16
 * bytes per load.  Three byte-class masks over the 48 bytes in hand (byte <= 1: a length of at most
 * 16 MiB starts with 0x00 or 0x01; byte == 0; byte == 0xff) say, with a few shifts and ANDs for all
 * 16 positions of a load at once, where a header could start at all: a request has the three high
 * bytes of a small opcode eight bytes in, a reply the four bytes of err 0 - or the three 0xff of a
 * negative err - sixteen bytes in.  Only those few positions get the local test, and what passes it
 * is checked by following the chain.  (Requests
 * with a negative opcode are not proposed: a granule that starts with one takes the next frame and is
 * put right by the fix-up.) */
__device__ uint64_t guess_entry(const SplitArgs &a, uint32_t s, uint64_t gs, uint64_t ge, uint64_t send)
{
	GuessCtx g;
	g.bytes = a.bytes;
	g.server = a.sess_flags && (a.sess_flags[s] & ZKB_SESS_SERVER);
	g.lim = ge < send ? ge : send;
	/* build knob: a guess may give up after ZK_GUESS_SPAN bytes (a frame that starts further on is then
	 * found by the fix-up pass).  Measured on config 4 with 2 KiB: the scan gets 4 x cheaper, but every
	 * megabyte frame that ends deep inside a granule costs a fix-up round, and those rounds run one after
	 * the other along a session - slower overall; the default scans the whole granule. */
	if (g.lim > gs + ZK_GUESS_SPAN) g.lim = gs + ZK_GUESS_SPAN;
	g.send = send;
	g.xw.on = a.xt_ops != NULL && !g.server;
	g.xw.base = g.xw.on ? (int64_t)__ldg(a.xt_base + s) : 0;
	g.xw.count = g.xw.on ? (int64_t)__ldg(a.xt_count + s) : 0;
	const uint64_t padend = (a.nbytes + 15) & ~15ull;
	g.padend = padend;
	const uint4 zero4 = make_uint4(0, 0, 0, 0);
	uint4 v0 = zero4, v1 = zero4, v2 = zero4;
	if (gs + 16 <= padend) v0 = __ldg(reinterpret_cast<const uint4 *>(a.bytes + gs));
	if (gs + 32 <= padend) v1 = __ldg(reinterpret_cast<const uint4 *>(a.bytes + gs + 16));
	uint32_t z0 = mask16_zero(v0), z1 = mask16_zero(v1);
	uint32_t f0 = mask16_ff(v0), f1 = mask16_ff(v1);
	/* replies: with a table whose xids stay below 2^24 the xid's first byte is 0x00, or 0xff for the
	 * special ones; the first byte of a zxid is one of the two as well (a miss only costs a fix-up) */
	const bool xid_small = g.server || !g.xw.on || (g.xw.base >= 0 && g.xw.base + g.xw.count <= (1 << 24));
	uint64_t blk = gs;            /* next 16 bytes to scan */
	uint64_t cand = 0, cbase = 0; /* positions of a 64-byte group that passed the masks and are not tested yet */
	for (;;) {
		/* scan on, 64 bytes at a time, until this lane has something to test; the lanes of a warp
		 * (which scan different granules) meet again below, so the tests - the local one and the far,
		 * DRAM-latency ones - run for all of them together */
		while (!cand && blk < g.lim) {
			cbase = blk;
#pragma unroll 1
			for (int q = 0; q < 4; ++q, blk += 16) {
				v2 = zero4;
				if (blk + 48 <= padend) v2 = __ldg(reinterpret_cast<const uint4 *>(a.bytes + blk + 32));
				const uint32_t z2 = mask16_zero(v2);
				const uint32_t f2 = mask16_ff(v2);
				if (blk < g.lim) {
					const uint64_t Z = (uint64_t)z0 | ((uint64_t)z1 << 16) | ((uint64_t)z2 << 32);
					const uint64_t F = (uint64_t)f0 | ((uint64_t)f1 << 16) | ((uint64_t)f2 << 32);
					uint32_t c;
					if (g.server) {
						/* opcode 0..255 eight bytes in; xid small or special */
						c = (uint32_t)(Z >> 8) & (uint32_t)(Z >> 9) & (uint32_t)(Z >> 10) & (uint32_t)((Z | F) >> 4);
					} else {
						const uint32_t e0 = (uint32_t)(Z >> 16) & (uint32_t)(Z >> 17) & (uint32_t)(Z >> 18) & (uint32_t)(Z >> 19);
						const uint32_t en = (uint32_t)(F >> 16) & (uint32_t)(F >> 17) & (uint32_t)(F >> 18);
						c = (e0 | en) & (uint32_t)((Z | F) >> 8);
						if (xid_small) c &= (uint32_t)((Z | F) >> 4);
					}
					c &= 0xffffu;
					if (c) c &= mask16_le1(v0);      /* the length's first byte: only looked at where the rest fits */
					cand |= (uint64_t)c << (16 * q);
				}
				v0 = v1; v1 = v2; z0 = z1; z1 = z2; f0 = f1; f1 = f2;
			}
		}
		if (!cand) break;
		/* every candidate of the group, in stream order */
		while (cand) {
			const int k = __ffsll((long long)cand) - 1;
			cand &= cand - 1;
			const uint64_t p = cbase + (uint64_t)k;
			if (!guess_local(g, p)) continue;
			const int32_t len = (int32_t)g_be32(a.bytes, p);
			if (chain_plausible(a.bytes, p + 4 + (uint64_t)len, send, g.server, g.server ? ZK_GUESS_HOPS_SERVER : ZK_GUESS_HOPS, g.xw))
				return p;
		}
	}
	return ZK_NONE;
}

/* ---- side-table rows a frame should produce ------------------------------------------
 * The walk has every frame header in hand, and for a well-formed frame the header (plus a
 * vector count right behind it) settles how many STAT / BUF / ACL rows the frame owns.  The
 * walk adds these up per granule, the scan turns them into each decode chunk's first row, and
 * k_decode<true> places its rows there without waiting for its predecessors; it compares what
 * it really found with the prediction, and a single mismatch (a malformed frame, a handshake
 * depth the walk cannot know) makes the host place the rest of the batch by the exact route.  So this is a
 * guess that costs time when wrong, never a wrong answer. */
struct RowCnt { uint32_t st, bf, ac; };

__device__ __forceinline__ int special_xid_op(int32_t xid)
{
	return xid == ZKB_XID_NOTIFICATION ? ZKB_OP_NOTIFICATION : xid == ZKB_XID_PING ? ZKB_OP_PING :
	    xid == ZKB_XID_AUTH ? ZKB_OP_AUTH : xid == ZKB_XID_SET_WATCHES ? ZKB_OP_SET_WATCHES : ZKB_OP_NONE;
}

/* The first 24 bytes at a frame start as six big-endian words: length prefix, then payload
 * words 0..4 (request: xid, type, path length; reply: xid, zxid, zxid, err, vector count).
 * Loaded together, one round trip; bytes past the end of the batch read as zero. */
struct Hdr { uint32_t len, w1, w2, w3, w4, w5; };   /* w1 = xid, w2 = request type, w3 = request path length, w4 = reply err, w5 = vector count */

__device__ __forceinline__ Hdr load_hdr(const uint8_t *bytes, uint64_t pos, uint64_t padend, bool server)
{
	Hdr h;
	const uint64_t base = pos & ~3ull;
	const uint32_t sel = 0x0123u + 0x1111u * (uint32_t)(pos & 3);    /* byte permute: four bytes, most significant first */
	const uint32_t *p = reinterpret_cast<const uint32_t *>(bytes + base);
	uint32_t x0 = 0, x1 = 0, x2 = 0, x3 = 0, x4 = 0, x5 = 0, x6 = 0;
	if (base + 28 <= padend) {
		/* a request is settled by its first 16 bytes (CREATE follows its buffers from there) */
		x0 = __ldg(p); x1 = __ldg(p + 1); x2 = __ldg(p + 2); x3 = __ldg(p + 3); x4 = __ldg(p + 4);
		if (!server) { x5 = __ldg(p + 5); x6 = __ldg(p + 6); }
	} else {
		if (base + 4 <= padend) x0 = __ldg(p);
		if (base + 8 <= padend) x1 = __ldg(p + 1);
		if (base + 12 <= padend) x2 = __ldg(p + 2);
		if (base + 16 <= padend) x3 = __ldg(p + 3);
		if (base + 20 <= padend) x4 = __ldg(p + 4);
		if (base + 24 <= padend) x5 = __ldg(p + 5);
	}
	h.len = __byte_perm(x0, x1, sel);
	h.w1 = __byte_perm(x1, x2, sel);
	h.w2 = __byte_perm(x2, x3, sel);
	h.w3 = __byte_perm(x3, x4, sel);
	h.w4 = __byte_perm(x4, x5, sel);
	h.w5 = __byte_perm(x5, x6, sel);
	return h;
}

/* per-session constants of the prediction */
struct SessCtx {
	bool     server;
	uint32_t hs;
	int64_t  xbase;
	int64_t  xcount;
	const int8_t *xops;     /* ops of this session, NULL = no table */
};

__device__ __forceinline__ SessCtx sess_ctx(const SplitArgs &a, uint32_t s)
{
	SessCtx c;
	c.server = a.sess_flags && (__ldg(a.sess_flags + s) & ZKB_SESS_SERVER);
	c.hs = a.sess_hs ? __ldg(a.sess_hs + s) : 1u;
	c.xops = NULL; c.xbase = 0; c.xcount = 0;
	if (a.xt_ops && !c.server) {
		c.xbase = (int64_t)__ldg(a.xt_base + s);
		c.xcount = (int64_t)__ldg(a.xt_count + s);
		c.xops = a.xt_ops + __ldg(a.xt_off + s);
	}
	return c;
}

__device__ void predict_rows(const SplitArgs &a, uint64_t pos, const Hdr &h, const SessCtx &sc, bool handshake,
    RowCnt &r)
{
	const uint64_t pay = pos + 4;
	const uint32_t L = h.len;
	if (handshake) {
		/* readConnectRequest / readConnectResponse (zk-buffer.js:22-30, 41-48): one buffer, passwd */
		if (L >= (sc.server ? 28u : 20u)) r.bf += 1;
		return;
	}
	if (sc.server) {                                /* readRequest zk-buffer.js:58-95 */
		if (L < 8) return;
		switch ((int32_t)h.w2) {
		case ZKB_OP_GET_CHILDREN: case ZKB_OP_GET_CHILDREN2: case ZKB_OP_GET_DATA: case ZKB_OP_EXISTS:
		case ZKB_OP_DELETE: case ZKB_OP_GET_ACL: case ZKB_OP_SYNC:
			r.bf += 1; return;
		case ZKB_OP_SET_DATA:
			r.bf += 2; return;
		case ZKB_OP_CREATE: {
			if (L < 12) return;
			const int32_t pl = (int32_t)h.w3;
			const uint64_t c1 = 12ull + (pl > 0 ? (uint32_t)pl : 0u);
			if (c1 + 4 > L) return;
			const int32_t dl = (int32_t)g_be32(a.bytes, pay + c1);
			const uint64_t c2 = c1 + 4 + (dl > 0 ? (uint32_t)dl : 0u);
			if (c2 + 4 > L) return;
			const int32_t n = (int32_t)g_be32(a.bytes, pay + c2);
			const uint32_t k = n > 0 ? (uint32_t)n : 0u;
			r.bf += 2 + 2 * k; r.ac += k;
			return;
		}
		case ZKB_OP_SET_WATCHES: {
			uint64_t cur = 16;
			uint32_t nb = 0;
			for (int v = 0; v < 3; ++v) {
				if (cur + 4 > L) return;
				const int32_t n = (int32_t)g_be32(a.bytes, pay + cur);
				cur += 4;
				for (int32_t i = 0; i < n; ++i) {
					if (cur + 4 > L) return;
					const int32_t l = (int32_t)g_be32(a.bytes, pay + cur);
					cur += 4ull + (l > 0 ? (uint32_t)l : 0u);
					nb++;
				}
			}
			r.bf += nb;
			return;
		}
		default: return;                            /* 'Unsupported opcode': no packet, no rows */
		}
	}
	/* readResponse zk-buffer.js:281-331 */
	if (L < 16) return;
	const int32_t xid = (int32_t)h.w1;
	int op = special_xid_op(xid);
	if (op == ZKB_OP_NONE && sc.xops) {
		const int64_t idx = (int64_t)xid - sc.xbase;
		if (idx >= 0 && idx < sc.xcount) op = __ldg(sc.xops + idx);
	}
	if (op == ZKB_OP_NONE) return;
	if (h.w4 != 0) return;                          /* only 'OK' replies have a body */
	switch (op) {
	case ZKB_OP_GET_CHILDREN: case ZKB_OP_GET_CHILDREN2: case ZKB_OP_GET_ACL: {
		if (L < 20) return;
		const int32_t n = (int32_t)h.w5;
		const uint32_t k = n > 0 ? (uint32_t)n : 0u;
		if (op == ZKB_OP_GET_ACL) { r.bf += 2 * k; r.ac += k; r.st += 1; }
		else { r.bf += k; if (op == ZKB_OP_GET_CHILDREN2) r.st += 1; }
		return;
	}
	case ZKB_OP_CREATE: case ZKB_OP_NOTIFICATION: r.bf += 1; return;
	case ZKB_OP_GET_DATA: r.bf += 1; r.st += 1; return;
	case ZKB_OP_EXISTS: case ZKB_OP_SET_DATA: r.st += 1; return;
	default: return;
	}
}

/* The loop of ZKDecodeStream._transform (zk-streams.js:47-64) restricted to the frames
 * that start before `ge`: counts them (and the rows they should produce) and returns where
 * the chain stands afterwards (ZK_NONE once the session's chain has ended: stream exhausted,
 * frame incomplete, or BAD_LENGTH).  `fidx` is the index within session `s` of the frame at
 * `pos`, or ZK_FIDX_UNKNOWN when the walk entered the session somewhere in the middle.  The
 * header of the next frame is requested before the current frame's rows are worked out, so
 * the table lookups of the prediction overlap that round trip.  The decode kernel repeats
 * this walk on the staged bytes to place frames. */
#define ZK_FIDX_UNKNOWN 0xFFFFFFFFu
__device__ uint64_t walk_chain(const SplitArgs &a, uint64_t pos, uint64_t send, uint64_t ge, uint32_t &count,
    uint64_t gs, uint16_t *row, uint32_t &last_sub, RowCnt &rows, uint32_t s, uint32_t fidx)
{
	const SessCtx sc = sess_ctx(a, s);
	const uint64_t padend = (a.nbytes + 15) & ~15ull;
	Hdr h = load_hdr(a.bytes, pos < ge ? pos : 0, padend, sc.server);
	for (;;) {
		if (pos < ge) {
			/* positions only grow, so each sub-range is recorded at most once */
			const uint32_t idx = (uint32_t)(pos - gs) >> 8;
			if (idx != last_sub) { row[idx] = (uint16_t)(pos - gs); last_sub = idx; }
		}
		if (pos == send) return ZK_NONE;
		if (pos >= ge) return pos;
		const uint64_t rem = send - pos;
		if (rem < 4) return ZK_NONE;                        /* zd_used < 4 */
		const int32_t len = (int32_t)h.len;
		if (!len_ok(len)) return ZK_NONE;                   /* :49-53 BAD_LENGTH */
		if (rem < 4 + (uint64_t)len) return ZK_NONE;        /* frame not complete yet */
		count++;
		const uint64_t next = pos + 4 + (uint64_t)len;
		Hdr hn;
		if (next < ge) hn = load_hdr(a.bytes, next, padend, sc.server);
		/* a walk that entered mid-session is past the handshake unless every frame is one */
#ifndef ZK_EXP_NO_ROWPRED
		predict_rows(a, pos, h, sc, fidx == ZK_FIDX_UNKNOWN ? sc.hs == 0xFFFFFFFFu : fidx < sc.hs, rows);
#endif
		if (fidx != ZK_FIDX_UNKNOWN) fidx++;
		pos = next;
		h = hn;
	}
}

/* One granule's walk: the chain that enters at `assumed` (if that lies inside the granule)
 * and then every session that starts inside it.  Fills the granule's row of sub-range
 * entries; `out` is where the last chain stands afterwards (ZK_INHERIT if nothing was walked). */
struct GranWalk { uint64_t out; uint32_t count; uint32_t open_local; RowCnt rows; };

__device__ GranWalk walk_granule(const SplitArgs &a, uint32_t g, uint32_t sf, uint64_t assumed, uint16_t *row)
{
	const uint64_t gs = (uint64_t)g << a.glog2;
	const bool last = (g == a.n_gran - 1);
	const uint64_t ge = last ? a.nbytes : gs + (1ull << a.glog2);
	const uint32_t subs = 1u << (a.glog2 - 8);
	for (uint32_t i = 0; i < subs; i += 8)
		*reinterpret_cast<uint4 *>(row + i) = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu);
	GranWalk w; w.count = 0; w.open_local = 0; w.rows.st = w.rows.bf = w.rows.ac = 0;
	uint64_t pos = ZK_NONE;
	bool touched = false;
	uint32_t last_sub = 0xFFFFFFFFu;
	if (__ldg(a.sess_off + sf) != gs && assumed >= gs && assumed < ge) {   /* inside session sf-1 */
		touched = true;
		pos = walk_chain(a, assumed, __ldg(a.sess_off + sf), ge, w.count, gs, row, last_sub, w.rows, sf - 1, ZK_FIDX_UNKNOWN);
	}
	for (uint32_t s = sf; s < a.nsess; ++s) {
		uint64_t so = __ldg(a.sess_off + s);
		if (!(so < ge || last)) break;
		touched = true;
		w.open_local = w.count;
		pos = walk_chain(a, so, __ldg(a.sess_off + s + 1), ge, w.count, gs, row, last_sub, w.rows, s, 0);
	}
	w.out = touched ? pos : ZK_INHERIT;
	return w;
}

/* A dirty granule's carried position can be trusted when the granule that produced it is
 * itself consistent and it does not point into an earlier granule. */
__device__ __forceinline__ bool fix_trusted(const SplitArgs &a, uint32_t g)
{
	const uint32_t src = a.g_src[g];
	const uint64_t carry = a.g_carry[g];
	return !(src && a.g_dirty[src - 1]) && !(carry != ZK_NONE && carry < ((uint64_t)g << a.glog2));
}

#ifndef ZK_FIX_EXT
#define ZK_FIX_EXT 64   /* granules one fix-up thread may carry its corrected chain through */
#endif

#ifndef ZK_SPLIT_MINBLOCKS
#define ZK_SPLIT_MINBLOCKS 1
#endif
template <int MODE>
__global__ void __launch_bounds__(128, ZK_SPLIT_MINBLOCKS) k_split(SplitArgs a)
{
	uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
	if (g >= a.n_gran) return;
	if (MODE == MODE_FIX && !a.g_dirty[g]) return;
	const uint64_t gs = (uint64_t)g << a.glog2;
	const bool last = (g == a.n_gran - 1);
	const uint64_t ge = last ? a.nbytes : gs + (1ull << a.glog2);
	const uint32_t subs = 1u << (a.glog2 - 8);

	if (MODE == MODE_SPEC) {
#ifdef ZK_PROFILE_PHASES
		const long long sp_t0 = clock64();
#endif
		const uint32_t sf = lower_bound_u64(a.sess_off, a.nsess + 1, gs);
		if (sf > a.nsess) return;   /* offsets do not cover the batch: k_check_sessions reports it */
		uint64_t assumed = ZK_NONE;
		if (__ldg(a.sess_off + sf) != gs)   /* granule starts inside session sf-1 */
			assumed = guess_entry(a, sf - 1, gs, ge, __ldg(a.sess_off + sf));
		__syncwarp();   /* lanes finish guessing at different times: walk together */
#ifdef ZK_PROFILE_PHASES
		const long long sp_t1 = clock64();
#endif
		uint16_t *row = a.g_sub + (size_t)g * subs;
		const GranWalk w = walk_granule(a, g, sf, assumed, row);
		uint16_t *row1 = a.g_sub + ((size_t)a.n_gran + g) * subs;
		for (uint32_t i = 0; i < subs; i += 8)
			*reinterpret_cast<uint4 *>(row1 + i) = *reinterpret_cast<const uint4 *>(row + i);
#ifdef ZK_PROFILE_PHASES
		__syncwarp();
		const long long sp_t2 = clock64();
		if ((threadIdx.x & 31) == 0) {
			atomicAdd(&a.ctrl->prof[0], (unsigned long long)(sp_t1 - sp_t0));
			atomicAdd(&a.ctrl->prof[1], (unsigned long long)(sp_t2 - sp_t1));
			atomicAdd(&a.ctrl->prof[2], 1ull);
		}
#endif
		a.g_assumed[g] = assumed; a.g_out[g] = w.out; a.g_cnt[g] = w.count; a.g_openlocal[g] = w.open_local;
		a.g_assumed[a.n_gran + g] = assumed; a.g_out[a.n_gran + g] = w.out; a.g_cnt[a.n_gran + g] = w.count;
		a.g_openlocal[a.n_gran + g] = w.open_local;
		a.g_rows[g] = a.g_rows[a.n_gran + g] = make_uint4(w.rows.st, w.rows.bf, w.rows.ac, 0);
		a.g_sel[g] = 0;
		a.g_sf[g] = sf;
		return;
	}

	/* FIX.  Only a dirty granule whose carried position can be trusted acts (the lowest dirty
	 * granule always qualifies, so every iteration makes progress); the others keep the walk
	 * from their own guess, whose exit is usually right even when its entry was not, so one
	 * bad guess cannot push an error wave down the stream.  The acting thread re-walks its
	 * granule into version 1 and then carries the corrected chain through the granules that
	 * follow - switching them to whichever recorded walk agrees with it, or re-walking them -
	 * until it reaches a granule that already agrees (the chains have merged) or another
	 * trusted dirty granule (which has its own thread).  The runs of two acting threads are
	 * disjoint, so nothing here is written twice.  The scan that follows re-verifies every
	 * granule, so a run cut short by ZK_FIX_EXT only costs another iteration. */
	if (!fix_trusted(a, g)) return;
	uint64_t carry = a.g_carry[g];
	{
		const GranWalk w = walk_granule(a, g, a.g_sf[g], carry, a.g_sub + ((size_t)a.n_gran + g) * subs);
		a.g_assumed[a.n_gran + g] = carry; a.g_out[a.n_gran + g] = w.out; a.g_cnt[a.n_gran + g] = w.count;
		a.g_openlocal[a.n_gran + g] = w.open_local;
		a.g_rows[a.n_gran + g] = make_uint4(w.rows.st, w.rows.bf, w.rows.ac, 0);
		a.g_sel[g] = 1;
		if (w.out != ZK_INHERIT) carry = w.out;
	}
	for (uint32_t h = g + 1, steps = 0; h < a.n_gran && steps < ZK_FIX_EXT; ++h, ++steps) {
		if (a.g_dirty[h] && fix_trusted(a, h)) break;
		const uint64_t hs = (uint64_t)h << a.glog2;
		const uint64_t he = (h == a.n_gran - 1) ? a.nbytes : hs + (1ull << a.glog2);
		const uint64_t nc = (carry >= hs && carry < he) ? carry : ZK_NONE;
		uint8_t sel = a.g_sel[h];
		const uint64_t as = a.g_assumed[(size_t)sel * a.n_gran + h];
		const uint64_t ao = a.g_assumed[(size_t)(sel ^ 1) * a.n_gran + h];
		bool changed = false;
		uint64_t hout;
		if (((as >= hs && as < he) ? as : ZK_NONE) == nc) {
			hout = a.g_out[(size_t)sel * a.n_gran + h];
		} else if (((ao >= hs && ao < he) ? ao : ZK_NONE) == nc) {
			sel ^= 1;
			a.g_sel[h] = sel;
			hout = a.g_out[(size_t)sel * a.n_gran + h];
			changed = true;
		} else {
			const GranWalk w = walk_granule(a, h, a.g_sf[h], nc, a.g_sub + ((size_t)a.n_gran + h) * subs);
			a.g_assumed[a.n_gran + h] = nc; a.g_out[a.n_gran + h] = w.out; a.g_cnt[a.n_gran + h] = w.count;
			a.g_openlocal[a.n_gran + h] = w.open_local;
			a.g_rows[a.n_gran + h] = make_uint4(w.rows.st, w.rows.bf, w.rows.ac, 0);
			a.g_sel[h] = 1;
			hout = w.out;
			changed = true;
		}
		if (hout == ZK_INHERIT) continue;          /* nothing starts in h: the carry passes through */
		carry = hout;
		if (changed) continue;                       /* what h hands on may differ from what was scanned */
		if (h + 1 < a.n_gran && a.g_dirty[h + 1] && !fix_trusted(a, h + 1)) continue;
		break;                                       /* merged with the chain the scan already has */
	}
}

/* ---- scan over granules: value = (last exit that is not INHERIT, sum of counts) ------ */

struct ScanVal { uint64_t v; uint32_t c; uint32_t src; uint32_t ssrc; uint32_t r0, r1, r2; };

__device__ __forceinline__ ScanVal sv_comb(ScanVal l, ScanVal r)
{
	ScanVal o;
	o.v = (r.v != ZK_INHERIT) ? r.v : l.v;
	o.src = (r.v != ZK_INHERIT) ? r.src : l.src;
	o.ssrc = r.ssrc ? r.ssrc : l.ssrc;
	o.c = l.c + r.c;
	o.r0 = l.r0 + r.r0; o.r1 = l.r1 + r.r1; o.r2 = l.r2 + r.r2;
	return o;
}
__device__ __forceinline__ ScanVal sv_shfl_up(ScanVal x, int d)
{
	ScanVal o;
	o.v = __shfl_up_sync(0xffffffffu, x.v, d);
	o.c = __shfl_up_sync(0xffffffffu, x.c, d);
	o.src = __shfl_up_sync(0xffffffffu, x.src, d);
	o.ssrc = __shfl_up_sync(0xffffffffu, x.ssrc, d);
	o.r0 = __shfl_up_sync(0xffffffffu, x.r0, d);
	o.r1 = __shfl_up_sync(0xffffffffu, x.r1, d);
	o.r2 = __shfl_up_sync(0xffffffffu, x.r2, d);
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
	ScanVal pre; pre.v = ZK_INHERIT; pre.c = 0; pre.src = 0; pre.ssrc = 0; pre.r0 = pre.r1 = pre.r2 = 0;
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

/* `gate`: the scan that follows the queued fix-up round has nothing to do when the first scan found
 * every granule consistent (request streams, ordinarily): NULL = always run, else run iff *gate != 0 */
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_reduce(SplitArgs a, ScanVal *blk, const unsigned int *gate)
{
	__shared__ ScanVal s_w[32];
	if (gate && *gate == 0) return;
	uint32_t g = blockIdx.x * SCAN_THREADS + threadIdx.x;
	ScanVal x; x.v = ZK_INHERIT; x.c = 0; x.src = 0; x.ssrc = 0; x.r0 = x.r1 = x.r2 = 0;
	if (g < a.n_gran) {
		const size_t i = (size_t)a.g_sel[g] * a.n_gran + g;
		x.v = a.g_out[i]; x.c = a.g_cnt[i]; x.src = g + 1;
		{ const uint4 rw = a.g_rows[i]; x.r0 = rw.x; x.r1 = rw.y; x.r2 = rw.z; }
		const uint32_t sf = a.g_sf[g];
		const uint64_t gend = (g == a.n_gran - 1) ? a.nbytes + 1 : ((uint64_t)(g + 1) << a.glog2);
		x.ssrc = (sf < a.nsess && __ldg(a.sess_off + sf) < gend) ? g + 1 : 0;
	}
	ScanVal tot;
	block_scan_incl(x, &tot, s_w);
	if (threadIdx.x == 0) blk[blockIdx.x] = tot;
}

/* exclusive scan of the block aggregates, in place, by one block */
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_blocks(ScanVal *blk, uint32_t nblk, const unsigned int *gate)
{
	__shared__ ScanVal s_w[32];
	if (gate && *gate == 0) return;
	ScanVal carry; carry.v = ZK_INHERIT; carry.c = 0; carry.src = 0; carry.ssrc = 0; carry.r0 = carry.r1 = carry.r2 = 0;
	for (uint32_t b0 = 0; b0 < nblk; b0 += SCAN_THREADS) {
		uint32_t i = b0 + threadIdx.x;
		ScanVal x; x.v = ZK_INHERIT; x.c = 0; x.src = 0; x.ssrc = 0; x.r0 = x.r1 = x.r2 = 0;
		if (i < nblk) x = blk[i];
		ScanVal tot;
		ScanVal inc = block_scan_incl(x, &tot, s_w);
		/* exclusive = carry (+) inclusive of the previous element */
		ScanVal prev = sv_shfl_up(inc, 1);
		__shared__ ScanVal s_last[32];
		if ((threadIdx.x & 31) == 31) s_last[threadIdx.x >> 5] = inc;
		__syncthreads();
		ScanVal ex;
		if (threadIdx.x == 0) { ex.v = ZK_INHERIT; ex.c = 0; ex.src = 0; ex.ssrc = 0; ex.r0 = ex.r1 = ex.r2 = 0; }
		else if ((threadIdx.x & 31) == 0) ex = s_last[(threadIdx.x >> 5) - 1];
		else ex = prev;
		if (i < nblk) blk[i] = sv_comb(carry, ex);
		carry = sv_comb(carry, tot);
		__syncthreads();
	}
}

__global__ void __launch_bounds__(SCAN_THREADS) k_scan_apply(SplitArgs a, const ScanVal *blk, const unsigned int *gate)
{
	__shared__ ScanVal s_w[32];
	__shared__ ScanVal s_last[32];
	if (gate && *gate == 0) return;
	uint32_t g = blockIdx.x * SCAN_THREADS + threadIdx.x;
	ScanVal x; x.v = ZK_INHERIT; x.c = 0; x.src = 0; x.ssrc = 0; x.r0 = x.r1 = x.r2 = 0;
	if (g < a.n_gran) {
		const size_t i = (size_t)a.g_sel[g] * a.n_gran + g;
		x.v = a.g_out[i]; x.c = a.g_cnt[i]; x.src = g + 1;
		{ const uint4 rw = a.g_rows[i]; x.r0 = rw.x; x.r1 = rw.y; x.r2 = rw.z; }
		const uint32_t sf = a.g_sf[g];
		const uint64_t gend = (g == a.n_gran - 1) ? a.nbytes + 1 : ((uint64_t)(g + 1) << a.glog2);
		x.ssrc = (sf < a.nsess && __ldg(a.sess_off + sf) < gend) ? g + 1 : 0;
	}
	ScanVal tot;
	ScanVal inc = block_scan_incl(x, &tot, s_w);
	ScanVal prev = sv_shfl_up(inc, 1);
	if ((threadIdx.x & 31) == 31) s_last[threadIdx.x >> 5] = inc;
	__syncthreads();
	ScanVal ex;
	if (threadIdx.x == 0) { ex.v = ZK_INHERIT; ex.c = 0; ex.src = 0; ex.ssrc = 0; ex.r0 = ex.r1 = ex.r2 = 0; }
	else if ((threadIdx.x & 31) == 0) ex = s_last[(threadIdx.x >> 5) - 1];
	else ex = prev;
	ex = sv_comb(blk[blockIdx.x], ex);
	if (g >= a.n_gran) return;
	const uint64_t gs = (uint64_t)g << a.glog2;
	const uint64_t ge = (g == a.n_gran - 1) ? a.nbytes : gs + (1ull << a.glog2);
	uint64_t carry = (ex.v == ZK_INHERIT) ? ZK_NONE : ex.v;
	a.g_carry[g] = carry;
	a.g_base[g] = ex.c;
	a.g_rbase[g] = make_uint4(ex.r0, ex.r1, ex.r2, 0);
	if (g == a.n_gran - 1) {
		a.g_rbase[a.n_gran] = make_uint4(ex.r0 + x.r0, ex.r1 + x.r1, ex.r2 + x.r2, 0);
		a.ctrl->pred_rows[0] = ex.r0 + x.r0; a.ctrl->pred_rows[1] = ex.r1 + x.r1; a.ctrl->pred_rows[2] = ex.r2 + x.r2;
	}
	a.g_src[g] = (ex.v == ZK_INHERIT) ? 0u : ex.src;
	a.g_ssrc[g] = ex.ssrc;
	/* a walk is consistent with the carry iff both name the same in-granule position, or
	 * neither names one; pick whichever version is, else the granule is dirty */
	const uint64_t nc = (carry >= gs && carry < ge) ? carry : ZK_NONE;
	const uint64_t a0 = a.g_assumed[g], a1 = a.g_assumed[a.n_gran + g];
	const uint64_t n0 = (a0 >= gs && a0 < ge) ? a0 : ZK_NONE;
	const uint64_t n1 = (a1 >= gs && a1 < ge) ? a1 : ZK_NONE;
	const uint8_t old = a.g_sel[g];
	uint8_t dirty = 0, sel = old;
	if ((old ? n1 : n0) == nc) sel = old;
	else if (n0 == nc) sel = 0;
	else if (n1 == nc) sel = 1;
	else dirty = 1;
	a.g_dirty[g] = dirty;
	a.g_sel[g] = sel;
	/* a changed selection changes what the scan saw: not converged yet */
	if (dirty || sel != old) atomicAdd(&a.ctrl->n_dirty, 1u);
	if (dirty) atomicAdd(&a.ctrl->n_rewalk, 1u);
	if (g == a.n_gran - 1) a.ctrl->n_frames = (unsigned long long)ex.c + x.c;
}

/* between the first scan of a split and its fix-up round: keep what the scan counted, start afresh */
__global__ void k_split_round(zk_ctrl *ctrl)
{
	ctrl->first_dirty[0] = ctrl->n_dirty; ctrl->first_dirty[1] = ctrl->n_rewalk;
	ctrl->n_dirty = 0; ctrl->n_rewalk = 0;
}

/* checks off[0] == 0, off[n] == nbytes, monotone */
__global__ void k_check_sessions(const uint64_t *off, const uint8_t *flags, uint32_t nsess, uint64_t nbytes,
    zk_ctrl *ctrl)
{
	uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i > nsess) return;
	if (i < nsess) {
		const unsigned int bit = (flags && (flags[i] & ZKB_SESS_SERVER)) ? 1u : 2u;
		if (!(ctrl->dir_mask & bit)) atomicOr(&ctrl->dir_mask, bit);
	}
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

/* a GET_CHILDREN(2) reply whose children vector is longer than a decode tile: what k_long_children
 * needs to walk it */
struct LongItem { uint64_t foff; uint32_t flen, frame, brow0, srow, with_stat, chunk; };

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
	const uint64_t *sess_off;
	uint32_t        nsess;
	const uint32_t *g_base;       /* per granule: first frame index */
	const uint32_t *g_sf;         /* per granule: first session starting at or after it */
	const uint32_t *g_ssrc;       /* per granule: 1 + last earlier granule with a session start */
	const uint32_t *g_openlocal;  /* [version][g] */
	const uint16_t *g_sub;        /* [version][g][granule/256] chain entry per 256-byte sub-range */
	const uint8_t  *g_sel;
	const uint4    *g_rbase;      /* [n_gran + 1] predicted first { stat, buf, acl } row per granule */
	uint32_t        glog2;
	uint32_t        n_gran;
	uint32_t        n_chunks;
	uint64_t        n_frames;
	zkb_decoded     o;
	const uint4    *meta0;        /* [n_chunks + 1] { first frame, first frame of the session the chunk starts in, first
	                                 session starting at or after the chunk, bytes to that session's start (saturated) } */
	const uint4    *meta1;        /* [n_chunks + 1] predicted first { stat, buf, acl } row of the chunk */
	uint4          *c_rows;       /* [n_chunks] { needs the exact body, stat rows, buf rows, acl rows } the chunk really has */
	uint4          *c_base;       /* [n_chunks] exclusive prefix of c_rows (redo passes) */
	uint32_t        chunk0;       /* first chunk of this launch (a redo covers only the tail of a batch) */
	uint32_t        force_exact;  /* every chunk takes the exact body (prediction switched off) */
	struct LongItem *long_items;  /* [n_chunks] first pass: long children vectors are listed here, not walked (NULL = walk them in place) */
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
			return __byte_perm(w[0], w[1], 0x0123u + 0x1111u * ((uint32_t)r & 3u));
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

/* what one frame contributes */
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

/* ---- per-warp chunk state in shared memory ------------------------------------------
 * A decode warp owns one 8 KiB chunk of the batch: the chunk plus a short overhang is
 * staged in `tile`, and every side-table row the chunk's frames produce is first recorded
 * as one bit in a bitmap indexed by the 4-byte slot that holds the field's first byte
 * (length prefix of a buffer/ustring, perms word of an ACL entry, first byte of a Stat).
 * Two fields never start in the same slot, and slot order is stream order, so compacting a
 * bitmap yields the chunk's rows already in (frame, read order) order.  sub0/sub1 keep the
 * low two bits of the byte position. */
#ifndef ZK_CHUNK_LOG2
#define ZK_CHUNK_LOG2  13
#endif
#define ZK_CHUNK       (1u << ZK_CHUNK_LOG2)
#ifndef ZK_OVER
#define ZK_OVER        768u
#endif
#define ZK_WIN         (ZK_CHUNK + ZK_OVER)
#define ZK_SLOTS       (ZK_WIN / 4)
#define ZK_BMW         (ZK_SLOTS / 32)
#define ZK_DEC_WARPS   4

/* deferred vector walk of the direct body: one entry per frame whose vector is walked later */
struct VEnt { uint32_t w0, w1, w2, w3; };   /* pr | vrel << 16;  buf row | acl row << 16 (chunk relative);  stat row | kind << 16;  frame index in chunk */
#define ZK_VL_CAP   64u
#define ZK_BM_BYTES (6u * ZK_BMW * 4u + (ZK_BMW + 2u) * 2u)
#define ZK_FREL_CAP ((ZK_BM_BYTES - ZK_VL_CAP * 16u) / 2u)

struct __align__(16) ChunkSm {
	uint8_t  tile[ZK_WIN + 16];
	union {
		struct {      /* exact body: one bit per 4-byte slot */
			uint32_t m_buf[ZK_BMW], m_acl[ZK_BMW], m_stat[ZK_BMW];
			uint32_t sub0[ZK_BMW], sub1[ZK_BMW];
			uint32_t fstart[ZK_BMW];       /* slots where a frame's length prefix starts */
			uint16_t fpre[ZK_BMW + 2];     /* exclusive prefix of popc(fstart[w]) */
		};
		struct {      /* direct body */
			VEnt     vl[ZK_VL_CAP];        /* frames whose vector is still to be walked */
			uint16_t frel[ZK_FREL_CAP];    /* tile offset of frame (window start + i) */
		};
	};
	/* direct body: the chunk's one frame that reaches past the tile, parsed by the lane that met it */
	Parsed   lp;
	uint64_t l_foff;
	int32_t  l_flen;
	uint32_t l_f;
	uint32_t have_long;
};

enum { WALK_COUNT = 0, WALK_EMIT = 1, WALK_MARK = 2 };
enum { FAST_MARK = 1, FAST_UNMARK = 2 };

__device__ __forceinline__ uint32_t sm_be32(const uint8_t *tile, uint32_t rel)
{
	/* one byte permute picks the four bytes, most significant first, out of the two aligned words */
	const uint32_t *w = reinterpret_cast<const uint32_t *>(tile + (rel & ~3u));
	return __byte_perm(w[0], w[1], 0x0123u + 0x1111u * (rel & 3u));
}

template <int M>
__device__ __forceinline__ void mark_pos(ChunkSm &c, uint32_t *bm, uint32_t prel)
{
	const uint32_t w = prel >> 7, bit = 1u << ((prel >> 2) & 31);
	if (M == FAST_MARK) {
		atomicOr(&bm[w], bit);
		if (prel & 1) atomicOr(&c.sub0[w], bit);
		if (prel & 2) atomicOr(&c.sub1[w], bit);
	} else {
		atomicAnd(&bm[w], ~bit);
		atomicAnd(&c.sub0[w], ~bit);
		atomicAnd(&c.sub1[w], ~bit);
	}
}

struct SideRows {
	zkb_decoded o;
	uint32_t frame;
	uint32_t buf_row, acl_row;
	ChunkSm *cs;        /* WALK_MARK: where to mark */
	uint32_t payrel;    /* WALK_MARK: payload byte 0 relative to the tile */
};

/* vector<ustring> (zk-buffer.js:334-337,238-252).  WALK_COUNT validates and counts,
 * WALK_EMIT writes BUF rows directly, WALK_MARK records them in the chunk bitmaps. */
template <int MODE>
__device__ bool walk_strvec(Cur &c, int32_t &n_out, uint32_t &n_buf, SideRows *sr)
{
	int32_t count;
	if (!c_int(c, ZKB_F_COUNT, count)) return false;
	n_out = count > 0 ? count : 0;
	for (int32_t i = 0; i < count; ++i) {
		uint32_t pos; int32_t len;
		if (!c_buf(c, ZKB_F_ELEM, pos, len)) return false;
		if (MODE == WALK_EMIT) {
			uint64_t row = sr->buf_row++;
			if (row < sr->o.cap_buf) {
				sr->o.buf_frame[row] = sr->frame;
				sr->o.buf_off[row] = c.pay + pos;
				sr->o.buf_len[row] = len;
			}
		} else if (MODE == WALK_MARK) {
			mark_pos<FAST_MARK>(*sr->cs, sr->cs->m_buf, sr->payrel + pos - 4);
		}
		n_buf++;
	}
	return true;
}

/* readACL (zk-buffer.js:372-382): count x { perms, scheme, id } */
template <int MODE>
__device__ bool walk_acl(Cur &c, int32_t &n_out, uint32_t &n_buf, uint32_t &n_acl, SideRows *sr)
{
	int32_t count;
	if (!c_int(c, ZKB_F_COUNT, count)) return false;
	n_out = count > 0 ? count : 0;
	for (int32_t i = 0; i < count; ++i) {
		int32_t perms; uint32_t p0, p1; int32_t l0, l1;
		const uint32_t at = c.cur;
		if (!c_int(c, ZKB_F_ACL_PERMS, perms)) return false;
		if (!c_buf(c, ZKB_F_ACL_SCHEME, p0, l0)) return false;
		if (!c_buf(c, ZKB_F_ACL_ID, p1, l1)) return false;
		if (MODE == WALK_EMIT) {
			uint64_t ar = sr->acl_row++;
			if (ar < sr->o.cap_acl) { sr->o.acl_frame[ar] = sr->frame; sr->o.acl_perms[ar] = perms; }
			uint64_t row = sr->buf_row; sr->buf_row += 2;
			if (row < sr->o.cap_buf) {
				sr->o.buf_frame[row] = sr->frame; sr->o.buf_off[row] = c.pay + p0; sr->o.buf_len[row] = l0;
			}
			if (row + 1 < sr->o.cap_buf) {
				sr->o.buf_frame[row + 1] = sr->frame; sr->o.buf_off[row + 1] = c.pay + p1; sr->o.buf_len[row + 1] = l1;
			}
		} else if (MODE == WALK_MARK) {
			mark_pos<FAST_MARK>(*sr->cs, sr->cs->m_acl, sr->payrel + at);
			mark_pos<FAST_MARK>(*sr->cs, sr->cs->m_buf, sr->payrel + p0 - 4);
			mark_pos<FAST_MARK>(*sr->cs, sr->cs->m_buf, sr->payrel + p1 - 4);
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
		if (!walk_strvec<WALK_COUNT>(c, p.a0, p.n_buf, nullptr)) return;
		if (op == ZKB_OP_GET_CHILDREN2) { if (!c_stat(c, p.stat_pos)) return; p.n_stat = 1; }
		break;
	case ZKB_OP_CREATE:
		fixed_buf(c, ZKB_F_PATH, p);
		break;
	case ZKB_OP_GET_ACL:
		p.vec_kind = VEC_ACL; p.vec_pos = c.cur;
		if (!walk_acl<WALK_COUNT>(c, p.a0, p.n_buf, p.n_acl, nullptr)) return;
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
		if (!walk_acl<WALK_COUNT>(c, p.a1, p.n_buf, p.n_acl, nullptr)) return;
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
		if (!walk_strvec<WALK_COUNT>(c, p.a0, p.n_buf, nullptr)) return;
		if (!walk_strvec<WALK_COUNT>(c, p.a1, p.n_buf, nullptr)) return;
		walk_strvec<WALK_COUNT>(c, p.err, p.n_buf, nullptr);
		break;
	default:
		c.fail = DETAIL(ZKB_CAUSE_UNSUPPORTED, ZKB_F_OPCODE);   /* :91-92 */
		break;
	}
}

/* handlePacket (zk-streams.js:66-99) for one frame; no side rows are written here */
/* DIR: 1 = every session of the batch is client->server, 2 = every one is server->client,
 * 0 = look at the session's flag.  The one-direction kernels carry only that direction's readers. */
template <int DIR>
__device__ void parse_frame(const Rd &rd, const DecArgs &a, uint64_t foff, int32_t flen,
    uint32_t s, uint32_t idx_in_session, Parsed &p)
{
	Cur c;
	c.rd = &rd; c.pay = foff + 4; c.size = (uint32_t)flen; c.cur = 0; c.fail = 0;
	p.xid = p.err = p.a0 = p.a1 = 0; p.zxid = 0;
	p.n_stat = p.n_buf = p.n_acl = 0; p.n_fixed = 0; p.vec_kind = VEC_NONE;
	p.stat_pos = 0; p.vec_pos = 0;
	p.fb0_pos = p.fb1_pos = 0; p.fb0_len = p.fb1_len = 0;
	const bool server = DIR == 1 ? true : DIR == 2 ? false : (a.sess_flags && (__ldg(a.sess_flags + s) & ZKB_SESS_SERVER));
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
	sr.o = o; sr.frame = frame; sr.buf_row = buf_row; sr.acl_row = acl_row; sr.cs = nullptr; sr.payrel = 0;
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
			walk_strvec<WALK_EMIT>(c, n, nb, &sr);
		} else if (p.vec_kind == VEC_ACL) {
			walk_acl<WALK_EMIT>(c, n, nb, na, &sr);
		} else {
			walk_strvec<WALK_EMIT>(c, n, nb, &sr);
			walk_strvec<WALK_EMIT>(c, n, nb, &sr);
			walk_strvec<WALK_EMIT>(c, n, nb, &sr);
		}
	}
}

/* ================================================================================== */
/* optimistic fast path: well-formed frames that lie completely inside the staged tile.
 * Same reads in the same order as the generic readers, but no exception bookkeeping: any
 * irregularity makes the function return false, the marks it made are undone by running it
 * again in FAST_UNMARK mode, and the frame is handed to parse_frame() for the exact answer. */

struct FrameCols {
	uint32_t meta;
	int32_t  xid, err, a0, a1;
	int64_t  zxid;
};

struct FC {
	ChunkSm *c;
	uint32_t pr;     /* payload byte 0, tile-relative */
	uint32_t size;
	uint32_t cur;
	__device__ __forceinline__ bool i32(int32_t &v)
	{
		if (cur + 4 > size) return false;
		v = (int32_t)sm_be32(c->tile, pr + cur);
		cur += 4;
		return true;
	}
	template <int M>
	__device__ __forceinline__ bool buf(uint32_t *bm)
	{
		if (cur + 4 > size) return false;
		const int32_t l = (int32_t)sm_be32(c->tile, pr + cur);
		const uint32_t ul = l > 0 ? (uint32_t)l : 0u;
		if (cur + 4 + ul > size) return false;
		mark_pos<M>(*c, bm, pr + cur);
		cur += 4 + ul;
		return true;
	}
	template <int M>
	__device__ __forceinline__ bool stat()
	{
		if (cur + 68 > size) return false;
		mark_pos<M>(*c, c->m_stat, pr + cur);
		cur += 68;
		return true;
	}
	/* bits gathered for one bitmap word go to shared memory with one atomic per bitmap */
	template <int M>
	__device__ __forceinline__ void flush(uint32_t w, uint32_t am, uint32_t a0, uint32_t a1)
	{
		if (!am) return;
		if (M == FAST_MARK) {
			atomicOr(&c->m_buf[w], am);
			if (a0) atomicOr(&c->sub0[w], a0);
			if (a1) atomicOr(&c->sub1[w], a1);
		} else {
			atomicAnd(&c->m_buf[w], ~am);
			atomicAnd(&c->sub0[w], ~am);
			atomicAnd(&c->sub1[w], ~am);
		}
	}
	/* vector<ustring>: neighbouring elements usually share a bitmap word (128 input bytes) */
	template <int M>
	__device__ __forceinline__ bool strvec(int32_t &n)
	{
		int32_t count;
		if (!i32(count)) return false;
		uint32_t cw = 0, am = 0, a0 = 0, a1 = 0;
		bool ok = true;
		for (int32_t i = 0; i < count; ++i) {
			if (cur + 4 > size) { ok = false; break; }
			const int32_t l = (int32_t)sm_be32(c->tile, pr + cur);
			const uint32_t ul = l > 0 ? (uint32_t)l : 0u;
			if (cur + 4 + ul > size) { ok = false; break; }
			const uint32_t prel = pr + cur;
			const uint32_t w = prel >> 7, bit = 1u << ((prel >> 2) & 31);
			if (w != cw) { flush<M>(cw, am, a0, a1); cw = w; am = a0 = a1 = 0; }
			am |= bit;
			if (prel & 1) a0 |= bit;
			if (prel & 2) a1 |= bit;
			cur += 4 + ul;
		}
		flush<M>(cw, am, a0, a1);
		if (!ok) return false;
		n = count > 0 ? count : 0;
		return true;
	}
	template <int M>
	__device__ __forceinline__ bool acl(int32_t &n)
	{
		int32_t count;
		if (!i32(count)) return false;
		for (int32_t i = 0; i < count; ++i) {
			if (cur + 4 > size) return false;
			const uint32_t at = pr + cur;
			cur += 4;
			if (!buf<M>(c->m_buf)) return false;
			if (!buf<M>(c->m_buf)) return false;
			mark_pos<M>(*c, c->m_acl, at);
		}
		n = count > 0 ? count : 0;
		return true;
	}
};

/* readResponse (zk-buffer.js:281-331), frames with a body the reference accepts */
/* returns 1 = decoded; 0 = not decodable here, marks may have been made; -1 = not decodable
 * here and nothing was marked */
template <int M>
__device__ int fast_response(ChunkSm &c, const DecArgs &a, uint32_t pr, uint32_t size, uint32_t s,
    FrameCols &o)
{
	if (size < 16) return -1;
	const uint8_t *t = c.tile;
	const int32_t xid = (int32_t)sm_be32(t, pr);
	const int32_t err = (int32_t)sm_be32(t, pr + 12);
	int op = special_xid(xid);
	if (op == ZKB_OP_NONE && a.xt_ops) {
		int64_t idx = (int64_t)xid - (int64_t)__ldg(a.xt_base + s);
		if (idx >= 0 && idx < (int64_t)__ldg(a.xt_count + s))
			op = __ldg(a.xt_ops + __ldg(a.xt_off + s) + (uint64_t)idx);
	}
	if (op == ZKB_OP_NONE) return -1;
	o.a0 = 0; o.a1 = 0;
	if (err == 0) {
		FC f; f.c = &c; f.pr = pr; f.size = size; f.cur = 16;
		switch (op) {
		case ZKB_OP_GET_CHILDREN:
			if (!f.strvec<M>(o.a0)) return false;
			break;
		case ZKB_OP_GET_CHILDREN2:
			if (!f.strvec<M>(o.a0)) return false;
			if (!f.stat<M>()) return false;
			break;
		case ZKB_OP_CREATE:
			if (!f.buf<M>(c.m_buf)) return false;
			break;
		case ZKB_OP_GET_ACL:
			if (!f.acl<M>(o.a0)) return false;
			if (!f.stat<M>()) return false;
			break;
		case ZKB_OP_GET_DATA:
			if (!f.buf<M>(c.m_buf)) return false;
			if (!f.stat<M>()) return false;
			break;
		case ZKB_OP_NOTIFICATION:
			if (!f.i32(o.a0)) return false;
			if (!f.i32(o.a1)) return false;
			if (!f.buf<M>(c.m_buf)) return false;
			break;
		case ZKB_OP_EXISTS:
		case ZKB_OP_SET_DATA:
			if (!f.stat<M>()) return false;
			break;
		case ZKB_OP_SET_WATCHES:
		case ZKB_OP_PING:
		case ZKB_OP_SYNC:
		case ZKB_OP_DELETE:
		case ZKB_OP_CLOSE_SESSION:
			break;
		default:
			return -1;
		}
	}
	o.xid = xid;
	o.err = err;
	o.zxid = (int64_t)(((uint64_t)sm_be32(t, pr + 4) << 32) | (uint64_t)sm_be32(t, pr + 8));
	o.meta = ZKB_ST_OK | ((uint32_t)(uint8_t)(int8_t)op << 8) | ((uint32_t)ZKB_KIND_RESPONSE << 16);
	return 1;
}

/* readRequest (zk-buffer.js:58-95,138-253), the opcodes it can read */
template <int M>
__device__ int fast_request(ChunkSm &c, uint32_t pr, uint32_t size, FrameCols &o)
{
	if (size < 8) return -1;
	const uint8_t *t = c.tile;
	const int32_t xid = (int32_t)sm_be32(t, pr);
	const int32_t op = (int32_t)sm_be32(t, pr + 4);
	FC f; f.c = &c; f.pr = pr; f.size = size; f.cur = 8;
	o.a0 = 0; o.a1 = 0; o.err = 0; o.zxid = 0;
	switch (op) {
	case ZKB_OP_GET_CHILDREN:
	case ZKB_OP_GET_CHILDREN2:
	case ZKB_OP_GET_DATA:
	case ZKB_OP_EXISTS: {
		if (!f.buf<M>(c.m_buf)) return false;
		if (!(f.cur < size)) return false;
		const uint32_t b = t[pr + f.cur];
		if (b > 1) return false;
		o.a0 = (int32_t)b;
		break;
	}
	case ZKB_OP_CREATE:
		if (!f.buf<M>(c.m_buf)) return false;
		if (!f.buf<M>(c.m_buf)) return false;
		if (!f.acl<M>(o.a1)) return false;
		if (!f.i32(o.a0)) return false;
		break;
	case ZKB_OP_DELETE:
		if (!f.buf<M>(c.m_buf)) return false;
		if (!f.i32(o.a0)) return false;
		break;
	case ZKB_OP_GET_ACL:
	case ZKB_OP_SYNC:
		if (!f.buf<M>(c.m_buf)) return false;
		break;
	case ZKB_OP_SET_DATA:
		if (!f.buf<M>(c.m_buf)) return false;
		if (!f.buf<M>(c.m_buf)) return false;
		if (!f.i32(o.a0)) return false;
		break;
	case ZKB_OP_SET_WATCHES:
		if (size < 16) return false;
		o.zxid = (int64_t)(((uint64_t)sm_be32(t, pr + 8) << 32) | (uint64_t)sm_be32(t, pr + 12));
		f.cur = 16;
		if (!f.strvec<M>(o.a0)) return false;
		if (!f.strvec<M>(o.a1)) return false;
		if (!f.strvec<M>(o.err)) return false;
		break;
	default: {
		/* readRequest throws 'Unsupported opcode' for everything else (zk-buffer.js:91-92),
		 * PING and CLOSE_SESSION included: same answer the generic parser gives */
		const int known = op_known(op) ? op : ZKB_OP_NONE;
		o.xid = xid;
		o.meta = ZKB_ST_BAD_DECODE | ((uint32_t)(uint8_t)(int8_t)known << 8) | ((uint32_t)ZKB_KIND_REQUEST << 16) |
		    (DETAIL(ZKB_CAUSE_UNSUPPORTED, ZKB_F_OPCODE) << 24);
		return 1;
	}
	}
	o.xid = xid;
	o.meta = ZKB_ST_OK | ((uint32_t)(uint8_t)(int8_t)op << 8) | ((uint32_t)ZKB_KIND_REQUEST << 16);
	return 1;
}

/* A frame the generic parser accepted and that lies inside the tile: record its rows */
__device__ void mark_from_parsed(ChunkSm &c, const Rd &rd, uint32_t payrel, uint64_t pay, int32_t flen,
    const Parsed &p)
{
	if (p.n_stat) mark_pos<FAST_MARK>(c, c.m_stat, payrel + p.stat_pos);
	if (p.n_fixed > 0) mark_pos<FAST_MARK>(c, c.m_buf, payrel + p.fb0_pos - 4);
	if (p.n_fixed > 1) mark_pos<FAST_MARK>(c, c.m_buf, payrel + p.fb1_pos - 4);
	if (p.vec_kind != VEC_NONE) {
		SideRows sr;
		sr.cs = &c; sr.payrel = payrel;
		Cur cu;
		cu.rd = &rd; cu.pay = pay; cu.size = (uint32_t)flen; cu.cur = p.vec_pos; cu.fail = 0;
		int32_t n; uint32_t nb = 0, na = 0;
		if (p.vec_kind == VEC_STR1) {
			walk_strvec<WALK_MARK>(cu, n, nb, &sr);
		} else if (p.vec_kind == VEC_ACL) {
			walk_acl<WALK_MARK>(cu, n, nb, na, &sr);
		} else {
			walk_strvec<WALK_MARK>(cu, n, nb, &sr);
			walk_strvec<WALK_MARK>(cu, n, nb, &sr);
			walk_strvec<WALK_MARK>(cu, n, nb, &sr);
		}
	}
}

__device__ __forceinline__ uint32_t warp_sum(uint32_t v)
{
	for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
	return v;
}
__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v, uint32_t lane)
{
	for (int d = 1; d < 32; d <<= 1) {
		uint32_t y = __shfl_up_sync(0xffffffffu, v, d);
		if (lane >= d) v += y;
	}
	return v;
}

/* Build with -DZK_PROFILE_PHASES to accumulate, per chunk, the SM cycles a decode warp spends
 * in each phase (ticket, staging wait, frame loop, counts, placement, emission) of the exact body into
 * zk_ctrl::prof; read them with zkb_debug_phase_cycles().  Off in normal builds. */
#ifdef ZK_PROFILE_PHASES
/* histogram of cycles from tile issue to AGG publish (1024-cycle buckets), chunks without /
 * with generic-parser frames */
__device__ unsigned long long zk_hist[3][64];
#define ZK_T(i) const long long zk_t##i = clock64()
#define ZK_ACC(slot, a, b) do { if (lane == 0) atomicAdd(&a_ctrl_prof[slot], (unsigned long long)((b) - (a))); } while (0)
#else
#define ZK_T(i) do { } while (0)
#define ZK_ACC(slot, a, b) do { } while (0)
#endif

/* Everything one warp does for one 8 KiB chunk.  `phase` is the parity of the warp's
 * mbarrier for this use (it flips every chunk). */
enum { XM_COUNT = 0, XM_EMIT = 1 };
template <int XM, int DIR>
__device__ __forceinline__ void decode_chunk_exact(const DecArgs &a, ChunkSm &c, uint64_t *bar, uint32_t phase,
    const uint32_t chunk, const uint32_t lane, const long long zk_t0)
{
#ifdef ZK_PROFILE_PHASES
	unsigned long long *a_ctrl_prof = a.ctrl->prof;
#endif
	(void)zk_t0;
	ZK_T(1);
	const uint64_t cs = (uint64_t)chunk << ZK_CHUNK_LOG2;
	const uint64_t avail = a.nbytes_pad - cs;
	const uint32_t wlen = avail < (uint64_t)ZK_WIN ? (uint32_t)avail : ZK_WIN;

	/* ---- stage the chunk: one TMA bulk copy (global -> shared, completion on an mbarrier)
	 * issued by lane 0; while it is in flight the warp clears its bitmaps and fetches what
	 * the splitter's scan says about the chunk, so those round trips overlap the copy ---- */
	if (lane == 0) {
		asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
		mbar_expect_tx(bar, wlen);
		bulk_g2s(c.tile, a.bytes + cs, wlen, bar);
		*reinterpret_cast<uint4 *>(c.tile + wlen) = make_uint4(0, 0, 0, 0);
	}
	for (uint32_t w = lane; w < ZK_BMW; w += 32) {
		c.m_buf[w] = 0; c.m_acl[w] = 0; c.m_stat[w] = 0; c.sub0[w] = 0; c.sub1[w] = 0; c.fstart[w] = 0;
	}
	Rd rd;
	rd.sm = c.tile; rd.lo = cs; rd.wlen = wlen; rd.g = a.bytes;

	const uint32_t gsh = ZK_CHUNK_LOG2 - a.glog2;
	const uint32_t g0 = chunk << gsh, g1 = (chunk + 1) << gsh;
	const uint32_t f0 = a.g_base[g0];
	const uint32_t f1 = (g1 < a.n_gran) ? a.g_base[g1] : (uint32_t)a.n_frames;

	const uint32_t nfr = f1 - f0;
	const bool last = (chunk == a.n_chunks - 1);
	const uint64_t ce = last ? a.nbytes : cs + ZK_CHUNK;
	const uint32_t sf = a.g_sf[g0];                  /* first session starting at or after cs */
	/* session boundaries from sf on, one per lane: boff[l] = sess_off[sf + l] */
	uint64_t boff = ZK_NONE;
	if (sf + lane <= a.nsess) boff = __ldg(a.sess_off + sf + lane);
	const uint64_t boff0 = __shfl_sync(0xffffffffu, boff, 0);
	const bool cont = (boff0 != cs);                 /* the chunk starts inside session sf-1 */
	/* first frame (global row) of the session the chunk starts in, from the splitter's scan */
	uint32_t sfirst = 0;
	if (cont) {
		const uint32_t src = a.g_ssrc[g0];
		if (src) sfirst = a.g_base[src - 1] + a.g_openlocal[(size_t)a.g_sel[src - 1] * a.n_gran + (src - 1)];
	}
	/* where the chain stands inside this lane's 256-byte sub-range, from the splitter */
	const uint64_t lo = cs + 256u * lane;
	const uint64_t hi = (lo + 256 < ce) ? lo + 256 : ce;
	uint64_t entry = ZK_NONE;
	if (lo < ce) {
		const uint32_t g = (uint32_t)(lo >> a.glog2);
		const uint64_t gstart = (uint64_t)g << a.glog2;
		const uint32_t subs = 1u << (a.glog2 - 8);
		const uint16_t v = a.g_sub[((size_t)a.g_sel[g] * a.n_gran + g) * subs + (uint32_t)((lo - gstart) >> 8)];
		if (v != 0xFFFFu) entry = gstart + v;
	}
	__syncwarp();
	ZK_T(2);
	mbar_wait(bar, phase);
	ZK_T(3);

	/* ---- frame starts: every lane repeats the loop of ZKDecodeStream._transform
	 * (zk-streams.js:47-64) on its own 256 bytes of the staged chunk, from where the splitter
	 * says the chain stands there; it marks frame starts in `fstart` (+ low position bits in
	 * sub0/sub1; a lane owns the two bitmap words of its sub-range) and writes the result of
	 * every session that ends under its hands ---- */
	{
		/* session context of the entry: jl = boundaries sess_off[sf + l], l < 32, strictly below it */
		const uint64_t e = (entry != ZK_NONE) ? entry : 0;
		uint32_t jl = 0;
#pragma unroll
		for (int step = 16; step > 0; step >>= 1) {
			const uint64_t bv = __shfl_sync(0xffffffffu, boff, (int)(jl + step - 1));
			if (bv < e) jl += step;
		}
		{
			const uint64_t bv = __shfl_sync(0xffffffffu, boff, 31);
			if (jl == 31 && bv < e) jl = 32;
		}
		uint64_t b_at = __shfl_sync(0xffffffffu, boff, (int)(jl < 32 ? jl : 31));       /* sess_off[sf + jl] */
		uint64_t b_next = __shfl_sync(0xffffffffu, boff, (int)(jl + 1 < 32 ? jl + 1 : 31)); /* sess_off[sf + jl + 1] */
		uint32_t cnt = 0;
		if (entry != ZK_NONE) {
			uint32_t sprev = sf - 1 + jl;          /* last session that starts strictly before the entry */
			if (jl >= 32) {                        /* > 32 session starts in this chunk: use the table */
				uint32_t l2 = sf + 31, h2 = a.nsess;
				while (h2 - l2 > 1) {
					const uint32_t mid = (l2 + h2) >> 1;
					if (__ldg(a.sess_off + mid) < entry) l2 = mid; else h2 = mid;
				}
				sprev = l2;
				b_at = __ldg(a.sess_off + sprev + 1);
			}
			if (jl + 1 >= 32 || jl >= 32) b_next = (sprev + 2 <= a.nsess) ? __ldg(a.sess_off + sprev + 2) : ZK_NONE;
			uint32_t s; uint64_t pos = entry, send;
			if (entry < b_at) { s = sprev; send = b_at; }          /* inside session sprev */
			else { s = sprev + 1; send = b_next; }                 /* the entry is where session sprev+1 starts */
			for (;;) {
				uint64_t bad = ZK_NONE;
				bool stop = false;
				for (;;) {
					if (pos == send) { stop = true; break; }
					if (pos >= hi) break;
					const uint64_t rem = send - pos;
					if (rem < 4) { stop = true; break; }                      /* zd_used < 4 */
					const uint32_t rel = (uint32_t)(pos - cs);
					const int32_t len = (int32_t)sm_be32(c.tile, rel);
					if (!len_ok(len)) { stop = true; bad = pos; break; }      /* :49-53 BAD_LENGTH */
					if (rem < 4 + (uint64_t)len) { stop = true; break; }      /* frame not complete yet */
					const uint32_t w = rel >> 7, bit = 1u << ((rel >> 2) & 31);
					c.fstart[w] |= bit;
					if (rel & 1) c.sub0[w] |= bit;
					if (rel & 2) c.sub1[w] |= bit;
					cnt++;
					pos += 4 + (uint64_t)len;
				}
				if (!stop) break;                 /* the chain leaves the sub-range alive */
				a.o.sess_consumed[s] = pos;
				a.o.sess_bad_length_at[s] = bad;
				/* next session, if it starts inside this sub-range */
				if (s + 1 >= a.nsess) break;
				const uint64_t so = send;
				if (!(so < hi)) break;
				s++;
				pos = so;
				send = __ldg(a.sess_off + s + 1);
			}
		}
		cnt = warp_sum(cnt);
		if (lane == 0 && cnt != nfr) atomicOr(&a.ctrl->error, ZK_DERR_COUNT);
		/* empty sessions at the very end of the batch start at nbytes, in no sub-range */
		if (last && lane == 0) {
			for (uint32_t sx = a.nsess; sx > 0 && __ldg(a.sess_off + sx - 1) == a.nbytes; --sx) {
				a.o.sess_consumed[sx - 1] = a.nbytes;
				a.o.sess_bad_length_at[sx - 1] = ZK_NONE;
			}
		}
	}
	__syncwarp();
	/* exclusive prefix of frame starts per bitmap word */
	{
		uint32_t run = 0;
		for (uint32_t w0 = 0; w0 < ZK_BMW; w0 += 32) {
			const uint32_t w = w0 + lane;
			const uint32_t fc = (w < ZK_BMW) ? __popc(c.fstart[w]) : 0u;
			const uint32_t inc = warp_incl_scan(fc, lane);
			if (w < ZK_BMW) c.fpre[w] = (uint16_t)(run + inc - fc);
			run += __shfl_sync(0xffffffffu, inc, 31);
		}
	}
	__syncwarp();
	/* rows before byte `relpos` of the chunk = frames that start below it */
	auto rank_below = [&](uint32_t relpos) -> uint32_t {
		if (relpos >= ZK_SLOTS * 4u) return nfr;
		const uint32_t w = relpos >> 7, b = (relpos >> 2) & 31;
		return (uint32_t)c.fpre[w] + __popc(c.fstart[w] & ((1u << b) - 1));
	};
	/* sess_first_frame for every session that starts inside the chunk (empty ones included) */
	for (uint32_t q = lane;; q += 32) {
		const uint32_t sq = sf + q;
		if (sq >= a.nsess) break;
		const uint64_t so = (q < 32) ? boff : __ldg(a.sess_off + sq);
		if (!(so < ce || last)) break;
		a.o.sess_first_frame[sq] = f0 + rank_below((uint32_t)(so - cs));
	}
	const bool one_session = cont && !(boff0 < ce || (last && sf < a.nsess));   /* no session starts inside */

	/* ---- frames: lane k takes the k-th, (k+32)-th ... frame start of the chunk ---- */
	Parsed lp;                       /* the chunk's long frame (reaches past the tile), if this lane has it */
	uint64_t l_foff = 0; int32_t l_flen = 0; uint32_t l_f = 0;
	lp.n_stat = lp.n_buf = lp.n_acl = 0;
	uint32_t n_slow = 0;
	for (uint32_t k0 = 0; k0 < nfr; k0 += 32) {
		const uint32_t k = k0 + lane;
		const bool valid = k < nfr;
		uint32_t rel = 0;
		if (valid) {
			/* word that holds the k-th frame start: last w with fpre[w] <= k */
			uint32_t lo = 0, hi = ZK_BMW;
			while (hi - lo > 1) {
				const uint32_t mid = (lo + hi) >> 1;
				if ((uint32_t)c.fpre[mid] <= k) lo = mid; else hi = mid;
			}
			const uint32_t b = __fns(c.fstart[lo], 0, (int)(k - c.fpre[lo]) + 1);
			rel = ((lo * 32 + b) << 2) | ((c.sub0[lo] >> b) & 1) | (((c.sub1[lo] >> b) & 1) << 1);
		}
		const uint32_t f = f0 + k;
		const uint64_t foff = cs + rel;
		const int32_t flen = valid ? (int32_t)sm_be32(c.tile, rel) : 0;
		/* which session, and where that session's frames begin */
		uint32_t s = sf - 1, first = sfirst;
		if (!one_session) {
			/* j = number of boundaries sess_off[sf + l], l < 32, at or below the frame */
			uint32_t j = 0;
#pragma unroll
			for (int step = 16; step > 0; step >>= 1) {
				const uint64_t bv = __shfl_sync(0xffffffffu, boff, (int)(j + step - 1));
				if (bv <= foff) j += step;
			}
			{
				const uint64_t bv = __shfl_sync(0xffffffffu, boff, 31);
				if (j == 31 && bv <= foff) j = 32;
			}
			uint64_t so = __shfl_sync(0xffffffffu, boff, (int)(j ? j - 1 : 0));
			if (j == 32 && valid) {
				/* more than 32 sessions start in this chunk before the frame: search the table */
				uint32_t lo = sf + 31, hi = a.nsess;          /* last s with sess_off[s] <= foff */
				while (hi - lo > 1) {
					const uint32_t mid = (lo + hi) >> 1;
					if (__ldg(a.sess_off + mid) <= foff) lo = mid; else hi = mid;
				}
				s = lo; so = __ldg(a.sess_off + lo);
				first = f0 + rank_below((uint32_t)(so - cs));
			} else if (j) {
				s = sf + j - 1;
				first = f0 + rank_below((uint32_t)(so - cs));
			}
		}
		if (!valid) continue;
		a.o.frame_off[f] = foff;
		a.o.frame_len[f] = flen;
		a.o.session[f] = s;
		const bool inwin = rel + 4u + (uint32_t)flen <= wlen;
		const bool server = DIR == 1 ? true : DIR == 2 ? false : (a.sess_flags && (__ldg(a.sess_flags + s) & ZKB_SESS_SERVER));
		const uint32_t hs = a.sess_hs ? __ldg(a.sess_hs + s) : 1u;
		const uint32_t idx = f - first;
		FrameCols o;
		int done = -1;
		if (inwin && idx >= hs) {
			if (server) {
				done = fast_request<FAST_MARK>(c, rel + 4, (uint32_t)flen, o);
				if (done == 0) fast_request<FAST_UNMARK>(c, rel + 4, (uint32_t)flen, o);
			} else {
				done = fast_response<FAST_MARK>(c, a, rel + 4, (uint32_t)flen, s, o);
				if (done == 0) fast_response<FAST_UNMARK>(c, a, rel + 4, (uint32_t)flen, s, o);
			}
		}
		if (done != 1) {
			Parsed p;
			n_slow++;
			parse_frame<DIR>(rd, a, foff, flen, s, idx, p);
			o.meta = p.meta; o.xid = p.xid; o.zxid = p.zxid; o.err = p.err; o.a0 = p.a0; o.a1 = p.a1;
			if (p.n_stat | p.n_buf | p.n_acl) {
				if (inwin) {
					mark_from_parsed(c, rd, rel + 4, foff + 4, flen, p);
				} else {
					lp = p; l_foff = foff; l_flen = flen; l_f = f;
				}
			}
		}
		a.o.meta[f] = o.meta;
		a.o.xid[f] = o.xid;
		a.o.zxid[f] = o.zxid;
		a.o.err[f] = o.err;
		a.o.a0[f] = o.a0;
		a.o.a1[f] = o.a1;
	}
	__syncwarp();
	ZK_T(4);
	n_slow = warp_sum(n_slow);
	if (lane == 0 && n_slow) atomicAdd(&a.ctrl->n_slow, (unsigned long long)n_slow);

	/* ---- side-table row counts of this chunk ---- */
	uint32_t nb = 0, na = 0, ns = 0;
	for (uint32_t w = lane; w < ZK_BMW; w += 32) {
		nb += __popc(c.m_buf[w]); na += __popc(c.m_acl[w]); ns += __popc(c.m_stat[w]);
	}
	nb = warp_sum(nb); na = warp_sum(na); ns = warp_sum(ns);
	const uint32_t tb = nb + warp_sum(lp.n_buf), ta = na + warp_sum(lp.n_acl), ts = ns + warp_sum(lp.n_stat);
	__syncwarp();

	/* ---- where the chunk's rows go.  The exact body never guesses: a first launch (XM_COUNT) only
	 * publishes what the chunk really has, a scan over all chunks turns that into every chunk's
	 * first rows, and a second launch (XM_EMIT) parses again and places them.  No chunk ever waits
	 * for another. ---- */
	ZK_T(5);
	if (XM == XM_COUNT) {
		if (lane == 0) a.c_rows[chunk] = make_uint4(1u, ts, tb, ta);
		return;
	}
	uint32_t e0, e1, e2;
	{
		const uint4 cb = a.c_base[chunk];
		e0 = cb.y; e1 = cb.z; e2 = cb.w;
	}
	ZK_T(6);
	ZK_ACC(0, zk_t0, zk_t1); ZK_ACC(1, zk_t1, zk_t2); ZK_ACC(2, zk_t2, zk_t3); ZK_ACC(3, zk_t3, zk_t4);
	ZK_ACC(4, zk_t4, zk_t5); ZK_ACC(5, zk_t5, zk_t6);
#ifdef ZK_PROFILE_PHASES
	if (lane == 0) {
		unsigned long long b = (unsigned long long)(zk_t5 - zk_t1) >> 11; if (b > 63) b = 63;
		atomicAdd(&zk_hist[n_slow ? 1 : 0][b], 1ull);
		b = (unsigned long long)(zk_t6 - zk_t5) >> 11; if (b > 63) b = 63;
		atomicAdd(&zk_hist[2][b], 1ull);
	}
#endif
	if (ts + tb + ta == 0) return;

	/* ---- emit: compact each bitmap, lane w owns bitmap word w (128 input bytes) ---- */
	const zkb_decoded &o = a.o;
	if (nb) {
		uint32_t run = e1;
		for (uint32_t w0 = 0; w0 < ZK_BMW; w0 += 32) {
			const uint32_t w = w0 + lane;
			uint32_t bits = (w < ZK_BMW) ? c.m_buf[w] : 0u;
			const uint32_t cnt = __popc(bits);
			const uint32_t inc = warp_incl_scan(cnt, lane);
			uint32_t row = run + inc - cnt;
			run += __shfl_sync(0xffffffffu, inc, 31);
			while (bits) {
				const uint32_t b = __ffs(bits) - 1;
				bits &= bits - 1;
				const uint32_t prel = ((w * 32 + b) << 2) | ((c.sub0[w] >> b) & 1) | (((c.sub1[w] >> b) & 1) << 1);
				const int32_t len = (int32_t)sm_be32(c.tile, prel);
				const uint32_t owner = f0 + c.fpre[w] + __popc(c.fstart[w] & ((1u << b) - 1)) - 1;
				if (row < o.cap_buf) {
					o.buf_frame[row] = owner;
					o.buf_off[row] = cs + prel + 4;
					o.buf_len[row] = len > 0 ? len : 0;
				}
				row++;
			}
		}
	}
	if (na) {
		uint32_t run = e2;
		for (uint32_t w0 = 0; w0 < ZK_BMW; w0 += 32) {
			const uint32_t w = w0 + lane;
			uint32_t bits = (w < ZK_BMW) ? c.m_acl[w] : 0u;
			const uint32_t cnt = __popc(bits);
			const uint32_t inc = warp_incl_scan(cnt, lane);
			uint32_t row = run + inc - cnt;
			run += __shfl_sync(0xffffffffu, inc, 31);
			while (bits) {
				const uint32_t b = __ffs(bits) - 1;
				bits &= bits - 1;
				const uint32_t prel = ((w * 32 + b) << 2) | ((c.sub0[w] >> b) & 1) | (((c.sub1[w] >> b) & 1) << 1);
				const uint32_t owner = f0 + c.fpre[w] + __popc(c.fstart[w] & ((1u << b) - 1)) - 1;
				if (row < o.cap_acl) {
					o.acl_frame[row] = owner;
					o.acl_perms[row] = (int32_t)sm_be32(c.tile, prel);
				}
				row++;
			}
		}
	}
	if (ns) {
		uint32_t run = e0;
		for (uint32_t w0 = 0; w0 < ZK_BMW; w0 += 32) {
			const uint32_t w = w0 + lane;
			uint32_t bits = (w < ZK_BMW) ? c.m_stat[w] : 0u;
			const uint32_t cnt = __popc(bits);
			const uint32_t inc = warp_incl_scan(cnt, lane);
			uint32_t row = run + inc - cnt;
			run += __shfl_sync(0xffffffffu, inc, 31);
			while (bits) {
				const uint32_t b = __ffs(bits) - 1;
				bits &= bits - 1;
				const uint32_t q = ((w * 32 + b) << 2) | ((c.sub0[w] >> b) & 1) | (((c.sub1[w] >> b) & 1) << 1);
				const uint32_t owner = f0 + c.fpre[w] + __popc(c.fstart[w] & ((1u << b) - 1)) - 1;
				if (row < o.cap_stat) {
					/* 68 bytes = 17 big-endian words from 18 aligned shared-memory words */
					const uint32_t *wp = reinterpret_cast<const uint32_t *>(c.tile + (q & ~3u));
					const uint32_t sh = (q & 3u) * 8;
					uint32_t x[18];
#pragma unroll
					for (int i = 0; i < 18; ++i) x[i] = wp[i];
					uint32_t v[17];
#pragma unroll
					for (int i = 0; i < 17; ++i) v[i] = bswap32(__funnelshift_r(x[i], x[i + 1], sh));
#define Q64(i) ((int64_t)(((uint64_t)v[i] << 32) | (uint64_t)v[(i) + 1]))
					o.stat_frame[row] = owner;
					o.czxid[row] = Q64(0); o.mzxid[row] = Q64(2); o.ctime[row] = Q64(4); o.mtime[row] = Q64(6);
					o.version[row] = (int32_t)v[8]; o.cversion[row] = (int32_t)v[9]; o.aversion[row] = (int32_t)v[10];
					o.ephemeralOwner[row] = Q64(11);
					o.dataLength[row] = (int32_t)v[13]; o.numChildren[row] = (int32_t)v[14];
					o.pzxid[row] = Q64(15);
#undef Q64
				}
				row++;
			}
		}
	}
	/* the long frame's rows come last: it is the chunk's last frame */
	if (lp.n_stat | lp.n_buf | lp.n_acl)
		emit_side(rd, a, l_foff, l_flen, lp, l_f, e0 + ns, e1 + nb, e2 + na);
	ZK_T(7);
	ZK_ACC(6, zk_t6, zk_t7);
#ifdef ZK_PROFILE_PHASES
	if (lane == 0) atomicAdd(&a_ctrl_prof[7], 1ull);
#endif
}

/* Everything the decode kernel wants to know about a chunk before it looks at the bytes, gathered
 * from the splitter's per-granule tables into two 16-byte records per chunk, so that a decode warp
 * gets it with one round of independent loads instead of a chain of dependent ones. */
__global__ void k_chunk_meta(DecArgs a, uint4 *meta0, uint4 *meta1)
{
	const uint32_t chunk = blockIdx.x * blockDim.x + threadIdx.x;
	if (chunk > a.n_chunks) return;
	if (chunk == a.n_chunks) {
		meta0[chunk] = make_uint4((uint32_t)a.n_frames, 0u, a.nsess, 0u);
		meta1[chunk] = a.g_rbase[a.n_gran];
		return;
	}
	const uint32_t g0 = chunk << (ZK_CHUNK_LOG2 - a.glog2);
	const uint64_t cs = (uint64_t)chunk << ZK_CHUNK_LOG2;
	const uint32_t sf = a.g_sf[g0];
	const uint64_t d = __ldg(a.sess_off + sf) - cs;
	uint32_t sfirst = 0;
	if (d != 0) {
		const uint32_t src = a.g_ssrc[g0];
		if (src) sfirst = a.g_base[src - 1] + a.g_openlocal[(size_t)a.g_sel[src - 1] * a.n_gran + (src - 1)];
	}
	meta0[chunk] = make_uint4(a.g_base[g0], sfirst, sf, d > 0xFFFFFFFFull ? 0xFFFFFFFFu : (uint32_t)d);
	meta1[chunk] = a.g_rbase[g0];
}

#include "zk_decode_direct.cuh"

/* One CTA = ZK_DEC_WARPS warps = that many consecutive chunks, chunk index from blockIdx.  The
 * warps of a CTA never wait for each other, and no chunk ever waits for another chunk.
 *   k_decode<DIR, false>   the ordinary pass: direct placement from the splitter's predicted rows
 *   k_decode<DIR, true>    the same body placing from scanned real counts (after a misprediction)
 *   k_decode_exact<DIR, XM_COUNT / XM_EMIT>   the exact body, for the chunks that need it
 * DIR: 1 = every session of the batch is client->server, 2 = every one is server->client, 0 = mixed. */
#ifndef ZK_DEC_MINBLOCKS
#define ZK_DEC_MINBLOCKS 5
#endif
template <int DIR, bool REDO>
__global__ void __launch_bounds__(ZK_DEC_WARPS * 32, ZK_DEC_MINBLOCKS) k_decode(DecArgs a)
{
	__shared__ __align__(16) ChunkSm sm_all[ZK_DEC_WARPS];
	__shared__ __align__(8) uint64_t s_bar[ZK_DEC_WARPS];
	const uint32_t lane = threadIdx.x & 31;
	ChunkSm &c = sm_all[threadIdx.x >> 5];
	uint64_t *bar = &s_bar[threadIdx.x >> 5];
	const uint32_t chunk = a.chunk0 + blockIdx.x * ZK_DEC_WARPS + (threadIdx.x >> 5);
	if (chunk >= a.n_chunks) return;
	if (lane == 0) mbar_init(bar, 1);
	__syncwarp();
	decode_chunk_direct<DIR, REDO>(a, c, bar, chunk, lane);
}

template <int DIR, int XM>
__global__ void __launch_bounds__(ZK_DEC_WARPS * 32, ZK_DEC_MINBLOCKS) k_decode_exact(DecArgs a)
{
	__shared__ __align__(16) ChunkSm sm_all[ZK_DEC_WARPS];
	__shared__ __align__(8) uint64_t s_bar[ZK_DEC_WARPS];
	const uint32_t lane = threadIdx.x & 31;
	ChunkSm &c = sm_all[threadIdx.x >> 5];
	uint64_t *bar = &s_bar[threadIdx.x >> 5];
#ifdef ZK_PROFILE_PHASES
	const long long t0 = clock64();
#else
	const long long t0 = 0;
#endif
	const uint32_t chunk = a.chunk0 + blockIdx.x * ZK_DEC_WARPS + (threadIdx.x >> 5);
	if (chunk >= a.n_chunks) return;
	if (!a.force_exact && !a.c_rows[chunk].x) return;      /* the direct body has this chunk */
	if (lane == 0) mbar_init(bar, 1);
	__syncwarp();
	decode_chunk_exact<XM, DIR>(a, c, bar, 0, chunk, lane, t0);
}

/* Exclusive prefix of the chunks' real row counts from chunk0 on, seeded with the rows before
 * chunk0 (all of which are where they belong); the totals go to the control block.  One block:
 * this only runs after a misprediction, or with the prediction switched off. */
/* One warp per listed children vector, one tile of shared memory per warp and nothing else, so that
 * many more of these walks are resident than the decode kernel's five CTAs per SM would allow: a
 * single walk is a chain of dependent tile loads and short loops, and only numbers hide that.  A
 * vector that does not keep what its header promised flags its chunk for the exact route, exactly
 * as the decode pass does for every other frame. */
#define ZK_LONG_WARPS 2
#define ZK_LONG_LIST  768u     /* element offsets of one tile (an 8 KiB tile of 8..24-byte names holds about 440) */
__global__ void __launch_bounds__(ZK_LONG_WARPS * 32) k_long_children(DecArgs a, const LongItem *items, uint32_t n)
{
	__shared__ __align__(128) uint8_t s_tile[ZK_LONG_WARPS][ZK_WIN + 16];
	__shared__ __align__(8) uint64_t s_bar[ZK_LONG_WARPS];
	__shared__ uint16_t s_list[ZK_LONG_WARPS][ZK_LONG_LIST];
	const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
	const uint32_t i = blockIdx.x * ZK_LONG_WARPS + w;
	if (i >= n) return;
	if (lane == 0) mbar_init(&s_bar[w], 1);
	__syncwarp();
	const LongItem it = items[i];
	const bool ok = long_children(a, s_tile[w], s_list[w], ZK_LONG_LIST, &s_bar[w], 0u, it.foff, it.flen, it.frame, it.brow0, it.srow,
	    it.with_stat != 0, lane);
	if (!ok && lane == 0) {
		a.c_rows[it.chunk].x = 1u;
		atomicOr(&a.ctrl->error, ZK_DERR_MISPREDICT);
		atomicMin(&a.ctrl->first_bad, it.chunk);
	}
}

__global__ void __launch_bounds__(1024) k_rows_scan(const uint4 *rows, uint4 *base, uint32_t chunk0, uint32_t n_chunks,
    uint4 seed, zk_ctrl *ctrl)
{
	__shared__ uint32_t s_tot[3][1024];
	const uint32_t n = n_chunks - chunk0;
	const uint32_t per = (n + 1023u) / 1024u;
	const uint32_t lo = chunk0 + threadIdx.x * per;
	const uint32_t hi = (lo + per < n_chunks) ? lo + per : n_chunks;
	uint32_t x0 = 0, x1 = 0, x2 = 0;
	for (uint32_t i = lo; i < hi; ++i) { const uint4 r = rows[i]; x0 += r.y; x1 += r.z; x2 += r.w; }
	s_tot[0][threadIdx.x] = x0; s_tot[1][threadIdx.x] = x1; s_tot[2][threadIdx.x] = x2;
	__syncthreads();
	/* Hillis-Steele over the 1024 partial sums */
	for (uint32_t d = 1; d < 1024; d <<= 1) {
		uint32_t y0 = 0, y1 = 0, y2 = 0;
		if (threadIdx.x >= d) { y0 = s_tot[0][threadIdx.x - d]; y1 = s_tot[1][threadIdx.x - d]; y2 = s_tot[2][threadIdx.x - d]; }
		__syncthreads();
		s_tot[0][threadIdx.x] += y0; s_tot[1][threadIdx.x] += y1; s_tot[2][threadIdx.x] += y2;
		__syncthreads();
	}
	uint32_t b0 = seed.x + s_tot[0][threadIdx.x] - x0, b1 = seed.y + s_tot[1][threadIdx.x] - x1,
	    b2 = seed.z + s_tot[2][threadIdx.x] - x2;
	for (uint32_t i = lo; i < hi; ++i) {
		const uint4 r = rows[i];
		base[i] = make_uint4(0, b0, b1, b2);
		b0 += r.y; b1 += r.z; b2 += r.w;
	}
	if (threadIdx.x == 1023) {
		ctrl->n_stat = (unsigned long long)seed.x + s_tot[0][1023];
		ctrl->n_buf = (unsigned long long)seed.y + s_tot[1][1023];
		ctrl->n_acl = (unsigned long long)seed.z + s_tot[2][1023];
	}
}

/* ================================================================================== */
/* xid -> opcode table from a decoded request stream                                   */

__global__ void k_xid_scatter(const zkb_decoded req, uint64_t n_frames, uint32_t nsess,
    const int32_t *base, const uint32_t *count, const uint64_t *off, unsigned long long *tmp, uint64_t nops)
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
	/* the layout arrays are the caller's device memory, which the host cannot inspect: an entry
	 * that points past the table (off + count > nops) is dropped, never written */
	const uint64_t slot = off[s] + (uint64_t)idx;
	if (slot >= nops) return;
	atomicMax(tmp + slot, key);
}

/* The same without atomics, for the ordinary case that no two requests of a connection share an
 * xid: every request stores its opcode byte, and the stores are counted.  If afterwards as many
 * slots hold an opcode as requests stored one, no slot was written twice and the table is exact;
 * otherwise the caller runs the ordered build above. */
__global__ void __launch_bounds__(256) k_xid_store(const zkb_decoded req, uint64_t n_frames, uint32_t nsess,
    const int32_t *base, const uint32_t *count, const uint64_t *off, int8_t *ops, uint64_t nops, zk_ctrl *ctrl)
{
	__shared__ uint32_t s_n;
	if (threadIdx.x == 0) s_n = 0;
	__syncthreads();
	uint32_t mine = 0;
	const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
	for (uint64_t f = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; f < n_frames; f += stride) {
		const uint32_t m = req.meta[f];
		if (ZKB_META_KIND(m) != ZKB_KIND_REQUEST || req.frame_len[f] < 8) continue;
		const uint32_t s = req.session[f];
		if (s >= nsess) continue;
		const int64_t idx = (int64_t)req.xid[f] - (int64_t)base[s];
		if (idx < 0 || idx >= (int64_t)count[s]) continue;
		const uint64_t slot = off[s] + (uint64_t)idx;
		if (slot >= nops) continue;       /* a layout entry that points past the table is dropped, as above */
		ops[slot] = (int8_t)((m >> 8) & 0xff);
		mine++;
	}
	for (int d = 16; d > 0; d >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, d);
	if ((threadIdx.x & 31) == 0 && mine) atomicAdd(&s_n, mine);
	__syncthreads();
	if (threadIdx.x == 0 && s_n) atomicAdd(&ctrl->xid_stored, (unsigned long long)s_n);
}

/* slots that hold an opcode (anything but ZKB_OP_NONE), 16 per thread */
__global__ void __launch_bounds__(256) k_xid_filled(const int8_t *ops, uint64_t nops, zk_ctrl *ctrl)
{
	__shared__ uint32_t s_n;
	if (threadIdx.x == 0) s_n = 0;
	__syncthreads();
	uint32_t c = 0;
	const uint64_t stride = (uint64_t)gridDim.x * blockDim.x * 16;
	const bool aligned = ((uintptr_t)ops & 15) == 0;
	for (uint64_t i = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) * 16; i < nops; i += stride) {
		if (aligned && i + 16 <= nops) {
			const uint4 v = *reinterpret_cast<const uint4 *>(ops + i);
			const uint32_t w[4] = { v.x, v.y, v.z, v.w };
#pragma unroll
			for (int k = 0; k < 4; ++k)
				c += 4u - (__popc(__vcmpeq4(w[k], 0x80808080u)) >> 3);     /* 0xff per byte that is ZKB_OP_NONE */
		} else {
			for (uint64_t k = i; k < nops && k < i + 16; ++k) c += ops[k] != (int8_t)ZKB_OP_NONE;
		}
	}
	for (int d = 16; d > 0; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
	if ((threadIdx.x & 31) == 0 && c) atomicAdd(&s_n, c);
	__syncthreads();
	if (threadIdx.x == 0 && s_n) atomicAdd(&ctrl->xid_filled, (unsigned long long)s_n);
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
	/* development knob: ZKBATCH_GLOG2=11..13 fixes the granule size */
	static int forced = -1;
	if (forced < 0) { const char *e = getenv("ZKBATCH_GLOG2"); forced = e ? atoi(e) : 0; }
	if (forced >= ZK_GRAN_MIN_LOG2 && forced <= ZK_GRAN_MAX_LOG2 && forced <= (int)ZK_CHUNK_LOG2) return (uint32_t)forced;
	/* aim for ~256K granules so the splitter has enough threads, within [2 KiB, 8 KiB] */
	uint64_t want = (nbytes + 262143) / 262144;
	uint32_t l = ZK_GRAN_MIN_LOG2;
	while (l < ZK_GRAN_MAX_LOG2 && l < ZK_CHUNK_LOG2 && (1ull << l) < want) l++;
	return l;
}

struct ScratchLayout {
	size_t g_assumed, g_out, g_carry, g_cnt, g_sel, g_base, g_dirty, g_src, g_sf, g_openlocal, g_ssrc, g_sub, g_rows, g_rbase, blk, c_rows, c_base, meta0, meta1, long_items, total;
	uint32_t glog2, n_gran, n_chunks, n_blk;
};

static ScratchLayout zk_layout(uint64_t nbytes)
{
	ScratchLayout L;
	memset(&L, 0, sizeof (L));
	L.glog2 = zk_pick_glog2(nbytes);
	L.n_gran = (uint32_t)((nbytes + (1ull << L.glog2) - 1) >> L.glog2);
	L.n_chunks = (uint32_t)((nbytes + ZK_CHUNK - 1) >> ZK_CHUNK_LOG2);
	L.n_blk = (L.n_gran + SCAN_THREADS - 1) / SCAN_THREADS;
	size_t o = 0;
	L.g_assumed = o; o += zk_align_up((size_t)L.n_gran * 16, 256);
	L.g_out = o;     o += zk_align_up((size_t)L.n_gran * 16, 256);
	L.g_carry = o;   o += zk_align_up((size_t)L.n_gran * 8, 256);
	L.g_cnt = o;     o += zk_align_up((size_t)L.n_gran * 8, 256);
	L.g_sel = o;     o += zk_align_up((size_t)L.n_gran, 256);
	L.g_base = o;    o += zk_align_up((size_t)(L.n_gran + 1) * 4, 256);
	L.g_dirty = o;   o += zk_align_up((size_t)L.n_gran, 256);
	L.g_src = o;     o += zk_align_up((size_t)L.n_gran * 4, 256);
	L.g_sf = o;      o += zk_align_up((size_t)L.n_gran * 4, 256);
	L.g_openlocal = o; o += zk_align_up((size_t)L.n_gran * 8, 256);
	L.g_ssrc = o;    o += zk_align_up((size_t)L.n_gran * 4, 256);
	L.g_sub = o;     o += zk_align_up((size_t)L.n_gran * 2 * ((size_t)1 << (L.glog2 - 8)) * 2, 256);
	L.g_rows = o;    o += zk_align_up((size_t)L.n_gran * 2 * 16, 256);
	L.g_rbase = o;   o += zk_align_up((size_t)(L.n_gran + 1) * 16, 256);
	L.blk = o;       o += zk_align_up((size_t)(L.n_blk + 1) * sizeof (ScanVal), 256);
	L.c_rows = o;    o += zk_align_up((size_t)L.n_chunks * 16, 256);
	L.c_base = o;    o += zk_align_up((size_t)L.n_chunks * 16, 256);
	L.meta0 = o;     o += zk_align_up((size_t)(L.n_chunks + 1) * 16, 256);
	L.meta1 = o;     o += zk_align_up((size_t)(L.n_chunks + 1) * 16, 256);
	L.long_items = o; o += zk_align_up((size_t)L.n_chunks * sizeof (LongItem), 256);
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
    const uint8_t *d_bytes, uint64_t nbytes, const zkb_sessions *sess, const zkb_xidtab *xid)
{
	uint8_t *S = (uint8_t *)ctx->d_scratch;
	memset(&sa, 0, sizeof (sa));
	sa.bytes = d_bytes; sa.nbytes = nbytes; sa.sess_off = sess->off; sa.nsess = sess->nsessions;
	sa.sess_flags = sess->flags; sa.sess_hs = sess->handshake_frames; sa.glog2 = L.glog2; sa.n_gran = L.n_gran;
	if (xid && xid->ops) { sa.xt_base = xid->base; sa.xt_count = xid->count; sa.xt_off = xid->off; sa.xt_ops = xid->ops; }
	sa.g_rows = (uint4 *)(S + L.g_rows); sa.g_rbase = (uint4 *)(S + L.g_rbase);
	sa.g_assumed = (uint64_t *)(S + L.g_assumed); sa.g_out = (uint64_t *)(S + L.g_out);
	sa.g_carry = (uint64_t *)(S + L.g_carry); sa.g_cnt = (uint32_t *)(S + L.g_cnt);
	sa.g_base = (uint32_t *)(S + L.g_base); sa.g_dirty = (uint8_t *)(S + L.g_dirty);
	sa.g_src = (uint32_t *)(S + L.g_src); sa.g_sel = (uint8_t *)(S + L.g_sel);
	sa.g_sf = (uint32_t *)(S + L.g_sf); sa.g_openlocal = (uint32_t *)(S + L.g_openlocal);
	sa.g_ssrc = (uint32_t *)(S + L.g_ssrc); sa.g_sub = (uint16_t *)(S + L.g_sub);
	sa.ctrl = ctx->d_ctrl;
	return ZKB_OK;
}

/* Phase A: find every frame boundary; on success *n_frames is exact. */
int zk_split_device(zkb_ctx *ctx, const uint8_t *d_bytes, uint64_t nbytes,
    const zkb_sessions *sess, const zkb_xidtab *xid, uint64_t *n_frames)
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
	k_check_sessions<<<(nsess + 1 + 255) / 256, 256, 0, ctx->stream>>>(sess->off, sess->flags, nsess, nbytes,
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
	zk_fill_split_args(ctx, sa, L, d_bytes, nbytes, sess, xid);
	ScanVal *blk = (ScanVal *)((uint8_t *)ctx->d_scratch + L.blk);

	const uint32_t split_blocks = (L.n_gran + 127) / 128;
	ZK_EV(ctx, 0);
	k_split<MODE_SPEC><<<split_blocks, 128, 0, ctx->stream>>>(sa);
	LAUNCH_CHECK(ctx);
	/* Ordinary traffic needs exactly one fix-up round, so SPEC, scan, FIX, scan are queued back to back
	 * and the host looks at the result once; FIX returns at once in granules that are not dirty. */
	for (uint32_t iter = 0;; ++iter) {
		/* the scan queued behind the first fix-up round runs only if the first scan left anything to fix */
		const unsigned int *gate = (iter == 1) ? &ctx->d_ctrl->first_dirty[0] : NULL;
		k_scan_reduce<<<L.n_blk, SCAN_THREADS, 0, ctx->stream>>>(sa, blk, gate);
		LAUNCH_CHECK(ctx);
		k_scan_blocks<<<1, SCAN_THREADS, 0, ctx->stream>>>(blk, L.n_blk, gate);
		LAUNCH_CHECK(ctx);
		k_scan_apply<<<L.n_blk, SCAN_THREADS, 0, ctx->stream>>>(sa, blk, gate);
		LAUNCH_CHECK(ctx);
		if (iter == 0) {
			k_split_round<<<1, 1, 0, ctx->stream>>>(ctx->d_ctrl);
			LAUNCH_CHECK(ctx);
			k_split<MODE_FIX><<<split_blocks, 128, 0, ctx->stream>>>(sa);
			LAUNCH_CHECK(ctx);
			continue;
		}
		ZK_EV(ctx, 1);
		if ((rc = zk_read_ctrl(ctx))) return rc;
		if (ctx->h_ctrl->error & ZK_DERR_SESSIONS) {
			snprintf(ctx->err, sizeof (ctx->err), "zkb_decode: session offsets invalid");
			return ZKB_EINVAL;
		}
		if (iter == 1) {
			ctx->split_iters[0][0] = ctx->h_ctrl->first_dirty[0];
			ctx->split_iters[0][1] = ctx->h_ctrl->first_dirty[1];
			if (ctx->h_ctrl->first_dirty[0]) ctx->split_fixups++;
		}
		if (iter < 32) {
			ctx->split_iters[iter][0] = ctx->h_ctrl->n_dirty;
			ctx->split_iters[iter][1] = ctx->h_ctrl->n_rewalk;
			if (iter < 31) ctx->split_iters[iter + 1][0] = ctx->split_iters[iter + 1][1] = 0;
		}
		if (ctx->h_ctrl->n_dirty == 0) break;
		if (iter > L.n_gran + 2) {
			snprintf(ctx->err, sizeof (ctx->err), "zkb_decode: splitter did not converge");
			return ZKB_EINTERNAL;
		}
		ctx->split_fixups++;
		const bool rewalk = ctx->h_ctrl->n_rewalk != 0;
		ZK_CUDA(ctx, cudaMemsetAsync(&ctx->d_ctrl->n_dirty, 0, 3 * sizeof (unsigned int), ctx->stream));
		if (rewalk) {
			k_split<MODE_FIX><<<split_blocks, 128, 0, ctx->stream>>>(sa);
			LAUNCH_CHECK(ctx);
		}
	}
	*n_frames = ctx->h_ctrl->n_frames;
	for (int k = 0; k < 3; ++k) ctx->pred_rows[k] = ctx->h_ctrl->pred_rows[k];
	ctx->dir = ctx->h_ctrl->dir_mask == 1u ? 1 : ctx->h_ctrl->dir_mask == 2u ? 2 : 0;
	memcpy(ctx->split_prof, ctx->h_ctrl->prof, sizeof (ctx->split_prof));
	return ZKB_OK;
}

#ifdef ZK_PROFILE_PHASES
extern "C" __attribute__((visibility("default"))) void zkb_debug_hist(unsigned long long out[192], int reset)
{
	cudaMemcpyFromSymbol(out, zk_hist, sizeof (zk_hist));
	if (reset) { static unsigned long long z[192]; cudaMemcpyToSymbol(zk_hist, z, sizeof (z)); }
}
#endif

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
	zk_fill_split_args(ctx, sa, L, d_bytes, nbytes, sess, xid);
	ZK_CUDA(ctx, cudaMemsetAsync(&ctx->d_ctrl->n_long, 0, sizeof (unsigned long long) * 10, ctx->stream));
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
	da.sess_off = sess->off; da.nsess = nsess;
	da.g_base = sa.g_base; da.g_sf = sa.g_sf; da.g_ssrc = sa.g_ssrc;
	da.g_openlocal = sa.g_openlocal; da.g_sub = sa.g_sub; da.g_sel = sa.g_sel;
	da.glog2 = L.glog2; da.n_gran = L.n_gran; da.n_chunks = L.n_chunks;
	da.n_frames = n_frames; da.o = *out;
	da.c_rows = (uint4 *)(S + L.c_rows); da.c_base = (uint4 *)(S + L.c_base); da.ctrl = ctx->d_ctrl;
	da.g_rbase = sa.g_rbase;
	const unsigned dec_blocks = (L.n_chunks + ZK_DEC_WARPS - 1) / ZK_DEC_WARPS;
	const unsigned dec_threads = ZK_DEC_WARPS * 32;
	ZK_EV(ctx, 4);
	{
		uint4 *m0 = (uint4 *)(S + L.meta0), *m1 = (uint4 *)(S + L.meta1);
		k_chunk_meta<<<(L.n_chunks + 1 + 255) / 256, 256, 0, ctx->stream>>>(da, m0, m1);
		LAUNCH_CHECK(ctx);
		da.meta0 = m0; da.meta1 = m1;
	}
	bool redo = ctx->no_predict != 0;
	unsigned long long slow_first = 0;
	bool have_slow = false;
	if (!redo) {
		/* the ordinary pass: every chunk places its rows where the splitter predicted */
		ZK_CUDA(ctx, cudaMemsetAsync(&ctx->d_ctrl->first_bad, 0xFF, sizeof (unsigned int), ctx->stream));
		da.long_items = (LongItem *)(S + L.long_items);
		if (ctx->dir == 1) k_decode<1, false><<<dec_blocks, dec_threads, 0, ctx->stream>>>(da);
		else if (ctx->dir == 2) k_decode<2, false><<<dec_blocks, dec_threads, 0, ctx->stream>>>(da);
		else k_decode<0, false><<<dec_blocks, dec_threads, 0, ctx->stream>>>(da);
		LAUNCH_CHECK(ctx);
		if ((rc = zk_read_ctrl(ctx))) return rc;
		if (ctx->h_ctrl->n_long) {
			/* children vectors longer than a tile were listed, not walked: their own launch */
			const uint32_t nl = (uint32_t)ctx->h_ctrl->n_long;
			k_long_children<<<(nl + ZK_LONG_WARPS - 1) / ZK_LONG_WARPS, ZK_LONG_WARPS * 32, 0, ctx->stream>>>(da,
			    da.long_items, nl);
			LAUNCH_CHECK(ctx);
			if ((rc = zk_read_ctrl(ctx))) return rc;
		}
		if (ctx->h_ctrl->error == ZK_DERR_MISPREDICT) {
			/* Some frame is not what its header promised.  Every chunk before the first one that
			 * noticed has its rows in the right place (its own and all earlier counts matched);
			 * the rest of the batch is placed again from the chunks' real counts. */
			ctx->mispredicts++;
			redo = true;
			da.chunk0 = ctx->h_ctrl->first_bad < L.n_chunks ? ctx->h_ctrl->first_bad : 0;
			slow_first = ctx->h_ctrl->n_slow;      /* the first pass already saw every frame */
			have_slow = true;
			ZK_CUDA(ctx, cudaMemsetAsync(&ctx->d_ctrl->error, 0, sizeof (unsigned int), ctx->stream));
		} else {
			/* the prediction held: its totals are the real ones */
			ctx->h_ctrl->n_stat = ctx->pred_rows[0]; ctx->h_ctrl->n_buf = ctx->pred_rows[1];
			ctx->h_ctrl->n_acl = ctx->pred_rows[2];
		}
	}
	da.long_items = NULL;       /* the passes below walk long vectors in place */
	if (redo) {
		/* exact route, no chunk waits for another: (1) the chunks the direct body could not vouch for
		 * (all of them when the prediction is switched off) count their rows with the exact body,
		 * (2) one scan turns every chunk's real counts into its first rows, (3) the direct body places
		 * the sound chunks again, the exact body the others. */
		const uint32_t todo = L.n_chunks - da.chunk0;
		const unsigned blocks = (todo + ZK_DEC_WARPS - 1) / ZK_DEC_WARPS;
		da.force_exact = ctx->no_predict ? 1u : 0u;
		uint4 seed = make_uint4(0, 0, 0, 0);
		if (da.chunk0) {
			uint4 rb;
			ZK_CUDA(ctx, cudaMemcpyAsync(&rb, sa.g_rbase + ((size_t)da.chunk0 << (ZK_CHUNK_LOG2 - L.glog2)), sizeof (rb),
			    cudaMemcpyDeviceToHost, ctx->stream));
			ZK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
			seed = rb;
		}
#define ZK_LAUNCH_EXACT(XM)                                                                              \
		do {                                                                                             \
			if (ctx->dir == 1) k_decode_exact<1, XM><<<blocks, dec_threads, 0, ctx->stream>>>(da);       \
			else if (ctx->dir == 2) k_decode_exact<2, XM><<<blocks, dec_threads, 0, ctx->stream>>>(da);  \
			else k_decode_exact<0, XM><<<blocks, dec_threads, 0, ctx->stream>>>(da);                     \
			LAUNCH_CHECK(ctx);                                                                           \
		} while (0)
		ZK_LAUNCH_EXACT(XM_COUNT);
		k_rows_scan<<<1, 1024, 0, ctx->stream>>>(da.c_rows, da.c_base, da.chunk0, L.n_chunks, seed, ctx->d_ctrl);
		LAUNCH_CHECK(ctx);
		if (!da.force_exact) {
			if (ctx->dir == 1) k_decode<1, true><<<blocks, dec_threads, 0, ctx->stream>>>(da);
			else if (ctx->dir == 2) k_decode<2, true><<<blocks, dec_threads, 0, ctx->stream>>>(da);
			else k_decode<0, true><<<blocks, dec_threads, 0, ctx->stream>>>(da);
			LAUNCH_CHECK(ctx);
		}
		ZK_LAUNCH_EXACT(XM_EMIT);
#undef ZK_LAUNCH_EXACT
	}
	ZK_EV(ctx, 5);
	if (redo && (rc = zk_read_ctrl(ctx))) return rc;
	if (have_slow) ctx->h_ctrl->n_slow = slow_first;
	if (ctx->timing) ZK_CUDA(ctx, cudaEventSynchronize(ctx->ev[5]));
	if (ctx->timing) {
		cudaEventElapsedTime(&ctx->phase_ms[0], ctx->ev[0], ctx->ev[1]);
		ctx->phase_ms[1] = 0.0f;   /* the frame rows are written by the decode kernel itself */
		cudaEventElapsedTime(&ctx->phase_ms[2], ctx->ev[4], ctx->ev[5]);
	}
	if (ctx->h_ctrl->error) {
		snprintf(ctx->err, sizeof (ctx->err), "zkb_decode: device error bits 0x%x", ctx->h_ctrl->error);
		return ZKB_EINTERNAL;
	}
	ctx->slow_frames = ctx->h_ctrl->n_slow;
	memcpy(ctx->prof, ctx->h_ctrl->prof, sizeof (ctx->prof));
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
	int rc = zk_split_device(ctx, d_bytes, nbytes, sess, xid, &n_frames);
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
	/* first without atomics; exact unless two requests of a connection share an xid */
	ZK_CUDA(ctx, cudaMemsetAsync(d_ops_out, (int)(uint8_t)(int8_t)ZKB_OP_NONE, n, ctx->stream));
	ZK_CUDA(ctx, cudaMemsetAsync(&ctx->d_ctrl->xid_stored, 0, 2 * sizeof (unsigned long long), ctx->stream));
	if (req->n_frames) {
		const uint64_t want_s = (req->n_frames + 255) / 256, want_f = (n + 4095) / 4096;
		k_xid_store<<<(unsigned)(want_s < 148 * 32 ? want_s : 148 * 32), 256, 0, ctx->stream>>>(*req,
		    req->n_frames, nsessions, layout->base, layout->count, layout->off, d_ops_out, n, ctx->d_ctrl);
		LAUNCH_CHECK(ctx);
		k_xid_filled<<<(unsigned)(want_f < 148 * 16 ? want_f : 148 * 16), 256, 0, ctx->stream>>>(d_ops_out, n, ctx->d_ctrl);
		LAUNCH_CHECK(ctx);
	}
	if ((rc = zk_read_ctrl(ctx))) return rc;
	if (ctx->h_ctrl->xid_stored == ctx->h_ctrl->xid_filled) return ZKB_OK;
	/* some slot was written more than once: the last request in stream order must win */
	if ((rc = zk_ensure(ctx, &ctx->d_scratch, &ctx->d_scratch_bytes, n * 8 + 256, 0))) return rc;
	unsigned long long *tmp = (unsigned long long *)ctx->d_scratch;
	ZK_CUDA(ctx, cudaMemsetAsync(tmp, 0, n * 8, ctx->stream));
	if (req->n_frames) {
		k_xid_scatter<<<(unsigned)((req->n_frames + 255) / 256), 256, 0, ctx->stream>>>(*req,
		    req->n_frames, nsessions, layout->base, layout->count, layout->off, tmp, n);
		LAUNCH_CHECK(ctx);
	}
	k_xid_finish<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(tmp, n, d_ops_out);
	LAUNCH_CHECK(ctx);
	ZK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
	return ZKB_OK;
}

/* ================================================================================== */
/* xid table build for callers that cannot allocate                                    */

/*
 * zkb_xidtab_build_dev borrows the context's scratch, which the context obtained with
 * cudaMalloc.  Two kinds of user of this library cannot go down that road: a host
 * process that has already given all device memory to its own pool (an inference server
 * or a database engine embedding the codec: a later cudaMalloc fails there, or breaks up
 * the pool), and a latency-bound service loop in which cudaMalloc/cudaFree are off limits
 * because they stall every stream on the device.  For them this entry point keeps the
 * one piece of scratch the build needs — a 64-bit slot per xid — in a fixed array that
 * is part of the library image.  Each GPU that loads libzkbatch.so carries these
 * ZKB_NOALLOC_MAX_XIDS * 8 bytes (8 MiB) until the process ends.
 */
static __device__ __align__(256) unsigned long long zk_noalloc_slots[ZKB_NOALLOC_MAX_XIDS];

#define ZK_NOALLOC_MAX_DEVICES 64
static unsigned long long *zk_noalloc_addr[ZK_NOALLOC_MAX_DEVICES];   /* device address per GPU, 0 = not looked up yet */

static int zk_noalloc_fail(char *errbuf, size_t errcap, int rc, const char *what, cudaError_t e)
{
	if (errbuf && errcap) {
		if (e != cudaSuccess) snprintf(errbuf, errcap, "zkb_xidtab_build_noalloc: %s: %s", what, cudaGetErrorString(e));
		else snprintf(errbuf, errcap, "zkb_xidtab_build_noalloc: %s", what);
	}
	return rc;
}

extern "C" __attribute__((visibility("default")))
int zkb_xidtab_build_noalloc(const zkb_decoded *req, uint32_t nsessions, const zkb_xidtab *layout,
    int8_t *ops_out, void *stream, char *errbuf, size_t errcap)
{
	if (errbuf && errcap) errbuf[0] = 0;
	if (!req || !layout) return zk_noalloc_fail(errbuf, errcap, ZKB_EINVAL, "NULL argument", cudaSuccess);
	const uint64_t n = layout->nops;
	if (n > ZKB_NOALLOC_MAX_XIDS) {
		if (errbuf && errcap)
			snprintf(errbuf, errcap, "zkb_xidtab_build_noalloc: table of %llu xids exceeds the fixed scratch of %u; "
			    "split the sessions over several calls or use zkb_xidtab_build_dev",
			    (unsigned long long)n, (unsigned)ZKB_NOALLOC_MAX_XIDS);
		return ZKB_ERANGE;
	}
	if (n == 0) return ZKB_OK;
	if (!ops_out) return zk_noalloc_fail(errbuf, errcap, ZKB_EINVAL, "NULL ops_out", cudaSuccess);

	/* The array's name means something only to device code.  Seen from the host it is a
	 * handle, so the first call on each GPU asks the runtime where the array really lives
	 * on that GPU and remembers the answer. */
	int dev = -1;
	cudaError_t e = cudaGetDevice(&dev);
	if (e != cudaSuccess) return zk_noalloc_fail(errbuf, errcap, ZKB_ECUDA, "cudaGetDevice", e);
	if (dev < 0 || dev >= ZK_NOALLOC_MAX_DEVICES)
		return zk_noalloc_fail(errbuf, errcap, ZKB_ERANGE, "device ordinal out of range", cudaSuccess);
	unsigned long long *slots = __atomic_load_n(&zk_noalloc_addr[dev], __ATOMIC_ACQUIRE);
	if (!slots) {
		void *p = NULL;
		e = cudaGetSymbolAddress(&p, zk_noalloc_slots);
		if (e != cudaSuccess || !p) return zk_noalloc_fail(errbuf, errcap, ZKB_ECUDA, "cudaGetSymbolAddress", e);
		slots = (unsigned long long *)p;
		__atomic_store_n(&zk_noalloc_addr[dev], slots, __ATOMIC_RELEASE);
	}

	cudaStream_t st = (cudaStream_t)stream;
	/* the slots still hold the previous call's keys: the part this call uses starts from zero */
	e = cudaMemsetAsync(slots, 0, n * sizeof (unsigned long long), st);
	if (e != cudaSuccess) return zk_noalloc_fail(errbuf, errcap, ZKB_ECUDA, "cudaMemsetAsync", e);
	if (req->n_frames) {
		k_xid_scatter<<<(unsigned)((req->n_frames + 255) / 256), 256, 0, st>>>(*req, req->n_frames, nsessions,
		    layout->base, layout->count, layout->off, slots, n);
		e = cudaGetLastError();
		if (e != cudaSuccess) return zk_noalloc_fail(errbuf, errcap, ZKB_ECUDA, "k_xid_scatter launch", e);
	}
	k_xid_finish<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(slots, n, ops_out);
	e = cudaGetLastError();
	if (e != cudaSuccess) return zk_noalloc_fail(errbuf, errcap, ZKB_ECUDA, "k_xid_finish launch", e);
	e = cudaStreamSynchronize(st);
	if (e != cudaSuccess) return zk_noalloc_fail(errbuf, errcap, ZKB_ECUDA, "cudaStreamSynchronize", e);
	return ZKB_OK;
}
