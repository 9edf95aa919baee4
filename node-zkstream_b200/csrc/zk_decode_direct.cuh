/*
 * zk_decode_direct.cuh — the decode kernel's ordinary body ("direct placement"); included by
 * zk_decode.cu inside its anonymous namespace, after the readers it shares with the exact body.
 *
 * One warp per 8 KiB chunk, as in the exact body, but nothing is recorded in bitmaps and nothing is
 * compacted afterwards: every side-table row is written where it belongs the moment its field is
 * read.  That is possible because
 *   - the chunk's first STAT / BUF / ACL rows are known before the kernel starts (the splitter's
 *     walk counted the rows each frame header promises, its scan carried the sums), and
 *   - inside the chunk the same header arithmetic gives every frame's row counts before any of its
 *     rows is written, so one warp scan per round of 32 frames turns them into each frame's first rows.
 * Frame headers and fixed fields are read by the frame's own lane in a few converged stages (path,
 * data, tail) instead of one branch per opcode; vectors (children, ACLs) are set aside and walked
 * together, one lane per vector frame, once enough have gathered or the chunk is done.
 *
 * What the header promised is verified as the fields are read.  A frame that breaks the promise (its
 * vector runs out of bytes) leaves rows it reserved unwritten: the chunk is flagged, and the host
 * runs the exact two-launch route from that chunk on (zk_fields_device).  A chunk whose total differs
 * from the splitter's prediction only raises the batch-wide flag: its rows are right relative to
 * each other and are placed again from the scanned bases.
 */

enum { VK_STR = 1, VK_STR_STAT = 2, VK_ACL_STAT = 3, VK_ACL_FLAGS = 4 };

/* The loop of ZKDecodeStream._transform (zk-streams.js:47-64) over one lane's 256-byte sub-range of
 * the staged chunk, from where the splitter says the chain stands there.
 * STORE == false: count the frames that start in the sub-range.
 * STORE == true:  store the tile offset of the lane's frames with chunk-local index in
 *                 [wlo, wlo + ZK_FREL_CAP) into c.frel; if `book`, also write consumed /
 *                 bad_length_at of sessions that end here and first_frame of sessions that start here
 *                 (`fbase` = batch row of the lane's first frame). */
template <bool STORE>
__device__ __forceinline__ uint32_t walk_subrange(const DecArgs &a, ChunkSm &c, const uint64_t cs, const uint64_t hi,
    const uint64_t entry, uint32_t s, uint64_t send, const uint32_t wlo, const uint32_t lbase, const bool book,
    const uint32_t fbase, const bool entry_is_start)
{
	uint32_t cnt = 0;
	if (entry == ZK_NONE) return 0;
	uint64_t pos = entry;
	if (STORE && book && entry_is_start) a.o.sess_first_frame[s] = fbase;
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
			if (STORE) {
				const uint32_t i = lbase + cnt - wlo;
				if (i < ZK_FREL_CAP) c.frel[i] = (uint16_t)rel;
			}
			cnt++;
			pos += 4 + (uint64_t)len;
		}
		if (!stop) break;                 /* the chain leaves the sub-range alive */
		if (STORE && book) {
			a.o.sess_consumed[s] = pos;
			a.o.sess_bad_length_at[s] = bad;
		}
		/* next session, if it starts inside this sub-range */
		if (s + 1 >= a.nsess) break;
		const uint64_t so = send;
		if (!(so < hi)) break;
		s++;
		pos = so;
		send = __ldg(a.sess_off + s + 1);
		if (STORE && book) a.o.sess_first_frame[s] = fbase + cnt;
	}
	return cnt;
}

/* The same loop for the ordinary case, in tile-relative 32-bit arithmetic: counts the frames that
 * start in [rel, rhi) and keeps the tile offsets of the first eight in four registers (16 bits
 * each).  Anything that is not a plain run of complete frames inside one session - the session's
 * bytes end here, a frame is incomplete, a length is bad - only sets `events`: that lane then goes
 * through walk_subrange, which also does the per-session bookkeeping. */
__device__ __forceinline__ uint32_t walk_count_fast(const uint8_t *tile, uint32_t rel, const uint32_t rhi,
    const uint32_t rsend, uint32_t (&pk)[4], bool &events)
{
	uint32_t cnt = 0;
	for (;;) {
		if (rel == rsend) { events = true; break; }
		if (rel >= rhi) break;
		const uint32_t rem = rsend - rel;
		if (rem < 4) { events = true; break; }
		const int32_t len = (int32_t)sm_be32(tile, rel);
		if (!len_ok(len) || rem < 4u + (uint32_t)len) { events = true; break; }
		const uint32_t v = rel << ((cnt & 1u) * 16u);
#pragma unroll
		for (int j = 0; j < 4; ++j)
			if ((cnt >> 1) == (uint32_t)j) pk[j] |= v;
		cnt++;
		rel += 4u + (uint32_t)len;
	}
	return cnt;
}

/* one BUF row */
__device__ __forceinline__ void put_buf(const zkb_decoded &o, uint32_t row, uint32_t frame, uint64_t off, int32_t len)
{
	if (row < o.cap_buf) {
		o.buf_frame[row] = frame;
		o.buf_off[row] = off;
		o.buf_len[row] = len;
	}
}

/* one STAT row from the 68 bytes at tile offset q (readStat, zk-buffer.js:428-442) */
__device__ __forceinline__ void put_stat(const zkb_decoded &o, const uint8_t *tile, uint32_t q, uint32_t row, uint32_t frame)
{
	if (row >= o.cap_stat) return;
	/* 68 bytes = 17 big-endian words from 18 aligned shared-memory words */
	const uint32_t *wp = reinterpret_cast<const uint32_t *>(tile + (q & ~3u));
	const uint32_t sel = 0x0123u + 0x1111u * (q & 3u);
	uint32_t x[18];
#pragma unroll
	for (int i = 0; i < 18; ++i) x[i] = wp[i];
	uint32_t v[17];
#pragma unroll
	for (int i = 0; i < 17; ++i) v[i] = __byte_perm(x[i], x[i + 1], sel);
#define Q64(i) ((int64_t)(((uint64_t)v[i] << 32) | (uint64_t)v[(i) + 1]))
	o.stat_frame[row] = frame;
	o.czxid[row] = Q64(0); o.mzxid[row] = Q64(2); o.ctime[row] = Q64(4); o.mtime[row] = Q64(6);
	o.version[row] = (int32_t)v[8]; o.cversion[row] = (int32_t)v[9]; o.aversion[row] = (int32_t)v[10];
	o.ephemeralOwner[row] = Q64(11);
	o.dataLength[row] = (int32_t)v[13]; o.numChildren[row] = (int32_t)v[14];
	o.pzxid[row] = Q64(15);
#undef Q64
}

/* Walk the vectors set aside so far, one lane per frame.  Returns false in a lane whose frame did
 * not keep what its header promised.
 *
 * LAST: the chunk's last drain.  The frame list is no longer needed then, so its memory takes the
 * tile offsets of the string elements: each lane walks its own vector reading lengths only, and all
 * lanes together write the rows afterwards (one element per lane, whichever vector it belongs to).
 * The entries of the batch at hand have been read into registers by then, so their memory holds
 * what the second phase needs to find an element's vector: one bit per element that opens a vector,
 * and { first row - first list slot, frame } of every vector that has elements. */
template <bool LAST>
__device__ __forceinline__ bool drain_vectors(const DecArgs &a, ChunkSm &c, const uint64_t cs, const uint32_t f0,
    const uint32_t e0, const uint32_t e1, const uint32_t e2, const uint32_t nv, const uint32_t lane)
{
	const zkb_decoded &o = a.o;
	uint16_t *list = c.frel;
	uint2 *vinfo = reinterpret_cast<uint2 *>(c.vl);                       /* 32 x 8 bytes */
	uint32_t *heads = reinterpret_cast<uint32_t *>(c.vl + 16);            /* ZK_FREL_CAP bits */
	bool good = true;
	for (uint32_t i0 = 0; i0 < nv; i0 += 32) {
		const uint32_t i = i0 + lane;
		const bool have = i < nv;
		VEnt e;
		e.w0 = 4u; e.w1 = e.w2 = e.w3 = 0u;
		if (have) e = c.vl[i];
		const uint32_t pr = e.w0 & 0xffffu, vrel = e.w0 >> 16;
		uint32_t brow = e1 + (e.w1 & 0xffffu), arow = e2 + (e.w1 >> 16);
		const uint32_t srow = e0 + (e.w2 & 0xffffu), kind = e.w2 >> 16;
		const uint32_t frame = f0 + e.w3;
		const uint32_t end = pr + sm_be32(c.tile, pr - 4);          /* tile offset one past the payload */
		const int32_t count = have ? (int32_t)sm_be32(c.tile, vrel) : 0;
		uint32_t cur = vrel + 4;
		bool ok = true;
		const uint64_t base = cs + 4;
		const uint64_t want = count > 0 ? (uint64_t)count : 0u;
		const bool str = have && (kind == VK_STR || kind == VK_STR_STAT);
		const bool fits = brow + want <= o.cap_buf;
		/* ---- which vectors go through the list ---- */
		bool listed = false;
		uint32_t slot = 0, total = 0;
		if (LAST) {
			const uint32_t n2 = (str && fits) ? (uint32_t)(want < ZK_FREL_CAP + 1u ? want : ZK_FREL_CAP + 1u) : 0u;
			const uint32_t inc = warp_incl_scan(n2, lane);
			total = __shfl_sync(0xffffffffu, inc, 31);
			if (total > ZK_FREL_CAP) total = 0;              /* too many elements for the list: lane by lane */
			const uint32_t nonempty = __ballot_sync(0xffffffffu, n2 != 0);
			__syncwarp();                                    /* every entry of the batch is in registers */
			if (total) {
				slot = inc - n2;
				listed = n2 != 0;
				if (lane < (ZK_FREL_CAP + 31u) / 32u) heads[lane] = 0;
				if (listed) vinfo[__popc(nonempty & ((1u << lane) - 1u))] = make_uint2(brow - slot, frame);
				__syncwarp();
				if (listed) atomicOr(&heads[slot >> 5], 1u << (slot & 31u));
			}
		}
		if (listed) {
			/* phase one: lengths only (vector<ustring>, zk-buffer.js:334-337).  An element's length is
			 * read even when fewer than four bytes are left (the tile is zero padded): the end test
			 * catches it. */
			uint16_t *lp = list + slot;
			uint16_t *const le = lp + (uint32_t)want;
			while (lp < le) {
				const int32_t l = (int32_t)sm_be32(c.tile, cur);
				const uint32_t nxt = cur + 4 + (uint32_t)max(l, 0);
				if (nxt > end) break;
				*lp++ = (uint16_t)cur;
				cur = nxt;
			}
			ok = lp == le;
			while (lp < le) *lp++ = 0xFFFFu;                     /* the count lied: no rows for the rest */
		} else if (str) {
			if (fits) {
				uint32_t *pf = o.buf_frame + brow; uint64_t *po = o.buf_off + brow; int32_t *pl = o.buf_len + brow;
				int32_t k = 0;
				for (; k < count; ++k) {
					const int32_t l = (int32_t)sm_be32(c.tile, cur);
					const uint32_t ul = l > 0 ? (uint32_t)l : 0u;
					const uint32_t nxt = cur + 4 + ul;
					if (nxt > end) break;
					pf[k] = frame; po[k] = base + cur; pl[k] = (int32_t)ul;
					cur = nxt;
				}
				ok = k >= count;
			} else {
				for (int32_t k = 0; k < count; ++k) {
					if (cur + 4 > end) { ok = false; break; }
					const int32_t l = (int32_t)sm_be32(c.tile, cur);
					const uint32_t ul = l > 0 ? (uint32_t)l : 0u;
					if (cur + 4 + ul > end) { ok = false; break; }
					put_buf(o, brow++, frame, base + cur, (int32_t)ul);
					cur += 4 + ul;
				}
			}
		} else if (have) {
			/* readACL (zk-buffer.js:372-382): count x { perms, scheme, id } */
			const bool afits = brow + 2 * want <= o.cap_buf && arow + want <= o.cap_acl;
			for (int32_t k = 0; k < count; ++k) {
				const int32_t perms = (int32_t)sm_be32(c.tile, cur);
				const int32_t l0 = (int32_t)sm_be32(c.tile, cur + 4);
				const uint32_t u0 = l0 > 0 ? (uint32_t)l0 : 0u;
				const uint32_t c1 = cur + 8 + u0;                  /* length prefix of the id */
				if (c1 + 4 > end) { ok = false; break; }
				const int32_t l1 = (int32_t)sm_be32(c.tile, c1);
				const uint32_t u1 = l1 > 0 ? (uint32_t)l1 : 0u;
				if (c1 + 4 + u1 > end) { ok = false; break; }
				if (afits) {
					o.acl_frame[arow] = frame; o.acl_perms[arow] = perms;
					o.buf_frame[brow] = frame; o.buf_off[brow] = base + cur + 4; o.buf_len[brow] = (int32_t)u0;
					o.buf_frame[brow + 1] = frame; o.buf_off[brow + 1] = base + c1; o.buf_len[brow + 1] = (int32_t)u1;
				} else {
					if (arow < o.cap_acl) { o.acl_frame[arow] = frame; o.acl_perms[arow] = perms; }
					put_buf(o, brow, frame, base + cur + 4, (int32_t)u0);
					put_buf(o, brow + 1, frame, base + c1, (int32_t)u1);
				}
				arow++; brow += 2;
				cur = c1 + 4 + u1;
			}
		}
		__syncwarp();
		if (LAST && total) {
			/* phase two: element j of the list belongs to the vector whose opening bit is the last one
			 * at or below j */
			uint32_t before = 0;
			for (uint32_t j0 = 0; j0 < total; j0 += 32) {
				const uint32_t hw = heads[j0 >> 5];
				const uint32_t j = j0 + lane;
				if (j < total) {
					const uint2 vi = vinfo[before + __popc(hw & (0xffffffffu >> (31u - lane))) - 1u];
					const uint32_t p = list[j];
					if (p != 0xFFFFu) {
						const int32_t l = (int32_t)sm_be32(c.tile, p);
						const uint32_t row = vi.x + j;
						o.buf_frame[row] = vi.y; o.buf_off[row] = base + p; o.buf_len[row] = l > 0 ? l : 0;
					}
				}
				before += __popc(hw);
			}
			__syncwarp();
		}
		/* what follows the vector, all lanes together */
		if (have && ok) {
			if (kind == VK_ACL_FLAGS) {                      /* CREATE request: flags follow the ACL */
				if (cur + 4 > end) ok = false;
				else a.o.a0[frame] = (int32_t)sm_be32(c.tile, cur);
			} else if (kind != VK_STR) {                     /* a Stat follows the vector */
				if (cur + 68 > end) ok = false;
				else put_stat(o, c.tile, cur, srow, frame);
			}
		}
		good = good && ok;
	}
	return good;
}

/* what one lane knows about its frame between the header stage and the row stage of a round */
struct Fr {
	uint32_t nst, nbf, nac;      /* rows the frame owns */
	uint32_t b0, b1;             /* tile offsets of the length prefixes of up to two scalar buffers (0 = none) */
	uint32_t st;                 /* tile offset of the Stat (0 = none) */
	uint32_t vrel, vkind;        /* vector to be walked later */
	uint32_t slow;               /* 1 = generic parser, 2 = generic parser and the frame reaches past the tile */
};

/* readRequest (zk-buffer.js:58-95,138-253) for a frame inside the tile: header, scalar fields and
 * the row counts; false = leave it to the generic parser.  `pr` = tile offset of payload byte 0. */
__device__ __forceinline__ bool direct_request(const uint8_t *t, const uint32_t pr, const uint32_t size, FrameCols &o, Fr &fr)
{
	if (size < 8) return false;
	const int32_t xid = (int32_t)sm_be32(t, pr);
	const int32_t op = (int32_t)sm_be32(t, pr + 4);
	o.xid = xid; o.zxid = 0; o.err = 0; o.a0 = 0; o.a1 = 0;
	/* opcodes whose body starts with a path: CREATE DELETE EXISTS GET_DATA SET_DATA GET_ACL GET_CHILDREN SYNC GET_CHILDREN2 */
	const bool has_path = op >= 0 && op <= 12 && ((0x137Eu >> op) & 1u);
	if (!has_path) {
		if (op == ZKB_OP_SET_WATCHES) return false;      /* three vectors: counted by the generic parser */
		/* 'Unsupported opcode' (zk-buffer.js:91-92), PING and CLOSE_SESSION included */
		const int known = op_known(op) ? op : ZKB_OP_NONE;
		o.meta = ZKB_ST_BAD_DECODE | ((uint32_t)(uint8_t)(int8_t)known << 8) | ((uint32_t)ZKB_KIND_REQUEST << 16) |
		    (DETAIL(ZKB_CAUSE_UNSUPPORTED, ZKB_F_OPCODE) << 24);
		return true;
	}
	/* path */
	if (12 > size) return false;
	const int32_t pl = (int32_t)sm_be32(t, pr + 8);
	uint32_t cur = 12u + (pl > 0 ? (uint32_t)pl : 0u);
	if (cur > size) return false;
	fr.b0 = pr + 8; fr.nbf = 1;
	/* data (CREATE, SET_DATA) */
	if (op == ZKB_OP_CREATE || op == ZKB_OP_SET_DATA) {
		if (cur + 4 > size) return false;
		const int32_t dl = (int32_t)sm_be32(t, pr + cur);
		fr.b1 = pr + cur;
		cur += 4u + (dl > 0 ? (uint32_t)dl : 0u);
		if (cur > size) return false;
		fr.nbf = 2;
	}
	/* tail */
	if (op == ZKB_OP_CREATE) {
		if (cur + 4 > size) return false;
		const int32_t n = (int32_t)sm_be32(t, pr + cur);
		const uint32_t k = n > 0 ? (uint32_t)n : 0u;
		fr.vrel = pr + cur; fr.vkind = VK_ACL_FLAGS;
		fr.nbf += 2 * k; fr.nac = k;
		o.a1 = (int32_t)k;                              /* a0 (flags) is written when the ACL has been walked */
	} else if (op == ZKB_OP_DELETE || op == ZKB_OP_SET_DATA) {
		if (cur + 4 > size) return false;
		o.a0 = (int32_t)sm_be32(t, pr + cur);
	} else if (op != ZKB_OP_GET_ACL && op != ZKB_OP_SYNC) {   /* watch flag (readBool asserts 0/1) */
		if (!(cur < size)) return false;
		const uint32_t b = t[pr + cur];
		if (b > 1) return false;
		o.a0 = (int32_t)b;
	}
	o.meta = ZKB_ST_OK | ((uint32_t)(uint8_t)(int8_t)op << 8) | ((uint32_t)ZKB_KIND_REQUEST << 16);
	return true;
}

/* readResponse (zk-buffer.js:281-331) for a frame inside the tile */
/* the xid table's entry of one session, fetched while the tile is on its way */
struct XidSess { uint32_t s; int32_t base; uint32_t count; uint64_t off; };

__device__ __forceinline__ bool direct_response(const uint8_t *t, const DecArgs &a, const uint32_t pr, const uint32_t size,
    const uint32_t s, const XidSess &xs, FrameCols &o, Fr &fr)
{
	if (size < 16) return false;
	const int32_t xid = (int32_t)sm_be32(t, pr);
	const int32_t err = (int32_t)sm_be32(t, pr + 12);
	int op = special_xid(xid);
	if (op == ZKB_OP_NONE && a.xt_ops) {
		const bool pre = (s == xs.s);
		const int64_t idx = (int64_t)xid - (int64_t)(pre ? xs.base : __ldg(a.xt_base + s));
		if (idx >= 0 && idx < (int64_t)(pre ? xs.count : __ldg(a.xt_count + s)))
			op = __ldg(a.xt_ops + (pre ? xs.off : __ldg(a.xt_off + s)) + (uint64_t)idx);
	}
	if (op == ZKB_OP_NONE) return false;                 /* 'reply packet must match a request' */
	o.xid = xid; o.err = err; o.a0 = 0; o.a1 = 0;
	o.zxid = (int64_t)(((uint64_t)sm_be32(t, pr + 4) << 32) | (uint64_t)sm_be32(t, pr + 8));
	o.meta = ZKB_ST_OK | ((uint32_t)(uint8_t)(int8_t)op << 8) | ((uint32_t)ZKB_KIND_RESPONSE << 16);
	if (err != 0) return true;                           /* only 'OK' replies have a body */
	if (op == ZKB_OP_GET_CHILDREN || op == ZKB_OP_GET_CHILDREN2 || op == ZKB_OP_GET_ACL) {
		if (20 > size) return false;
		const int32_t n = (int32_t)sm_be32(t, pr + 16);
		const uint32_t k = n > 0 ? (uint32_t)n : 0u;
		o.a0 = (int32_t)k;
		fr.vrel = pr + 16;
		if (op == ZKB_OP_GET_ACL) { fr.vkind = VK_ACL_STAT; fr.nbf = 2 * k; fr.nac = k; fr.nst = 1; }
		else if (op == ZKB_OP_GET_CHILDREN2) { fr.vkind = VK_STR_STAT; fr.nbf = k; fr.nst = 1; }
		else { fr.vkind = VK_STR; fr.nbf = k; }
		return true;
	}
	uint32_t cur = 16;
	if (op == ZKB_OP_NOTIFICATION) {
		if (24 > size) return false;
		o.a0 = (int32_t)sm_be32(t, pr + 16);
		o.a1 = (int32_t)sm_be32(t, pr + 20);
		cur = 24;
	}
	if (op == ZKB_OP_NOTIFICATION || op == ZKB_OP_CREATE || op == ZKB_OP_GET_DATA) {
		if (cur + 4 > size) return false;
		const int32_t l = (int32_t)sm_be32(t, pr + cur);
		fr.b0 = pr + cur; fr.nbf = 1;
		cur += 4u + (l > 0 ? (uint32_t)l : 0u);
		if (cur > size) return false;
	}
	if (op == ZKB_OP_GET_DATA || op == ZKB_OP_EXISTS || op == ZKB_OP_SET_DATA) {
		if (cur + 68 > size) return false;
		fr.st = pr + cur; fr.nst = 1;
		return true;
	}
	if (op == ZKB_OP_NOTIFICATION || op == ZKB_OP_CREATE) return true;
	/* header only; everything else is 'Unsupported opcode' (zk-buffer.js:326-327): generic parser */
	return op == ZKB_OP_SET_WATCHES || op == ZKB_OP_PING || op == ZKB_OP_SYNC || op == ZKB_OP_DELETE ||
	    op == ZKB_OP_CLOSE_SESSION;
}

/* A GET_CHILDREN / GET_CHILDREN2 reply that reaches past the tile (10^4-10^5 children, config 4):
 * header and child count are inside the tile, so the FRAMES row is written like any other; the
 * children are walked by the whole warp once the chunk's other frames are done (long_children).
 * Returns false for everything else, which the generic parser takes. */
__device__ __forceinline__ bool direct_response_long(ChunkSm &c, const DecArgs &a, const uint32_t pr, const uint32_t size,
    const uint32_t wlen, const uint32_t s, FrameCols &o)
{
	if (size < 20 || pr + 20 > wlen) return false;
	const uint8_t *t = c.tile;
	const int32_t xid = (int32_t)sm_be32(t, pr);
	if (special_xid(xid) != ZKB_OP_NONE || !a.xt_ops) return false;
	const int64_t idx = (int64_t)xid - (int64_t)__ldg(a.xt_base + s);
	if (idx < 0 || idx >= (int64_t)__ldg(a.xt_count + s)) return false;
	const int op = __ldg(a.xt_ops + __ldg(a.xt_off + s) + (uint64_t)idx);
	if (op != ZKB_OP_GET_CHILDREN && op != ZKB_OP_GET_CHILDREN2) return false;
	const int32_t err = (int32_t)sm_be32(t, pr + 12);
	if (err != 0) return false;
	const int32_t n = (int32_t)sm_be32(t, pr + 16);
	o.xid = xid; o.err = 0; o.a1 = 0; o.a0 = n > 0 ? n : 0;
	o.zxid = (int64_t)(((uint64_t)sm_be32(t, pr + 4) << 32) | (uint64_t)sm_be32(t, pr + 8));
	o.meta = ZKB_ST_OK | ((uint32_t)(uint8_t)(int8_t)op << 8) | ((uint32_t)ZKB_KIND_RESPONSE << 16);
	c.lp.n_buf = (uint32_t)o.a0; c.lp.n_stat = (op == ZKB_OP_GET_CHILDREN2) ? 1u : 0u; c.lp.n_acl = 0;
	return true;
}

/* ---- children vectors longer than the tile ------------------------------------------------
 * The warp stages the vector through its tile 8 KiB at a time.  Inside a tile the elements are a
 * chain of length prefixes like the frames of a stream, and they are found the same way: every lane
 * takes a 280-byte sub-range, guesses where the chain stands in it (the first position that looks
 * like the prefix of a short, non-empty string and from which three more such prefixes follow),
 * and walks to the end of its sub-range; then the lanes compare notes from the left - the chain's
 * true entry into the tile is known - and a lane whose guess is not where its left neighbour's
 * walk ended walks again from there.  A wrong guess costs a repair round, never a wrong row. */
#define ZK_LSUB 280u
static_assert(ZK_LSUB * 32u >= ZK_WIN, "the 32 sub-ranges must cover the tile");

/* elements that start in [p, hi): count them (EMIT == false) or write their BUF rows (EMIT == true,
 * at most `maxn`); returns where the chain stands afterwards.  A prefix that cannot be read inside
 * the tile, or an element that runs past the frame (`vlim`), ends the walk. */
template <bool EMIT>
__device__ __forceinline__ uint32_t walk_elems(const uint8_t *tile, uint32_t p, const uint32_t hi, const uint32_t tlen,
    const uint32_t vlim, uint32_t &n, const uint32_t maxn, const zkb_decoded &o, uint32_t row, const uint32_t frame,
    const uint64_t tb)
{
	n = 0;
	while (p < hi && p + 4 <= tlen && (!EMIT || n < maxn)) {
		const int32_t l = (int32_t)sm_be32(tile, p);
		const uint32_t ul = l > 0 ? (uint32_t)l : 0u;
		if (p + 4 + ul > vlim) break;
		if (EMIT) put_buf(o, row++, frame, tb + p + 4, (int32_t)ul);
		n++;
		p += 4 + ul;
	}
	return p;
}

/* bit k set iff byte k (in address order) of the word is zero */
__device__ __forceinline__ uint32_t zero_nib(const uint32_t w)
{
	const uint32_t t = ((w & 0x7f7f7f7fu) + 0x7f7f7f7fu) | w;       /* bit 7 of a byte stays clear iff the byte is zero */
	return (((~t & 0x80808080u) >> 7) * 0x01020408u) >> 24;        /* bit 8k -> bit 24 + k, no two products collide */
}

/* positions b .. b+31 (restricted to [lo, hi)) where a prefix of a string of 1..255 bytes may start:
 * three zero bytes and a non-zero one.  b is a multiple of four. */
__device__ __forceinline__ uint32_t elem_mask32(const uint8_t *tile, const uint32_t b, const uint32_t lo, const uint32_t hi)
{
	const uint32_t *wp = reinterpret_cast<const uint32_t *>(tile);
	uint64_t Z = 0;
#pragma unroll
	for (int k = 0; k < 9; ++k) {
		const uint32_t wi = (b >> 2) + (uint32_t)k;
		const uint32_t w = wi < (ZK_WIN + 16u) / 4u ? wp[wi] : 0xffffffffu;
		Z |= (uint64_t)zero_nib(w) << (4 * k);
	}
	uint32_t cand = (uint32_t)(Z & (Z >> 1) & (Z >> 2) & ~(Z >> 3));     /* bit i: position b + i */
	if (b < lo) cand &= 0xffffffffu << (lo - b);
	if (hi - b < 32u) cand &= (1u << (hi - b)) - 1u;
	return cand;
}

/* For every lane with `need`: the first position in its [lo, hi) that looks like an element prefix
 * (see above), or 0xFFFFFFFF.  Called by the whole warp: the lanes look for their next candidate each
 * on its own (a few mask words), then follow the three prefixes behind their candidates together -
 * done lane by lane as each found one, that check ran at under two lanes per instruction and was
 * nearly half of the whole walk. */
__device__ __forceinline__ uint32_t guess_elem_warp(const uint8_t *tile, const uint32_t lo, const uint32_t hi,
    const uint32_t tlen, const bool need)
{
	uint32_t b = lo & ~3u, cb = 0, cand = 0, found = 0xFFFFFFFFu;
	bool scanning = need;
	for (;;) {
		uint32_t q = 0xFFFFFFFFu;
		while (scanning && q == 0xFFFFFFFFu) {
			if (!cand) {
				if (b >= hi) { scanning = false; break; }
				cand = elem_mask32(tile, b, lo, hi);
				cb = b;
				b += 32;
				if (!cand) continue;
			}
			q = cb + (uint32_t)(__ffs((int)cand) - 1);
			cand &= cand - 1;
			if (q + 4 > tlen) { q = 0xFFFFFFFFu; scanning = false; }
		}
		__syncwarp();
		if (!__any_sync(0xffffffffu, q != 0xFFFFFFFFu)) break;
		bool good = q != 0xFFFFFFFFu, stop = false;
		uint32_t r = good ? q + 4 + sm_be32(tile, q) : 0u;
#pragma unroll
		for (int h = 0; h < 3; ++h) {
			if (good && !stop) {
				if (r + 4 > tlen) stop = true;
				else {
					const uint32_t w2 = sm_be32(tile, r);
					if (w2 > 255u) good = false;
					r += 4 + w2;
				}
			}
		}
		if (q != 0xFFFFFFFFu && good) { found = q; scanning = false; }
	}
	return found;
}

/* `phase` is the parity of the warp's mbarrier for its next use.  Returns false if the vector or
 * the Stat behind it runs out of bytes. */
/* `list`: shared memory for `list_cap` tile offsets.  The rows of a tile are written from it, one
 * element per lane in row order, so a warp's stores fall into a few sectors (written by each lane
 * as it walks, every store instruction touched 32 sectors and the walks of a full GPU queued up
 * behind the store path). */
__device__ __forceinline__ bool long_children(const DecArgs &a, uint8_t *tile, uint16_t *list, const uint32_t list_cap,
    uint64_t *bar, uint32_t phase,
    const uint64_t foff, const uint32_t flen, const uint32_t frame, const uint32_t brow0, const uint32_t srow,
    const bool with_stat, const uint32_t lane)
{
	const zkb_decoded &o = a.o;
	const uint64_t fend = foff + 4 + (uint64_t)flen;
	uint64_t pos = foff + 4 + 20;                         /* first element */
	int64_t remaining = (int64_t)(int32_t)g_be32(a.bytes, foff + 4 + 16);
	uint32_t row = brow0;
	bool ok = true;
	while (remaining > 0 && ok) {
		const uint64_t tb = pos & ~15ull;
		const uint64_t avail = a.nbytes_pad - tb;
		const uint32_t tlen = avail < (uint64_t)ZK_WIN ? (uint32_t)avail : ZK_WIN;
		const uint32_t vlim = fend - tb > 0x40000000ull ? 0x40000000u : (uint32_t)(fend - tb);
		__syncwarp();
		if (lane == 0) {
			asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
			mbar_expect_tx(bar, tlen);
			bulk_g2s(tile, a.bytes + tb, tlen, bar);
		}
		__syncwarp();
		mbar_wait(bar, phase);
		phase ^= 1u;
		const uint32_t start = (uint32_t)(pos - tb);
		const uint32_t lo = lane * ZK_LSUB;
		const uint32_t hi = (lo + ZK_LSUB < tlen) ? lo + ZK_LSUB : tlen;
		/* first try: the lane that holds the chain's entry walks from it, the lanes after it guess */
		uint32_t in, out, n = 0;
		const bool idle = hi <= start || lo >= tlen;
		const uint32_t guessed = guess_elem_warp(tile, lo, hi, tlen, !idle && lo > start);
		if (idle) { in = out = start; }
		else {
			in = (lo <= start) ? start : guessed;
			out = (in == 0xFFFFFFFFu) ? in : walk_elems<false>(tile, in, hi, tlen, vlim, n, 0, o, 0, 0, 0);
		}
		/* repair from the left until every lane starts where its neighbour ended */
		for (;;) {
			uint32_t prev = __shfl_up_sync(0xffffffffu, out, 1);
			if (lane == 0) prev = start;
			const bool active = !(hi <= start || lo >= tlen);
			const uint32_t wrong = __ballot_sync(0xffffffffu, active && in != prev);
			if (!wrong) break;
			if (lane == (uint32_t)(__ffs(wrong) - 1)) {
				in = prev;
				out = walk_elems<false>(tile, in, hi, tlen, vlim, n, 0, o, 0, 0, 0);
			}
		}
		/* rows: lane l's elements follow those of the lanes before it */
		const uint32_t inc = warp_incl_scan(n, lane);
		const uint32_t total = __shfl_sync(0xffffffffu, inc, 31);
		const uint32_t before = inc - n;
		const uint32_t left = remaining > (int64_t)0x7fffffff ? 0x7fffffffu : (uint32_t)remaining;
		const uint32_t mine = before >= left ? 0u : (left - before < n ? left - before : n);
		uint32_t wrote = 0;
		uint32_t endp;
		if (total <= list_cap) {
			/* the same walk once more, leaving the offsets of this lane's elements in the list */
			uint32_t p = in, k = 0;
			while (p < hi && p + 4 <= tlen && k < mine) {
				const int32_t l = (int32_t)sm_be32(tile, p);
				const uint32_t ul = l > 0 ? (uint32_t)l : 0u;
				if (p + 4 + ul > vlim) break;
				list[before + k] = (uint16_t)p;
				k++;
				p += 4 + ul;
			}
			endp = p;
			__syncwarp();
			const uint32_t emit_n = total < left ? total : left;
			for (uint32_t j = lane; j < emit_n; j += 32) {
				const uint32_t q = list[j];
				const int32_t l = (int32_t)sm_be32(tile, q);
				put_buf(o, row + j, frame, tb + q + 4, l > 0 ? l : 0);
			}
		} else {
			endp = walk_elems<true>(tile, in, hi, tlen, vlim, wrote, mine, o, row + before, frame, tb);
		}
		/* where the chain stands after the tile: what the last lane that has bytes says */
		const uint32_t act = __ballot_sync(0xffffffffu, lo < tlen);
		uint32_t last_out = __shfl_sync(0xffffffffu, out, 31 - __clz((int)act));
		if (total >= left) {
			/* the vector ends in this tile: the lane that wrote its last element knows where */
			const uint32_t who = __ballot_sync(0xffffffffu, mine > 0 && before + mine == left);
			last_out = __shfl_sync(0xffffffffu, endp, who ? __ffs(who) - 1 : 0);
			row += left; remaining = 0;
		} else {
			if (total == 0) ok = false;              /* no readable element where one must be */
			row += total; remaining -= total;
		}
		pos = tb + last_out;
	}
	if (ok && with_stat) {
		if (pos + 68 > fend) ok = false;
		else if (lane == 0 && srow < o.cap_stat) {
			const uint8_t *g = a.bytes;
#define G64(q) ((int64_t)(((uint64_t)g_be32(g, (q)) << 32) | (uint64_t)g_be32(g, (q) + 4)))
			o.stat_frame[srow] = frame;
			o.czxid[srow] = G64(pos); o.mzxid[srow] = G64(pos + 8); o.ctime[srow] = G64(pos + 16); o.mtime[srow] = G64(pos + 24);
			o.version[srow] = (int32_t)g_be32(g, pos + 32); o.cversion[srow] = (int32_t)g_be32(g, pos + 36);
			o.aversion[srow] = (int32_t)g_be32(g, pos + 40); o.ephemeralOwner[srow] = G64(pos + 44);
			o.dataLength[srow] = (int32_t)g_be32(g, pos + 52); o.numChildren[srow] = (int32_t)g_be32(g, pos + 56);
			o.pzxid[srow] = G64(pos + 60);
#undef G64
		}
	}
	return ok;
}

/* Everything one warp does for one 8 KiB chunk.  REDO == false: first rows from the splitter's
 * prediction (a.g_rbase), verified; REDO == true: first rows from the scan over the chunks' real
 * counts (a.c_base), chunks that need the exact body are skipped (they have their own launch). */
template <int DIR, bool REDO>
__device__ __forceinline__ void decode_chunk_direct(const DecArgs &a, ChunkSm &c, uint64_t *bar, const uint32_t chunk,
    const uint32_t lane)
{
	if (REDO && a.c_rows[chunk].x) return;
	const uint64_t cs = (uint64_t)chunk << ZK_CHUNK_LOG2;
	const uint64_t avail = a.nbytes_pad - cs;
	const uint32_t wlen = avail < (uint64_t)ZK_WIN ? (uint32_t)avail : ZK_WIN;

	/* ---- stage the chunk: one TMA bulk copy, completion on an mbarrier; the splitter's records of
	 * the chunk are fetched while it is in flight ---- */
	if (lane == 0) {
		asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
		mbar_expect_tx(bar, wlen);
		bulk_g2s(c.tile, a.bytes + cs, wlen, bar);
		*reinterpret_cast<uint4 *>(c.tile + wlen) = make_uint4(0, 0, 0, 0);
		c.have_long = 0;
	}
	Rd rd;
	rd.sm = c.tile; rd.lo = cs; rd.wlen = wlen; rd.g = a.bytes;

	/* the chunk's records and this lane's sub-range entry: independent loads, one round trip */
	const uint4 m0 = a.meta0[chunk];
	const uint32_t f1 = a.meta0[chunk + 1].x;
	uint32_t e0, e1, e2, p0 = 0, p1 = 0, p2 = 0;     /* first rows of the chunk; rows the splitter predicted for it */
	if (REDO) {
		const uint4 cb = a.c_base[chunk];
		e0 = cb.y; e1 = cb.z; e2 = cb.w;
	} else {
		const uint4 rb0 = a.meta1[chunk], rb1 = a.meta1[chunk + 1];
		e0 = rb0.x; e1 = rb0.y; e2 = rb0.z;
		p0 = rb1.x - rb0.x; p1 = rb1.y - rb0.y; p2 = rb1.z - rb0.z;
	}
	const bool last = (chunk == a.n_chunks - 1);
	const uint64_t ce = last ? a.nbytes : cs + ZK_CHUNK;
	/* where the chain stands inside this lane's 256-byte sub-range, from the splitter (both recorded
	 * walks are fetched, the selection picks one) */
	const uint64_t lo = cs + 256u * lane;
	const uint64_t hi = (lo + 256 < ce) ? lo + 256 : ce;
	uint64_t entry = ZK_NONE;
	if (lo < ce) {
		const uint32_t g = (uint32_t)(lo >> a.glog2);
		const uint64_t gstart = (uint64_t)g << a.glog2;
		const uint32_t subs = 1u << (a.glog2 - 8);
		const uint32_t si = (uint32_t)((lo - gstart) >> 8);
		const uint16_t v0 = a.g_sub[(size_t)g * subs + si];
		const uint16_t v1 = a.g_sub[((size_t)a.n_gran + g) * subs + si];
		const uint16_t v = a.g_sel[g] ? v1 : v0;
		if (v != 0xFFFFu) entry = gstart + v;
	}
	const uint32_t f0 = m0.x, sfirst = m0.y, sf = m0.z;
	const uint32_t nfr = f1 - f0;
	const bool cont = (m0.w != 0);                   /* the chunk starts inside session sf-1 */
	/* no session starts inside the chunk: then nothing but the records above is needed */
	const bool one_session = cont && !((uint64_t)m0.w < ce - cs || (last && sf < a.nsess));
	uint64_t boff = ZK_NONE;                         /* boff[l] = sess_off[sf + l] */
	if (!one_session || m0.w == 0xFFFFFFFFu) { if (sf + lane <= a.nsess) boff = __ldg(a.sess_off + sf + lane); }
	const uint64_t boff0 = (one_session && m0.w != 0xFFFFFFFFu) ? cs + m0.w : __shfl_sync(0xffffffffu, boff, 0);
	/* session the entry lies in */
	uint32_t es = sf - 1; uint64_t esend = boff0; bool e_start = false;
	if (!one_session) {
		/* jl = boundaries sess_off[sf + l], l < 32, strictly below the entry */
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
		uint64_t b_at = __shfl_sync(0xffffffffu, boff, (int)(jl < 32 ? jl : 31));
		uint64_t b_next = __shfl_sync(0xffffffffu, boff, (int)(jl + 1 < 32 ? jl + 1 : 31));
		if (entry != ZK_NONE) {
			uint32_t sprev = sf - 1 + jl;              /* last session that starts strictly before the entry */
			if (jl >= 32) {                            /* > 32 session starts in this chunk: use the table */
				uint32_t l2 = sf + 31, h2 = a.nsess;
				while (h2 - l2 > 1) {
					const uint32_t mid = (l2 + h2) >> 1;
					if (__ldg(a.sess_off + mid) < entry) l2 = mid; else h2 = mid;
				}
				sprev = l2;
				b_at = __ldg(a.sess_off + sprev + 1);
			}
			if (jl + 1 >= 32 || jl >= 32) b_next = (sprev + 2 <= a.nsess) ? __ldg(a.sess_off + sprev + 2) : ZK_NONE;
			if (entry < b_at) { es = sprev; esend = b_at; }
			else { es = sprev + 1; esend = b_next; e_start = true; }
		}
	}
	/* replies: the xid table's entry of the session the chunk starts in */
	XidSess xs;
	xs.s = 0xFFFFFFFFu; xs.base = 0; xs.count = 0; xs.off = 0;
	if (DIR != 1 && a.xt_ops && cont) {
		xs.s = sf - 1;
		xs.base = __ldg(a.xt_base + xs.s); xs.count = __ldg(a.xt_count + xs.s); xs.off = __ldg(a.xt_off + xs.s);
	}
	__syncwarp();
	mbar_wait(bar, 0);
	/* replies: ask for the xid table's bytes now - the first frames of every eighth sub-range span the
	 * chunk's xids closely enough - so the header stage finds them in cache (a wrong or special xid asks
	 * for nothing) */
	if (DIR != 1 && (lane & 7u) == 0 && xs.s != 0xFFFFFFFFu && entry != ZK_NONE && (uint32_t)(entry - cs) + 8u <= wlen) {
		const int64_t idx = (int64_t)(int32_t)sm_be32(c.tile, (uint32_t)(entry - cs) + 4u) - (int64_t)xs.base;
		if (idx >= 0 && idx < (int64_t)xs.count)
			asm volatile("prefetch.global.L1 [%0];" :: "l"(a.xt_ops + xs.off + (uint64_t)idx));
	}

	/* ---- frame starts: count per lane, prefix over lanes.  A lane whose sub-range is a plain run of
	 * frames keeps their offsets in registers; the others (a session ends or starts there, a bad or
	 * incomplete frame, more than eight frames) walk again below, with the bookkeeping ---- */
	uint32_t pk[4] = { 0u, 0u, 0u, 0u };
	bool events = e_start;
	uint32_t mycnt = 0;
	if (entry != ZK_NONE) {
		const uint64_t d = esend - cs;
		const uint32_t rsend = d > 0x7fffffffull ? 0x7fffffffu : (uint32_t)d;
		mycnt = walk_count_fast(c.tile, (uint32_t)(entry - cs), (uint32_t)(hi - cs), rsend, pk, events);
		if (events) mycnt = walk_subrange<false>(a, c, cs, hi, entry, es, esend, 0, 0, false, 0, false);
	}
	const uint32_t lbase = warp_incl_scan(mycnt, lane) - mycnt;
	if (lane == 31 && lbase + mycnt != nfr) atomicOr(&a.ctrl->error, ZK_DERR_COUNT);
	const bool simple = nfr <= ZK_FREL_CAP;      /* one window: offsets can come straight from the registers */
	if (last && lane == 0) {
		/* empty sessions at the very end of the batch start at nbytes, in no sub-range */
		for (uint32_t sx = a.nsess; sx > 0 && __ldg(a.sess_off + sx - 1) == a.nbytes; --sx) {
			a.o.sess_consumed[sx - 1] = a.nbytes;
			a.o.sess_bad_length_at[sx - 1] = ZK_NONE;
			a.o.sess_first_frame[sx - 1] = (uint32_t)a.n_frames;
		}
	}
	uint32_t rs = 0, rb = 0, ra = 0;       /* rows of the chunk placed so far (stat, buf, acl) */
	uint32_t nv = 0;                       /* vectors set aside */
	bool good = true;
	uint32_t n_slow = 0;

	uint32_t wlo = 0;
	do {
		/* ---- tile offsets of frames wlo .. wlo + ZK_FREL_CAP - 1; the first window also does the
		 * per-session bookkeeping ---- */
		if (simple && !events) {
#pragma unroll
			for (int i = 0; i < 8; ++i)
				if ((uint32_t)i < mycnt) c.frel[lbase + i] = (uint16_t)(pk[i >> 1] >> ((i & 1) * 16));
			if (mycnt > 8) {
				/* more than eight frames in 256 bytes (ping replies, notifications): a plain run all the
				 * same, followed on from the eighth */
				uint32_t rel = pk[3] >> 16;
				for (uint32_t i = 8; i < mycnt; ++i) {
					rel += 4u + sm_be32(c.tile, rel);
					c.frel[lbase + i] = (uint16_t)rel;
				}
			}
		} else {
			walk_subrange<true>(a, c, cs, hi, entry, es, esend, wlo, lbase, wlo == 0, f0 + lbase, e_start);
		}
		__syncwarp();
		const uint32_t wend = (nfr - wlo < ZK_FREL_CAP) ? nfr : wlo + ZK_FREL_CAP;
		for (uint32_t k0 = wlo; k0 < wend; k0 += 32) {
			if (nv > ZK_VL_CAP - 32) {
				__syncwarp();
				good = drain_vectors<false>(a, c, cs, f0, e0, e1, e2, nv, lane) && good;
				nv = 0;
				__syncwarp();
			}
			const uint32_t k = k0 + lane;
			const bool valid = k < wend;
			const uint32_t rel = valid ? c.frel[k - wlo] : 0u;
			const uint32_t f = f0 + k;
			const uint64_t foff = cs + rel;
			const int32_t flen = valid ? (int32_t)sm_be32(c.tile, rel) : 0;
			/* which session, and where that session's frames begin */
			uint32_t s = sf - 1, first = sfirst;
			if (!one_session) {
				uint32_t j = 0;            /* boundaries sess_off[sf + l], l < 32, at or below the frame */
#pragma unroll
				for (int step = 16; step > 0; step >>= 1) {
					const uint64_t bv = __shfl_sync(0xffffffffu, boff, (int)(j + step - 1));
					if (bv <= foff) j += step;
				}
				{
					const uint64_t bv = __shfl_sync(0xffffffffu, boff, 31);
					if (j == 31 && bv <= foff) j = 32;
				}
				if (j == 32 && valid) {
					uint32_t l2 = sf + 31, h2 = a.nsess;          /* last s with sess_off[s] <= foff */
					while (h2 - l2 > 1) {
						const uint32_t mid = (l2 + h2) >> 1;
						if (__ldg(a.sess_off + mid) <= foff) l2 = mid; else h2 = mid;
					}
					s = l2;
				} else if (j) {
					s = sf + j - 1;
				}
				/* a session that starts inside the chunk had its first frame written by the walk above
				 * (same warp, ordered by the __syncwarp) */
				if (valid && s != sf - 1) first = __ldcg(a.o.sess_first_frame + s);
			}
			Fr fr;
			fr.nst = fr.nbf = fr.nac = 0; fr.b0 = fr.b1 = fr.st = 0; fr.vrel = 0; fr.vkind = 0; fr.slow = 0;
			if (valid) {
				a.o.frame_off[f] = foff;
				a.o.frame_len[f] = flen;
				a.o.session[f] = s;
				const bool inwin = rel + 4u + (uint32_t)flen <= wlen;
				const bool server = DIR == 1 ? true : DIR == 2 ? false : (a.sess_flags && (__ldg(a.sess_flags + s) & ZKB_SESS_SERVER));
				const uint32_t hs = a.sess_hs ? __ldg(a.sess_hs + s) : 1u;
				const uint32_t idx = f - first;
				FrameCols o;
				bool done = false;
				if (inwin && idx >= hs)
					done = server ? direct_request(c.tile, rel + 4, (uint32_t)flen, o, fr)
					              : direct_response(c.tile, a, rel + 4, (uint32_t)flen, s, xs, o, fr);
				if (!done && !inwin && !server && idx >= hs &&
				    direct_response_long(c, a, rel + 4, (uint32_t)flen, wlen, s, o)) {
					done = true;
					c.l_foff = foff; c.l_flen = flen; c.l_f = f; c.have_long = 2;
				}
				if (!done) {
					/* handshake frames, malformed frames, frames longer than the tile: exact reader */
					Parsed p;
					n_slow++;
					parse_frame<DIR>(rd, a, foff, flen, s, idx, p);
					o.meta = p.meta; o.xid = p.xid; o.zxid = p.zxid; o.err = p.err; o.a0 = p.a0; o.a1 = p.a1;
					fr.nst = fr.nbf = fr.nac = 0; fr.b0 = fr.b1 = fr.st = 0; fr.vrel = 0; fr.vkind = 0;
					if (p.n_stat | p.n_buf | p.n_acl) {
						if (inwin) { fr.slow = 1; fr.nst = p.n_stat; fr.nbf = p.n_buf; fr.nac = p.n_acl; }
						else { fr.slow = 2; c.lp = p; c.l_foff = foff; c.l_flen = flen; c.l_f = f; c.have_long = 1; }
					}
				}
				a.o.meta[f] = o.meta;
				a.o.xid[f] = o.xid;
				a.o.zxid[f] = o.zxid;
				a.o.err[f] = o.err;
				a.o.a0[f] = o.a0;
				a.o.a1[f] = o.a1;
			}
			/* ---- every frame's first rows: one scan over the round (buf and acl counts share a word;
			 * at most one Stat per frame) ---- */
			const uint32_t packed = fr.nbf | (fr.nac << 17);
			const uint32_t inc = warp_incl_scan(packed, lane);
			const uint32_t stm = __ballot_sync(0xffffffffu, fr.nst != 0);
			const uint32_t ex = inc - packed;
			const uint32_t brow = e1 + rb + (ex & 0x1ffffu), arow = e2 + ra + (ex >> 17);
			const uint32_t srow = e0 + rs + __popc(stm & ((1u << lane) - 1u));
			const uint32_t tot = __shfl_sync(0xffffffffu, inc, 31);
			/* ---- rows ---- */
			if (fr.slow == 1) {
				Parsed p;
				const bool server = DIR == 1 ? true : DIR == 2 ? false : (a.sess_flags && (__ldg(a.sess_flags + s) & ZKB_SESS_SERVER));
				(void)server;
				parse_frame<DIR>(rd, a, foff, flen, s, f - first, p);
				emit_side(rd, a, foff, flen, p, f, srow, brow, arow);
			} else {
				if (fr.b0) put_buf(a.o, brow, f, cs + fr.b0 + 4, max((int32_t)sm_be32(c.tile, fr.b0), 0));
				if (fr.b1) put_buf(a.o, brow + 1, f, cs + fr.b1 + 4, max((int32_t)sm_be32(c.tile, fr.b1), 0));
				if (fr.st) put_stat(a.o, c.tile, fr.st, srow, f);
			}
			/* vectors wait for the drain */
			const uint32_t vm = __ballot_sync(0xffffffffu, fr.vkind != 0);
			if (fr.vkind) {
				VEnt e;
				e.w0 = (rel + 4) | (fr.vrel << 16);
				e.w1 = (brow + (fr.b0 ? 1u : 0u) + (fr.b1 ? 1u : 0u) - e1) | ((arow - e2) << 16);
				e.w2 = (srow - e0) | (fr.vkind << 16);
				e.w3 = k;
				c.vl[nv + __popc(vm & ((1u << lane) - 1u))] = e;
			}
			nv += __popc(vm);
			rb += tot & 0x1ffffu; ra += tot >> 17; rs += __popc(stm);
		}
		wlo += ZK_FREL_CAP;
		__syncwarp();
	} while (wlo < nfr);
	__syncwarp();
		/* request streams carry no string vectors (the ACL of a CREATE is walked lane by lane) */
#ifndef ZK_DRAIN_LIST
#define ZK_DRAIN_LIST (DIR != 1)
#endif
	if (nv) good = drain_vectors<ZK_DRAIN_LIST>(a, c, cs, f0, e0, e1, e2, nv, lane) && good;
	good = __all_sync(0xffffffffu, good);
	n_slow = warp_sum(n_slow);
	if (lane == 0 && n_slow) atomicAdd(&a.ctrl->n_slow, (unsigned long long)n_slow);

	/* the long frame's rows come last: it is the chunk's last frame */
	__syncwarp();
	uint32_t ts = rs, tb = rb, ta = ra;
	const uint32_t have_long = c.have_long;
	if (have_long) {
		ts += c.lp.n_stat; tb += c.lp.n_buf; ta += c.lp.n_acl;
		if (have_long == 2) {
			const uint64_t lf = c.l_foff; const uint32_t ll = (uint32_t)c.l_flen, lfr = c.l_f;
			const bool st = c.lp.n_stat != 0;
			__syncwarp();
			if (!REDO && a.long_items && ll > 4u * ZK_WIN) {
				/* many tiles long: listed for k_long_children (at most one per chunk, so the list cannot
				 * overflow); a frame that merely crosses the end of the tile is walked here */
				if (lane == 0) {
					LongItem it;
					it.foff = lf; it.flen = ll; it.frame = lfr; it.brow0 = e1 + rb; it.srow = e0 + rs;
					it.with_stat = st ? 1u : 0u; it.chunk = chunk;
					a.long_items[atomicAdd(&a.ctrl->n_long, 1ull)] = it;
				}
			} else {
				/* the work list and the frame list are free by now: they hold the element offsets */
				good = long_children(a, c.tile, reinterpret_cast<uint16_t *>(c.vl), ZK_BM_BYTES / 2u, bar, 1u, lf, ll, lfr,
				    e1 + rb, e0 + rs, st, lane) && good;
			}
		} else if (lane == 0) {
			const Parsed p = c.lp;
			emit_side(rd, a, c.l_foff, c.l_flen, p, c.l_f, e0 + rs, e1 + rb, e2 + ra);
		}
	}
	if (!REDO && lane == 0) {
		/* what the chunk really has (a frame that broke its promise owns no rows: the exact body
		 * counts them when the chunk is done again) */
		a.c_rows[chunk] = make_uint4(good ? 0u : 1u, ts, tb, ta);
		if (!good || ts != p0 || tb != p1 || ta != p2) {
			atomicOr(&a.ctrl->error, ZK_DERR_MISPREDICT);
			atomicMin(&a.ctrl->first_bad, chunk);
		}
	}
}
