/*
 * zkoracle.c — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * A plain-C, single-frame-at-a-time CPU restatement of node-zkstream's codec hot
 * path, used only as the checker for the CUDA library (tests/, __graft_entry__.smoke)
 * and as the timed CPU baseline of bench.py.  Nothing in the product links, loads
 * or calls this file; libzkbatch.so has no CPU path.
 *
 * Parity status: PINNED.  tests/test_oracle_golden.py checks this file against the
 * reference's own golden vectors (test/streams.test.js:21-56 CAPTURE1/DECODED1,
 * test/nasty.test.js:125-126,148,170 bad length prefixes), against known-answer
 * vectors produced by executing the reference's unmodified JavaScript under
 * oracle/_ref/js (tests/golden/ JSON files, generator: oracle/make_golden.py), and —
 * where /root/reference is present — against that JavaScript run live on fuzzed
 * corpora (tests/test_oracle_vs_reference_js.py).
 *
 * Each function cites the reference lines it follows (paths relative to the
 * node-zkstream tree).  Output layout: include/zkbatch.h (zkb_decoded).
 */
#define _GNU_SOURCE
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "../include/zkbatch.h"

/* ------------------------------------------------------------------------------ */
/* constant tables: lib/zk-consts.js                                              */

/* OP_CODE_LOOKUP (zk-consts.js:84-109): is this wire integer a named opcode? */
static int op_known(int32_t v)
{
	switch (v) {
	case 0: case 1: case 2: case 3: case 4: case 5: case 6: case 7: case 8: case 9:
	case 11: case 12: case 13: case 14: case 100: case 101: case 102:
	case -10: case -11: case -1:
		return 1;
	default:
		return 0;
	}
}

/* ERR_LOOKUP[v] === 'OK' (zk-consts.js:26-51) */
static int err_is_ok(int32_t v) { return v == 0; }

/* ------------------------------------------------------------------------------ */
/* growable columns                                                               */

typedef struct {
	void  *p;
	size_t n, cap, esz;
} col_t;

static void col_init(col_t *c, size_t esz) { c->p = NULL; c->n = 0; c->cap = 0; c->esz = esz; }
static void *col_push(col_t *c)
{
	if (c->n == c->cap) {
		size_t nc = c->cap ? c->cap * 2 : 1024;
		void *np = realloc(c->p, nc * c->esz);
		if (!np) { fprintf(stderr, "zkoracle: out of memory\n"); abort(); }
		c->p = np; c->cap = nc;
	}
	return (char *)c->p + (c->n++) * c->esz;
}
#define PUSH(col, type, v) (*(type *)col_push(&(col)) = (v))

typedef struct zko_out {
	/* frames */
	col_t frame_off, frame_len, session, meta, xid, zxid, err, a0, a1;
	/* stat */
	col_t stat_frame, czxid, mzxid, ctime, mtime, version, cversion, aversion,
	    ephemeralOwner, dataLength, numChildren, pzxid;
	/* buf */
	col_t buf_frame, buf_off, buf_len;
	/* acl */
	col_t acl_frame, acl_perms;
	uint64_t *sess_consumed, *sess_bad_length_at;
	uint32_t *sess_first_frame;
	uint32_t nsessions;
} zko_out;

static void out_init(zko_out *o, uint32_t nsessions)
{
	col_init(&o->frame_off, 8); col_init(&o->frame_len, 4); col_init(&o->session, 4);
	col_init(&o->meta, 4); col_init(&o->xid, 4); col_init(&o->zxid, 8); col_init(&o->err, 4);
	col_init(&o->a0, 4); col_init(&o->a1, 4);
	col_init(&o->stat_frame, 4);
	col_init(&o->czxid, 8); col_init(&o->mzxid, 8); col_init(&o->ctime, 8); col_init(&o->mtime, 8);
	col_init(&o->version, 4); col_init(&o->cversion, 4); col_init(&o->aversion, 4);
	col_init(&o->ephemeralOwner, 8); col_init(&o->dataLength, 4); col_init(&o->numChildren, 4);
	col_init(&o->pzxid, 8);
	col_init(&o->buf_frame, 4); col_init(&o->buf_off, 8); col_init(&o->buf_len, 4);
	col_init(&o->acl_frame, 4); col_init(&o->acl_perms, 4);
	o->nsessions = nsessions;
	o->sess_consumed = calloc(nsessions + 1, 8);
	o->sess_bad_length_at = calloc(nsessions + 1, 8);
	o->sess_first_frame = calloc(nsessions + 1, 4);
}

static void out_free(zko_out *o)
{
	col_t *cols[] = { &o->frame_off, &o->frame_len, &o->session, &o->meta, &o->xid, &o->zxid,
	    &o->err, &o->a0, &o->a1, &o->stat_frame, &o->czxid, &o->mzxid, &o->ctime, &o->mtime,
	    &o->version, &o->cversion, &o->aversion, &o->ephemeralOwner, &o->dataLength,
	    &o->numChildren, &o->pzxid, &o->buf_frame, &o->buf_off, &o->buf_len, &o->acl_frame,
	    &o->acl_perms };
	for (size_t i = 0; i < sizeof (cols) / sizeof (cols[0]); i++) free(cols[i]->p);
	free(o->sess_consumed); free(o->sess_bad_length_at); free(o->sess_first_frame);
}

/* ------------------------------------------------------------------------------ */
/* JuteBuffer read side: lib/jute-buffer.js:14-21,46-109                           */

typedef struct {
	const uint8_t *buf;   /* jb_buffer: the frame payload */
	uint64_t abs;         /* where buf[0] sits in the batch (for offsets in the BUF table) */
	int64_t size;         /* jb_size */
	int64_t off;          /* jb_offset */
	int cause, field;     /* first exception thrown, if any */
} jute_t;

static void jb_throw(jute_t *jb, int cause, int field)
{
	if (!jb->cause) { jb->cause = cause; jb->field = field; }
}

/* readInt, jute-buffer.js:58-62: Buffer.readInt32BE throws RangeError past the end */
static int32_t jb_read_int(jute_t *jb, int field)
{
	if (jb->cause) return 0;
	if (jb->off + 4 > jb->size) { jb_throw(jb, ZKB_CAUSE_RANGE, field); return 0; }
	const uint8_t *p = jb->buf + jb->off;
	uint32_t v = ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3];
	jb->off += 4;
	return (int32_t)v;
}

/* readLongBuffer, jute-buffer.js:64-69: assert.ok(off + 8 <= size); 8 raw bytes */
static int64_t jb_read_long(jute_t *jb, int field)
{
	if (jb->cause) return 0;
	if (!(jb->off + 8 <= jb->size)) { jb_throw(jb, ZKB_CAUSE_ASSERT, field); return 0; }
	const uint8_t *p = jb->buf + jb->off;
	uint64_t v = 0;
	for (int i = 0; i < 8; i++) v = (v << 8) | p[i];
	jb->off += 8;
	return (int64_t)v;
}

/* readBool, jute-buffer.js:51-56 */
static int32_t jb_read_bool(jute_t *jb, int field)
{
	if (jb->cause) return 0;
	if (!(jb->off < jb->size)) { jb_throw(jb, ZKB_CAUSE_ASSERT, field); return 0; }
	uint8_t v = jb->buf[jb->off++];
	if (!(v == 0 || v == 1)) { jb_throw(jb, ZKB_CAUSE_ASSERT, field); return 0; }
	return v == 1;
}

/* readBuffer / readUString, jute-buffer.js:97-109: negative length reads as empty.
 * The bytes are not copied or utf-8 decoded: the row records where they are. */
typedef struct { uint64_t off; int32_t len; } bufref_t;
static bufref_t jb_read_buffer(jute_t *jb, int field)
{
	bufref_t r = { 0, 0 };
	int32_t len = jb_read_int(jb, field);
	if (jb->cause) return r;
	if (len < 0) len = 0;
	if (!(jb->off + len <= jb->size)) { jb_throw(jb, ZKB_CAUSE_ASSERT, field); return r; }
	r.off = jb->abs + (uint64_t)jb->off;
	r.len = len;
	jb->off += len;
	return r;
}

/* ------------------------------------------------------------------------------ */
/* one frame's worth of pending rows (committed only if the reader did not throw) */

#define MAXPEND 64
typedef struct {
	int32_t xid, err, a0, a1;
	int64_t zxid;
	int opcode;
	int has_stat;
	int64_t st64[6];  /* czxid mzxid ctime mtime ephemeralOwner pzxid */
	int32_t st32[5];  /* version cversion aversion dataLength numChildren */
	bufref_t *bufs; size_t nbufs, capbufs;
	int32_t *perms; size_t nperms, capperms;
} pend_t;

static void pend_reset(pend_t *p)
{
	p->xid = p->err = p->a0 = p->a1 = 0; p->zxid = 0; p->opcode = ZKB_OP_NONE;
	p->has_stat = 0; p->nbufs = 0; p->nperms = 0;
}
static void pend_buf(pend_t *p, bufref_t r)
{
	if (p->nbufs == p->capbufs) {
		p->capbufs = p->capbufs ? p->capbufs * 2 : MAXPEND;
		p->bufs = realloc(p->bufs, p->capbufs * sizeof (bufref_t));
	}
	p->bufs[p->nbufs++] = r;
}
static void pend_perm(pend_t *p, int32_t v)
{
	if (p->nperms == p->capperms) {
		p->capperms = p->capperms ? p->capperms * 2 : MAXPEND;
		p->perms = realloc(p->perms, p->capperms * sizeof (int32_t));
	}
	p->perms[p->nperms++] = v;
}

/* ------------------------------------------------------------------------------ */
/* ZKBuffer readers: lib/zk-buffer.js                                              */

/* readStat, zk-buffer.js:428-442 */
static void read_stat(jute_t *jb, pend_t *p)
{
	p->st64[0] = jb_read_long(jb, ZKB_F_STAT);  /* czxid */
	p->st64[1] = jb_read_long(jb, ZKB_F_STAT);  /* mzxid */
	p->st64[2] = jb_read_long(jb, ZKB_F_STAT);  /* ctime (LongDate over the same 8 bytes) */
	p->st64[3] = jb_read_long(jb, ZKB_F_STAT);  /* mtime */
	p->st32[0] = jb_read_int(jb, ZKB_F_STAT);   /* version */
	p->st32[1] = jb_read_int(jb, ZKB_F_STAT);   /* cversion */
	p->st32[2] = jb_read_int(jb, ZKB_F_STAT);   /* aversion */
	p->st64[4] = jb_read_long(jb, ZKB_F_STAT);  /* ephemeralOwner */
	p->st32[3] = jb_read_int(jb, ZKB_F_STAT);   /* dataLength */
	p->st32[4] = jb_read_int(jb, ZKB_F_STAT);   /* numChildren */
	p->st64[5] = jb_read_long(jb, ZKB_F_STAT);  /* pzxid */
	if (!jb->cause) p->has_stat = 1;
}

/* readACL + readPerms + readID, zk-buffer.js:372-382,395-403,416-421.
 * perms stay raw; the `val & MASK != 0` quirk is applied by the binding. */
static void read_acl(jute_t *jb, pend_t *p, int32_t *n_out)
{
	int32_t count = jb_read_int(jb, ZKB_F_COUNT);
	if (jb->cause) return;
	*n_out = count > 0 ? count : 0;   /* count column is set as soon as the count is read */
	for (int32_t i = 0; i < count; ++i) {
		int32_t perms = jb_read_int(jb, ZKB_F_ACL_PERMS);
		bufref_t scheme = jb_read_buffer(jb, ZKB_F_ACL_SCHEME);
		bufref_t id = jb_read_buffer(jb, ZKB_F_ACL_ID);
		if (jb->cause) return;
		pend_perm(p, perms); pend_buf(p, scheme); pend_buf(p, id);
	}
}

/* a vector<ustring>: zk-buffer.js:334-337 and 238-252; count <= 0 reads as empty */
static void read_strvec(jute_t *jb, pend_t *p, int32_t *n_out)
{
	int32_t count = jb_read_int(jb, ZKB_F_COUNT);
	if (jb->cause) return;
	*n_out = count > 0 ? count : 0;
	for (int32_t i = 0; i < count; ++i) {
		bufref_t e = jb_read_buffer(jb, ZKB_F_ELEM);
		if (jb->cause) return;
		pend_buf(p, e);
	}
}

/* SPECIAL_XIDS, zk-buffer.js:275-279 */
static int special_xid(int32_t xid)
{
	switch (xid) {
	case ZKB_XID_NOTIFICATION: return ZKB_OP_NOTIFICATION;
	case ZKB_XID_PING: return ZKB_OP_PING;
	case ZKB_XID_AUTH: return ZKB_OP_AUTH;
	case ZKB_XID_SET_WATCHES: return ZKB_OP_SET_WATCHES;
	default: return ZKB_OP_NONE;
	}
}

static int xid_lookup(const zkb_xidtab *t, uint32_t s, int32_t xid)
{
	if (!t || !t->ops) return ZKB_OP_NONE;
	int64_t idx = (int64_t)xid - (int64_t)t->base[s];
	if (idx < 0 || idx >= (int64_t)t->count[s]) return ZKB_OP_NONE;
	return t->ops[t->off[s] + (uint64_t)idx];
}

/* readResponse, zk-buffer.js:281-331 */
static void read_response(jute_t *jb, pend_t *p, const zkb_xidtab *t, uint32_t s)
{
	p->xid = jb_read_int(jb, ZKB_F_XID);
	if (jb->cause) return;
	p->zxid = jb_read_long(jb, ZKB_F_ZXID);
	if (jb->cause) return;
	p->err = jb_read_int(jb, ZKB_F_ERR);
	if (jb->cause) return;
	/* :288-291 — opcode is resolved, and asserted, before err is looked at */
	int op = special_xid(p->xid);
	if (op == ZKB_OP_NONE) op = xid_lookup(t, s, p->xid);
	p->opcode = op;
	if (op == ZKB_OP_NONE) { jb_throw(jb, ZKB_CAUSE_NO_MATCH, ZKB_F_OPCODE); return; }
	if (!err_is_ok(p->err)) return;   /* :292 — unknown codes are also "not OK" */
	switch (op) {
	case ZKB_OP_GET_CHILDREN:
	case ZKB_OP_GET_CHILDREN2:            /* readGetChildrenResponse :333-340 */
		read_strvec(jb, p, &p->a0);
		if (jb->cause) return;
		if (op == ZKB_OP_GET_CHILDREN2) read_stat(jb, p);
		break;
	case ZKB_OP_CREATE:                   /* readCreateResponse :342-344 */
		{ bufref_t r = jb_read_buffer(jb, ZKB_F_PATH); if (!jb->cause) pend_buf(p, r); }
		break;
	case ZKB_OP_GET_ACL:                  /* readGetACLResponse :354-357 */
		read_acl(jb, p, &p->a0);
		if (jb->cause) return;
		read_stat(jb, p);
		break;
	case ZKB_OP_GET_DATA:                 /* readGetDataResponse :359-362 */
		{ bufref_t r = jb_read_buffer(jb, ZKB_F_DATA); if (!jb->cause) pend_buf(p, r); }
		read_stat(jb, p);
		break;
	case ZKB_OP_NOTIFICATION:             /* readNotification :364-370 */
		p->a0 = jb_read_int(jb, ZKB_F_NTYPE);
		if (jb->cause) return;
		p->a1 = jb_read_int(jb, ZKB_F_NSTATE);
		if (jb->cause) return;
		{ bufref_t r = jb_read_buffer(jb, ZKB_F_PATH); if (!jb->cause) pend_buf(p, r); }
		break;
	case ZKB_OP_EXISTS:                   /* readExistsResponse :346-348 */
	case ZKB_OP_SET_DATA:                 /* readSetDataResponse :350-352 */
		read_stat(jb, p);
		break;
	case ZKB_OP_SET_WATCHES:
	case ZKB_OP_PING:
	case ZKB_OP_SYNC:
	case ZKB_OP_DELETE:
	case ZKB_OP_CLOSE_SESSION:            /* :316-325 header only */
		break;
	default:                              /* :326-327 */
		jb_throw(jb, ZKB_CAUSE_UNSUPPORTED, ZKB_F_OPCODE);
		break;
	}
}

/* readRequest + sub-readers, zk-buffer.js:58-95,138-253 */
static void read_request(jute_t *jb, pend_t *p)
{
	p->xid = jb_read_int(jb, ZKB_F_XID);
	if (jb->cause) return;
	int32_t type = jb_read_int(jb, ZKB_F_OPCODE);
	if (jb->cause) return;
	int op = op_known(type) ? type : ZKB_OP_NONE;   /* OP_CODE_LOOKUP[...] may be undefined */
	p->opcode = op;
	bufref_t r;
	switch (op) {
	case ZKB_OP_GET_CHILDREN:
	case ZKB_OP_GET_CHILDREN2:            /* :138-141 */
	case ZKB_OP_GET_DATA:                 /* :193-196 */
	case ZKB_OP_EXISTS:                   /* :203-206 */
		r = jb_read_buffer(jb, ZKB_F_PATH); if (jb->cause) return; pend_buf(p, r);
		p->a0 = jb_read_bool(jb, ZKB_F_WATCH);
		break;
	case ZKB_OP_CREATE:                   /* :148-159 */
		r = jb_read_buffer(jb, ZKB_F_PATH); if (jb->cause) return; pend_buf(p, r);
		r = jb_read_buffer(jb, ZKB_F_DATA); if (jb->cause) return; pend_buf(p, r);
		read_acl(jb, p, &p->a1); if (jb->cause) return;
		p->a0 = jb_read_int(jb, ZKB_F_FLAGS);
		break;
	case ZKB_OP_DELETE:                   /* :175-178 */
		r = jb_read_buffer(jb, ZKB_F_PATH); if (jb->cause) return; pend_buf(p, r);
		p->a0 = jb_read_int(jb, ZKB_F_VERSION);
		break;
	case ZKB_OP_GET_ACL:                  /* :185-187 */
	case ZKB_OP_SYNC:                     /* :225-227 */
		r = jb_read_buffer(jb, ZKB_F_PATH); if (jb->cause) return; pend_buf(p, r);
		break;
	case ZKB_OP_SET_DATA:                 /* :213-217 */
		r = jb_read_buffer(jb, ZKB_F_PATH); if (jb->cause) return; pend_buf(p, r);
		r = jb_read_buffer(jb, ZKB_F_DATA); if (jb->cause) return; pend_buf(p, r);
		p->a0 = jb_read_int(jb, ZKB_F_VERSION);
		break;
	case ZKB_OP_SET_WATCHES:              /* :233-253 */
		p->zxid = jb_read_long(jb, ZKB_F_REL_ZXID); if (jb->cause) return;
		read_strvec(jb, p, &p->a0); if (jb->cause) return;
		read_strvec(jb, p, &p->a1); if (jb->cause) return;
		read_strvec(jb, p, &p->err);
		break;
	default:                              /* :91-92 — PING, CLOSE_SESSION, everything else */
		jb_throw(jb, ZKB_CAUSE_UNSUPPORTED, ZKB_F_OPCODE);
		break;
	}
}

/* readConnectResponse, zk-buffer.js:41-48 */
static void read_connect_response(jute_t *jb, pend_t *p)
{
	p->xid = jb_read_int(jb, ZKB_F_PROTOCOL_VERSION); if (jb->cause) return;
	p->err = jb_read_int(jb, ZKB_F_TIMEOUT); if (jb->cause) return;
	p->zxid = jb_read_long(jb, ZKB_F_SESSION_ID); if (jb->cause) return;
	bufref_t r = jb_read_buffer(jb, ZKB_F_PASSWD); if (jb->cause) return;
	pend_buf(p, r);
}

/* readConnectRequest, zk-buffer.js:22-30 */
static void read_connect_request(jute_t *jb, pend_t *p)
{
	p->xid = jb_read_int(jb, ZKB_F_PROTOCOL_VERSION); if (jb->cause) return;
	int64_t last = jb_read_long(jb, ZKB_F_LAST_ZXID); if (jb->cause) return;
	p->a0 = (int32_t)((uint64_t)last >> 32);
	p->a1 = (int32_t)((uint64_t)last & 0xffffffffu);
	p->err = jb_read_int(jb, ZKB_F_TIMEOUT); if (jb->cause) return;
	p->zxid = jb_read_long(jb, ZKB_F_SESSION_ID); if (jb->cause) return;
	bufref_t r = jb_read_buffer(jb, ZKB_F_PASSWD); if (jb->cause) return;
	pend_buf(p, r);
}

/* ------------------------------------------------------------------------------ */
/* handlePacket, zk-streams.js:66-99                                               */

static void handle_packet(zko_out *o, pend_t *p, const uint8_t *pkt, uint64_t abs_prefix,
    int32_t len, uint32_t s, int is_server, int handshaking, const zkb_xidtab *t)
{
	jute_t jb = { pkt, abs_prefix + 4, len, 0, 0, 0 };
	int kind;
	pend_reset(p);
	if (handshaking) {
		if (is_server) { kind = ZKB_KIND_CONNECT_REQUEST; read_connect_request(&jb, p); }
		else { kind = ZKB_KIND_CONNECT_RESPONSE; read_connect_response(&jb, p); }
	} else {
		if (is_server) { kind = ZKB_KIND_REQUEST; read_request(&jb, p); }
		else { kind = ZKB_KIND_RESPONSE; read_response(&jb, p, t, s); }
	}
	uint32_t row = (uint32_t)o->frame_off.n;
	uint32_t status = jb.cause ? ZKB_ST_BAD_DECODE : ZKB_ST_OK;
	uint32_t detail = jb.cause ? (uint32_t)(jb.cause | (jb.field << 3)) : 0;
	uint32_t meta = status | ((uint32_t)(uint8_t)(int8_t)p->opcode << 8) | ((uint32_t)kind << 16) |
	    (detail << 24);
	PUSH(o->frame_off, uint64_t, abs_prefix);
	PUSH(o->frame_len, int32_t, len);
	PUSH(o->session, uint32_t, s);
	PUSH(o->meta, uint32_t, meta);
	PUSH(o->xid, int32_t, p->xid);
	PUSH(o->zxid, int64_t, p->zxid);
	PUSH(o->err, int32_t, p->err);
	PUSH(o->a0, int32_t, p->a0);
	PUSH(o->a1, int32_t, p->a1);
	if (jb.cause) return;   /* the packet is never pushed: it owns no side rows */
	if (p->has_stat) {
		PUSH(o->stat_frame, uint32_t, row);
		PUSH(o->czxid, int64_t, p->st64[0]); PUSH(o->mzxid, int64_t, p->st64[1]);
		PUSH(o->ctime, int64_t, p->st64[2]); PUSH(o->mtime, int64_t, p->st64[3]);
		PUSH(o->version, int32_t, p->st32[0]); PUSH(o->cversion, int32_t, p->st32[1]);
		PUSH(o->aversion, int32_t, p->st32[2]); PUSH(o->ephemeralOwner, int64_t, p->st64[4]);
		PUSH(o->dataLength, int32_t, p->st32[3]); PUSH(o->numChildren, int32_t, p->st32[4]);
		PUSH(o->pzxid, int64_t, p->st64[5]);
	}
	for (size_t i = 0; i < p->nbufs; i++) {
		PUSH(o->buf_frame, uint32_t, row);
		PUSH(o->buf_off, uint64_t, p->bufs[i].off);
		PUSH(o->buf_len, int32_t, p->bufs[i].len);
	}
	for (size_t i = 0; i < p->nperms; i++) {
		PUSH(o->acl_frame, uint32_t, row);
		PUSH(o->acl_perms, int32_t, p->perms[i]);
	}
}

static int32_t rd_be32(const uint8_t *p)
{
	return (int32_t)(((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]);
}

/* ------------------------------------------------------------------------------ */
/* ZKDecodeStream._transform, zk-streams.js:39-65                                  */

/*
 * mode 0 ("in place"): walk the session's bytes once, no copies.  Same results as the
 * reference for any chunking, because the loop at :47-64 only ever looks at the front
 * of zd_buffer.
 * mode 1 ("faithful"): does what the reference does byte for byte — chunks of `chunk`
 * bytes are appended to a doubling staging buffer (:43-46,102-106), each frame is copied
 * into a fresh allocation (:55-56) and the rest of the staging buffer is moved down to
 * offset 0 (:57-58).  Used for the timed CPU baseline.
 */
static void decode_session(zko_out *o, pend_t *p, const uint8_t *bytes, uint64_t lo, uint64_t hi,
    uint32_t s, int is_server, uint32_t hs, const zkb_xidtab *t, int mode, uint64_t chunk)
{
	uint32_t nframes = 0;
	o->sess_first_frame[s] = (uint32_t)o->frame_off.n;
	o->sess_bad_length_at[s] = ZKB_NO_OFFSET;
	if (mode == 0) {
		uint64_t pos = lo;
		while (hi - pos >= 4) {
			int32_t len = rd_be32(bytes + pos);
			if (len < 0 || len > ZKB_MAX_PACKET) {     /* :49-53 emit BAD_LENGTH, return */
				o->sess_bad_length_at[s] = pos;
				break;
			}
			if (hi - pos >= 4 + (uint64_t)len) {
				handle_packet(o, p, bytes + pos + 4, pos, len, s, is_server, nframes < hs, t);
				nframes++;
				pos += 4 + (uint64_t)len;
			} else {
				break;
			}
		}
		o->sess_consumed[s] = pos;
		return;
	}
	/* faithful */
	uint64_t zd_size = 4096, zd_used = 0, consumed = lo;
	uint8_t *zd_buffer = malloc(zd_size);
	int dead = 0;
	if (chunk == 0) chunk = 65536;
	for (uint64_t c = lo; c < hi && !dead; c += chunk) {
		uint64_t clen = (hi - c < chunk) ? hi - c : chunk;
		while (zd_size < zd_used + clen) {             /* expandBuffer :102-106 */
			uint8_t *nb = malloc(zd_size * 2);
			memcpy(nb, zd_buffer, zd_size);
			free(zd_buffer); zd_buffer = nb; zd_size *= 2;
		}
		memcpy(zd_buffer + zd_used, bytes + c, clen);
		zd_used += clen;
		while (zd_used >= 4) {
			int32_t len = rd_be32(zd_buffer);
			if (len < 0 || len > ZKB_MAX_PACKET) {
				o->sess_bad_length_at[s] = consumed;
				dead = 1;
				break;
			}
			if (zd_used >= 4 + (uint64_t)len) {
				uint8_t *pkt = malloc(len ? len : 1);      /* new Buffer(len) */
				memcpy(pkt, zd_buffer + 4, len);
				memmove(zd_buffer, zd_buffer + 4 + len, zd_used - (4 + (uint64_t)len));
				zd_used -= 4 + (uint64_t)len;
				handle_packet(o, p, pkt, consumed, len, s, is_server, nframes < hs, t);
				free(pkt);
				nframes++;
				consumed += 4 + (uint64_t)len;
			} else {
				break;
			}
		}
	}
	free(zd_buffer);
	o->sess_consumed[s] = consumed;
}

static int check_sessions(uint64_t nbytes, const zkb_sessions *sess)
{
	if (!sess || !sess->off) return ZKB_EINVAL;
	if (sess->off[0] != 0 || sess->off[sess->nsessions] != nbytes) return ZKB_EINVAL;
	for (uint32_t s = 0; s < sess->nsessions; s++)
		if (sess->off[s] > sess->off[s + 1]) return ZKB_EINVAL;
	return 0;
}

static void decode_range(zko_out *o, const uint8_t *bytes, const zkb_sessions *sess,
    const zkb_xidtab *t, uint32_t s0, uint32_t s1, int mode, uint64_t chunk)
{
	pend_t p; memset(&p, 0, sizeof (p));
	for (uint32_t s = s0; s < s1; s++) {
		int is_server = sess->flags ? (sess->flags[s] & ZKB_SESS_SERVER) != 0 : 0;
		uint32_t hs = sess->handshake_frames ? sess->handshake_frames[s] : 1;
		decode_session(o, &p, bytes, sess->off[s], sess->off[s + 1], s, is_server, hs, t, mode,
		    chunk);
	}
	free(p.bufs); free(p.perms);
}

/* ---- public oracle API ------------------------------------------------------------- */

/* Decode a whole batch on one thread.  Fills `out` with pointers into *handle;
 * release with zko_free(handle). */
int zko_decode(const uint8_t *bytes, uint64_t nbytes, const zkb_sessions *sess,
    const zkb_xidtab *t, int mode, uint64_t chunk, zkb_decoded *out, void **handle)
{
	int rc = check_sessions(nbytes, sess);
	if (rc) return rc;
	zko_out *o = calloc(1, sizeof (*o));
	out_init(o, sess->nsessions);
	decode_range(o, bytes, sess, t, 0, sess->nsessions, mode, chunk);
	o->sess_first_frame[sess->nsessions] = (uint32_t)o->frame_off.n;
	memset(out, 0, sizeof (*out));
	out->n_frames = o->frame_off.n; out->n_stat = o->stat_frame.n;
	out->n_buf = o->buf_frame.n; out->n_acl = o->acl_frame.n;
	out->frame_off = o->frame_off.p; out->frame_len = o->frame_len.p; out->session = o->session.p;
	out->meta = o->meta.p; out->xid = o->xid.p; out->zxid = o->zxid.p; out->err = o->err.p;
	out->a0 = o->a0.p; out->a1 = o->a1.p;
	out->stat_frame = o->stat_frame.p; out->czxid = o->czxid.p; out->mzxid = o->mzxid.p;
	out->ctime = o->ctime.p; out->mtime = o->mtime.p; out->version = o->version.p;
	out->cversion = o->cversion.p; out->aversion = o->aversion.p;
	out->ephemeralOwner = o->ephemeralOwner.p; out->dataLength = o->dataLength.p;
	out->numChildren = o->numChildren.p; out->pzxid = o->pzxid.p;
	out->buf_frame = o->buf_frame.p; out->buf_off = o->buf_off.p; out->buf_len = o->buf_len.p;
	out->acl_frame = o->acl_frame.p; out->acl_perms = o->acl_perms.p;
	out->sess_consumed = o->sess_consumed; out->sess_bad_length_at = o->sess_bad_length_at;
	out->sess_first_frame = o->sess_first_frame;
	*handle = o;
	return 0;
}

void zko_free(void *handle)
{
	if (!handle) return;
	out_free((zko_out *)handle);
	free(handle);
}

/* ---- timed CPU baseline: sessions sharded over threads ---------------------------------- */

typedef struct {
	const uint8_t *bytes; const zkb_sessions *sess; const zkb_xidtab *t;
	uint32_t s0, s1; int mode; uint64_t chunk;
	uint64_t frames, ok, checksum;
} job_t;

static void *bench_thread(void *arg)
{
	job_t *j = arg;
	zko_out o;
	/* decode in groups of sessions so the result columns stay cache sized, as a streaming
	 * consumer would; the columns are folded into a checksum so the work cannot be elided */
	uint32_t s = j->s0;
	while (s < j->s1) {
		uint32_t e = s;
		uint64_t b = 0;
		while (e < j->s1 && b < (64u << 20)) { b += j->sess->off[e + 1] - j->sess->off[e]; e++; }
		out_init(&o, j->sess->nsessions);
		decode_range(&o, j->bytes, j->sess, j->t, s, e, j->mode, j->chunk);
		j->frames += o.frame_off.n;
		const uint32_t *meta = o.meta.p; const int64_t *zx = o.zxid.p;
		for (size_t i = 0; i < o.frame_off.n; i++) {
			j->ok += (meta[i] & 0xff) == 0;
			j->checksum += (uint64_t)zx[i] ^ meta[i];
		}
		out_free(&o);
		s = e;
	}
	return NULL;
}

/* Returns seconds of wall time; *frames gets the number of frames decoded. */
double zko_bench_decode(const uint8_t *bytes, uint64_t nbytes, const zkb_sessions *sess,
    const zkb_xidtab *t, int mode, uint64_t chunk, int nthreads, uint64_t *frames,
    uint64_t *checksum)
{
	if (check_sessions(nbytes, sess)) return -1.0;
	if (nthreads < 1) nthreads = 1;
	if ((uint32_t)nthreads > sess->nsessions && sess->nsessions) nthreads = sess->nsessions;
	job_t *jobs = calloc(nthreads, sizeof (job_t));
	pthread_t *th = calloc(nthreads, sizeof (pthread_t));
	/* contiguous session ranges balanced by bytes */
	uint32_t s = 0;
	for (int i = 0; i < nthreads; i++) {
		uint64_t target = nbytes / nthreads * (i + 1);
		uint32_t e = s;
		if (i == nthreads - 1) e = sess->nsessions;
		else while (e < sess->nsessions && sess->off[e + 1] <= target) e++;
		if (e == s && s < sess->nsessions && i < nthreads - 1) e = s + 1;
		jobs[i] = (job_t){ bytes, sess, t, s, e, mode, chunk, 0, 0, 0 };
		s = e;
	}
	struct timespec t0, t1;
	clock_gettime(CLOCK_MONOTONIC, &t0);
	for (int i = 0; i < nthreads; i++) pthread_create(&th[i], NULL, bench_thread, &jobs[i]);
	for (int i = 0; i < nthreads; i++) pthread_join(th[i], NULL);
	clock_gettime(CLOCK_MONOTONIC, &t1);
	uint64_t f = 0, c = 0;
	for (int i = 0; i < nthreads; i++) { f += jobs[i].frames; c += jobs[i].checksum + jobs[i].ok; }
	if (frames) *frames = f;
	if (checksum) *checksum = c;
	free(jobs); free(th);
	return (t1.tv_sec - t0.tv_sec) + (t1.tv_nsec - t0.tv_nsec) * 1e-9;
}

/* ---- whole-batch digest (bench.py's parity gate on the corpus it times) ------------------
 *
 * Decodes the batch on `nthreads` threads (contiguous session ranges, in place) and folds every
 * output column into one order-sensitive 64-bit digest:
 *     D(col) = sum over rows i of mix64(value_i zero-extended to 64 bits) * (i + 1)   (mod 2^64)
 * bench.py computes the same sums over the GPU's columns with integer tensor arithmetic; equal
 * digests for all 29 columns = same rows in the same places.  out_digest order: the FRAMES, STAT,
 * BUF, ACL columns as declared in zkb_decoded, then sess_consumed, sess_bad_length_at,
 * sess_first_frame[nsessions + 1].  out_counts = { frames, stat, buf, acl }.
 */
static inline uint64_t mix64(uint64_t v)
{
	v ^= v >> 33; v *= 0xff51afd7ed558ccdULL; v ^= v >> 29; v *= 0xc4ceb9fe1a85ec53ULL; v ^= v >> 32;
	return v;
}

static uint64_t digest_col(const col_t *c, uint64_t row_base, uint64_t val_base)
{
	uint64_t d = 0;
	if (c->esz == 8) {
		const uint64_t *p = c->p;
		for (size_t i = 0; i < c->n; i++) d += mix64(p[i] + val_base) * (row_base + i + 1);
	} else {
		const uint32_t *p = c->p;
		for (size_t i = 0; i < c->n; i++) d += mix64((uint64_t)(uint32_t)(p[i] + (uint32_t)val_base)) * (row_base + i + 1);
	}
	return d;
}

typedef struct {
	const uint8_t *bytes; const zkb_sessions *sess; const zkb_xidtab *t;
	uint32_t s0, s1;
	zko_out o;
	uint64_t base[4];      /* first FRAMES / STAT / BUF / ACL row of this thread's range */
	uint64_t dig[29];
} djob_t;

static void *digest_decode_thread(void *arg)
{
	djob_t *j = arg;
	out_init(&j->o, j->sess->nsessions);
	decode_range(&j->o, j->bytes, j->sess, j->t, j->s0, j->s1, 0, 0);
	return NULL;
}

static void *digest_fold_thread(void *arg)
{
	djob_t *j = arg;
	zko_out *o = &j->o;
	const col_t *cols[26] = { &o->frame_off, &o->frame_len, &o->session, &o->meta, &o->xid, &o->zxid,
	    &o->err, &o->a0, &o->a1, &o->stat_frame, &o->czxid, &o->mzxid, &o->ctime, &o->mtime,
	    &o->version, &o->cversion, &o->aversion, &o->ephemeralOwner, &o->dataLength,
	    &o->numChildren, &o->pzxid, &o->buf_frame, &o->buf_off, &o->buf_len, &o->acl_frame,
	    &o->acl_perms };
	for (int c = 0; c < 26; c++) {
		const int table = c < 9 ? 0 : c < 21 ? 1 : c < 24 ? 2 : 3;
		/* the *_frame columns hold rows of this thread's own FRAMES table: shift to batch rows */
		const uint64_t vb = (c == 9 || c == 21 || c == 24) ? j->base[0] : 0;
		j->dig[c] = digest_col(cols[c], j->base[table], vb);
	}
	uint64_t d0 = 0, d1 = 0, d2 = 0;
	for (uint32_t s = j->s0; s < j->s1; s++) {
		d0 += mix64(o->sess_consumed[s]) * ((uint64_t)s + 1);
		d1 += mix64(o->sess_bad_length_at[s]) * ((uint64_t)s + 1);
		d2 += mix64((uint64_t)(uint32_t)(o->sess_first_frame[s] + (uint32_t)j->base[0])) * ((uint64_t)s + 1);
	}
	j->dig[26] = d0; j->dig[27] = d1; j->dig[28] = d2;
	return NULL;
}

int zko_digest(const uint8_t *bytes, uint64_t nbytes, const zkb_sessions *sess, const zkb_xidtab *t,
    int nthreads, uint64_t out_digest[29], uint64_t out_counts[4])
{
	int rc = check_sessions(nbytes, sess);
	if (rc) return rc;
	if (nthreads < 1) nthreads = 1;
	if ((uint32_t)nthreads > sess->nsessions) nthreads = sess->nsessions ? (int)sess->nsessions : 1;
	djob_t *jobs = calloc(nthreads, sizeof (djob_t));
	pthread_t *th = calloc(nthreads, sizeof (pthread_t));
	uint32_t s = 0;
	for (int i = 0; i < nthreads; i++) {
		uint64_t target = nbytes / nthreads * (i + 1);
		uint32_t e = s;
		if (i == nthreads - 1) e = sess->nsessions;
		else while (e < sess->nsessions && sess->off[e + 1] <= target) e++;
		if (e == s && s < sess->nsessions && i < nthreads - 1) e = s + 1;
		jobs[i].bytes = bytes; jobs[i].sess = sess; jobs[i].t = t; jobs[i].s0 = s; jobs[i].s1 = e;
		s = e;
	}
	for (int i = 0; i < nthreads; i++) pthread_create(&th[i], NULL, digest_decode_thread, &jobs[i]);
	for (int i = 0; i < nthreads; i++) pthread_join(th[i], NULL);
	uint64_t run[4] = { 0, 0, 0, 0 };
	for (int i = 0; i < nthreads; i++) {
		memcpy(jobs[i].base, run, sizeof (run));
		run[0] += jobs[i].o.frame_off.n; run[1] += jobs[i].o.stat_frame.n;
		run[2] += jobs[i].o.buf_frame.n; run[3] += jobs[i].o.acl_frame.n;
	}
	for (int i = 0; i < nthreads; i++) pthread_create(&th[i], NULL, digest_fold_thread, &jobs[i]);
	for (int i = 0; i < nthreads; i++) pthread_join(th[i], NULL);
	for (int c = 0; c < 29; c++) out_digest[c] = 0;
	for (int i = 0; i < nthreads; i++) {
		for (int c = 0; c < 29; c++) out_digest[c] += jobs[i].dig[c];
		out_free(&jobs[i].o);
	}
	/* sess_first_frame has one more entry: [nsessions] = n_frames */
	out_digest[28] += mix64((uint64_t)(uint32_t)run[0]) * ((uint64_t)sess->nsessions + 1);
	memcpy(out_counts, run, sizeof (run));
	free(jobs); free(th);
	return 0;
}

/* ------------------------------------------------------------------------------ */
/* xid map: zk-streams.js:145 — zcf_xidMap[pkt.xid] = pkt.opcode, last write wins  */

int zko_xidtab_build(const zkb_decoded *req, uint32_t nsessions, const zkb_xidtab *layout,
    int8_t *ops_out)
{
	for (uint64_t i = 0; i < layout->nops; i++) ops_out[i] = ZKB_OP_NONE;
	for (uint64_t f = 0; f < req->n_frames; f++) {
		uint32_t m = req->meta[f];
		if (ZKB_META_KIND(m) != ZKB_KIND_REQUEST) continue;
		/* a frame too short to hold xid + type was never a request the encoder wrote */
		if (req->frame_len[f] < 8) continue;
		uint32_t s = req->session[f];
		if (s >= nsessions) continue;
		int64_t idx = (int64_t)req->xid[f] - (int64_t)layout->base[s];
		if (idx < 0 || idx >= (int64_t)layout->count[s]) continue;
		ops_out[layout->off[s] + (uint64_t)idx] = ZKB_META_OPCODE(m);
	}
	return 0;
}

/* ------------------------------------------------------------------------------ */
/* ZKSession's per-packet folds (lib/zk-session.js:227-236, :389-419), packet by packet */

int zko_session_stats(const zkb_decoded *dec, uint32_t nsessions, uint32_t flags, zkb_session_stats *out)
{
	uint8_t *dead = calloc(nsessions ? nsessions : 1, 1);   /* connection torn down after a decoder error */
	for (uint32_t s = 0; s < nsessions; s++) {
		out->max_zxid[s] = INT64_MIN;
		out->n_replies[s] = 0;
		for (int k = 0; k < ZKB_NOTIFY_SLOTS; k++) out->notifications[(size_t)s * ZKB_NOTIFY_SLOTS + k] = 0;
	}
	for (uint64_t f = 0; f < dec->n_frames; f++) {
		uint32_t m = dec->meta[f];
		uint32_t s = dec->session[f];
		if (s >= nsessions) continue;
		if (ZKB_META_STATUS(m) != ZKB_ST_OK) {
			/* 'error' on the decoder: connection-fsm.js:234-239 moves the connection to 'error' */
			if (flags & ZKB_STATS_UNTIL_ERROR) dead[s] = 1;
			continue;
		}
		if (dead[s]) continue;
		/* only packets read with readResponse reach the attached session's handler; the
		 * connect response is consumed by state_attaching (zk-session.js:170-190) */
		if (ZKB_META_KIND(m) != ZKB_KIND_RESPONSE) continue;
		if (ZKB_META_OPCODE(m) != ZKB_OP_NOTIFICATION) {
			/* :229-235 — BigInteger(pkt.zxid).compareTo(zs_lastZxid) > 0, signed */
			if (out->n_replies[s] == 0 || dec->zxid[f] > out->max_zxid[s]) out->max_zxid[s] = dec->zxid[f];
			out->n_replies[s]++;
			continue;
		}
		/* processNotification :389-419.  readResponse reads the body only when err is OK
		 * (zk-buffer.js:296), so an error reply has no pkt.state */
		int slot;
		if (dec->err[f] != 0 || dec->a1[f] != 3 /* STATE.SYNC_CONNECTED, zk-consts.js:124 */)
			slot = ZKB_NOTIFY_IGNORED;
		else if (dec->a0[f] >= 1 && dec->a0[f] <= 4)   /* NOTIFICATION_TYPE, zk-consts.js:111-116 */
			slot = dec->a0[f];
		else
			slot = ZKB_NOTIFY_BAD_TYPE;                 /* pkt.type undefined: :402 throws */
		out->notifications[(size_t)s * ZKB_NOTIFY_SLOTS + slot]++;
	}
	free(dead);
	return 0;
}

/* ------------------------------------------------------------------------------ */
/* JuteBuffer write side + ZKBuffer writers                                       */

typedef struct { uint8_t *p; uint64_t n, cap; } wbuf_t;

static void wb_need(wbuf_t *w, uint64_t extra)
{
	if (w->n + extra > w->cap) {
		uint64_t nc = w->cap ? w->cap : 1024;       /* jute-buffer.js:18,39-44: 1 KiB, doubling */
		while (nc < w->n + extra) nc *= 2;
		w->p = realloc(w->p, nc); w->cap = nc;
	}
}
/* writeInt, jute-buffer.js:141-147 */
static void wb_int(wbuf_t *w, int32_t v)
{
	wb_need(w, 4);
	w->p[w->n++] = (uint8_t)((uint32_t)v >> 24); w->p[w->n++] = (uint8_t)((uint32_t)v >> 16);
	w->p[w->n++] = (uint8_t)((uint32_t)v >> 8);  w->p[w->n++] = (uint8_t)v;
}
/* writeLong with an 8-byte Buffer argument, jute-buffer.js:149-165 */
static void wb_long(wbuf_t *w, int64_t v)
{
	wb_need(w, 8);
	for (int i = 7; i >= 0; i--) w->p[w->n++] = (uint8_t)((uint64_t)v >> (8 * i));
}
/* writeByte / writeBool, jute-buffer.js:111-121 */
static void wb_byte(wbuf_t *w, int v) { wb_need(w, 1); w->p[w->n++] = (uint8_t)v; }
/* writeBuffer / writeUString, jute-buffer.js:123-139: empty encodes as length -1 */
static void wb_buffer(wbuf_t *w, const uint8_t *arena, uint64_t off, int32_t len)
{
	if (len <= 0) { wb_int(w, -1); return; }
	wb_int(w, len);
	wb_need(w, (uint64_t)len);
	memcpy(w->p + w->n, arena + off, (size_t)len);
	w->n += (uint64_t)len;
}

/* writeRequest and friends (zk-buffer.js:97-136,143-273,384-426), writeConnectRequest
 * (:32-39), writeConnectResponse (:50-56), wrapped by writeLengthPrefixed
 * (jute-buffer.js:181-189) as ZKEncodeStream._transform does (zk-streams.js:121-148). */
static void encode_range(const zkb_requests *r, uint64_t lo, uint64_t hi, wbuf_t *wp, uint64_t *frame_off,
    uint8_t *status)
{
	wbuf_t w = *wp;
	for (uint64_t i = lo; i < hi; i++) {
		uint64_t start = w.n;
		int op = r->op[i];
		int bad = 0;
		if (frame_off) frame_off[i] = start;
		wb_int(&w, 0);   /* length back-patched below */
		if (op == ZKB_ENC_CONNECT_REQUEST) {
			wb_int(&w, r->a0[i]); wb_long(&w, r->zxid2[i]); wb_int(&w, r->a1[i]);
			wb_long(&w, r->zxid[i]);
			wb_buffer(&w, r->arena, r->data_off[i], r->data_len[i]);
		} else if (op == ZKB_ENC_CONNECT_RESPONSE) {
			wb_int(&w, r->a0[i]); wb_int(&w, r->a1[i]); wb_long(&w, r->zxid[i]);
			wb_buffer(&w, r->arena, r->data_off[i], r->data_len[i]);
		} else {
			wb_int(&w, r->xid[i]);
			wb_int(&w, op);
			switch (op) {
			case ZKB_OP_GET_CHILDREN: case ZKB_OP_GET_CHILDREN2:
			case ZKB_OP_GET_DATA: case ZKB_OP_EXISTS:
				wb_buffer(&w, r->arena, r->path_off[i], r->path_len[i]);
				wb_byte(&w, r->a0[i] ? 1 : 0);
				break;
			case ZKB_OP_CREATE:
				wb_buffer(&w, r->arena, r->path_off[i], r->path_len[i]);
				wb_buffer(&w, r->arena, r->data_off[i], r->data_len[i]);
				wb_int(&w, r->vec_count[i]);          /* writeACL :384-393 */
				for (int32_t k = 0; k < r->vec_count[i]; k++) {
					uint64_t a = (uint64_t)r->vec_first[i] + (uint64_t)k;
					wb_int(&w, r->acl_perms[a]);
					wb_buffer(&w, r->arena, r->acl_scheme_off[a], r->acl_scheme_len[a]);
					wb_buffer(&w, r->arena, r->acl_id_off[a], r->acl_id_len[a]);
				}
				wb_int(&w, r->a0[i]);
				break;
			case ZKB_OP_DELETE:
				wb_buffer(&w, r->arena, r->path_off[i], r->path_len[i]);
				wb_int(&w, r->a0[i]);
				break;
			case ZKB_OP_GET_ACL: case ZKB_OP_SYNC:
				wb_buffer(&w, r->arena, r->path_off[i], r->path_len[i]);
				break;
			case ZKB_OP_SET_DATA:
				wb_buffer(&w, r->arena, r->path_off[i], r->path_len[i]);
				wb_buffer(&w, r->arena, r->data_off[i], r->data_len[i]);
				wb_int(&w, r->a0[i]);
				break;
			case ZKB_OP_SET_WATCHES: {
				int32_t n[3] = { r->n0[i], r->n1[i], r->vec_count[i] - r->n0[i] - r->n1[i] };
				uint64_t e = r->vec_first[i];
				wb_long(&w, r->zxid[i]);
				for (int v = 0; v < 3; v++) {
					wb_int(&w, n[v]);
					for (int32_t k = 0; k < n[v]; k++, e++)
						wb_buffer(&w, r->arena, r->elem_off[e], r->elem_len[e]);
				}
				break;
			}
			case ZKB_OP_CLOSE_SESSION: case ZKB_OP_PING:
				break;
			default:
				bad = 1;
				break;
			}
		}
		if (bad) {
			w.n = start;                       /* the reference throws: nothing is pushed */
			if (status) status[i] = ZKB_ENC_UNSUPPORTED;
			continue;
		}
		if (status) status[i] = ZKB_ENC_OK;
		uint32_t len = (uint32_t)(w.n - start - 4);
		w.p[start] = (uint8_t)(len >> 24); w.p[start + 1] = (uint8_t)(len >> 16);
		w.p[start + 2] = (uint8_t)(len >> 8); w.p[start + 3] = (uint8_t)len;
	}
	*wp = w;
}

int zko_encode(const zkb_requests *r, uint8_t **out, uint64_t *out_len, uint64_t *frame_off,
    uint8_t *status)
{
	wbuf_t w = { 0, 0, 0 };
	encode_range(r, 0, r->n, &w, frame_off, status);
	frame_off[r->n] = w.n;
	*out = w.p; *out_len = w.n;
	return 0;
}

/* ---- timed CPU baseline of the encoder: packets sharded over threads, each thread grows its own
 * output buffer as ZKEncodeStream's per-packet ZKBuffer would (zk-streams.js:121-148) ---------- */
typedef struct { const zkb_requests *r; uint64_t lo, hi, bytes, chk; } ejob_t;

static void *bench_encode_thread(void *arg)
{
	ejob_t *j = arg;
	/* in groups of packets, so the frames stay cache sized as a socket writer would see them */
	for (uint64_t lo = j->lo; lo < j->hi; lo += 65536) {
		uint64_t hi = lo + 65536 < j->hi ? lo + 65536 : j->hi;
		wbuf_t w = { 0, 0, 0 };
		encode_range(j->r, lo, hi, &w, NULL, NULL);
		j->bytes += w.n;
		for (uint64_t k = 0; k < w.n; k += 64) j->chk += w.p[k];
		free(w.p);
	}
	return NULL;
}

double zko_bench_encode(const zkb_requests *r, int nthreads, uint64_t *bytes, uint64_t *checksum)
{
	if (nthreads < 1) nthreads = 1;
	ejob_t *jobs = calloc(nthreads, sizeof (ejob_t));
	pthread_t *th = calloc(nthreads, sizeof (pthread_t));
	for (int i = 0; i < nthreads; i++) {
		jobs[i].r = r;
		jobs[i].lo = r->n * (uint64_t)i / nthreads;
		jobs[i].hi = r->n * (uint64_t)(i + 1) / nthreads;
	}
	struct timespec t0, t1;
	clock_gettime(CLOCK_MONOTONIC, &t0);
	for (int i = 0; i < nthreads; i++) pthread_create(&th[i], NULL, bench_encode_thread, &jobs[i]);
	for (int i = 0; i < nthreads; i++) pthread_join(th[i], NULL);
	clock_gettime(CLOCK_MONOTONIC, &t1);
	uint64_t b = 0, c = 0;
	for (int i = 0; i < nthreads; i++) { b += jobs[i].bytes; c += jobs[i].chk; }
	if (bytes) *bytes = b;
	if (checksum) *checksum = c;
	free(jobs); free(th);
	return (t1.tv_sec - t0.tv_sec) + (t1.tv_nsec - t0.tv_nsec) * 1e-9;
}

void zko_buf_free(uint8_t *p) { free(p); }
