/*
 * zkgen.c — seeded synthetic ZooKeeper traffic at benchmark scale (host utility, plain C).
 *
 * Builds libzkgen.so.  It only WRITES wire bytes (it contains no decoder), so it is neither
 * part of the codec nor of the oracle: bench.py and the large-size tests use it to make the
 * multi-gigabyte corpora of BASELINE.json's configs, which the Python generator
 * (synth.py, same frame shapes) would take hours to produce.
 *
 * Config 2: per session a paired request stream and response stream;
 * frame 0 of each direction is the connect handshake; ops CREATE / GET_CHILDREN2 /
 * SET_DATA / EXISTS / PING 20 % each; path 16-64 ASCII bytes; data 0-1024 bytes; children
 * 0-32 x 8-24 bytes; 5 % error replies (NO_NODE / NODE_EXISTS / BAD_VERSION); xids dense
 * from 0 per session; PING uses xid -2.
 * Config 1: GET_DATA replies with a fixed payload + Stat.
 * Config 3: notifications / ping replies after a connect response.
 */
#define _GNU_SOURCE
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct { uint64_t s; } rng_t;

static inline uint64_t rnd(rng_t *r)
{
	uint64_t z = (r->s += 0x9E3779B97F4A7C15ull);
	z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
	z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
	return z ^ (z >> 31);
}
static inline uint32_t rnd_below(rng_t *r, uint32_t n) { return (uint32_t)((rnd(r) >> 32) * (uint64_t)n >> 32); }

/* a writer that only counts when buf == NULL */
typedef struct { uint8_t *buf; uint64_t n; } w_t;

static inline void w_u32(w_t *w, uint32_t v)
{
	if (w->buf) { uint8_t *p = w->buf + w->n; p[0] = v >> 24; p[1] = v >> 16; p[2] = v >> 8; p[3] = v; }
	w->n += 4;
}
static inline void w_u64(w_t *w, uint64_t v) { w_u32(w, (uint32_t)(v >> 32)); w_u32(w, (uint32_t)v); }
static inline void w_u8(w_t *w, uint8_t v) { if (w->buf) w->buf[w->n] = v; w->n += 1; }
static void w_random(w_t *w, rng_t *r, uint32_t len)
{
	if (w->buf) {
		uint8_t *p = w->buf + w->n;
		uint32_t i = 0;
		for (; i + 8 <= len; i += 8) { uint64_t v = rnd(r); memcpy(p + i, &v, 8); }
		if (i < len) { uint64_t v = rnd(r); memcpy(p + i, &v, len - i); }
	} else {
		/* keep the generator state identical between the sizing and the filling pass */
		for (uint32_t i = 0; i < len; i += 8) (void)rnd(r);
	}
	w->n += len;
}
static void w_ascii(w_t *w, rng_t *r, uint32_t len, int leading_slash)
{
	static const char A[] = "abcdefghijklmnopqrstuvwxyz0123456789-_";
	for (uint32_t i = 0; i < len; i += 8) {
		uint64_t v = rnd(r);
		for (uint32_t k = 0; k < 8 && i + k < len; k++) {
			uint8_t c = (uint8_t)A[(v >> (8 * k)) % 38];
			if (leading_slash && i + k == 0) c = '/';
			if (w->buf) w->buf[w->n + i + k] = c;
		}
	}
	w->n += len;
}
static void w_stat(w_t *w, rng_t *r, uint32_t data_len, uint32_t nchildren)
{
	w_u64(w, rnd(r) >> 24); w_u64(w, rnd(r) >> 24);
	w_u64(w, 1400000000000ull + (rnd(r) >> 28)); w_u64(w, 1400000000000ull + (rnd(r) >> 28));
	w_u32(w, rnd_below(r, 1000)); w_u32(w, rnd_below(r, 1000)); w_u32(w, rnd_below(r, 10));
	w_u64(w, (rnd(r) & 1) ? rnd(r) >> 8 : 0);
	w_u32(w, data_len); w_u32(w, nchildren);
	w_u64(w, rnd(r) >> 24);
}
/* reserve the length prefix; returns where it is so it can be patched */
static inline uint64_t frame_begin(w_t *w) { uint64_t at = w->n; w_u32(w, 0); return at; }
static inline void frame_end(w_t *w, uint64_t at)
{
	uint32_t len = (uint32_t)(w->n - at - 4);
	if (w->buf) { uint8_t *p = w->buf + at; p[0] = len >> 24; p[1] = len >> 16; p[2] = len >> 8; p[3] = len; }
}
static void connect_request(w_t *w, rng_t *r)
{
	uint64_t at = frame_begin(w);
	w_u32(w, 0); w_u64(w, rnd(r) >> 24); w_u32(w, 30000); w_u64(w, rnd(r) >> 2);
	w_u32(w, 16); for (int i = 0; i < 16; i++) w_u8(w, 0);
	w_u8(w, 0);
	frame_end(w, at);
}
static void connect_response(w_t *w, rng_t *r)
{
	uint64_t at = frame_begin(w);
	w_u32(w, 0); w_u32(w, 30000); w_u64(w, rnd(r) >> 2);
	w_u32(w, 16); w_random(w, r, 16);
	w_u8(w, 0);
	frame_end(w, at);
}

enum { OP_CREATE = 1, OP_EXISTS = 3, OP_GET_DATA = 4, OP_SET_DATA = 5, OP_PING = 11, OP_GET_CHILDREN2 = 12 };

/* one connection, both directions; xid_ops (may be NULL) gets the opcode per xid */
static void cfg2_session(uint64_t seed, uint32_t s, uint32_t frames, w_t *rq, w_t *rs, int8_t *xid_ops)
{
	rng_t r = { seed ^ (0xD1B54A32D192ED03ull * (uint64_t)(s + 1)) };
	static const int OPS[5] = { OP_CREATE, OP_GET_CHILDREN2, OP_SET_DATA, OP_EXISTS, OP_PING };
	static const int32_t ERRS[3] = { -101, -110, -103 };
	connect_request(rq, &r);
	connect_response(rs, &r);
	int32_t xid = 0;
	uint64_t zxid = 0x100000000ull;
	for (uint32_t i = 0; i < frames; i++) {
		int op = OPS[rnd_below(&r, 5)];
		uint64_t a, b;
		if (op == OP_PING) {
			a = frame_begin(rq); w_u32(rq, (uint32_t)-2); w_u32(rq, OP_PING); frame_end(rq, a);
			b = frame_begin(rs); w_u32(rs, (uint32_t)-2); w_u64(rs, zxid); w_u32(rs, 0); frame_end(rs, b);
			continue;
		}
		uint32_t plen = 16 + rnd_below(&r, 49);
		a = frame_begin(rq);
		w_u32(rq, (uint32_t)xid); w_u32(rq, (uint32_t)op);
		w_u32(rq, plen); w_ascii(rq, &r, plen, 1);
		if (op == OP_CREATE || op == OP_SET_DATA) {
			uint32_t dlen = rnd_below(&r, 1025);
			w_u32(rq, dlen); w_random(rq, &r, dlen);
			if (op == OP_CREATE) {
				w_u32(rq, 1); w_u32(rq, 31);
				w_u32(rq, 5); if (rq->buf) memcpy(rq->buf + rq->n, "world", 5); rq->n += 5;
				w_u32(rq, 6); if (rq->buf) memcpy(rq->buf + rq->n, "anyone", 6); rq->n += 6;
				w_u32(rq, rnd_below(&r, 4));
			} else {
				w_u32(rq, (uint32_t)((int32_t)rnd_below(&r, 51) - 1));
			}
		} else {
			w_u8(rq, (uint8_t)(rnd(&r) & 1));
		}
		frame_end(rq, a);
		zxid++;
		b = frame_begin(rs);
		w_u32(rs, (uint32_t)xid); w_u64(rs, zxid);
		if (rnd_below(&r, 100) < 5) {
			w_u32(rs, (uint32_t)ERRS[rnd_below(&r, 3)]);
		} else {
			w_u32(rs, 0);
			if (op == OP_CREATE) {
				w_u32(rs, plen); w_ascii(rs, &r, plen, 1);
			} else if (op == OP_GET_CHILDREN2) {
				uint32_t n = rnd_below(&r, 33);
				w_u32(rs, n);
				for (uint32_t k = 0; k < n; k++) {
					uint32_t cl = 8 + rnd_below(&r, 17);
					w_u32(rs, cl); w_ascii(rs, &r, cl, 0);
				}
				w_stat(rs, &r, 0, n);
			} else {
				w_stat(rs, &r, rnd_below(&r, 1025), 0);
			}
		}
		frame_end(rs, b);
		if (xid_ops) xid_ops[xid] = (int8_t)op;
		xid++;
	}
}

typedef struct {
	uint64_t seed; uint32_t s0, s1, frames;
	uint64_t *rq_off, *rs_off;   /* [nsessions+1]: in the size pass entry s+1 receives the LENGTH */
	uint8_t *rq, *rs; int8_t *ops;
	int fill;
} job_t;

static void *cfg2_thread(void *arg)
{
	job_t *j = arg;
	for (uint32_t s = j->s0; s < j->s1; s++) {
		w_t rq = { NULL, 0 }, rs = { NULL, 0 };
		if (j->fill) { rq.buf = j->rq + j->rq_off[s]; rs.buf = j->rs + j->rs_off[s]; }
		cfg2_session(j->seed, s, j->frames, &rq, &rs,
		    (j->fill && j->ops) ? j->ops + (uint64_t)s * j->frames : NULL);
		if (!j->fill) { j->rq_off[s + 1] = rq.n; j->rs_off[s + 1] = rs.n; }
	}
	return NULL;
}

static void run_jobs(job_t proto, uint32_t nsessions, int nthreads)
{
	if (nthreads < 1) nthreads = 1;
	if ((uint32_t)nthreads > nsessions) nthreads = nsessions ? (int)nsessions : 1;
	pthread_t *th = calloc(nthreads, sizeof (*th));
	job_t *jobs = calloc(nthreads, sizeof (*jobs));
	for (int i = 0; i < nthreads; i++) {
		jobs[i] = proto;
		jobs[i].s0 = (uint32_t)((uint64_t)nsessions * i / nthreads);
		jobs[i].s1 = (uint32_t)((uint64_t)nsessions * (i + 1) / nthreads);
		pthread_create(&th[i], NULL, cfg2_thread, &jobs[i]);
	}
	for (int i = 0; i < nthreads; i++) pthread_join(th[i], NULL);
	free(th); free(jobs);
}

/* Pass 1: fills rq_off / rs_off (nsessions+1 entries each) with stream offsets. */
void zkgen_cfg2_sizes(uint64_t seed, uint32_t nsessions, uint32_t frames, uint64_t *rq_off,
    uint64_t *rs_off, int nthreads)
{
	job_t p = { seed, 0, 0, frames, rq_off, rs_off, NULL, NULL, NULL, 0 };
	rq_off[0] = rs_off[0] = 0;
	run_jobs(p, nsessions, nthreads);
	for (uint32_t s = 0; s < nsessions; s++) { rq_off[s + 1] += rq_off[s]; rs_off[s + 1] += rs_off[s]; }
}

/* Pass 2: writes the bytes; ops (may be NULL) is [nsessions][frames] int8, opcode per xid
 * (entries past a session's last xid are left untouched: pre-fill with -128). */
void zkgen_cfg2_fill(uint64_t seed, uint32_t nsessions, uint32_t frames, const uint64_t *rq_off,
    const uint64_t *rs_off, uint8_t *rq, uint8_t *rs, int8_t *ops, int nthreads)
{
	job_t p = { seed, 0, 0, frames, (uint64_t *)rq_off, (uint64_t *)rs_off, rq, rs, ops, 1 };
	run_jobs(p, nsessions, nthreads);
}

/* Config 1: one stream, ConnectResponse then n GET_DATA replies with `payload` data bytes.
 * Returns the byte count; writes when out != NULL. */
uint64_t zkgen_cfg1(uint64_t seed, uint64_t n, uint32_t payload, uint8_t *out)
{
	rng_t r = { seed };
	w_t w = { out, 0 };
	connect_response(&w, &r);
	for (uint64_t i = 0; i < n; i++) {
		uint64_t at = frame_begin(&w);
		w_u32(&w, (uint32_t)i); w_u64(&w, 0x100000000ull + i); w_u32(&w, 0);
		w_u32(&w, payload); w_random(&w, &r, payload);
		w_stat(&w, &r, payload, 0);
		frame_end(&w, at);
	}
	return w.n;
}

/* Config 3: one stream, ConnectResponse then n tiny frames: 2/3 notifications
 * (xid -1, zxid -1, type 1-4, state 3, path 4-12 bytes), 1/3 ping replies (20 bytes). */
uint64_t zkgen_cfg3(uint64_t seed, uint64_t n, uint8_t *out)
{
	rng_t r = { seed };
	w_t w = { out, 0 };
	connect_response(&w, &r);
	for (uint64_t i = 0; i < n; i++) {
		uint64_t at = frame_begin(&w);
		if (rnd_below(&r, 3) < 2) {
			uint32_t pl = 4 + rnd_below(&r, 9);
			w_u32(&w, (uint32_t)-1); w_u64(&w, (uint64_t)-1); w_u32(&w, 0);
			w_u32(&w, 1 + rnd_below(&r, 4)); w_u32(&w, 3);
			w_u32(&w, pl); w_ascii(&w, &r, pl, 1);
		} else {
			w_u32(&w, (uint32_t)-2); w_u64(&w, rnd(&r) >> 24); w_u32(&w, 0);
		}
		frame_end(&w, at);
	}
	return w.n;
}
