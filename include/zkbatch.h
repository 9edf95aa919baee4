/*
 * zkbatch.h — C ABI of libzkbatch.so, the B200-native batch codec for ZooKeeper
 * jute frames (the data-parallel hot path of node-zkstream).
 *
 * Everything here is plain C: pointers, sizes, fixed-width integers.  No CUDA,
 * torch or C++ types appear in any signature, so the library can be bound from
 * an N-API addon, ctypes, cgo, JNI ...
 *
 * Reference interfaces replaced (paths relative to the node-zkstream tree):
 *   lib/zk-streams.js:39-100   ZKDecodeStream._transform + handlePacket  -> zkb_decode_*
 *   lib/zk-buffer.js:281-331   ZKBuffer.readResponse                     -> zkb_decode_* (kind RESPONSE)
 *   lib/zk-buffer.js:58-95     ZKBuffer.readRequest                      -> zkb_decode_* (kind REQUEST)
 *   lib/zk-buffer.js:22-30,41-48 readConnectRequest/readConnectResponse  -> zkb_decode_* (kind CONNECT_*)
 *   lib/zk-streams.js:121-148  ZKEncodeStream._transform                 -> zkb_encode_*
 *   lib/zk-buffer.js:97-136    ZKBuffer.writeRequest (+ sub-writers)     -> zkb_encode_*
 *   lib/zk-buffer.js:32-39     ZKBuffer.writeConnectRequest              -> zkb_encode_* (ZKB_ENC_CONNECT_REQUEST)
 *   lib/zk-buffer.js:50-56     ZKBuffer.writeConnectResponse             -> zkb_encode_* (ZKB_ENC_CONNECT_RESPONSE)
 *   lib/zk-streams.js:145      fsm.zcf_xidMap[xid] = opcode              -> zkb_xidtab_build_*
 *   lib/jute-buffer.js:46-189  JuteBuffer primitives                     -> device functions inside the kernels
 *
 * There is NO CPU implementation behind these entry points: if no CUDA device
 * is usable, zkb_ctx_create fails with ZKB_ENODEV and nothing else can be called.
 */
#ifndef ZKBATCH_H
#define ZKBATCH_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ZKB_VERSION 1

/* ---- return codes (0 = success, negative = failure) ------------------------ */
#define ZKB_OK          0
#define ZKB_EINVAL    (-22)  /* bad argument (NULL pointer, offsets not monotone, ...) */
#define ZKB_ENOMEM    (-12)  /* host or device allocation failed */
#define ZKB_ENODEV    (-19)  /* no usable CUDA device */
#define ZKB_ECUDA      (-5)  /* a CUDA call or kernel failed; see zkb_last_error() */
#define ZKB_EOVERFLOW (-75)  /* caller-provided capacity too small; required row counts are in the result */
#define ZKB_ERANGE    (-34)  /* input larger than the documented limits */
#define ZKB_EINTERNAL (-71)  /* internal consistency check failed (bug) */

/* ---- wire constants (lib/zk-streams.js:23, lib/zk-consts.js) ----------------- */
#define ZKB_MAX_PACKET (16 * 1024 * 1024)

/* opcodes, lib/zk-consts.js:84-105; values are the wire integers */
#define ZKB_OP_NOTIFICATION     0
#define ZKB_OP_CREATE           1
#define ZKB_OP_DELETE           2
#define ZKB_OP_EXISTS           3
#define ZKB_OP_GET_DATA         4
#define ZKB_OP_SET_DATA         5
#define ZKB_OP_GET_ACL          6
#define ZKB_OP_SET_ACL          7
#define ZKB_OP_GET_CHILDREN     8
#define ZKB_OP_SYNC             9
#define ZKB_OP_PING            11
#define ZKB_OP_GET_CHILDREN2   12
#define ZKB_OP_CHECK           13
#define ZKB_OP_MULTI           14
#define ZKB_OP_AUTH           100
#define ZKB_OP_SET_WATCHES    101
#define ZKB_OP_SASL           102
#define ZKB_OP_CREATE_SESSION (-10)
#define ZKB_OP_CLOSE_SESSION  (-11)
#define ZKB_OP_ERROR           (-1)
#define ZKB_OP_NONE          (-128) /* "undefined": xid not in the map / unknown wire opcode */

/* special xids, lib/zk-consts.js:135-138 and lib/zk-buffer.js:275-279 */
#define ZKB_XID_NOTIFICATION (-1)
#define ZKB_XID_PING         (-2)
#define ZKB_XID_AUTH         (-4)
#define ZKB_XID_SET_WATCHES  (-8)

/* ---- per-frame meta word ------------------------------------------------------
 * meta = status | (uint8_t)opcode << 8 | kind << 16 | detail << 24
 */
#define ZKB_META_STATUS(m) ((uint8_t)((m) & 0xff))
#define ZKB_META_OPCODE(m) ((int8_t)(((m) >> 8) & 0xff))
#define ZKB_META_KIND(m)   ((uint8_t)(((m) >> 16) & 0xff))
#define ZKB_META_DETAIL(m) ((uint8_t)(((m) >> 24) & 0xff))

/* status: what ZKDecodeStream did with the frame (lib/zk-streams.js:66-99) */
#define ZKB_ST_OK          0 /* self.push(pkt) */
#define ZKB_ST_BAD_DECODE  1 /* emit('error', ZKProtocolError('BAD_DECODE', ...)); later frames still decoded */

/* kind: which reader ran, chosen from isServer x handshaking (lib/zk-streams.js:68-97) */
#define ZKB_KIND_RESPONSE          0 /* readResponse(xidMap) */
#define ZKB_KIND_REQUEST           1 /* readRequest() */
#define ZKB_KIND_CONNECT_RESPONSE  2 /* readConnectResponse() */
#define ZKB_KIND_CONNECT_REQUEST   3 /* readConnectRequest() */

/* detail (only when status != OK) = cause | field << 3.
 * cause says which exception the reference's reader threw, so that a binding can
 * rebuild the 'Failed to decode ...: ' + e.name + ': ' + e.message text. */
#define ZKB_DETAIL_CAUSE(d) ((d) & 7)
#define ZKB_DETAIL_FIELD(d) ((d) >> 3)
#define ZKB_CAUSE_RANGE        1 /* Buffer RangeError out of readInt32BE (jute-buffer.js:59) */
#define ZKB_CAUSE_ASSERT       2 /* assert.ok(...) failed (jute-buffer.js:52,54,65,101) */
#define ZKB_CAUSE_NO_MATCH     3 /* 'reply packet must match a request' (zk-buffer.js:291) */
#define ZKB_CAUSE_UNSUPPORTED  4 /* 'Unsupported opcode X' (zk-buffer.js:92,327) */

/* field ids: which read failed */
enum {
	ZKB_F_NONE = 0,
	ZKB_F_XID, ZKB_F_ZXID, ZKB_F_ERR, ZKB_F_OPCODE,
	ZKB_F_PATH, ZKB_F_DATA, ZKB_F_WATCH, ZKB_F_VERSION, ZKB_F_FLAGS,
	ZKB_F_COUNT, ZKB_F_ELEM, ZKB_F_ACL_PERMS, ZKB_F_ACL_SCHEME, ZKB_F_ACL_ID,
	ZKB_F_STAT, ZKB_F_NTYPE, ZKB_F_NSTATE,
	ZKB_F_PROTOCOL_VERSION, ZKB_F_LAST_ZXID, ZKB_F_TIMEOUT, ZKB_F_SESSION_ID, ZKB_F_PASSWD,
	ZKB_F_REL_ZXID
};

/* session flags */
#define ZKB_SESS_SERVER 1u /* isServer:true — the stream holds client->server bytes */

#define ZKB_NO_OFFSET UINT64_MAX

/* ---- inputs -------------------------------------------------------------------- */

/*
 * A batch is the concatenation of `nsessions` independent byte streams (one
 * direction of one connection each; what a ZKDecodeStream instance would be fed,
 * in any chunking).  off[0] must be 0 and off[nsessions] the total byte count.
 */
typedef struct zkb_sessions {
	uint32_t        nsessions;
	const uint64_t *off;              /* [nsessions + 1], monotone */
	const uint8_t  *flags;            /* [nsessions], ZKB_SESS_*; NULL = all 0 (client side) */
	const uint32_t *handshake_frames; /* [nsessions] leading frames read in 'handshaking' state
	                                     (zk-streams.js:68); NULL = 1 each, as with the real FSM */
} zkb_sessions;

/*
 * The batched form of fsm.zcf_xidMap (connection-fsm.js:39, zk-streams.js:145):
 * for session s, xid x maps to ops[off[s] + (x - base[s])] when
 * 0 <= x - base[s] < count[s], else to ZKB_OP_NONE.  The four special xids
 * win over the table exactly as SPECIAL_XIDS does (zk-buffer.js:288-290).
 * NULL table = every non-special xid is unknown.
 */
typedef struct zkb_xidtab {
	const int32_t  *base;   /* [nsessions] */
	const uint32_t *count;  /* [nsessions] */
	const uint64_t *off;    /* [nsessions] */
	const int8_t   *ops;    /* wire opcode or ZKB_OP_NONE */
	uint64_t        nops;   /* total entries in ops */
} zkb_xidtab;

/* ---- decoded output (struct of arrays) ------------------------------------------
 *
 * FRAMES table: one row per complete frame the reference would have handed to
 * handlePacket, in (session, stream position) order.
 *
 *  frame_off  offset of the 4-byte length prefix in the batch buffer
 *  frame_len  L (payload is bytes [frame_off+4, frame_off+4+L))
 *  session    index into zkb_sessions
 *  meta       see ZKB_META_*
 *  xid, zxid, err, a0, a1 — by kind:
 *    RESPONSE          xid, zxid, err = header (err is the raw int32; the binding maps it through
 *                      ERR_LOOKUP).  NOTIFICATION: a0 = type, a1 = state (raw).  GET_CHILDREN,
 *                      GET_CHILDREN2: a0 = number of children.  GET_ACL: a0 = number of ACL entries.
 *    REQUEST           xid.  a0 = watch (0/1) | version | flags (raw).  CREATE: a1 = number of ACL
 *                      entries.  SET_WATCHES: zxid = relZxid, a0/a1/err = number of dataChanged /
 *                      createdOrDestroyed / childrenChanged paths.
 *    CONNECT_RESPONSE  xid = protocolVersion, err = timeOut, zxid = sessionId.
 *    CONNECT_REQUEST   as CONNECT_RESPONSE, plus lastZxidSeen = (int64)a0 << 32 | (uint32)a1.
 *  Every 64-bit column is the field's 8 wire bytes read as a big-endian two's complement
 *  integer (jute-buffer.js:64-69 returns those 8 bytes); it converts back losslessly.
 *  Columns not listed for a kind are 0.  For a BAD_DECODE row, columns hold whatever had
 *  been read before the failing field (0 elsewhere) and the row owns no side-table rows.
 *
 * STAT table: one row per OK frame that carries a Stat (zk-buffer.js:428-442), frame order.
 * BUF table: one row per buffer/ustring field of OK frames, in read order within the frame:
 *  path, data, passwd, every child, every watch path, and for each ACL entry scheme then id.
 *  buf_off is absolute in the batch buffer (payload bytes are never copied), buf_len >= 0
 *  (negative wire lengths read as empty, jute-buffer.js:98-100).
 * ACL table: one row per ACL entry of OK frames: raw perms (the binding applies readPerms,
 *  including its operator-precedence quirk, zk-buffer.js:399).
 * Each side table has a `*_frame` column: the FRAMES row that owns the side row.
 *
 * Per session (all offsets are absolute in the batch buffer): consumed = end of the last
 * complete frame (the caller carries bytes [consumed, off[s+1]) into the next batch, as
 * zd_buffer does); bad_length_at = offset of the length prefix that raised BAD_LENGTH
 * (nothing after it in that session is decoded, zk-streams.js:49-53) or ZKB_NO_OFFSET;
 * first_frame[s] = FRAMES row of the session's first frame, first_frame[nsessions] = n_frames.
 */
typedef struct zkb_decoded {
	/* in (zkb_decode_dev only): rows each table can hold */
	uint64_t cap_frames, cap_stat, cap_buf, cap_acl;
	/* out: rows produced (with ZKB_EOVERFLOW: rows required) */
	uint64_t n_frames, n_stat, n_buf, n_acl;

	uint64_t *frame_off;
	int32_t  *frame_len;
	uint32_t *session;
	uint32_t *meta;
	int32_t  *xid;
	int64_t  *zxid;
	int32_t  *err;
	int32_t  *a0;
	int32_t  *a1;

	uint32_t *stat_frame;
	int64_t  *czxid, *mzxid, *ctime, *mtime;
	int32_t  *version, *cversion, *aversion;
	int64_t  *ephemeralOwner;
	int32_t  *dataLength, *numChildren;
	int64_t  *pzxid;

	uint32_t *buf_frame;
	uint64_t *buf_off;
	int32_t  *buf_len;

	uint32_t *acl_frame;
	int32_t  *acl_perms;

	uint64_t *sess_consumed;      /* [nsessions] batch-absolute end of the last complete frame */
	uint64_t *sess_bad_length_at; /* [nsessions] */
	uint32_t *sess_first_frame;   /* [nsessions + 1] */

	/* compact replay results only (zkb_ctx_set_replay_compact), NULL otherwise: per BUF row the
	 * offset of the buffer's first byte from the length prefix of the frame that owns it */
	uint32_t *buf_rel;
} zkb_decoded;

/* ---- encode input (struct of arrays) ----------------------------------------------
 *
 * One row per packet handed to ZKEncodeStream.write().  `op` is the wire opcode
 * (writeRequest, zk-buffer.js:97-136) or one of the two connect pseudo-ops.
 * Strings and buffers live in `arena`; a field is the bytes [off, off+len).
 *   path      ustring for every op with a path; data: CREATE/SET_DATA data, connect passwd
 *   a0        watch (0/1) | version | flags | connect protocolVersion
 *   a1        connect timeOut
 *   zxid      SET_WATCHES relZxid | connect sessionId      (8 wire bytes as big-endian int64)
 *   zxid2     ConnectRequest lastZxidSeen
 *   vec_first/vec_count  CREATE: rows of the acl_* arrays; SET_WATCHES: rows of the elem_* arrays
 *             holding dataChanged, createdOrDestroyed, childrenChanged back to back, with
 *             n0 / n1 / vec_count - n0 - n1 entries.
 * A zero-length buffer or string is written as length -1 (jute-buffer.js:127-130).
 */
#define ZKB_ENC_CONNECT_REQUEST   ((int8_t)-100)
#define ZKB_ENC_CONNECT_RESPONSE  ((int8_t)-101)

typedef struct zkb_requests {
	uint64_t        n;
	const int32_t  *xid;
	const int8_t   *op;
	const uint8_t  *arena;
	uint64_t        arena_len;
	const uint64_t *path_off;  const int32_t *path_len;
	const uint64_t *data_off;  const int32_t *data_len;
	const int32_t  *a0;
	const int32_t  *a1;
	const int64_t  *zxid;
	const int64_t  *zxid2;
	const uint32_t *vec_first; const int32_t *vec_count;
	const int32_t  *n0;        const int32_t *n1;
	/* element tables */
	uint64_t        n_elem;
	const uint64_t *elem_off;  const int32_t *elem_len;
	uint64_t        n_acl;
	const int32_t  *acl_perms;
	const uint64_t *acl_scheme_off; const int32_t *acl_scheme_len;
	const uint64_t *acl_id_off;     const int32_t *acl_id_len;
} zkb_requests;

/* Encode status per packet: 0 = frame written; else no bytes were produced for it
 * (frame_off[i+1] == frame_off[i]) and the reference would have thrown. */
#define ZKB_ENC_OK           0
#define ZKB_ENC_UNSUPPORTED  1 /* 'Unsupported opcode' (zk-buffer.js:133-134) */

/* ---- context ---------------------------------------------------------------------- */
typedef struct zkb_ctx zkb_ctx;

/* Create a codec context bound to CUDA device `device_id`.  Returns ZKB_ENODEV when CUDA
 * or the device is missing — there is no CPU path to fall back to. */
int  zkb_ctx_create(int device_id, zkb_ctx **out);
void zkb_ctx_destroy(zkb_ctx *ctx);
/* Text of the last failure seen by this context (never NULL). */
const char *zkb_last_error(const zkb_ctx *ctx);
/* Number of CUDA kernels this context has launched so far (for bench accounting). */
uint64_t zkb_kernel_launches(const zkb_ctx *ctx);
/* Number of splitter fix-up rounds (entry-point guesses that had to be corrected) this context
 * has needed so far; one per batch is normal, more on payloads that look like frames.  Diagnostic only. */
uint64_t zkb_split_fixups(const zkb_ctx *ctx);
/* Decode passes this context has had to run twice.  The splitter predicts from the frame
 * headers how many side-table rows every chunk will produce, so that the decode kernel can
 * place rows without waiting for the chunks before it; a frame that does not keep what its
 * header promises (malformed body, handshake deeper than the splitter can see) is noticed by
 * the kernel, and the batch is decoded again from that frame's chunk onwards by the exact
 * route (count the rows of every chunk, one scan, place) — same result, at most one more pass over
 * the rest of the batch, and no chunk ever waits for another.  0 on well-formed traffic.
 * ZKBATCH_NO_PREDICT=1 in the environment (read at zkb_ctx_create) turns the prediction off: every
 * batch then takes the exact route.  Diagnostic only. */
uint64_t zkb_row_mispredicts(const zkb_ctx *ctx);
/* Frames of the last decode call that were not well-formed (or reached past the staged
 * tile) and therefore went through the exact, slower reader.  Diagnostic only. */
uint64_t zkb_last_slow_frames(const zkb_ctx *ctx);
/* Per-phase device timing of decode calls.  With timing on, each zkb_decode_* records CUDA
 * events on the context's stream around its three phases; zkb_last_phase_ms returns the
 * milliseconds the last call spent in { frame split (first pass), frame-row emit, field
 * decode kernel }.  Off by default. */
int zkb_ctx_set_timing(zkb_ctx *ctx, int on);
int zkb_last_phase_ms(const zkb_ctx *ctx, float out[3]);
/* The cudaStream_t the context launches on, as an opaque pointer (for event timing). */
void *zkb_ctx_stream(zkb_ctx *ctx);
int  zkb_ctx_sync(zkb_ctx *ctx);
/* Bytes of device scratch a decode of `nbytes` input bytes over `nsessions` sessions needs. */
uint64_t zkb_decode_scratch_bytes(uint64_t nbytes, uint32_t nsessions);

/* ---- decode ------------------------------------------------------------------------- */

/*
 * Device-resident decode.  Every pointer in `sess`, `xid`, `out` and `bytes` is a DEVICE
 * pointer (bytes must be 16-byte aligned); `out` itself is a host struct whose cap_* fields
 * say how many rows the caller allocated.  Asynchronous work is queued on the context's
 * stream and the call returns after it has completed (the row counts are read back).
 * Replaces ZKDecodeStream._transform over the whole batch (zk-streams.js:39-100).
 */
int zkb_decode_dev(zkb_ctx *ctx, const uint8_t *bytes, uint64_t nbytes,
    const zkb_sessions *sess, const zkb_xidtab *xid, zkb_decoded *out);

/*
 * Host-buffer decode: same contract, but every input pointer is HOST memory.  The context
 * copies the batch to the device, decodes, and copies the columns back into buffers it
 * owns; on return `out`'s pointers address that (pinned) host memory, valid until the next
 * zkb_decode_host / zkb_ctx_destroy on this context.  cap_* are ignored.
 */
int zkb_decode_host(zkb_ctx *ctx, const uint8_t *bytes, uint64_t nbytes,
    const zkb_sessions *sess, const zkb_xidtab *xid, zkb_decoded *out);

/* ---- paired replay, pipelined over PCIe ------------------------------------------------ */

/*
 * What a capture replay does with one ZKEncodeStream/ZKDecodeStream pair per connection
 * (lib/zk-streams.js:145 feeding lib/zk-buffer.js:288-290), for a whole capture in host
 * memory: decode the client->server streams, derive each connection's xid -> opcode map from
 * them, decode the server->client streams with that map.  Session i of `req` pairs with
 * session i of `rsp`; xids [xid_base[i], xid_base[i] + xid_count[i]) are tracked per session.
 *
 * The sessions are cut into slices of about `slice_bytes` (0 = default, 512 MiB) and the
 * slices flow through three CUDA streams — upload, decode, download — so the PCIe transfers of
 * neighbouring slices overlap the kernels.  Results are handed to `cb` one slice at a time, in
 * session order, as context-owned pinned host columns that stay valid until `cb` returns:
 *   - `session` columns and all byte offsets (frame_off, buf_off, sess_consumed,
 *     sess_bad_length_at) are global: they index the caller's session arrays and byte buffers;
 *   - row references (stat_frame, buf_frame, acl_frame, sess_first_frame) are local to the
 *     slice's own tables; sess_* arrays hold the slice's `nsessions` sessions.
 * A non-zero return from `cb` stops the replay and is returned.  Pin the input buffers
 * (cudaHostRegister / cudaHostAlloc) for full PCIe speed; pageable memory works, slower.
 */
typedef int (*zkb_slice_cb)(void *user, uint32_t first_session, uint32_t nsessions,
    const zkb_decoded *req, const zkb_decoded *rsp);

/*
 * Compact replay results.  With several GPUs on one host the download of the result columns,
 * not the upload of the capture, is what the host's PCIe complex runs out of first
 * (profiles/pcie_peak_n*.json), and more than a third of those bytes can be worked out from the
 * rest.  With compact mode on, zkb_replay_host leaves these columns out (their pointers are NULL
 * in the structs handed to the callback):
 *   frame_off    frames of a session lie back to back from the session's first byte:
 *                frame_off[f] = off[s] + sum over the session's earlier frames of (4 + frame_len)
 *   session      from sess_first_frame
 *   stat_frame, buf_frame, acl_frame   the rows of a frame follow those of the frame before it, and how
 *                many rows a frame owns follows from meta / err / a0 / a1 (see the column list above)
 *   buf_off      replaced by buf_rel (32 bits): buf_off = frame_off[owner] + buf_rel
 * Everything else is delivered as usual.  node-zkstream_b200/columns.py (Columns.restore_derived)
 * is the reference implementation of putting the columns back.  Off by default.
 */
int zkb_ctx_set_replay_compact(zkb_ctx *ctx, int on);

int zkb_replay_host(zkb_ctx *ctx,
    const uint8_t *req_bytes, uint64_t req_nbytes, const zkb_sessions *req_sess,
    const uint8_t *rsp_bytes, uint64_t rsp_nbytes, const zkb_sessions *rsp_sess,
    const int32_t *xid_base, const uint32_t *xid_count, uint64_t slice_bytes,
    zkb_slice_cb cb, void *user);

/* ---- xid table ------------------------------------------------------------------------ */

/*
 * Build the xid -> opcode table the paired response stream needs from a decoded REQUEST
 * stream (what ZKEncodeStream records at zk-streams.js:145: zcf_xidMap[pkt.xid] = pkt.opcode,
 * last write wins).  `req` is the result of decoding the client->server direction with
 * ZKB_SESS_SERVER; request session i pairs with response session i.  Only xids in
 * [base[s], base[s]+count[s]) are recorded.  Host pointers in, host pointers out
 * (tab->ops is written; base/count/off are read).
 */
int zkb_xidtab_build_host(zkb_ctx *ctx, const zkb_decoded *req, uint32_t nsessions,
    const zkb_xidtab *layout, int8_t *ops_out);
/* Device-pointer variant: req columns, layout arrays and ops_out are device memory. */
int zkb_xidtab_build_dev(zkb_ctx *ctx, const zkb_decoded *req, uint32_t nsessions,
    const zkb_xidtab *layout, int8_t *ops_out);

/*
 * The same table build for callers that must not allocate.  (No counterpart in the
 * reference: it exists for the embedding, not for the protocol.)  Who needs it: a host
 * whose own pool already owns all device memory, where a cudaMalloc made behind its back
 * fails or fragments the pool; and a service loop that may not call cudaMalloc/cudaFree
 * at all because those calls stall every stream on the device.  This entry point takes no
 * context and calls no allocator: the per-xid scratch it needs is a fixed array inside the
 * library image, ZKB_NOALLOC_MAX_XIDS slots of 8 bytes (8 MiB), present on every GPU that
 * has loaded the library for as long as the process lives.
 *
 * All pointers — the req columns (meta, frame_len, session, xid), the layout arrays and
 * ops_out — are DEVICE pointers on the current device; `stream` is a cudaStream_t (NULL =
 * the default stream) on which the work is queued, and the call returns once it has
 * finished.  Results are identical to zkb_xidtab_build_dev.
 *   - layout->nops above ZKB_NOALLOC_MAX_XIDS is refused with ZKB_ERANGE (text in errbuf);
 *     cover more xids by building the table for a few sessions per call.
 *   - The call waits for its own work (it ends with a stream synchronize), so it can never
 *     be part of a stream capture; the first call on each GPU also looks up where the array
 *     lives on that GPU.  With lazy module loading that first call is also the moment the
 *     8 MiB are claimed on that GPU: make one warm-up call per GPU at start-up, before the
 *     pool takes the rest of the memory.  The address is remembered per device ordinal for
 *     the life of the process; after cudaDeviceReset the process must be restarted.
 *   - A layout entry that points past the table (off[s] + count[s] > nops) is ignored.
 *   - Every call uses the same array (a call clears the slots it is about to use), so two
 *     calls whose GPU work overlaps, from any threads or streams, corrupt each other: run
 *     them one after the other.
 * errbuf (may be NULL) receives a message when the result is not ZKB_OK.
 */
#define ZKB_NOALLOC_MAX_XIDS (1u << 20)
int zkb_xidtab_build_noalloc(const zkb_decoded *req, uint32_t nsessions, const zkb_xidtab *layout,
    int8_t *ops_out, void *stream, char *errbuf, size_t errcap);

/* ---- session-level reductions --------------------------------------------------------- */

/*
 * What ZKSession folds out of the decoded server->client packets one at a time while a
 * connection is attached (lib/zk-session.js:227-236 and processNotification, :389-419),
 * as per-session reductions over a decoded RESPONSE stream:
 *
 *  max_zxid[s], n_replies[s]   every OK frame of kind RESPONSE whose opcode is not NOTIFICATION
 *      takes the zxid branch (:229-235): n_replies counts them, max_zxid is the greatest zxid
 *      among them under BigInteger.compareTo, i.e. as signed 64-bit integers (INT64_MIN when
 *      there were none).  The session's zs_lastZxid, which starts at 0 (:57), becomes
 *      max(zs_lastZxid, max_zxid[s]) — the binding does that fold, so a session carried across
 *      batches keeps its value.  jsbn is not vendored with the reference, so the signedness of
 *      this compare rests on jsbn's documented byte-array constructor, not on a reference test.
 *  notifications[s * ZKB_NOTIFY_SLOTS + k]   NOTIFICATION frames by what processNotification
 *      does with them: k = ZKB_NOTIFY_IGNORED when pkt.state is not 'SYNC_CONNECTED' (logged and
 *      dropped, :391-400 — that includes error replies, whose body is never read); k = 1..4 the
 *      event counted at :413-415, by NOTIFICATION_TYPE value (CREATED, DELETED, DATA_CHANGED,
 *      CHILDREN_CHANGED, zk-consts.js:111-116); k = ZKB_NOTIFY_BAD_TYPE when the type is outside
 *      that table (pkt.type is undefined and :402 throws a TypeError).
 *
 * Frames that are not OK (BAD_DECODE) never became packets and are skipped.  With
 * ZKB_STATS_UNTIL_ERROR only the frames before a session's first BAD_DECODE frame count: the
 * reference tears the connection down on the decoder's first 'error' (connection-fsm.js:234-239),
 * so nothing later would have reached the session.
 */
#define ZKB_NOTIFY_SLOTS     6
#define ZKB_NOTIFY_IGNORED   0
#define ZKB_NOTIFY_BAD_TYPE  5
#define ZKB_STATS_UNTIL_ERROR 1u

typedef struct zkb_session_stats {
	int64_t  *max_zxid;       /* [nsessions] */
	uint32_t *n_replies;      /* [nsessions] */
	uint32_t *notifications;  /* [nsessions * ZKB_NOTIFY_SLOTS] */
} zkb_session_stats;

/* Device pointers in `dec` (meta, session, zxid, a0, a1 and n_frames are read) and in `out`. */
int zkb_session_stats_dev(zkb_ctx *ctx, const zkb_decoded *dec, uint32_t nsessions, uint32_t flags,
    zkb_session_stats *out);
/* Host pointers in `dec` and `out`. */
int zkb_session_stats_host(zkb_ctx *ctx, const zkb_decoded *dec, uint32_t nsessions, uint32_t flags,
    zkb_session_stats *out);

/* ---- encode ----------------------------------------------------------------------------- */

/*
 * Batched ZKEncodeStream (zk-streams.js:121-148): frame i = 4-byte length + writeRequest /
 * writeConnectRequest / writeConnectResponse body.  frame_off must hold n+1 entries; status n.
 * Device variant: all pointers device memory, out_cap = bytes available in `out`; the
 * required size is returned in *out_len (ZKB_EOVERFLOW if larger than out_cap).
 */
int zkb_encode_dev(zkb_ctx *ctx, const zkb_requests *req, uint8_t *out, uint64_t out_cap,
    uint64_t *frame_off, uint8_t *status, uint64_t *out_len);
/* Host variant: pointers are host memory; *out receives a context-owned pinned buffer. */
int zkb_encode_host(zkb_ctx *ctx, const zkb_requests *req, const uint8_t **out,
    const uint64_t **frame_off, const uint8_t **status, uint64_t *out_len);

#ifdef __cplusplus
}
#endif
#endif /* ZKBATCH_H */
