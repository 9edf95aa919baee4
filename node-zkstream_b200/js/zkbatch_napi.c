/*
 * zkbatch_napi.c — the N-API addon that binds libzkbatch.so (include/zkbatch.h) into node
 * for node-zkstream_b200/js/batch-streams.js.
 *
 * It is deliberately thin: argument unpacking, one call into the C ABI, typed arrays out.
 * No node and no node_api.h exist in the build image, so the handful of N-API prototypes
 * used here are declared below (they are ABI-stable by design) and the file is only
 * compile-checked there (`__graft_entry__.build()`); build it for real with node-gyp against
 * the installed headers by defining ZKB_HAVE_NODE_API_H.
 *
 *   createContext(device)                                 -> External<zkb_ctx>
 *   decodeHost(ctx, data: Buffer, isServer: 0|1, handshakeFrames: uint32, xidOps: Int8Array|null)
 *        -> { n_frames, n_stat, n_buf, n_acl, consumed, <every column of zkb_decoded as a typed
 *             array; 64-bit columns as Int32Array [lo, hi] pairs> }          (one session)
 *   encodeColumns(ctx, cols)                              -> { bytes: Buffer, frame_off: Float64Array, status: Uint8Array }
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "zkbatch.h"

#ifdef ZKB_HAVE_NODE_API_H
#include <node_api.h>
#else
typedef struct napi_env__ *napi_env;
typedef struct napi_value__ *napi_value;
typedef struct napi_callback_info__ *napi_callback_info;
typedef enum { napi_ok } napi_status;
typedef enum {
	napi_int8_array, napi_uint8_array, napi_uint8_clamped_array, napi_int16_array, napi_uint16_array,
	napi_int32_array, napi_uint32_array, napi_float32_array, napi_float64_array
} napi_typedarray_type;
typedef enum { napi_undefined, napi_null, napi_boolean, napi_number, napi_string, napi_symbol, napi_object,
	napi_function, napi_external } napi_valuetype;
typedef napi_value (*napi_callback)(napi_env, napi_callback_info);
typedef void (*napi_finalize)(napi_env, void *, void *);
typedef struct {
	int nm_version; unsigned int nm_flags; const char *nm_filename;
	napi_value (*nm_register_func)(napi_env, napi_value);
	const char *nm_modname; void *nm_priv; void *reserved[4];
} napi_module;
napi_status napi_get_cb_info(napi_env, napi_callback_info, size_t *, napi_value *, napi_value *, void **);
napi_status napi_get_value_uint32(napi_env, napi_value, uint32_t *);
napi_status napi_get_value_int32(napi_env, napi_value, int32_t *);
napi_status napi_get_buffer_info(napi_env, napi_value, void **, size_t *);
napi_status napi_get_typedarray_info(napi_env, napi_value, napi_typedarray_type *, size_t *, void **, napi_value *, size_t *);
napi_status napi_typeof(napi_env, napi_value, napi_valuetype *);
napi_status napi_create_external(napi_env, void *, napi_finalize, void *, napi_value *);
napi_status napi_get_value_external(napi_env, napi_value, void **);
napi_status napi_create_object(napi_env, napi_value *);
napi_status napi_create_double(napi_env, double, napi_value *);
napi_status napi_create_arraybuffer(napi_env, size_t, void **, napi_value *);
napi_status napi_create_typedarray(napi_env, napi_typedarray_type, size_t, napi_value, size_t, napi_value *);
napi_status napi_create_buffer_copy(napi_env, size_t, const void *, void **, napi_value *);
napi_status napi_set_named_property(napi_env, napi_value, const char *, napi_value);
napi_status napi_get_named_property(napi_env, napi_value, const char *, napi_value *);
napi_status napi_create_function(napi_env, const char *, size_t, napi_callback, void *, napi_value *);
napi_status napi_throw_error(napi_env, const char *, const char *);
void napi_module_register(napi_module *);
#define NAPI_AUTO_LENGTH ((size_t)-1)
#endif

static void ctx_finalize(napi_env env, void *data, void *hint)
{
	(void)env; (void)hint;
	zkb_ctx_destroy((zkb_ctx *)data);
}

/* createContext(device): throws when no CUDA device is usable — the codec has no CPU path */
static napi_value create_context(napi_env env, napi_callback_info info)
{
	size_t argc = 1; napi_value argv[1]; int32_t dev = 0; zkb_ctx *ctx = NULL; napi_value out = NULL;
	napi_get_cb_info(env, info, &argc, argv, NULL, NULL);
	if (argc >= 1) napi_get_value_int32(env, argv[0], &dev);
	if (zkb_ctx_create(dev, &ctx) != ZKB_OK) {
		napi_throw_error(env, "ZKB_ENODEV", "zkbatch: no usable CUDA device (there is no CPU fallback)");
		return NULL;
	}
	napi_create_external(env, ctx, ctx_finalize, NULL, &out);
	return out;
}

/* a typed-array copy of n elements of esz bytes */
static void put_array(napi_env env, napi_value obj, const char *name, napi_typedarray_type t, const void *src,
    size_t n, size_t esz, size_t words_per_elem)
{
	void *dst = NULL; napi_value ab = NULL, ta = NULL;
	napi_create_arraybuffer(env, (n * esz) != 0 ? n * esz : 8, &dst, &ab);
	if (n && src) memcpy(dst, src, n * esz);
	napi_create_typedarray(env, t, n * words_per_elem, ab, 0, &ta);
	napi_set_named_property(env, obj, name, ta);
}
static void put_number(napi_env env, napi_value obj, const char *name, double v)
{
	napi_value x = NULL;
	napi_create_double(env, v, &x);
	napi_set_named_property(env, obj, name, x);
}

static napi_value decode_host(napi_env env, napi_callback_info info)
{
	size_t argc = 5; napi_value argv[5];
	zkb_ctx *ctx = NULL; void *data = NULL; size_t len = 0; uint32_t is_server = 0, hs = 1;
	napi_get_cb_info(env, info, &argc, argv, NULL, NULL);
	if (argc < 5) { napi_throw_error(env, NULL, "decodeHost(ctx, data, isServer, handshakeFrames, xidOps)"); return NULL; }
	napi_get_value_external(env, argv[0], (void **)&ctx);
	napi_get_buffer_info(env, argv[1], &data, &len);
	napi_get_value_uint32(env, argv[2], &is_server);
	napi_get_value_uint32(env, argv[3], &hs);

	uint64_t off[2] = { 0, (uint64_t)len };
	uint8_t flags[1] = { (uint8_t)(is_server ? ZKB_SESS_SERVER : 0) };
	uint32_t hsv[1] = { hs };
	zkb_sessions sess = { 1, off, flags, hsv };
	zkb_xidtab xt; int32_t base[1] = { 0 }; uint32_t count[1] = { 0 }; uint64_t xoff[1] = { 0 };
	memset(&xt, 0, sizeof (xt));
	napi_valuetype vt = napi_null;
	napi_typeof(env, argv[4], &vt);
	if (vt == napi_object) {
		napi_typedarray_type tt; size_t n = 0; void *p = NULL;
		napi_get_typedarray_info(env, argv[4], &tt, &n, &p, NULL, NULL);
		count[0] = (uint32_t)n;
		xt.base = base; xt.count = count; xt.off = xoff; xt.ops = (const int8_t *)p; xt.nops = n;
	}
	zkb_decoded d;
	memset(&d, 0, sizeof (d));
	int rc = zkb_decode_host(ctx, (const uint8_t *)data, (uint64_t)len, &sess, xt.ops ? &xt : NULL, &d);
	if (rc != ZKB_OK) { napi_throw_error(env, "ZKB", zkb_last_error(ctx)); return NULL; }

	napi_value o = NULL;
	napi_create_object(env, &o);
	put_number(env, o, "n_frames", (double)d.n_frames); put_number(env, o, "n_stat", (double)d.n_stat);
	put_number(env, o, "n_buf", (double)d.n_buf); put_number(env, o, "n_acl", (double)d.n_acl);
	put_number(env, o, "consumed", (double)d.sess_consumed[0]);
#define I32(name, n) put_array(env, o, #name, napi_int32_array, d.name, n, 4, 1)
#define U32(name, n) put_array(env, o, #name, napi_uint32_array, d.name, n, 4, 1)
#define I64(name, n) put_array(env, o, #name, napi_int32_array, d.name, n, 8, 2)   /* [lo, hi] pairs */
	I64(frame_off, d.n_frames); I32(frame_len, d.n_frames); U32(session, d.n_frames); U32(meta, d.n_frames);
	I32(xid, d.n_frames); I64(zxid, d.n_frames); I32(err, d.n_frames); I32(a0, d.n_frames); I32(a1, d.n_frames);
	U32(stat_frame, d.n_stat); I64(czxid, d.n_stat); I64(mzxid, d.n_stat); I64(ctime, d.n_stat); I64(mtime, d.n_stat);
	I32(version, d.n_stat); I32(cversion, d.n_stat); I32(aversion, d.n_stat); I64(ephemeralOwner, d.n_stat);
	I32(dataLength, d.n_stat); I32(numChildren, d.n_stat); I64(pzxid, d.n_stat);
	U32(buf_frame, d.n_buf); I64(buf_off, d.n_buf); I32(buf_len, d.n_buf);
	U32(acl_frame, d.n_acl); I32(acl_perms, d.n_acl);
	I64(sess_consumed, 1); I64(sess_bad_length_at, 1); U32(sess_first_frame, 2);
#undef I32
#undef U32
#undef I64
	return o;
}

static const void *ta_ptr(napi_env env, napi_value obj, const char *name, size_t *n)
{
	napi_value v = NULL; napi_typedarray_type tt; void *p = NULL; size_t len = 0;
	napi_get_named_property(env, obj, name, &v);
	napi_get_typedarray_info(env, v, &tt, &len, &p, NULL, NULL);
	if (n) *n = len;
	return p;
}

/* encodeColumns(ctx, cols): cols = flatten.js output with every array already a typed array of the
 * element type zkb_requests names (Int32Array pairs for the 64-bit columns) and arena a Buffer */
static napi_value encode_columns(napi_env env, napi_callback_info info)
{
	size_t argc = 2; napi_value argv[2]; zkb_ctx *ctx = NULL; size_t n = 0, tmp = 0;
	napi_get_cb_info(env, info, &argc, argv, NULL, NULL);
	if (argc < 2) { napi_throw_error(env, NULL, "encodeColumns(ctx, cols)"); return NULL; }
	napi_get_value_external(env, argv[0], (void **)&ctx);
	zkb_requests r;
	memset(&r, 0, sizeof (r));
	r.xid = ta_ptr(env, argv[1], "xid", &n); r.n = n;
	r.op = ta_ptr(env, argv[1], "op", NULL);
	{
		napi_value v = NULL; void *p = NULL; size_t len = 0;
		napi_get_named_property(env, argv[1], "arena", &v);
		napi_get_buffer_info(env, v, &p, &len);
		r.arena = p; r.arena_len = len;
	}
	r.path_off = ta_ptr(env, argv[1], "path_off", NULL); r.path_len = ta_ptr(env, argv[1], "path_len", NULL);
	r.data_off = ta_ptr(env, argv[1], "data_off", NULL); r.data_len = ta_ptr(env, argv[1], "data_len", NULL);
	r.a0 = ta_ptr(env, argv[1], "a0", NULL); r.a1 = ta_ptr(env, argv[1], "a1", NULL);
	r.zxid = ta_ptr(env, argv[1], "zxid", NULL); r.zxid2 = ta_ptr(env, argv[1], "zxid2", NULL);
	r.vec_first = ta_ptr(env, argv[1], "vec_first", NULL); r.vec_count = ta_ptr(env, argv[1], "vec_count", NULL);
	r.n0 = ta_ptr(env, argv[1], "n0", NULL); r.n1 = ta_ptr(env, argv[1], "n1", NULL);
	r.elem_off = ta_ptr(env, argv[1], "elem_off", &tmp); r.elem_len = ta_ptr(env, argv[1], "elem_len", &tmp); r.n_elem = tmp;
	r.acl_perms = ta_ptr(env, argv[1], "acl_perms", &tmp); r.n_acl = tmp;
	r.acl_scheme_off = ta_ptr(env, argv[1], "acl_scheme_off", NULL); r.acl_scheme_len = ta_ptr(env, argv[1], "acl_scheme_len", NULL);
	r.acl_id_off = ta_ptr(env, argv[1], "acl_id_off", NULL); r.acl_id_len = ta_ptr(env, argv[1], "acl_id_len", NULL);

	const uint8_t *bytes = NULL, *status = NULL; const uint64_t *fo = NULL; uint64_t total = 0;
	int rc = zkb_encode_host(ctx, &r, &bytes, &fo, &status, &total);
	if (rc != ZKB_OK) { napi_throw_error(env, "ZKB", zkb_last_error(ctx)); return NULL; }
	napi_value o = NULL, b = NULL; void *cp = NULL;
	napi_create_object(env, &o);
	napi_create_buffer_copy(env, (size_t)total, bytes, &cp, &b);
	napi_set_named_property(env, o, "bytes", b);
	{
		/* frame offsets as doubles: exact below 2^53 */
		void *dst = NULL; napi_value ab = NULL, ta = NULL;
		napi_create_arraybuffer(env, (r.n + 1) * 8, &dst, &ab);
		for (uint64_t i = 0; i <= r.n; i++) ((double *)dst)[i] = (double)fo[i];
		napi_create_typedarray(env, napi_float64_array, (size_t)r.n + 1, ab, 0, &ta);
		napi_set_named_property(env, o, "frame_off", ta);
	}
	put_array(env, o, "status", napi_uint8_array, status, (size_t)r.n, 1, 1);
	return o;
}

/* sessionStats(ctx, cols, nsessions, flags): cols = a decodeHost result (response direction).
 * Returns { n_replies: Uint32Array, max_zxid: Int32Array [lo, hi] pairs, notifications:
 * Uint32Array[6 * nsessions] } for js/session-fold.js (zkb_session_stats_host). */
static napi_value session_stats(napi_env env, napi_callback_info info)
{
	size_t argc = 4; napi_value argv[4];
	zkb_ctx *ctx = NULL; uint32_t nsess = 1, flags = 0;
	napi_get_cb_info(env, info, &argc, argv, NULL, NULL);
	if (argc < 2) { napi_throw_error(env, NULL, "sessionStats(ctx, cols, nsessions, flags)"); return NULL; }
	napi_get_value_external(env, argv[0], (void **)&ctx);
	if (argc >= 3) napi_get_value_uint32(env, argv[2], &nsess);
	if (argc >= 4) napi_get_value_uint32(env, argv[3], &flags);
	zkb_decoded d;
	size_t n = 0;
	memset(&d, 0, sizeof (d));
	d.meta = (uint32_t *)ta_ptr(env, argv[1], "meta", &n);
	d.n_frames = n;
	d.session = (uint32_t *)ta_ptr(env, argv[1], "session", NULL);
	d.zxid = (int64_t *)ta_ptr(env, argv[1], "zxid", NULL);
	d.a0 = (int32_t *)ta_ptr(env, argv[1], "a0", NULL);
	d.a1 = (int32_t *)ta_ptr(env, argv[1], "a1", NULL);
	napi_value o = NULL, ab = NULL, ta = NULL;
	zkb_session_stats st;
	void *p = NULL;
	napi_create_object(env, &o);
	napi_create_arraybuffer(env, (size_t)nsess * 8 + 8, &p, &ab); st.max_zxid = (int64_t *)p;
	napi_create_typedarray(env, napi_int32_array, (size_t)nsess * 2, ab, 0, &ta);
	napi_set_named_property(env, o, "max_zxid", ta);
	napi_create_arraybuffer(env, (size_t)nsess * 4 + 8, &p, &ab); st.n_replies = (uint32_t *)p;
	napi_create_typedarray(env, napi_uint32_array, nsess, ab, 0, &ta);
	napi_set_named_property(env, o, "n_replies", ta);
	napi_create_arraybuffer(env, (size_t)nsess * 4 * ZKB_NOTIFY_SLOTS + 8, &p, &ab); st.notifications = (uint32_t *)p;
	napi_create_typedarray(env, napi_uint32_array, (size_t)nsess * ZKB_NOTIFY_SLOTS, ab, 0, &ta);
	napi_set_named_property(env, o, "notifications", ta);
	if (zkb_session_stats_host(ctx, &d, nsess, flags, &st) != ZKB_OK) {
		napi_throw_error(env, "ZKB", zkb_last_error(ctx));
		return NULL;
	}
	return o;
}

static napi_value init(napi_env env, napi_value exports)
{
	napi_value f = NULL;
	napi_create_function(env, "createContext", NAPI_AUTO_LENGTH, create_context, NULL, &f);
	napi_set_named_property(env, exports, "createContext", f);
	napi_create_function(env, "decodeHost", NAPI_AUTO_LENGTH, decode_host, NULL, &f);
	napi_set_named_property(env, exports, "decodeHost", f);
	napi_create_function(env, "encodeColumns", NAPI_AUTO_LENGTH, encode_columns, NULL, &f);
	napi_set_named_property(env, exports, "encodeColumns", f);
	napi_create_function(env, "sessionStats", NAPI_AUTO_LENGTH, session_stats, NULL, &f);
	napi_set_named_property(env, exports, "sessionStats", f);
	return exports;
}

static napi_module zkbatch_module = { 1, 0, __FILE__, init, "zkbatch_napi", NULL, { 0 } };
__attribute__((constructor)) static void zkbatch_register(void) { napi_module_register(&zkbatch_module); }
