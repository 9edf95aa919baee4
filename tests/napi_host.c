/*
 * napi_host.c — TEST INFRASTRUCTURE: a minimal stand-in for the part of node's N-API that
 * node-zkstream_b200/js/zkbatch_napi.c uses, so that the addon can be executed (not merely
 * compiled) on a box without node.  Values are tagged structs; there is no JavaScript engine
 * behind them.  tests/test_gpu_napi_addon.py links this file with the addon object and
 * libzkbatch.so and drives the addon's four functions through the nh_* helpers below.
 *
 * Only what the addon calls is implemented, with node's documented semantics: typed arrays view
 * an ArrayBuffer, napi_create_buffer_copy copies, a thrown error makes the pending-exception slot
 * non-empty and the native function's return value is ignored.
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

typedef struct napi_env__ *napi_env;
typedef struct nh_value *napi_value;
typedef struct nh_cbinfo *napi_callback_info;
typedef enum { napi_ok, napi_invalid_arg } napi_status;
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

enum { NH_UNDEFINED, NH_NULL, NH_NUMBER, NH_OBJECT, NH_ARRAYBUFFER, NH_TYPEDARRAY, NH_BUFFER, NH_FUNCTION, NH_EXTERNAL };

struct nh_prop { char *name; struct nh_value *v; struct nh_prop *next; };
struct nh_value {
	int kind;
	double num;
	void *data; size_t len;                 /* ARRAYBUFFER / BUFFER bytes; TYPEDARRAY: first element, len = elements */
	int owns;
	napi_typedarray_type tt;
	struct nh_value *ab;                    /* TYPEDARRAY: its ArrayBuffer */
	struct nh_prop *props;                  /* OBJECT */
	napi_callback fn;                       /* FUNCTION */
	void *ext; napi_finalize fin;           /* EXTERNAL */
	struct nh_value *all_next;
};
struct nh_cbinfo { size_t argc; struct nh_value **argv; };

static struct nh_value *nh_all = NULL;
static napi_module *nh_module = NULL;
static char nh_error[512];
static int nh_pending = 0;
static struct napi_env__ { int unused; } nh_env_obj;

static struct nh_value *nh_new(int kind)
{
	struct nh_value *v = calloc(1, sizeof (*v));
	v->kind = kind; v->all_next = nh_all; nh_all = v;
	return v;
}
static size_t esize(napi_typedarray_type t)
{
	switch (t) {
	case napi_int8_array: case napi_uint8_array: case napi_uint8_clamped_array: return 1;
	case napi_int16_array: case napi_uint16_array: return 2;
	case napi_float64_array: return 8;
	default: return 4;
	}
}

/* ---- the N-API functions the addon declares (zkbatch_napi.c:43-61) ---------------------- */
void napi_module_register(napi_module *m) { nh_module = m; }
napi_status napi_get_cb_info(napi_env e, napi_callback_info i, size_t *argc, napi_value *argv, napi_value *this_arg, void **data)
{
	(void)e; (void)data;
	size_t want = argc ? *argc : 0;
	for (size_t k = 0; k < want; k++) argv[k] = k < i->argc ? i->argv[k] : nh_new(NH_UNDEFINED);
	if (argc) *argc = i->argc;
	if (this_arg) *this_arg = NULL;
	return napi_ok;
}
napi_status napi_get_value_uint32(napi_env e, napi_value v, uint32_t *out)
{
	(void)e;
	if (!v || v->kind != NH_NUMBER) return napi_invalid_arg;
	*out = (uint32_t)(int64_t)v->num;
	return napi_ok;
}
napi_status napi_get_value_int32(napi_env e, napi_value v, int32_t *out)
{
	(void)e;
	if (!v || v->kind != NH_NUMBER) return napi_invalid_arg;
	*out = (int32_t)(int64_t)v->num;
	return napi_ok;
}
napi_status napi_get_buffer_info(napi_env e, napi_value v, void **data, size_t *len)
{
	(void)e;
	if (!v || v->kind != NH_BUFFER) return napi_invalid_arg;
	if (data) *data = v->data;
	if (len) *len = v->len;
	return napi_ok;
}
napi_status napi_get_typedarray_info(napi_env e, napi_value v, napi_typedarray_type *t, size_t *len, void **data,
    napi_value *ab, size_t *off)
{
	(void)e;
	if (!v || v->kind != NH_TYPEDARRAY) return napi_invalid_arg;
	if (t) *t = v->tt;
	if (len) *len = v->len;
	if (data) *data = v->data;
	if (ab) *ab = v->ab;
	if (off) *off = 0;
	return napi_ok;
}
napi_status napi_typeof(napi_env e, napi_value v, napi_valuetype *out)
{
	(void)e;
	switch (v ? v->kind : NH_UNDEFINED) {
	case NH_NULL: *out = napi_null; break;
	case NH_NUMBER: *out = napi_number; break;
	case NH_OBJECT: case NH_ARRAYBUFFER: case NH_TYPEDARRAY: case NH_BUFFER: *out = napi_object; break;
	case NH_FUNCTION: *out = napi_function; break;
	case NH_EXTERNAL: *out = napi_external; break;
	default: *out = napi_undefined; break;
	}
	return napi_ok;
}
napi_status napi_create_external(napi_env e, void *data, napi_finalize fin, void *hint, napi_value *out)
{
	(void)e; (void)hint;
	struct nh_value *v = nh_new(NH_EXTERNAL);
	v->ext = data; v->fin = fin;
	*out = v;
	return napi_ok;
}
napi_status napi_get_value_external(napi_env e, napi_value v, void **out)
{
	(void)e;
	if (!v || v->kind != NH_EXTERNAL) return napi_invalid_arg;
	*out = v->ext;
	return napi_ok;
}
napi_status napi_create_object(napi_env e, napi_value *out) { (void)e; *out = nh_new(NH_OBJECT); return napi_ok; }
napi_status napi_create_double(napi_env e, double d, napi_value *out)
{
	(void)e;
	*out = nh_new(NH_NUMBER); (*out)->num = d;
	return napi_ok;
}
napi_status napi_create_arraybuffer(napi_env e, size_t n, void **data, napi_value *out)
{
	(void)e;
	struct nh_value *v = nh_new(NH_ARRAYBUFFER);
	v->data = calloc(1, n ? n : 1); v->len = n; v->owns = 1;
	if (data) *data = v->data;
	*out = v;
	return napi_ok;
}
napi_status napi_create_typedarray(napi_env e, napi_typedarray_type t, size_t n, napi_value ab, size_t off, napi_value *out)
{
	(void)e;
	if (!ab || ab->kind != NH_ARRAYBUFFER || off + n * esize(t) > ab->len + 8) return napi_invalid_arg;
	struct nh_value *v = nh_new(NH_TYPEDARRAY);
	v->tt = t; v->len = n; v->data = (char *)ab->data + off; v->ab = ab;
	*out = v;
	return napi_ok;
}
napi_status napi_create_buffer_copy(napi_env e, size_t n, const void *src, void **data, napi_value *out)
{
	(void)e;
	struct nh_value *v = nh_new(NH_BUFFER);
	v->data = malloc(n ? n : 1); v->len = n; v->owns = 1;
	if (n) memcpy(v->data, src, n);
	if (data) *data = v->data;
	*out = v;
	return napi_ok;
}
napi_status napi_set_named_property(napi_env e, napi_value o, const char *name, napi_value val)
{
	(void)e;
	if (!o || o->kind != NH_OBJECT) return napi_invalid_arg;
	for (struct nh_prop *p = o->props; p; p = p->next)
		if (!strcmp(p->name, name)) { p->v = val; return napi_ok; }
	struct nh_prop *p = calloc(1, sizeof (*p));
	p->name = strdup(name); p->v = val; p->next = o->props; o->props = p;
	return napi_ok;
}
napi_status napi_get_named_property(napi_env e, napi_value o, const char *name, napi_value *out)
{
	(void)e;
	*out = NULL;
	if (!o || o->kind != NH_OBJECT) return napi_invalid_arg;
	for (struct nh_prop *p = o->props; p; p = p->next)
		if (!strcmp(p->name, name)) { *out = p->v; return napi_ok; }
	*out = nh_new(NH_UNDEFINED);
	return napi_ok;
}
napi_status napi_create_function(napi_env e, const char *name, size_t len, napi_callback cb, void *data, napi_value *out)
{
	(void)e; (void)name; (void)len; (void)data;
	*out = nh_new(NH_FUNCTION); (*out)->fn = cb;
	return napi_ok;
}
napi_status napi_throw_error(napi_env e, const char *code, const char *msg)
{
	(void)e;
	snprintf(nh_error, sizeof (nh_error), "%s%s%s", code ? code : "", code ? ": " : "", msg ? msg : "");
	nh_pending = 1;
	return napi_ok;
}

/* ---- what the test driver calls ------------------------------------------------------------ */
static struct nh_value *nh_exports_obj = NULL;
struct nh_value *nh_exports(void)
{
	if (!nh_exports_obj && nh_module) {
		nh_exports_obj = nh_new(NH_OBJECT);
		nh_module->nm_register_func(&nh_env_obj, nh_exports_obj);
	}
	return nh_exports_obj;
}
const char *nh_module_name(void) { return nh_module ? nh_module->nm_modname : NULL; }
struct nh_value *nh_number(double d) { struct nh_value *v = nh_new(NH_NUMBER); v->num = d; return v; }
struct nh_value *nh_null(void) { return nh_new(NH_NULL); }
struct nh_value *nh_object(void) { return nh_new(NH_OBJECT); }
/* Buffer / typed array over memory the caller keeps alive */
struct nh_value *nh_buffer(void *p, size_t n) { struct nh_value *v = nh_new(NH_BUFFER); v->data = p; v->len = n; return v; }
struct nh_value *nh_typedarray(int type, void *p, size_t n)
{
	struct nh_value *ab = nh_new(NH_ARRAYBUFFER);
	ab->data = p; ab->len = n * esize((napi_typedarray_type)type);
	struct nh_value *v = nh_new(NH_TYPEDARRAY);
	v->tt = (napi_typedarray_type)type; v->len = n; v->data = p; v->ab = ab;
	return v;
}
int nh_set(struct nh_value *o, const char *name, struct nh_value *v) { return napi_set_named_property(&nh_env_obj, o, name, v); }
struct nh_value *nh_get(struct nh_value *o, const char *name)
{
	struct nh_value *v = NULL;
	napi_get_named_property(&nh_env_obj, o, name, &v);
	return v;
}
int nh_kind(struct nh_value *v) { return v ? v->kind : NH_UNDEFINED; }
double nh_num(struct nh_value *v) { return v && v->kind == NH_NUMBER ? v->num : 0.0; }
void *nh_data(struct nh_value *v) { return v ? v->data : NULL; }
size_t nh_len(struct nh_value *v) { return v ? v->len : 0; }
int nh_array_type(struct nh_value *v) { return v && v->kind == NH_TYPEDARRAY ? (int)v->tt : -1; }
/* call exports[name](argv...): NULL + nh_last_error() if the function threw */
struct nh_value *nh_call(const char *name, size_t argc, struct nh_value **argv)
{
	struct nh_value *f = nh_get(nh_exports(), name);
	if (!f || f->kind != NH_FUNCTION) { snprintf(nh_error, sizeof (nh_error), "no such export: %s", name); return NULL; }
	struct nh_cbinfo info = { argc, argv };
	nh_pending = 0; nh_error[0] = 0;
	struct nh_value *r = f->fn(&nh_env_obj, &info);
	return nh_pending ? NULL : r;
}
const char *nh_last_error(void) { return nh_error; }
/* drop every value created so far; externals are finalised as the garbage collector would */
void nh_reset(void)
{
	struct nh_value *v = nh_all;
	while (v) {
		struct nh_value *n = v->all_next;
		if (v->kind == NH_EXTERNAL && v->fin) v->fin(&nh_env_obj, v->ext, NULL);
		if (v->owns) free(v->data);
		for (struct nh_prop *p = v->props; p;) { struct nh_prop *q = p->next; free(p->name); free(p); p = q; }
		free(v);
		v = n;
	}
	nh_all = NULL;
	nh_exports_obj = NULL;
}
