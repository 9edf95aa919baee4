"""TEST INFRASTRUCTURE — runs the reference's UNMODIFIED JavaScript codec.

The reference (node-zkstream) is pure JavaScript and this image has no node, but
the reference vendors a SpiderMonkey 1.x C engine (a dependency of its lint tool);
oracle/build_ref.sh compiles that into oracle/_ref/js.  This module wraps the
reference's own lib/zk-consts.js, lib/jute-buffer.js, lib/zk-buffer.js,
lib/errors.js and lib/zk-streams.js — read at run time from /root/reference, never
copied into this repository — in Node-style module functions, puts a small shim
(Buffer, assert-plus, util.inherits, stream.Transform, jsbn stub, ES5 polyfills)
in front of them and runs a driver script.  It exists to pin oracle/zkoracle.c and
to generate tests/golden/*.json (oracle/make_golden.py).  It only works where
/root/reference exists; callers skip otherwise.  Never imported by the product.

Strings: the shim's Buffer.toString('utf-8') maps each byte to one char code
(and back in new Buffer(str, 'utf-8')), so ustrings cross this harness as raw
bytes.  UTF-8 decoding is left to the JavaScript side of the real binding
(SURVEY.md §8a quirk 12); parity is on bytes.
"""
import json
import os
import subprocess
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("ZK_REFERENCE", "/root/reference")
JS = os.path.join(HERE, "_ref", "js")
REF_FILES = ["zk-consts", "errors", "jute-buffer", "zk-buffer", "zk-streams"]


def available():
    return os.path.isdir(os.path.join(REF, "lib")) and os.path.exists(JS)


SHIM_JS = r"""
/* ---- ES5 polyfills the 1.x engine lacks ---- */
if (!Object.keys) Object.keys = function (o) { var r = []; for (var k in o) if (o.hasOwnProperty(k)) r.push(k); return r; };
if (!Object.create) Object.create = function (p) { function F() {} F.prototype = p; return new F(); };
if (!Array.isArray) Array.isArray = function (a) { return a instanceof Array; };
if (!Function.prototype.bind) Function.prototype.bind = function (t) { var f = this; return function () { return f.apply(t, arguments); }; };

/* ---- Buffer: bytes as indexed own properties ---- */
function Buffer(a, enc) {
	var i;
	if (typeof (a) === 'number') {
		this.length = a;
		for (i = 0; i < a; ++i) this[i] = 0;
	} else if (typeof (a) === 'string') {
		var bytes = [];
		if (enc === 'hex') {
			for (i = 0; i + 1 < a.length; i += 2) bytes.push(parseInt(a.substr(i, 2), 16));
		} else if (enc === 'base64') {
			var T = 'ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz0123456789+/';
			var acc = 0, nb = 0;
			for (i = 0; i < a.length; ++i) {
				var c = T.indexOf(a.charAt(i));
				if (c < 0) continue;
				acc = (acc << 6) | c; nb += 6;
				if (nb >= 8) { nb -= 8; bytes.push((acc >> nb) & 0xff); }
			}
		} else {
			for (i = 0; i < a.length; ++i) bytes.push(a.charCodeAt(i) & 0xff);
		}
		this.length = bytes.length;
		for (i = 0; i < bytes.length; ++i) this[i] = bytes[i];
	} else {
		this.length = a.length;
		for (i = 0; i < a.length; ++i) this[i] = a[i] & 0xff;
	}
}
Buffer.isBuffer = function (b) { return (b instanceof Buffer); };
Buffer.prototype._chk = function (off, n) {
	if (typeof (off) !== 'number' || off < 0 || off + n > this.length)
		throw new RangeError('index out of range');
};
Buffer.prototype.readUInt32BE = function (off) {
	this._chk(off, 4);
	return (this[off] * 16777216 + (this[off + 1] << 16) + (this[off + 2] << 8) + this[off + 3]);
};
Buffer.prototype.readInt32BE = function (off) {
	this._chk(off, 4);
	return ((this[off] << 24) | (this[off + 1] << 16) | (this[off + 2] << 8) | this[off + 3]);
};
Buffer.prototype.readInt8 = function (off) {
	this._chk(off, 1);
	var v = this[off];
	return (v & 0x80) ? v - 256 : v;
};
Buffer.prototype.writeUInt32BE = function (v, off) {
	this._chk(off, 4);
	this[off] = (v >>> 24) & 0xff; this[off + 1] = (v >>> 16) & 0xff;
	this[off + 2] = (v >>> 8) & 0xff; this[off + 3] = v & 0xff;
};
Buffer.prototype.writeInt32BE = function (v, off) {
	this._chk(off, 4);
	this[off] = (v >>> 24) & 0xff; this[off + 1] = (v >>> 16) & 0xff;
	this[off + 2] = (v >>> 8) & 0xff; this[off + 3] = v & 0xff;
};
Buffer.prototype.writeInt8 = function (v, off) { this._chk(off, 1); this[off] = v & 0xff; };
Buffer.prototype.slice = function (s, e) {
	if (s === undefined) s = 0;
	if (e === undefined || e > this.length) e = this.length;
	if (s > e) s = e;
	var b = new Buffer(e - s);
	for (var i = s; i < e; ++i) b[i - s] = this[i];
	return b;
};
Buffer.prototype.copy = function (t, ts, ss, se) {
	if (ts === undefined) ts = 0;
	if (ss === undefined) ss = 0;
	if (se === undefined || se > this.length) se = this.length;
	var n = se - ss, i;
	if (t.length - ts < n) n = t.length - ts;
	if (t === this && ts > ss) { for (i = n - 1; i >= 0; --i) t[ts + i] = this[ss + i]; }
	else { for (i = 0; i < n; ++i) t[ts + i] = this[ss + i]; }
	return n;
};
Buffer.prototype.fill = function (v, s, e) {
	if (s === undefined) s = 0;
	if (e === undefined) e = this.length;
	for (var i = s; i < e; ++i) this[i] = v & 0xff;
};
Buffer.prototype.toString = function (enc) {
	var s = '', i;
	if (enc === 'hex') {
		for (i = 0; i < this.length; ++i) s += (this[i] < 16 ? '0' : '') + this[i].toString(16);
		return s;
	}
	for (i = 0; i < this.length; ++i) s += String.fromCharCode(this[i]);
	return s;
};

/* ---- assert-plus subset ---- */
function AssertionError(msg) { this.name = 'AssertionError'; this.message = msg; }
AssertionError.prototype = new Error();
var mod_assert_plus = (function () {
	var a = {};
	function mk(name, test) {
		a[name] = function (v, n) { if (!test(v)) throw new AssertionError((n || 'value') + ' (' + name + ') is required'); };
		var on = 'optional' + name.charAt(0).toUpperCase() + name.substr(1);
		a[on] = function (v, n) { if (v !== undefined && v !== null) a[name](v, n); };
		var an = 'arrayOf' + name.charAt(0).toUpperCase() + name.substr(1);
		a[an] = function (v, n) {
			if (!Array.isArray(v)) throw new AssertionError((n || 'value') + ' ([' + name + ']) is required');
			for (var i = 0; i < v.length; ++i) a[name](v[i], n);
		};
	}
	mk('object', function (v) { return (typeof (v) === 'object' && v !== null); });
	mk('buffer', function (v) { return Buffer.isBuffer(v); });
	mk('number', function (v) { return (typeof (v) === 'number' && !isNaN(v)); });
	mk('bool', function (v) { return (typeof (v) === 'boolean'); });
	mk('string', function (v) { return (typeof (v) === 'string'); });
	mk('func', function (v) { return (typeof (v) === 'function'); });
	a.ok = function (v, msg) { if (!v) throw new AssertionError(msg || (String(v) + ' == true')); };
	return a;
})();

var mod_util_shim = {
	inherits: function (ctor, sup) {
		ctor.super_ = sup;
		function F() {} F.prototype = sup.prototype;
		ctor.prototype = new F(); ctor.prototype.constructor = ctor;
	}
};

/* ---- stream.Transform: push() collects, emit() records ---- */
function Transform(opts) { this._pushed = []; this._events = []; }
Transform.prototype.push = function (x) { this._pushed.push(x); this._events.push({ push: x }); };
Transform.prototype.emit = function (ev, arg) { this._events.push({ ev: ev, arg: arg }); };
var mod_stream_shim = { Transform: Transform };

/* jsbn stand-in (the real one is not vendored): construction from a byte array or Buffer
 * (big-endian two's complement, as jsbn documents) or a small decimal literal, toByteArray
 * for the write path, and the signed compareTo / equals ZKSession uses on zxids. */
function BigInteger(x) {
	if (typeof (x) === 'string') {
		var v = parseInt(x, 10), b = [], i;
		for (i = 7; i >= 0; --i) { b[i] = v & 0xff; v = Math.floor(v / 256); }
		this._b = b;
	} else {
		this._b = x;
	}
}
BigInteger.prototype.toByteArray = function () { var r = []; for (var i = 0; i < this._b.length; ++i) r.push(this._b[i]); return r; };
BigInteger.prototype._parts = function () {
	/* [signed high word, unsigned low word] of the value sign-extended to 8 bytes */
	var n = this._b.length, neg = (n > 0 && (this._b[0] & 0x80) !== 0), hi = 0, lo = 0, i, byte;
	for (i = 0; i < 8; ++i) {
		byte = (i < 8 - n) ? (neg ? 0xff : 0) : (this._b[i - (8 - n)] & 0xff);
		if (i < 4) hi = hi * 256 + byte; else lo = lo * 256 + byte;
	}
	if (hi >= 2147483648) hi -= 4294967296;
	return [hi, lo];
};
BigInteger.prototype.compareTo = function (o) {
	var a = this._parts(), b = o._parts();
	if (a[0] !== b[0]) return (a[0] < b[0] ? -1 : 1);
	if (a[1] !== b[1]) return (a[1] < b[1] ? -1 : 1);
	return 0;
};
BigInteger.prototype.equals = function (o) { return this.compareTo(o) === 0; };
BigInteger.ZERO = new BigInteger('0');
function EventEmitterStub() {}
function FSMStub() {}

var DEFS = {};
var CACHE = {};
function requireShim(name) {
	if (name === 'assert-plus') return mod_assert_plus;
	if (name === 'util') return mod_util_shim;
	if (name === 'stream') return mod_stream_shim;
	if (name === 'jsbn') return { BigInteger: BigInteger };
	if (name === 'events') return { EventEmitter: EventEmitterStub };
	if (name === 'mooremachine') return { FSM: FSMStub };
	if (name === 'bunyan' || name === 'verror' || name === 'vasync') return {};
	var key = name.replace(/^\.\//, '');
	if (CACHE[key]) return CACHE[key].exports;
	if (!DEFS[key]) throw new Error('no module ' + name);
	var m = { exports: {} };
	CACHE[key] = m;
	DEFS[key](m, m.exports, requireShim);
	return m.exports;
}

/* ---- canonical JSON out ---- */
function jstr(s) {
	var r = '"', i, c;
	for (i = 0; i < s.length; ++i) {
		c = s.charCodeAt(i);
		if (c === 34) r += '\\"'; else if (c === 92) r += '\\\\';
		else if (c < 32 || c > 126) { var h = c.toString(16); while (h.length < 4) h = '0' + h; r += '\\u' + h; }
		else r += s.charAt(i);
	}
	return r + '"';
}
function isLongDate(v) { return (v && typeof (v) === 'object' && v.ld_buffer !== undefined); }
function canon(v) {
	var i, parts;
	if (v === undefined) return '{"$undefined":1}';
	if (v === null) return 'null';
	if (typeof (v) === 'number') return String(v);
	if (typeof (v) === 'boolean') return v ? 'true' : 'false';
	if (typeof (v) === 'string') return '{"$str":"' + new Buffer(v, 'utf-8').toString('hex') + '"}';
	if (Buffer.isBuffer(v)) return '{"$buf":"' + v.toString('hex') + '"}';
	if (isLongDate(v)) return '{"$ld":"' + v.ld_buffer.toString('hex') + '"}';
	if (Array.isArray(v)) {
		parts = [];
		for (i = 0; i < v.length; ++i) parts.push(canon(v[i]));
		return '[' + parts.join(',') + ']';
	}
	parts = [];
	for (var k in v) if (v.hasOwnProperty(k)) parts.push(jstr(k) + ':' + canon(v[k]));
	return '{' + parts.join(',') + '}';
}
function hex(s) { return new Buffer(s, 'hex'); }
"""

DRIVER_COMMON = r"""
var zkstreams = requireShim('./zk-streams');
var zkbuffer = requireShim('./zk-buffer');
var consts = requireShim('./zk-consts');

/* One ZKDecodeStream per session; fsm is the duck-typed stand-in test/nasty.test.js:298-310
 * uses, with 'handshaking' true for the first `hs` frames handled. */
function decodeSession(sess) {
	var handled = 0;
	var fsm = {
		zcf_xidMap: sess.xidMap || {},
		isInState: function (st) { return (st === 'handshaking' && handled < sess.hs); }
	};
	var ds = new zkstreams.ZKDecodeStream({ fsm: fsm, isServer: sess.isServer });
	var out = [], seen = 0, dead = false;
	function drain() {
		while (seen < ds._events.length) {
			var e = ds._events[seen++];
			if (e.push !== undefined) { out.push('{"pkt":' + canon(e.push) + '}'); handled++; }
			else if (e.ev === 'error') {
				out.push('{"error":{"code":' + jstr(e.arg.code) + ',"message":' + jstr(e.arg.message) + '}}');
				if (e.arg.code === 'BAD_LENGTH') dead = true; else handled++;
			}
		}
	}
	/* the real Transform calls handlePacket synchronously, so `handled` must advance between
	 * frames of one chunk: patch push/emit to drain eagerly */
	var opush = ds.push, oemit = ds.emit;
	ds.push = function (x) { opush.call(ds, x); drain(); };
	ds.emit = function (ev, a) { oemit.call(ds, ev, a); drain(); };
	for (var i = 0; i < sess.chunks.length && !dead; ++i) {
		var called = false;
		ds._transform(hex(sess.chunks[i]), null, function () { called = true; });
		drain();
		if (!called) dead = true;   /* BAD_LENGTH returns without cb(): the stream is stuck */
	}
	return '{"events":[' + out.join(',') + '],"used":' + ds.zd_used + '}';
}

function encodeSession(sess) {
	var hsleft = sess.hs;
	var fsm = { zcf_xidMap: {}, isInState: function (st) { return (st === 'handshaking' && hsleft > 0); } };
	var es = new zkstreams.ZKEncodeStream({ fsm: fsm, isServer: sess.isServer });
	var out = [];
	for (var i = 0; i < sess.pkts.length; ++i) {
		var n0 = es._pushed.length;
		try {
			es._transform(sess.pkts[i], null, function () {});
			out.push('{"frame":"' + es._pushed[n0].toString('hex') + '"}');
		} catch (e) {
			out.push('{"throw":' + jstr(String(e.name) + ': ' + String(e.message)) + '}');
		}
		if (hsleft > 0) hsleft--;
	}
	var xm = [];
	for (var k in fsm.zcf_xidMap) xm.push(jstr(k) + ':' + canon(fsm.zcf_xidMap[k]));
	return '{"frames":[' + out.join(',') + '],"xidMap":{' + xm.join(',') + '}}';
}
"""


def bundle(files=None):
    parts = [SHIM_JS]
    for name in (files or REF_FILES):
        with open(os.path.join(REF, "lib", name + ".js")) as f:
            src = f.read()
        parts.append("DEFS[%s] = function (module, exports, require) {\n%s\n};\n" % (json.dumps(name), src))
    parts.append(DRIVER_COMMON)
    return "\n".join(parts)


def run_js(driver, timeout=600, files=None):
    """Run bundle + driver under oracle/_ref/js; returns stdout."""
    if not available():
        raise RuntimeError("reference JS engine unavailable (need %s and %s)" % (REF, JS))
    with tempfile.NamedTemporaryFile("w", suffix=".js", delete=False) as f:
        f.write(bundle(files))
        f.write("\n")
        f.write(driver)
        path = f.name
    try:
        r = subprocess.run([JS, path], capture_output=True, text=True, timeout=timeout)
    finally:
        os.unlink(path)
    if r.returncode != 0 or r.stderr.strip():
        raise RuntimeError("reference JS failed: rc=%d\n%s\n%s" % (r.returncode, r.stderr[-2000:], r.stdout[-500:]))
    return r.stdout


def _decanon(v):
    """Turn the shim's canonical JSON back into Python: bytes for $buf/$str/$ld."""
    if isinstance(v, dict):
        if "$buf" in v:
            return bytes.fromhex(v["$buf"])
        if "$str" in v:
            return ("str", bytes.fromhex(v["$str"]))
        if "$ld" in v:
            return ("longdate", bytes.fromhex(v["$ld"]))
        if "$undefined" in v:
            return ("undefined",)
        return {k: _decanon(x) for k, x in v.items()}
    if isinstance(v, list):
        return [_decanon(x) for x in v]
    return v


def decode_sessions(sessions):
    """sessions: list of dict(chunks=[bytes,...], isServer=bool, hs=int, xidMap={xid:name}).
    Returns per session: dict(events=[{'pkt':..}|{'error':{code,message}}], used=int)."""
    lines = ["var SESS = ["]
    for s in sessions:
        xm = ",".join("%s:%s" % (json.dumps(str(k)), json.dumps(v)) for k, v in s.get("xidMap", {}).items())
        chunks = ",".join(json.dumps(bytes(c).hex()) for c in s["chunks"])
        lines.append("{isServer:%s,hs:%d,xidMap:{%s},chunks:[%s]}," % (
            "true" if s.get("isServer") else "false", s.get("hs", 1), xm, chunks))
    lines.append("];\nfor (var i = 0; i < SESS.length; ++i) print(decodeSession(SESS[i]));")
    out = run_js("\n".join(lines))
    return [_decanon(json.loads(l)) for l in out.splitlines() if l.strip()]


SESSION_DRIVER = r"""
var zksession = requireShim('./zk-session');

/* Decode one server->client stream with the reference decoder and hand every packet that
 * comes out after the handshake to the handler ZKSession.state_attached registers for its
 * connection's 'packet' event (lib/zk-session.js:227-236), on a bare object that has just
 * the members that handler and processNotification (:389-419) touch. */
function sessionFold(sess, untilError) {
	var handled = 0;
	var fsm = {
		zcf_xidMap: sess.xidMap || {},
		isInState: function (st) { return (st === 'handshaking' && handled < sess.hs); }
	};
	var ds = new zkstreams.ZKDecodeStream({ fsm: fsm, isServer: false });
	var counts = {}, ignored = 0, threw = 0, replies = 0, stopped = false;
	var self = {
		zs_lastZxid: undefined,
		zs_conn: {}, zs_expiryTimer: {},
		zs_log: { warn: function () { ignored++; }, trace: function () {} },
		zs_watchers: {},
		zs_collector: { getCollector: function () {
			return ({ increment: function (l) { counts[l.event] = (counts[l.event] || 0) + 1; } });
		} },
		resetExpiryTimer: function () {},
		processNotification: zksession.ZKSession.prototype.processNotification
	};
	var handler;
	var S = { on: function (em, ev, fn) { if (em === self.zs_conn && ev === 'packet') handler = fn; },
	    gotoState: function () {} };
	zksession.ZKSession.prototype.state_attached.call(self, S);
	var seen = 0, dead = false;
	function drain() {
		while (seen < ds._events.length) {
			var e = ds._events[seen++];
			if (e.push !== undefined) {
				var wasHandshake = (handled < sess.hs);
				handled++;
				if (wasHandshake || stopped) continue;
				var before = self.zs_lastZxid;
				if (e.push.opcode !== 'NOTIFICATION') replies++;
				try { handler(e.push); } catch (ex) { threw++; }
			} else if (e.ev === 'error') {
				if (e.arg.code === 'BAD_LENGTH') dead = true; else handled++;
				if (untilError) stopped = true;
			}
		}
	}
	var opush = ds.push, oemit = ds.emit;
	ds.push = function (x) { opush.call(ds, x); drain(); };
	ds.emit = function (ev, a) { oemit.call(ds, ev, a); drain(); };
	for (var i = 0; i < sess.chunks.length && !dead; ++i) {
		var called = false;
		ds._transform(hex(sess.chunks[i]), null, function () { called = true; });
		drain();
		if (!called) dead = true;
	}
	var z = self.zs_lastZxid, zhex = null;
	if (z !== undefined) { zhex = ''; for (var k = 0; k < z._b.length; ++k) { var h = (z._b[k] & 0xff).toString(16); zhex += (h.length < 2 ? '0' : '') + h; } }
	var cs = [];
	for (var ev in counts) cs.push(jstr(ev) + ':' + counts[ev]);
	return '{"maxZxid":' + (zhex === null ? 'null' : '"' + zhex + '"') + ',"replies":' + replies +
	    ',"counts":{' + cs.join(',') + '},"ignored":' + ignored + ',"threw":' + threw + '}';
}
"""


def session_folds(sessions, until_error=False):
    """Run the reference's own ZKSession packet handler over client-side sessions (as for
    decode_sessions).  zs_lastZxid starts undefined here (not 0) so the result is the plain
    maximum.  Returns per session dict(maxZxid=hex|None, replies, counts={event:n}, ignored, threw)."""
    lines = [SESSION_DRIVER, "var SESS = ["]
    for s in sessions:
        xm = ",".join("%s:%s" % (json.dumps(str(k)), json.dumps(v)) for k, v in s.get("xidMap", {}).items())
        chunks = ",".join(json.dumps(bytes(c).hex()) for c in s["chunks"])
        lines.append("{hs:%d,xidMap:{%s},chunks:[%s]}," % (s.get("hs", 1), xm, chunks))
    lines.append("];\nfor (var i = 0; i < SESS.length; ++i) print(sessionFold(SESS[i], %s));" % (
        "true" if until_error else "false"))
    out = run_js("\n".join(lines), files=REF_FILES + ["zk-session"])
    return [json.loads(l) for l in out.splitlines() if l.strip()]


def js_literal(v):
    """Python packet description -> JS source.  bytes -> Buffer, ('str', bytes) or str -> string."""
    if isinstance(v, bool):
        return "true" if v else "false"
    if isinstance(v, int):
        return str(v)
    if isinstance(v, (bytes, bytearray)):
        return "hex(%s)" % json.dumps(bytes(v).hex())
    if isinstance(v, str):
        return "hex(%s).toString('utf-8')" % json.dumps(v.encode("utf-8").hex())
    if isinstance(v, tuple) and v and v[0] == "str":
        return "hex(%s).toString('utf-8')" % json.dumps(bytes(v[1]).hex())
    if isinstance(v, list):
        return "[" + ",".join(js_literal(x) for x in v) + "]"
    if isinstance(v, dict):
        return "{" + ",".join("%s:%s" % (json.dumps(k), js_literal(x)) for k, x in v.items()) + "}"
    if v is None:
        return "undefined"
    raise TypeError(type(v))


def encode_sessions(sessions):
    """sessions: list of dict(pkts=[dict,...], isServer=bool, hs=int).
    Returns per session dict(frames=[{'frame':hex}|{'throw':msg}], xidMap={...})."""
    lines = ["var SESS = ["]
    for s in sessions:
        lines.append("{isServer:%s,hs:%d,pkts:%s}," % (
            "true" if s.get("isServer") else "false", s.get("hs", 0), js_literal(s["pkts"])))
    lines.append("];\nfor (var i = 0; i < SESS.length; ++i) print(encodeSession(SESS[i]));")
    out = run_js("\n".join(lines))
    return [_decanon(json.loads(l)) for l in out.splitlines() if l.strip()]
