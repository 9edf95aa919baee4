"""The JavaScript half of the drop-in (node-zkstream_b200/js/materialise.js) against the
reference, in the same JavaScript engine: columns from the CPU oracle are handed to the shim,
which must rebuild exactly the packet objects (and error objects) that the reference's
ZKDecodeStream produced for the same bytes (tests/golden/ref_decode.json).  Needs the
SpiderMonkey shell built from the reference tree; skipped where that is absent."""
import json
import os

import numpy as np
import pytest

import ref_js
from _canon import _decanon_guard  # noqa: F401  (import check below)

pytestmark = pytest.mark.skipif(not ref_js.available(), reason="reference JS engine not available here")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


def js_array(a):
    return "[" + ",".join(str(int(x)) for x in a) + "]"


def js_array64(a):
    """64-bit column -> interleaved [lo, hi] int32 words, as an Int32Array view would show it."""
    w = np.ascontiguousarray(a).view(np.int32)
    return js_array(w)


def test_js_materialiser_reproduces_reference_packets(oracle, synth, zk):
    from _canon import canon_events
    g = json.load(open(os.path.join(GOLD, "ref_decode.json")))
    sessions = [s for s in g["sessions"] if s["corpus"] in ("capture1+nasty", "fuzz0", "fuzz1", "fuzz2", "cfg2", "cfg3")]
    c = synth.Corpus()
    for s in sessions:
        c.add(bytes.fromhex(s["hex"]), server=s["isServer"], hs=s["hs"], xidmap={int(k): v for k, v in s["xidMap"].items()})
    d, off, fl, hs = c.arrays()
    cols = oracle.decode(d, off, fl, hs, c.xidtable())
    parts = ["var cols = {"]
    for name, dt in zk._abi.FRAME_COLS + zk._abi.STAT_COLS + zk._abi.BUF_COLS + zk._abi.ACL_COLS + zk._abi.SESS_COLS:
        a = getattr(cols, name)
        parts.append("%s: %s," % (name, js_array64(a) if dt in ("i8", "u8") else js_array(a)))
    parts.append("n_frames: %d, n_stat: %d, n_buf: %d, n_acl: %d};" % (cols.n_frames, cols.n_stat, cols.n_buf, cols.n_acl))
    parts.append("var data = hex(%s);" % json.dumps(c.data.hex()))
    parts.append(open(os.path.join(ROOT, "node-zkstream_b200", "js", "materialise.js")).read())
    parts.append(r"""
var jute = requireShim('./jute-buffer');
var LongDate = (new jute.JuteBuffer({ buffer: new Buffer(8) })).readLongDate().constructor;
var M = makeMaterialiser({ consts: requireShim('./zk-consts'), LongDate: LongDate,
    ZKProtocolError: requireShim('./errors').ZKProtocolError, Buffer: Buffer });
var m = new M(cols, data);
for (var s = 0; s < %d; ++s) {
	var out = [];
	m.forEach(s, function (pkt) { out.push('{"pkt":' + canon(pkt) + '}'); },
	    function (e) { out.push('{"error":{"code":' + jstr(e.code) + ',"message":' + jstr(e.message) + '}}'); });
	print('{"events":[' + out.join(',') + ']}');
}
""" % len(sessions))
    out = ref_js.run_js("\n".join(parts))
    lines = [l for l in out.splitlines() if l.strip()]
    assert len(lines) == len(sessions)
    for i, (line, s) in enumerate(zip(lines, sessions)):
        got = canon_events(ref_js._decanon(json.loads(line))["events"])
        assert got == s["events"], (s["corpus"], i)


def test_js_flatten_matches_python_request_batch(zk):
    """flatten.js (packets -> encode descriptors) must lay packets out exactly as encode.py does."""
    from test_oracle_golden import unjson
    g = json.load(open(os.path.join(GOLD, "ref_encode.json")))
    src = open(os.path.join(ROOT, "node-zkstream_b200", "js", "flatten.js")).read()
    for p in g["packets"]:
        pkt = unjson(p["pkt"])
        batch = zk.RequestBatch.from_packets([pkt], p["handshaking"], p["isServer"])
        driver = src + r"""
var fl = makeFlatten({ consts: requireShim('./zk-consts'), assert: requireShim('assert-plus'), Buffer: Buffer });
var c = fl([%s], %s, %s);
var ar = '';
for (var i = 0; i < c.arena.length; ++i) ar += c.arena[i].toString('hex');
var o = [];
for (var k in c) if (k !== 'arena') o.push(jstr(k) + ':' + (typeof (c[k]) === 'number' ? c[k] : '[' + c[k].join(',') + ']'));
print('{' + o.join(',') + ',"arena":"' + ar + '"}');
""" % (ref_js.js_literal(pkt), "true" if p["handshaking"] else "false", "true" if p["isServer"] else "false")
        got = json.loads(ref_js.run_js(driver))
        assert got["arena"] == batch.arena.tobytes().hex(), pkt
        for name, dt in zk._abi.REQ_COLS + zk._abi.ELEM_COLS + zk._abi.REQ_ACL_COLS:
            want = np.ascontiguousarray(batch.cols[name], dtype=dt)
            if name in ("zxid", "zxid2"):
                want = want.view(np.int32)
            assert [int(x) for x in want] == got[name], (name, pkt)


def test_js_session_fold_matches_reference_handler(oracle, synth, zk):
    """js/session-fold.js applied to the oracle's reductions leaves a session where the
    reference's own packet handler left it (tests/golden/ref_session_folds.json)."""
    g = json.load(open(os.path.join(GOLD, "ref_session_folds.json")))["sessions"]
    c = synth.Corpus()
    for s in g:
        c.add(bytes.fromhex(s["hex"]), hs=s["hs"], xidmap={int(k): v for k, v in s["xidMap"].items()})
    d, off, fl, hs = c.arrays()
    cols = oracle.decode(d, off, fl, hs, c.xidtable())
    st = oracle.session_stats(cols, len(g), 0)
    parts = ["var stats = {n_replies: %s, max_zxid: %s, notifications: %s};" % (
        js_array(st.n_replies), js_array64(st.max_zxid), js_array(st.notifications.reshape(-1)))]
    parts.append(open(os.path.join(ROOT, "node-zkstream_b200", "js", "session-fold.js")).read())
    parts.append(r"""
var fold = makeSessionFold({ BigInteger: BigInteger, consts: requireShim('./zk-consts') });
for (var s = 0; s < %d; ++s) {
	var counts = {};
	var session = { zs_lastZxid: undefined, zs_collector: { getCollector: function () {
		return ({ increment: function (l) { counts[l.event] = (counts[l.event] || 0) + 1; } }); } } };
	fold(session, stats, s);
	var z = session.zs_lastZxid, zhex = null;
	if (z !== undefined) { zhex = ''; for (var k = 0; k < 8; ++k) { var h = (z._b[k] & 0xff).toString(16); zhex += (h.length < 2 ? '0' : '') + h; } }
	var cs = [];
	for (var ev in counts) cs.push(jstr(ev) + ':' + counts[ev]);
	print('{"maxZxid":' + (zhex === null ? 'null' : '"' + zhex + '"') + ',"counts":{' + cs.join(',') + '}}');
}
""" % len(g))
    out = ref_js.run_js("\n".join(parts))
    lines = [json.loads(l) for l in out.splitlines() if l.strip()]
    assert len(lines) == len(g)
    for got, ref in zip(lines, g):
        assert got["maxZxid"] == ref["all"]["maxZxid"]
        assert got["counts"] == ref["all"]["counts"]


def _cols_js(cols, zk, consumed):
    parts = []
    for name, dt in zk._abi.FRAME_COLS + zk._abi.STAT_COLS + zk._abi.BUF_COLS + zk._abi.ACL_COLS + zk._abi.SESS_COLS:
        a = getattr(cols, name)
        parts.append("%s:%s" % (name, js_array64(a) if dt in ("i8", "u8") else js_array(a)))
    parts.append("n_frames:%d,n_stat:%d,n_buf:%d,n_acl:%d,consumed:%d" % (cols.n_frames, cols.n_stat, cols.n_buf,
                                                                       cols.n_acl, consumed))
    return "{" + ",".join(parts) + "}"


def test_js_batch_decode_stream_with_recorded_addon(oracle, synth, zk):
    """js/batch-streams.js (the drop-in ZKDecodeStream) run in the reference's engine against a
    recorded addon: every decodeHost call the stream makes is answered with the columns the
    oracle gives for exactly those bytes, so what is tested is the stream's own logic - carrying
    incomplete frames over, the handshake state per call, the xid table it builds from
    fsm.zcf_xidMap, packets pushed and errors emitted in order, getting stuck after BAD_LENGTH -
    against what the reference's ZKDecodeStream did with the same bytes (ref_decode.json)."""
    import random
    from _canon import canon_events
    g = json.load(open(os.path.join(GOLD, "ref_decode.json")))["sessions"]
    rng = random.Random(5)
    chosen = []
    for s in g:
        ev = s["events"]
        data = bytes.fromhex(s["hex"])
        if s["hs"] != 1 or len(ev) < 2 or "pkt" not in ev[0] or len(data) < 8:
            continue
        first = 4 + int.from_bytes(data[:4], "big")
        if first >= len(data):
            continue
        chosen.append((s, data, first))
    chosen = chosen[:60]
    assert len(chosen) >= 30
    table, sess_js = {}, []
    for s, data, first in chosen:
        cuts = sorted(set([first] + [rng.randrange(first, len(data)) for _ in range(rng.randrange(0, 4))]))
        chunks = [data[a:b] for a, b in zip([0] + cuts, cuts + [len(data)]) if b > a]
        xm = {int(k): v for k, v in s["xidMap"].items()}
        xt = zk.XidTable.from_maps([xm]) if not s["isServer"] else None
        flags = np.array([zk._abi.SESS_SERVER if s["isServer"] else 0], dtype=np.uint8)
        carry, stuck = b"", False
        for i, ch in enumerate(chunks):
            if stuck:
                break
            pend = carry + ch
            hsv = np.array([0xFFFFFFFF if i == 0 else 0], dtype=np.uint32)
            cols = oracle.decode(np.frombuffer(pend, dtype=np.uint8), None, flags, hsv, xt)
            consumed = int(cols.sess_consumed[0])
            table[pend.hex() + ("S" if s["isServer"] else "C") + "h%d" % int(hsv[0])] = _cols_js(cols, zk, consumed)
            stuck = int(cols.sess_bad_length_at[0]) != zk._abi.NO_OFFSET
            carry = pend[consumed:]
        # the batching replay: the whole stream in one GPU call, told that its first frame is the handshake
        whole = oracle.decode(np.frombuffer(data, dtype=np.uint8), None, flags, np.array([1], dtype=np.uint32), xt)
        table[data.hex() + ("S" if s["isServer"] else "C") + "h1"] = _cols_js(whole, zk, int(whole.sess_consumed[0]))
        names = {str(k): zk.consts.OP_CODE_LOOKUP[v] for k, v in xm.items()}
        sess_js.append("{isServer:%s,xidMap:%s,chunks:[%s]}" % (
            "true" if s["isServer"] else "false", json.dumps(names), ",".join(json.dumps(c.hex()) for c in chunks)))
    src = {n: open(os.path.join(ROOT, "node-zkstream_b200", "js", n)).read()
           for n in ("materialise.js", "flatten.js", "batch-streams.js")}
    driver = "\n".join([
        "var TABLE = {" + ",".join("%s:%s" % (json.dumps(k), v) for k, v in table.items()) + "};",
        "function Int8Array(n) { var a = []; for (var i = 0; i < n; ++i) a.push(0); return a; }",
        "Buffer.concat = function (list, total) { var h = ''; for (var i = 0; i < list.length; ++i) h += list[i].toString('hex'); return new Buffer(h, 'hex'); };",
        src["materialise.js"], src["flatten.js"],
        "var CUR = { server: false, hs: true };",
        "var fakeAddon = { createContext: function () { return {}; }, decodeHost: function (ctx, data, isServer, hs, ops) {",
        "  var key = data.toString('hex') + (isServer ? 'S' : 'C') + 'h' + hs;",
        "  if (!TABLE[key]) throw new Error('unrecorded decodeHost call ' + key.slice(-40));",
        "  CUR.ops = ops; return TABLE[key]; } };",
        "DEFS['batch-streams'] = function (module, exports, require) {\n" + src["batch-streams.js"] + "\n};",
        r"""
function req2(name) {
	if (name === 'zkstream/lib/zk-consts') return requireShim('./zk-consts');
	if (name === 'zkstream/lib/errors') return requireShim('./errors');
	if (name === 'zkstream/lib/jute-buffer') return requireShim('./jute-buffer');
	if (name === './build/Release/zkbatch_napi.node') return fakeAddon;
	if (name === './materialise') return { makeMaterialiser: makeMaterialiser };
	if (name === './flatten') return { makeFlatten: makeFlatten };
	return requireShim(name);
}
var bm = { exports: {} };
DEFS['batch-streams'](bm, bm.exports, req2);
var BatchDecode = bm.exports.ZKDecodeStream;
var consts = requireShim('./zk-consts');
for (var si = 0; si < SESS.length; ++si) {
	var sess = SESS[si], calls = 0;
	var fsm = { zcf_xidMap: sess.xidMap, isInState: function (st) { return (st === 'handshaking' && calls === 0); } };
	var ds = new BatchDecode({ fsm: fsm, isServer: sess.isServer });
	var out = [], opsOk = true;
	for (var i = 0; i < sess.chunks.length; ++i) {
		var called = false;
		ds._transform(hex(sess.chunks[i]), null, function () { called = true; });
		calls++;
		if (!sess.isServer) {
			/* the table handed to the addon must say what fsm.zcf_xidMap says */
			for (var k in sess.xidMap) if (parseInt(k, 10) >= 0 && CUR.ops[parseInt(k, 10)] !== consts.OP_CODES[sess.xidMap[k]]) opsOk = false;
		}
		if (!called) break;            /* BAD_LENGTH: stuck, as the reference is */
	}
	for (var e = 0; e < ds._events.length; ++e) {
		var ev = ds._events[e];
		if (ev.push !== undefined) out.push('{"pkt":' + canon(ev.push) + '}');
		else if (ev.ev === 'error') out.push('{"error":{"code":' + jstr(ev.arg.code) + ',"message":' + jstr(ev.arg.message) + '}}');
	}
	print('{"opsOk":' + (opsOk ? 'true' : 'false') + ',"events":[' + out.join(',') + ']}');
	/* the same stream as a batching replay: nothing is decoded until the flush, the fsm is not in its
	 * handshaking state any more, and handshakeFrames says that the first frame is the connect frame */
	var fsm2 = { zcf_xidMap: sess.xidMap, isInState: function (st) { return (false); } };
	var bs = new BatchDecode({ fsm: fsm2, isServer: sess.isServer, batchBytes: 1073741824, handshakeFrames: 1 });
	var cbs = 0;
	for (var i2 = 0; i2 < sess.chunks.length; ++i2)
		bs._transform(hex(sess.chunks[i2]), null, function () { cbs++; });
	var early = bs._events.length;
	bs._flush(function () { });
	var out2 = [];
	for (var e2 = 0; e2 < bs._events.length; ++e2) {
		var ev2 = bs._events[e2];
		if (ev2.push !== undefined) out2.push('{"pkt":' + canon(ev2.push) + '}');
		else if (ev2.ev === 'error') out2.push('{"error":{"code":' + jstr(ev2.arg.code) + ',"message":' + jstr(ev2.arg.message) + '}}');
	}
	print('{"batched":true,"early":' + early + ',"cbs":' + cbs + ',"chunks":' + sess.chunks.length + ',"events":[' + out2.join(',') + ']}');
}
"""])
    driver = "var SESS = [" + ",".join(sess_js) + "];\n" + driver
    out = ref_js.run_js(driver)
    lines = [l for l in out.splitlines() if l.strip()]
    assert len(lines) == 2 * len(chosen)
    for line, line2, (s, _, _) in zip(lines[0::2], lines[1::2], chosen):
        got = ref_js._decanon(json.loads(line))
        assert got["opsOk"] is True
        assert canon_events(got["events"]) == s["events"], s["corpus"]
        # batchBytes > 0 with handshakeFrames: 1 — every write acknowledged, nothing pushed before the
        # flush, and then exactly what the reference's stream produced, connect frame included
        got2 = ref_js._decanon(json.loads(line2))
        assert got2["batched"] is True and got2["early"] == 0 and got2["cbs"] == got2["chunks"]
        assert canon_events(got2["events"]) == s["events"], s["corpus"]


def test_js_batch_encode_stream_with_recorded_addon(oracle, zk):
    """js/batch-streams.js ZKEncodeStream in the reference's engine: the packet goes through
    flatten.js, the recorded addon returns what the oracle encodes for the same descriptors, and the
    stream must push the reference's frame, record xid -> opcode in fsm.zcf_xidMap, or throw the
    reference's 'Unsupported opcode' error (tests/golden/ref_encode.json)."""
    from test_oracle_golden import unjson
    g = json.load(open(os.path.join(GOLD, "ref_encode.json")))["packets"]
    recs = []
    for p in g:
        pkt = unjson(p["pkt"])
        batch = zk.RequestBatch.from_packets([pkt], p["handshaking"], p["isServer"])
        data, fo, st = oracle.encode(batch)
        recs.append("{bytes:hex(%s),frame_off:%s,status:%s}" % (json.dumps(data.tobytes().hex()), js_array(fo), js_array(st)))
    src = {n: open(os.path.join(ROOT, "node-zkstream_b200", "js", n)).read()
           for n in ("materialise.js", "flatten.js", "batch-streams.js")}
    cases = ",".join("{hs:%s,server:%s,pkt:%s}" % ("true" if p["handshaking"] else "false",
                                                   "true" if p["isServer"] else "false",
                                                   ref_js.js_literal(unjson(p["pkt"]))) for p in g)
    driver = "\n".join([
        "var RECS = [" + ",".join(recs) + "]; var CASES = [" + cases + "]; var NEXT = 0;",
        "function Int8Array(n) { var a = []; for (var i = 0; i < n; ++i) a.push(0); return a; }",
        "Buffer.concat = function (list, total) { var h = ''; for (var i = 0; i < list.length; ++i) h += list[i].toString('hex'); return new Buffer(h, 'hex'); };",
        src["materialise.js"], src["flatten.js"],
        "var fakeAddon = { createContext: function () { return {}; }, encodeColumns: function (ctx, cols) { return RECS[NEXT]; } };",
        "DEFS['batch-streams'] = function (module, exports, require) {\n" + src["batch-streams.js"] + "\n};",
        r"""
function req2(name) {
	if (name === 'zkstream/lib/zk-consts') return requireShim('./zk-consts');
	if (name === 'zkstream/lib/errors') return requireShim('./errors');
	if (name === 'zkstream/lib/jute-buffer') return requireShim('./jute-buffer');
	if (name === './build/Release/zkbatch_napi.node') return fakeAddon;
	if (name === './materialise') return { makeMaterialiser: makeMaterialiser };
	if (name === './flatten') return { makeFlatten: makeFlatten };
	return requireShim(name);
}
var bm = { exports: {} };
DEFS['batch-streams'](bm, bm.exports, req2);
for (NEXT = 0; NEXT < CASES.length; ++NEXT) {
	var cs = CASES[NEXT];
	var fsm = { zcf_xidMap: {}, isInState: function (st) { return (st === 'handshaking' && cs.hs); } };
	var es = new bm.exports.ZKEncodeStream({ fsm: fsm, isServer: cs.server });
	var res;
	try {
		es._transform(cs.pkt, null, function () {});
		res = '{"frame":"' + es._pushed[0].toString('hex') + '"}';
	} catch (e) {
		res = '{"throw":' + jstr(String(e.name) + ': ' + String(e.message)) + '}';
	}
	var xm = [];
	for (var k in fsm.zcf_xidMap) xm.push(jstr(k) + ':' + jstr(String(fsm.zcf_xidMap[k])));
	print('{"result":' + res + ',"xidMap":{' + xm.join(',') + '}}');
}
"""])
    out = ref_js.run_js(driver)
    lines = [json.loads(l) for l in out.splitlines() if l.strip()]
    assert len(lines) == len(g)
    for got, p in zip(lines, g):
        assert got["result"] == p["result"], p["pkt"]
        assert got["xidMap"] == {k: v for k, v in p["xidMap"].items() if v is not None}, p["pkt"]
