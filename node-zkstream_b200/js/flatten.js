/*
 * flatten.js — packets -> the struct-of-arrays encode descriptors of include/zkbatch.h
 * (zkb_requests).  JavaScript twin of node-zkstream_b200/encode.py: it performs the
 * argument checks the reference's writers make with assert-plus (lib/zk-buffer.js:143-273,
 * 384-426) and leaves the byte layout to the GPU.  ES3, plain arrays out; the addon wraps
 * them in typed arrays.
 */
function makeFlatten(deps) {
	var consts = deps.consts, assert = deps.assert, BufferCtor = deps.Buffer;
	var ENC_CONNECT_REQUEST = -100, ENC_CONNECT_RESPONSE = -101;
	var WRITABLE = { GET_CHILDREN: 1, GET_CHILDREN2: 1, CREATE: 1, DELETE: 1, GET_ACL: 1, SET_WATCHES: 1,
	    GET_DATA: 1, EXISTS: 1, SET_DATA: 1, SYNC: 1, CLOSE_SESSION: 1, PING: 1 };

	/* writeLong's argument (lib/jute-buffer.js:149-165) -> [hi, lo] int32 words, left-padded with zeros */
	function long2(v) {
		if (!BufferCtor.isBuffer(v)) {
			if (v && v.ld_buffer !== undefined) v = v.toBuffer();
			else v = new BufferCtor(v.toByteArray());
		}
		assert.ok(v.length <= 8);
		var b = new BufferCtor(8);
		b.fill(0);
		v.copy(b, 8 - v.length);
		return ([b.readInt32BE(0), b.readInt32BE(4)]);
	}

	return (function flatten(pkts, handshaking, isServer) {
		var n = pkts.length, i;
		var c = { n: n, xid: [], op: [], path_off: [], path_len: [], data_off: [], data_len: [], a0: [], a1: [],
		    zxid: [], zxid2: [], vec_first: [], vec_count: [], n0: [], n1: [],
		    elem_off: [], elem_len: [], acl_perms: [], acl_scheme_off: [], acl_scheme_len: [],
		    acl_id_off: [], acl_id_len: [] };
		var arena = [], alen = 0;
		function put(buf) { var o = alen; arena.push(buf); alen += buf.length; return ([o, buf.length]); }
		function ustr(v, what) { assert.string(v, what); return (put(new BufferCtor(v, 'utf-8'))); }
		function buf(v, what) { assert.buffer(v, what); return (put(v)); }
		for (i = 0; i < n; ++i) {
			var pkt = pkts[i], r;
			var hs = (typeof (handshaking) === 'boolean') ? handshaking : handshaking[i];
			assert.object(pkt, 'packet');
			var row = { xid: 0, op: 0, path: [0, 0], data: [0, 0], a0: 0, a1: 0, zxid: [0, 0], zxid2: [0, 0],
			    vec_first: 0, vec_count: 0, n0: 0, n1: 0 };
			if (hs) {
				row.op = isServer ? ENC_CONNECT_RESPONSE : ENC_CONNECT_REQUEST;
				assert.number(pkt.protocolVersion, 'protocolVersion');
				assert.number(pkt.timeOut, 'timeOut');
				row.a0 = pkt.protocolVersion;
				row.a1 = pkt.timeOut;
				row.zxid = long2(pkt.sessionId);
				if (!isServer) row.zxid2 = long2(pkt.lastZxidSeen);
				row.data = buf(pkt.passwd, 'passwd');
			} else {
				assert.number(pkt.xid, 'xid');
				assert.optionalString(pkt.opcode, 'opcode');
				row.xid = pkt.xid;
				var code = consts.OP_CODES[pkt.opcode];
				if (!WRITABLE[pkt.opcode]) {
					row.op = (code !== undefined && code > -100 && code < 128) ? code : 127;
				} else {
					row.op = code;
					switch (pkt.opcode) {
					case 'GET_CHILDREN': case 'GET_CHILDREN2': case 'GET_DATA': case 'EXISTS':
						row.path = ustr(pkt.path, 'path');
						assert.bool(pkt.watch, 'watch');
						row.a0 = pkt.watch ? 1 : 0;
						break;
					case 'CREATE':
						row.path = ustr(pkt.path, 'path');
						row.data = buf(pkt.data, 'data');
						assert.arrayOfObject(pkt.acl, 'acl');
						row.vec_first = c.acl_perms.length;
						row.vec_count = pkt.acl.length;
						pkt.acl.forEach(function (line) {
							assert.arrayOfString(line.perms, 'permissions');
							var val = 0;
							line.perms.forEach(function (k) {
								var mask = consts.PERM_MASKS[k.toUpperCase()];
								assert.number(mask, 'permission mask for ' + k);
								val |= mask;
							});
							var s = ustr(line.id.scheme, 'id.scheme'), d = ustr(line.id.id, 'id');
							c.acl_perms.push(val);
							c.acl_scheme_off.push(s[0]); c.acl_scheme_len.push(s[1]);
							c.acl_id_off.push(d[0]); c.acl_id_len.push(d[1]);
						});
						assert.arrayOfString(pkt.flags, 'flags');
						pkt.flags.forEach(function (k) {
							var mask = consts.CREATE_FLAGS[k];
							assert.number(mask, 'unknown flag ' + k);
							row.a0 |= mask;
						});
						break;
					case 'DELETE':
						row.path = ustr(pkt.path, 'name');
						assert.number(pkt.version, 'version');
						row.a0 = pkt.version;
						break;
					case 'GET_ACL': case 'SYNC':
						row.path = ustr(pkt.path, 'path');
						break;
					case 'SET_DATA':
						row.path = ustr(pkt.path, 'path');
						row.data = buf(pkt.data, 'data');
						assert.number(pkt.version, 'version');
						row.a0 = pkt.version;
						break;
					case 'SET_WATCHES':
						row.zxid = long2(pkt.relZxid);
						row.vec_first = c.elem_off.length;
						var counts = [];
						['dataChanged', 'createdOrDestroyed', 'childrenChanged'].forEach(function (key) {
							var lst = pkt.events[key] || [];
							counts.push(lst.length);
							lst.forEach(function (p, j) {
								var e = ustr(p, key + '[' + j + ']');
								c.elem_off.push(e[0]); c.elem_len.push(e[1]);
							});
						});
						row.n0 = counts[0]; row.n1 = counts[1];
						row.vec_count = counts[0] + counts[1] + counts[2];
						break;
					default:
						break;
					}
				}
			}
			c.xid.push(row.xid); c.op.push(row.op);
			c.path_off.push(row.path[0]); c.path_len.push(row.path[1]);
			c.data_off.push(row.data[0]); c.data_len.push(row.data[1]);
			c.a0.push(row.a0); c.a1.push(row.a1);
			c.zxid.push(row.zxid[1], row.zxid[0]);      /* [lo, hi] per packet: Int32 view of int64 */
			c.zxid2.push(row.zxid2[1], row.zxid2[0]);
			c.vec_first.push(row.vec_first); c.vec_count.push(row.vec_count);
			c.n0.push(row.n0); c.n1.push(row.n1);
		}
		c.arena = arena;
		c.arena_len = alen;
		return (c);
	});
}

if (typeof (module) !== 'undefined')
	module.exports = { makeFlatten: makeFlatten };
