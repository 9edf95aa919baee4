/*
 * materialise.js — columnar decode result -> the packet objects node-zkstream's
 * ZKDecodeStream pushes (lib/zk-streams.js:81,97).
 *
 * Runs on the JavaScript side of the N-API addon (zkbatch_napi.c).  `cols` holds the typed
 * arrays of include/zkbatch.h's zkb_decoded; every 64-bit column arrives as an Int32Array
 * over the same memory (little endian: [2*i] = low word, [2*i+1] = high word), so no BigInt
 * is needed and old runtimes work.  `data` is the Buffer that was decoded: payload fields
 * are slices of it, exactly as the reference's fields are slices of the frame buffer
 * (lib/jute-buffer.js:66,102).
 *
 * Written in ES3 so that it also loads under the SpiderMonkey 1.x shell the tests use to
 * compare it with the reference (tests/test_js_shim.py).
 */
var ST_OK = 0;
var KIND_RESPONSE = 0, KIND_REQUEST = 1, KIND_CONNECT_RESPONSE = 2, KIND_CONNECT_REQUEST = 3;
var CAUSE_RANGE = 1, CAUSE_ASSERT = 2, CAUSE_NO_MATCH = 3, CAUSE_UNSUPPORTED = 4;

function makeMaterialiser(deps) {
	var consts = deps.consts;          /* the reference's lib/zk-consts.js */
	var LongDate = deps.LongDate;      /* function (buffer) -> LongDate, lib/jute-buffer.js:79 */
	var ZKProtocolError = deps.ZKProtocolError;
	var BufferCtor = deps.Buffer;

	function long8(col32, i) {
		var b = new BufferCtor(8);
		b.writeInt32BE(col32[2 * i + 1], 0);
		b.writeInt32BE(col32[2 * i], 4);
		return (b);
	}
	function long8parts(hi, lo) {
		var b = new BufferCtor(8);
		b.writeInt32BE(hi, 0);
		b.writeInt32BE(lo, 4);
		return (b);
	}
	/* 64-bit offsets stay below 2^53 for any batch the codec accepts (< 16 GiB) */
	function off64(col32, i) {
		return ((col32[2 * i + 1] >>> 0) * 4294967296 + (col32[2 * i] >>> 0));
	}

	function Materialiser(cols, data) {
		this.c = cols;
		this.data = data;
	}
	Materialiser.prototype.bytes = function (row) {
		var o = off64(this.c.buf_off, row);
		return (this.data.slice(o, o + this.c.buf_len[row]));
	};
	Materialiser.prototype.str = function (row) {
		return (this.bytes(row).toString('utf-8'));
	};
	Materialiser.prototype.stat = function (r) {
		var c = this.c;
		return ({
			czxid: long8(c.czxid, r), mzxid: long8(c.mzxid, r),
			ctime: new LongDate(long8(c.ctime, r)), mtime: new LongDate(long8(c.mtime, r)),
			version: c.version[r], cversion: c.cversion[r], aversion: c.aversion[r],
			ephemeralOwner: long8(c.ephemeralOwner, r),
			dataLength: c.dataLength[r], numChildren: c.numChildren[r], pzxid: long8(c.pzxid, r)
		});
	};
	/* readPerms exactly as written, operator precedence included (lib/zk-buffer.js:395-403) */
	function perms(val) {
		var out = [];
		Object.keys(consts.PERM_MASKS).forEach(function (k) {
			if (val & consts.PERM_MASKS[k] != 0)
				out.push(k);
		});
		return (out);
	}
	Materialiser.prototype.acl = function (n, b, a) {
		var out = [];
		for (var i = 0; i < n; ++i) {
			out.push({ perms: perms(this.c.acl_perms[a + i]),
			    id: { scheme: this.str(b + 2 * i), id: this.str(b + 2 * i + 1) } });
		}
		return (out);
	};
	Materialiser.prototype.strs = function (b, n) {
		var out = [];
		for (var i = 0; i < n; ++i)
			out.push(this.str(b + i));
		return (out);
	};

	/* the packet of FRAMES row f; s/b/a = first STAT/BUF/ACL row the frame owns */
	Materialiser.prototype.packet = function (f, s, b, a) {
		var c = this.c, self = this;
		var m = c.meta[f];
		var kind = (m >>> 16) & 0xff;
		var op = (m >>> 8) & 0xff;
		if (op >= 128) op -= 256;
		var a0 = c.a0[f], a1 = c.a1[f];
		var pkt = {};
		if (kind === KIND_CONNECT_RESPONSE) {
			pkt.protocolVersion = c.xid[f];
			pkt.timeOut = c.err[f];
			pkt.sessionId = long8(c.zxid, f);
			pkt.passwd = this.bytes(b);
			return (pkt);
		}
		if (kind === KIND_CONNECT_REQUEST) {
			pkt.protocolVersion = c.xid[f];
			pkt.lastZxidSeen = long8parts(a0, a1);
			pkt.timeOut = c.err[f];
			pkt.sessionId = long8(c.zxid, f);
			pkt.passwd = this.bytes(b);
			return (pkt);
		}
		var name = consts.OP_CODE_LOOKUP[op];
		if (kind === KIND_REQUEST) {
			pkt.xid = c.xid[f];
			pkt.opcode = name;
			switch (name) {
			case 'GET_CHILDREN': case 'GET_CHILDREN2': case 'GET_DATA': case 'EXISTS':
				pkt.path = this.str(b);
				pkt.watch = (a0 === 1);
				break;
			case 'CREATE':
				pkt.path = this.str(b);
				pkt.data = this.bytes(b + 1);
				pkt.acl = this.acl(a1, b + 2, a);
				pkt.flags = [];
				Object.keys(consts.CREATE_FLAGS).forEach(function (k) {
					var mask = consts.CREATE_FLAGS[k];
					if ((a0 & mask) === mask)
						pkt.flags.push(k);
				});
				break;
			case 'DELETE':
				pkt.path = this.str(b);
				pkt.version = a0;
				break;
			case 'GET_ACL': case 'SYNC':
				pkt.path = this.str(b);
				break;
			case 'SET_DATA':
				pkt.path = this.str(b);
				pkt.data = this.bytes(b + 1);
				pkt.version = a0;
				break;
			case 'SET_WATCHES':
				pkt.relZxid = long8(c.zxid, f);
				pkt.events = {};
				pkt.events.dataChanged = this.strs(b, a0);
				pkt.events.createdOrDestroyed = this.strs(b + a0, a1);
				pkt.events.childrenChanged = this.strs(b + a0 + a1, c.err[f]);
				break;
			default:
				break;
			}
			return (pkt);
		}
		pkt.xid = c.xid[f];
		pkt.zxid = long8(c.zxid, f);
		pkt.err = consts.ERR_LOOKUP[c.err[f]];
		pkt.opcode = name;
		if (pkt.err !== 'OK')
			return (pkt);
		switch (name) {
		case 'GET_CHILDREN': case 'GET_CHILDREN2':
			pkt.children = this.strs(b, a0);
			if (name === 'GET_CHILDREN2')
				pkt.stat = this.stat(s);
			break;
		case 'CREATE':
			pkt.path = this.str(b);
			break;
		case 'GET_ACL':
			pkt.acl = this.acl(a0, b, a);
			pkt.stat = this.stat(s);
			break;
		case 'GET_DATA':
			pkt.data = this.bytes(b);
			pkt.stat = this.stat(s);
			break;
		case 'NOTIFICATION':
			pkt.type = consts.NOTIFICATION_TYPE_LOOKUP[a0];
			pkt.state = consts.STATE_LOOKUP[a1];
			pkt.path = this.str(b);
			break;
		case 'EXISTS': case 'SET_DATA':
			pkt.stat = this.stat(s);
			break;
		default:
			break;
		}
		return (pkt);
	};

	/* 'Failed to decode ...: ' + e.name + ': ' + e.message (lib/zk-streams.js:74-79,90-95) */
	function decodeError(m) {
		var kind = (m >>> 16) & 0xff;
		var op = (m >>> 8) & 0xff;
		if (op >= 128) op -= 256;
		var cause = (m >>> 24) & 7;
		var txt;
		if (cause === CAUSE_RANGE) txt = 'RangeError: index out of range';
		else if (cause === CAUSE_ASSERT) txt = 'AssertionError: false == true';
		else if (cause === CAUSE_NO_MATCH) txt = 'AssertionError: reply packet must match a request';
		else txt = 'Error: Unsupported opcode ' + consts.OP_CODE_LOOKUP[op];
		var what = (kind === KIND_CONNECT_REQUEST || kind === KIND_CONNECT_RESPONSE) ?
		    'ConnectResponse' : 'Response';
		return (new ZKProtocolError('BAD_DECODE', 'Failed to decode ' + what + ': ' + txt));
	}

	/* In stream order for session s: onPacket(pkt) for pushed packets, onError(err) for
	 * BAD_DECODE frames and for the BAD_LENGTH that ends the session. */
	Materialiser.prototype.forEach = function (sess, onPacket, onError) {
		var c = this.c;
		var f0 = c.sess_first_frame[sess], f1 = c.sess_first_frame[sess + 1];
		var s = lowerBound(c.stat_frame, c.n_stat, f0);
		var b = lowerBound(c.buf_frame, c.n_buf, f0);
		var a = lowerBound(c.acl_frame, c.n_acl, f0);
		for (var f = f0; f < f1; ++f) {
			var m = c.meta[f];
			if ((m & 0xff) === ST_OK) {
				onPacket(this.packet(f, s, b, a));
				while (s < c.n_stat && c.stat_frame[s] === f) ++s;
				while (b < c.n_buf && c.buf_frame[b] === f) ++b;
				while (a < c.n_acl && c.acl_frame[a] === f) ++a;
			} else {
				onError(decodeError(m));
			}
		}
		var bl = c.sess_bad_length_at;
		if (!(bl[2 * sess] === -1 && bl[2 * sess + 1] === -1))
			onError(new ZKProtocolError('BAD_LENGTH', 'Invalid ZK packet length'));
	};

	function lowerBound(arr, n, key) {
		var lo = 0, hi = n;
		while (lo < hi) {
			var mid = (lo + hi) >>> 1;
			if (arr[mid] < key) lo = mid + 1; else hi = mid;
		}
		return (lo);
	}

	return (Materialiser);
}

if (typeof (module) !== 'undefined')
	module.exports = { makeMaterialiser: makeMaterialiser };
