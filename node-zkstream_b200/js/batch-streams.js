/*
 * batch-streams.js — drop-in ZKDecodeStream / ZKEncodeStream for node-zkstream that decode
 * and encode in batches on the GPU through the N-API addon (zkbatch_napi.c -> libzkbatch.so).
 *
 * Same constructor options and the same observable behaviour as lib/zk-streams.js:25-148:
 * Buffers in / packet objects out (and the reverse), 'error' events carrying
 * ZKProtocolError BAD_LENGTH / BAD_DECODE, fsm.zcf_xidMap updated per encoded request.
 * `require` this file from lib/connection-fsm.js:86-91 instead of './zk-streams' to switch
 * a client over; pass {batchBytes: n} to collect n bytes before each GPU call when
 * replaying captures (0, the default, decodes every chunk as it arrives, like the
 * reference).  The reference asks fsm.isInState('handshaking') once per FRAME
 * (lib/zk-streams.js:68); one GPU call covers many frames, so a batching replay says how many
 * leading frames of the stream are handshake frames with {handshakeFrames: n} (1 for a
 * connection captured from its start) instead of being asked per call.  With batchBytes > 0 the
 * xid map is read when the batch is decoded, not per frame: a capture that re-uses an xid inside
 * one batch resolves it to the later opcode (cut such captures into quiescent segments, see
 * include/zkcapture.h).
 *
 * Needs a node with N-API (>= 8.x) and the built addon; neither exists in the build image,
 * so both streams are exercised in the reference's own JavaScript engine against a recorded
 * addon (tests/test_js_shim.py).
 */
var mod_stream = require('stream');
var mod_util = require('util');
var mod_assert = require('assert-plus');
var mod_consts = require('zkstream/lib/zk-consts');
var mod_errors = require('zkstream/lib/errors');
var mod_jute = require('zkstream/lib/jute-buffer');
var addon = require('./build/Release/zkbatch_napi.node');
var makeMaterialiser = require('./materialise').makeMaterialiser;
var flatten = require('./flatten').makeFlatten({ consts: mod_consts, assert: mod_assert, Buffer: Buffer });

/* LongDate is not exported by jute-buffer.js: take its constructor from an instance */
var LongDate = (new mod_jute.JuteBuffer({ buffer: new Buffer(8) })).readLongDate().constructor;
var Materialiser = makeMaterialiser({ consts: mod_consts, LongDate: LongDate,
    ZKProtocolError: mod_errors.ZKProtocolError, Buffer: Buffer });

var ctx = null;
function context() {
	if (ctx === null)
		ctx = addon.createContext(0);    /* throws when no CUDA device is usable: no CPU path */
	return (ctx);
}

/* fsm.zcf_xidMap ({xid: opcode name}) -> dense Int8Array starting at xid 0.  The map only ever
 * grows (the reference never deletes from it), so the table is kept between calls and only keys
 * that are new, or whose opcode changed, are written; it is rebuilt when the map object changes. */
function XidTable() {
	this.map = null;
	this.seen = {};
	this.nseen = 0;
	this.ops = new Int8Array(0);
}
XidTable.prototype.update = function (map) {
	var self = this;
	if (map !== this.map) { this.map = map; this.seen = {}; this.nseen = 0; this.ops = new Int8Array(0); }
	var keys = Object.keys(map);
	var max = this.ops.length - 1;
	keys.forEach(function (k) { k = parseInt(k, 10); if (k > max) max = k; });
	if (max + 1 > this.ops.length) {
		var cap = Math.max(max + 1, this.ops.length * 2);
		var grown = new Int8Array(cap);
		for (var i = 0; i < cap; ++i) grown[i] = (i < this.ops.length) ? this.ops[i] : -128;
		this.ops = grown;
	}
	keys.forEach(function (k) {
		if (self.seen[k] === map[k]) return;
		self.seen[k] = map[k];
		var code = mod_consts.OP_CODES[map[k]];
		var x = parseInt(k, 10);
		if (x >= 0) self.ops[x] = (code !== undefined) ? code : -128;
	});
	return (this.ops);
};

function ZKDecodeStream(opts) {
	mod_assert.object(opts, 'options');
	mod_assert.object(opts.fsm, 'options.fsm');
	mod_assert.optionalBool(opts.isServer, 'options.isServer');
	this.zd_fsm = opts.fsm;
	this.zd_server = (opts.isServer === true);
	this.zd_pending = [];
	this.zd_used = 0;
	this.zd_batch = opts.batchBytes || 0;
	mod_assert.optionalNumber(opts.handshakeFrames, 'options.handshakeFrames');
	this.zd_hsFrames = (opts.handshakeFrames === undefined) ? null : opts.handshakeFrames;
	this.zd_xids = new XidTable();
	opts.readableObjectMode = true;
	opts.writableObjectMode = false;
	mod_stream.Transform.call(this, opts);
}
mod_util.inherits(ZKDecodeStream, mod_stream.Transform);

ZKDecodeStream.prototype._transform = function (chunk, enc, cb) {
	mod_assert.buffer(chunk);
	this.zd_pending.push(chunk);
	this.zd_used += chunk.length;
	if (this.zd_used < Math.max(this.zd_batch, 1)) {
		cb();
		return;
	}
	if (this.decodePending())
		cb();
	/* after BAD_LENGTH the reference never calls cb(): the stream stays stuck */
};

ZKDecodeStream.prototype._flush = function (cb) {
	if (this.zd_used > 0)
		this.decodePending();
	cb();
};

ZKDecodeStream.prototype.decodePending = function () {
	var self = this;
	var data = Buffer.concat(this.zd_pending, this.zd_used);
	/* handshake depth: as many leading frames as the caller said are still owed, else what the fsm
	 * says now for the whole call (exact when every call holds one frame, as on a live socket) */
	var hs = (this.zd_hsFrames !== null) ? this.zd_hsFrames :
	    (this.zd_fsm.isInState('handshaking') ? 0xffffffff : 0);
	var cols = addon.decodeHost(context(), data, this.zd_server ? 1 : 0, hs,
	    this.zd_server ? null : this.zd_xids.update(this.zd_fsm.zcf_xidMap));
	if (this.zd_hsFrames !== null)
		this.zd_hsFrames = Math.max(0, this.zd_hsFrames - cols.n_frames);
	var ok = true;
	new Materialiser(cols, data).forEach(0, function (pkt) {
		self.push(pkt);
	}, function (err) {
		if (err.code === 'BAD_LENGTH') ok = false;
		self.emit('error', err);
	});
	var rest = data.slice(cols.consumed);
	this.zd_pending = rest.length ? [rest] : [];
	this.zd_used = rest.length;
	return (ok);
};

function ZKEncodeStream(opts) {
	mod_assert.object(opts, 'options');
	mod_assert.object(opts.fsm, 'options.fsm');
	mod_assert.optionalBool(opts.isServer, 'options.isServer');
	this.ze_fsm = opts.fsm;
	this.ze_server = (opts.isServer === true);
	opts.readableObjectMode = false;
	opts.writableObjectMode = true;
	mod_stream.Transform.call(this, opts);
}
mod_util.inherits(ZKEncodeStream, mod_stream.Transform);

ZKEncodeStream.prototype._transform = function (pkt, enc, cb) {
	mod_assert.object(pkt);
	var hs = this.ze_fsm.isInState('handshaking');
	if (!hs) {
		mod_assert.number(pkt.xid, 'xid');
		mod_assert.optionalString(pkt.opcode, 'opcode');
	}
	/* field checks as in the reference's writers, then one GPU call for the batch */
	var cols = flatten([pkt], hs, this.ze_server);
	cols.arena = Buffer.concat(cols.arena, cols.arena_len);
	var res = addon.encodeColumns(context(), cols);     /* { bytes: Buffer, frame_off: [...], status: [...] } */
	if (res.status[0] !== 0)
		throw (new Error('Unsupported opcode ' + pkt.opcode));
	this.push(res.bytes.slice(res.frame_off[0], res.frame_off[1]));
	if (!hs)
		this.ze_fsm.zcf_xidMap[pkt.xid] = pkt.opcode;
	cb();
};

module.exports = { ZKDecodeStream: ZKDecodeStream, ZKEncodeStream: ZKEncodeStream };
