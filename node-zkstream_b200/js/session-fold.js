/*
 * session-fold.js — applies one batch's per-session reductions (zkb_session_stats_host /
 * _dev, include/zkbatch.h) to a ZKSession, with the outcome its own 'packet' handler reaches
 * by looking at the packets one at a time (lib/zk-session.js:227-236 and
 * processNotification, :389-419): zs_lastZxid moves up to the greatest reply zxid, and the
 * zookeeper_notifications counter grows by the number of notifications of each event.
 *
 * Waking watchers (watcher.notify, :417-418) needs each notification's path and stays with
 * the per-packet route: hand the materialised NOTIFICATION packets to processNotification as
 * before when the session has watchers.  This fold is for replaying captures, where only the
 * session state and the metrics are wanted.
 *
 *   var fold = makeSessionFold({ BigInteger: require('jsbn').BigInteger,
 *       consts: require('zkstream/lib/zk-consts') });
 *   fold(session, stats, s);
 *
 * stats = { n_replies: Uint32Array[n], max_zxid: Int32Array[2n] (the int64 column viewed as
 * little-endian word pairs), notifications: Uint32Array[6n] }.
 */
function makeSessionFold(deps) {
	var SLOTS = 6;
	var names = {};
	var lookup = deps.consts.NOTIFICATION_TYPE_LOOKUP;
	for (var v in lookup) {
		if (lookup.hasOwnProperty(v)) {
			/* the expression at zk-session.js:402-405 */
			names[v] = lookup[v].toLowerCase().replace(/_[a-z]/g, function (s) {
				return (s.slice(1).toUpperCase());
			});
		}
	}

	function zxidBytes(stats, s) {
		var lo = stats.max_zxid[2 * s], hi = stats.max_zxid[2 * s + 1], b = [], i;
		for (i = 3; i >= 0; --i) { b[i] = hi & 0xff; hi >>= 8; }
		for (i = 7; i >= 4; --i) { b[i] = lo & 0xff; lo >>= 8; }
		/* jsbn reads a byte array as big-endian two's complement with signed bytes */
		for (i = 0; i < 8; ++i) if (b[i] > 127) b[i] -= 256;
		return (b);
	}

	return (function fold(session, stats, s) {
		if (stats.n_replies[s] > 0) {
			var zxid = new deps.BigInteger(zxidBytes(stats, s));
			if (session.zs_lastZxid === undefined ||
			    zxid.compareTo(session.zs_lastZxid) > 0) {
				session.zs_lastZxid = zxid;
			}
		}
		var counter = null;
		for (var t = 1; t <= 4; ++t) {
			var n = stats.notifications[s * SLOTS + t];
			if (n === 0)
				continue;
			if (counter === null)
				counter = session.zs_collector.getCollector('zookeeper_notifications');
			if (typeof (counter.add) === 'function') {
				counter.add(n, { event: names[t] });
			} else {
				for (var i = 0; i < n; ++i)
					counter.increment({ event: names[t] });
			}
		}
		return (session);
	});
}

if (typeof (module) !== 'undefined')
	module.exports = { makeSessionFold: makeSessionFold };
