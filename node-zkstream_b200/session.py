"""Session-level folds over decoded response columns.

The reference's ZKSession looks at every decoded packet of an attached connection
(lib/zk-session.js:227-236): anything that is not a NOTIFICATION advances zs_lastZxid if its
zxid is greater, a NOTIFICATION goes to processNotification (:389-419), which drops it unless
its state is SYNC_CONNECTED and otherwise counts it under its event name and wakes the
watcher of that path.  zkb_session_stats_* computes the reductions for a whole batch on the
GPU; this module holds the result type and the fold into reference-shaped session state.
"""
import ctypes as C

import numpy as np

from . import _abi, consts

INT64_MIN = -(1 << 63)

# NOTIFICATION_TYPE value -> the event name ZKSession.processNotification derives from it
# (zk-session.js:402-405: lower-case, '_x' -> 'X'); the same names watchers are notified with
EVENT_NAMES = {}
for _v, _name in consts.NOTIFICATION_TYPE_LOOKUP.items():
    _parts = _name.lower().split("_")
    EVENT_NAMES[_v] = _parts[0] + "".join(p[:1].upper() + p[1:] for p in _parts[1:])


class SessionStats:
    """max_zxid int64[n], n_replies uint32[n], notifications uint32[n, NOTIFY_SLOTS]."""

    def __init__(self, max_zxid, n_replies, notifications):
        self.max_zxid = max_zxid
        self.n_replies = n_replies
        self.notifications = notifications

    @classmethod
    def empty(cls, nsessions):
        return cls(np.full(nsessions, INT64_MIN, dtype=np.int64), np.zeros(nsessions, dtype=np.uint32),
                   np.zeros((nsessions, _abi.NOTIFY_SLOTS), dtype=np.uint32))

    @property
    def nsessions(self):
        return len(self.n_replies)

    def struct(self):
        st = _abi.SessionStats()
        st.max_zxid = self.max_zxid.ctypes.data if self.max_zxid.size else None
        st.n_replies = self.n_replies.ctypes.data if self.n_replies.size else None
        st.notifications = self.notifications.ctypes.data if self.notifications.size else None
        return st

    def diff(self, other):
        out = []
        for name in ("max_zxid", "n_replies", "notifications"):
            a, b = getattr(self, name), getattr(other, name)
            if a.shape != b.shape or not np.array_equal(a, b):
                out.append(name)
        return out

    def event_counts(self, s):
        """{event name: count} for session s, as the zookeeper_notifications counter would hold
        (zk-session.js:411-415); ignored and untyped notifications are not in it."""
        row = self.notifications[s]
        return {EVENT_NAMES[t]: int(row[t]) for t in sorted(EVENT_NAMES) if row[t]}


class SessionState:
    """The two pieces of ZKSession state the packet handler maintains: zs_lastZxid (starts at
    0, zk-session.js:57) and the per-event notification counter.  fold() applies one batch's
    reductions exactly as handling those packets one by one would have."""

    def __init__(self):
        self.last_zxid = 0
        self.notifications = {}

    def fold(self, stats, s):
        if stats.n_replies[s] and int(stats.max_zxid[s]) > self.last_zxid:
            self.last_zxid = int(stats.max_zxid[s])
        for ev, n in stats.event_counts(s).items():
            self.notifications[ev] = self.notifications.get(ev, 0) + n
        return self
