"""Session-level folds: max zxid and notification counts per session.

The CPU oracle's zko_session_stats is pinned to what the reference's own ZKSession packet
handler (lib/zk-session.js:227-236, :389-419) made of the same bytes — tests/golden/
ref_session_folds.json, and live when the reference is present — and the CUDA reductions are
compared with the oracle bit for bit, then at bench scale with plain numpy segment reductions.
"""
import json
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(__file__), "golden", "ref_session_folds.json")
OPS = None


def _xidmap(zk, names):
    return {int(k): v for k, v in names.items()}


def _corpus_from_golden(synth, sessions):
    c = synth.Corpus()
    for s in sessions:
        c.add(bytes.fromhex(s["hex"]), hs=s["hs"], xidmap={int(k): v for k, v in s["xidMap"].items()})
    return c


def _expect_row(zk, ref):
    """(n_replies, max_zxid or None, notifications[6]) from one reference result."""
    inv = {v: k for k, v in zk.session.EVENT_NAMES.items()}
    row = [0] * zk._abi.NOTIFY_SLOTS
    row[zk._abi.NOTIFY_IGNORED] = ref["ignored"]
    row[zk._abi.NOTIFY_BAD_TYPE] = ref["threw"]
    for ev, n in ref["counts"].items():
        row[inv[ev]] = n
    mz = None if ref["maxZxid"] is None else int.from_bytes(bytes.fromhex(ref["maxZxid"]), "big", signed=True)
    return ref["replies"], mz, row


def _check_against_reference(zk, stats, refs):
    assert stats.nsessions == len(refs)
    for s, ref in enumerate(refs):
        nr, mz, row = _expect_row(zk, ref)
        assert int(stats.n_replies[s]) == nr, (s, ref)
        assert (mz is None) == (nr == 0)
        if mz is None:
            assert int(stats.max_zxid[s]) == zk.session.INT64_MIN
        else:
            assert int(stats.max_zxid[s]) == mz, (s, ref)
        assert [int(x) for x in stats.notifications[s]] == row, (s, ref)


def _golden():
    with open(GOLD) as f:
        return json.load(f)["sessions"]


def test_event_names_are_the_references(zk):
    # zk-session.js:402-405 applied to NOTIFICATION_TYPE (zk-consts.js:111-116); README's watcher events
    assert zk.session.EVENT_NAMES == {1: "created", 2: "deleted", 3: "dataChanged", 4: "childrenChanged"}


def test_oracle_folds_match_reference_golden(zk, oracle, synth):
    sessions = _golden()
    assert len(sessions) >= 80
    c = _corpus_from_golden(synth, sessions)
    d, off, fl, hs = c.arrays()
    cols = oracle.decode(d, off, fl, hs, c.xidtable())
    _check_against_reference(zk, oracle.session_stats(cols, len(sessions), 0), [s["all"] for s in sessions])
    _check_against_reference(zk, oracle.session_stats(cols, len(sessions), zk._abi.STATS_UNTIL_ERROR),
                             [s["until_error"] for s in sessions])
    # the fixture exercises every outcome
    tot = oracle.session_stats(cols, len(sessions), 0).notifications.sum(axis=0)
    assert (tot > 0).all()
    assert any(s["all"] != s["until_error"] for s in sessions)


def test_oracle_folds_match_live_reference(zk, oracle, synth):
    import ref_js
    if not ref_js.available():
        pytest.skip("reference JS engine not available here")
    from _canon import xidmap_names
    c = synth.session_corpus(101, nsessions=10, max_frames=40)
    sessions = [dict(chunks=[c.chunks[i]], hs=c.hs[i], xidMap=xidmap_names(c.xidmaps[i])) for i in range(len(c.chunks))]
    d, off, fl, hs = c.arrays()
    cols = oracle.decode(d, off, fl, hs, c.xidtable())
    for flag in (0, zk._abi.STATS_UNTIL_ERROR):
        refs = ref_js.session_folds(sessions, until_error=bool(flag))
        _check_against_reference(zk, oracle.session_stats(cols, len(sessions), flag), refs)


def test_session_state_fold(zk):
    st = zk.SessionStats.empty(2)
    st.n_replies[:] = [3, 0]
    st.max_zxid[0] = -7
    st.notifications[0, 3] = 2
    st.notifications[0, 0] = 9
    s = zk.SessionState().fold(st, 0)
    assert s.last_zxid == 0 and s.notifications == {"dataChanged": 2}      # zs_lastZxid starts at 0 (:57)
    st.max_zxid[0] = 41
    assert s.fold(st, 0).last_zxid == 41 and s.notifications == {"dataChanged": 4}
    assert zk.SessionState().fold(st, 1).last_zxid == 0


# ---- CUDA ------------------------------------------------------------------------------------

def _both_dirs(ctx, oracle, c, nsess):
    d, off, fl, hs = c.arrays()
    xt = c.xidtable()
    cols_o = oracle.decode(d, off, fl, hs, xt)
    cols_g = ctx.decode_host(d, off, fl, hs, xt)
    assert cols_g.diff(cols_o) == []
    return cols_g, cols_o


@pytest.mark.gpu
@pytest.mark.parametrize("flags", [0, 1])
def test_gpu_folds_match_oracle_and_golden(ctx, zk, oracle, synth, flags):
    sessions = _golden()
    c = _corpus_from_golden(synth, sessions)
    cols_g, cols_o = _both_dirs(ctx, oracle, c, len(sessions))
    got = ctx.session_stats_host(cols_g, flags=flags)
    assert got.diff(oracle.session_stats(cols_o, len(sessions), flags)) == []
    _check_against_reference(zk, got, [s["until_error" if flags else "all"] for s in sessions])


@pytest.mark.gpu
@pytest.mark.parametrize("seed", range(4))
def test_gpu_folds_fuzz(ctx, zk, oracle, synth, seed):
    c = synth.session_corpus(200 + seed, nsessions=300, max_frames=90)
    # mixed in: server-side and damaged sessions, which must contribute nothing / stop early
    f = synth.fuzz_corpus(300 + seed, nsessions=40, max_frames=30)
    for i in range(len(f.chunks)):
        c.add(f.chunks[i], server=bool(f.flags[i]), hs=f.hs[i], xidmap=f.xidmaps[i])
    n = len(c.chunks)
    cols_g, cols_o = _both_dirs(ctx, oracle, c, n)
    for flags in (0, 1):
        assert ctx.session_stats_host(cols_g, flags=flags).diff(oracle.session_stats(cols_o, n, flags)) == []


@pytest.mark.gpu
def test_gpu_folds_empty_and_bad_arguments(ctx, zk, oracle, synth):
    c = synth.Corpus()
    c.add(b"")
    c.add(b"")
    cols_g, cols_o = _both_dirs(ctx, oracle, c, 2)
    got = ctx.session_stats_host(cols_g)
    assert got.diff(oracle.session_stats(cols_o, 2, 0)) == []
    assert list(got.n_replies) == [0, 0] and int(got.max_zxid[0]) == zk.session.INT64_MIN
    with pytest.raises(zk.ZKBatchError):
        ctx.session_stats_host(cols_g, flags=8)


@pytest.mark.gpu
def test_gpu_folds_device_resident_at_scale(ctx, zk, synth):
    """~25 M frames: the device-resident entry point against numpy segment reductions."""
    import importlib
    import torch
    dev = importlib.import_module("node-zkstream_b200.device")
    pc = synth.native_cfg2(9, 64, 20000)
    n = pc.nsessions
    d = dev.to_device(pc.rsp, "cuda:0", pad=64)
    ds = dev.DeviceSessions(pc.rsp.size, pc.rsp_off, pc.rsp_flags, None, "cuda:0")
    dx = dev.DeviceXidTable(pc.xidtable(), "cuda:0")
    dc = dev.decode_dev_autosize(ctx, d, int(pc.rsp.size), ds, dx)
    got = dev.session_stats_dev(ctx, dc, n)
    cols = dc.to_host()
    st = cols.meta & 0xFF
    kind = (cols.meta >> 16) & 0xFF
    op = ((cols.meta >> 8) & 0xFF).astype(np.int8)
    ok = (st == 0) & (kind == 0)
    rep = ok & (op != 0)
    first = cols.sess_first_frame.astype(np.int64)
    assert (np.diff(first) > 0).all()
    z = np.where(rep, cols.zxid, np.iinfo(np.int64).min)
    assert np.array_equal(got.max_zxid, np.maximum.reduceat(z, first[:-1]))
    assert np.array_equal(got.n_replies, np.add.reduceat(rep.astype(np.int64), first[:-1]).astype(np.uint32))
    note = ok & (op == 0)
    slot = np.where(cols.a1 != 3, 0, np.where((cols.a0 >= 1) & (cols.a0 <= 4), cols.a0, 5))
    for k in range(zk._abi.NOTIFY_SLOTS):
        want = np.add.reduceat((note & (slot == k)).astype(np.int64), first[:-1]).astype(np.uint32)
        assert np.array_equal(got.notifications[:, k], want)
    assert int(got.n_replies.sum()) > 1_000_000
    del d, dc
    torch.cuda.empty_cache()
