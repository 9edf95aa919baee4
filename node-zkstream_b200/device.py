"""Device-resident use of the C ABI: inputs and output columns live in HBM (torch tensors
are used purely as device allocations; every pointer handed to libzkbatch.so is raw).

This is what bench.py times for the device-resident number and what the GPU tests use to
exercise zkb_decode_dev / zkb_xidtab_build_dev / zkb_encode_dev.
"""
import ctypes as C

import numpy as np
import torch

from . import _abi
from ._lib import ZKBatchError
from .columns import Columns

_TORCH_DT = {"u1": torch.uint8, "i1": torch.int8, "i4": torch.int32, "u4": torch.int32,
             "i8": torch.int64, "u8": torch.int64}
_NP_DT = {"u1": np.uint8, "i1": np.int8, "i4": np.int32, "u4": np.uint32, "i8": np.int64, "u8": np.uint64}


def _dev_empty(n, dt, device):
    return torch.empty(max(int(n), 1), dtype=_TORCH_DT[dt], device=device)


def to_device(a, device, pad=0):
    """numpy array -> device tensor of the same bytes (+ `pad` spare bytes for uint8 inputs)."""
    a = np.ascontiguousarray(a)
    raw = torch.from_numpy(a.view(np.uint8).reshape(-1))
    t = torch.empty(raw.numel() + pad + 16, dtype=torch.uint8, device=device)
    t[: raw.numel()].copy_(raw)
    return t


class DeviceSessions:
    def __init__(self, nbytes, sess_off, flags=None, hs=None, device="cuda:0"):
        off = np.ascontiguousarray(sess_off, dtype=np.uint64)
        self.n = len(off) - 1
        self.off = to_device(off, device)
        self.flags = to_device(np.ascontiguousarray(flags, dtype=np.uint8), device) if flags is not None else None
        self.hs = to_device(np.ascontiguousarray(hs, dtype=np.uint32), device) if hs is not None else None
        s = _abi.Sessions()
        s.nsessions = self.n
        s.off = self.off.data_ptr()
        s.flags = self.flags.data_ptr() if self.flags is not None else None
        s.handshake_frames = self.hs.data_ptr() if self.hs is not None else None
        self.struct = s


class DeviceXidTable:
    def __init__(self, xt, device="cuda:0"):
        """xt: _lib.XidTable (host).  ops may be overwritten on device by build()."""
        self.host = xt
        self.base = to_device(xt.base, device)
        self.count = to_device(xt.count, device)
        self.off = to_device(xt.off, device)
        self.ops = to_device(xt.ops, device)
        t = _abi.XidTab()
        t.base, t.count, t.off, t.ops = (self.base.data_ptr(), self.count.data_ptr(), self.off.data_ptr(),
                                         self.ops.data_ptr())
        t.nops = xt.nops
        self.struct = t


class DeviceColumns:
    """Output columns of one decode call, allocated in HBM with given capacities."""

    def __init__(self, nsess, cap_frames, cap_stat, cap_buf, cap_acl, device="cuda:0"):
        self.nsess = nsess
        self.device = device
        self.t = {}
        d = _abi.Decoded()
        d.cap_frames, d.cap_stat, d.cap_buf, d.cap_acl = cap_frames, cap_stat, cap_buf, cap_acl
        for cap, cols in ((cap_frames, _abi.FRAME_COLS), (cap_stat, _abi.STAT_COLS),
                          (cap_buf, _abi.BUF_COLS), (cap_acl, _abi.ACL_COLS)):
            for name, dt in cols:
                self.t[name] = _dev_empty(cap, dt, device)
                setattr(d, name, self.t[name].data_ptr())
        for name, dt in _abi.SESS_COLS:
            n = nsess + 1 if name == "sess_first_frame" else nsess
            self.t[name] = _dev_empty(n, dt, device)
            setattr(d, name, self.t[name].data_ptr())
        self.struct = d

    def nbytes_capacity(self):
        return sum(t.numel() * t.element_size() for t in self.t.values())

    def to_host(self):
        """Columns (numpy) with the rows the last decode produced."""
        d = self.struct
        c = Columns()
        c.nsessions = self.nsess
        for cnt, cols in (("n_frames", _abi.FRAME_COLS), ("n_stat", _abi.STAT_COLS),
                          ("n_buf", _abi.BUF_COLS), ("n_acl", _abi.ACL_COLS)):
            n = int(getattr(d, cnt))
            setattr(c, cnt, n)
            for name, dt in cols:
                setattr(c, name, self.t[name][:n].cpu().numpy().view(_NP_DT[dt]).copy())
        for name, dt in _abi.SESS_COLS:
            n = self.nsess + 1 if name == "sess_first_frame" else self.nsess
            setattr(c, name, self.t[name][:n].cpu().numpy().view(_NP_DT[dt]).copy())
        return c


def decode_dev(ctx, d_bytes, nbytes, dsess, dxid, dcols):
    """zkb_decode_dev; raises on anything but OK / EOVERFLOW, returns the rc."""
    rc = ctx.decode_dev(d_bytes.data_ptr(), nbytes, dsess.struct, dxid.struct if dxid is not None else None,
                        dcols.struct)
    if rc not in (0, _abi.EOVERFLOW):
        raise ZKBatchError(rc, ctx._lib.zkb_last_error(ctx._h).decode("utf-8", "replace"))
    return rc


def decode_dev_autosize(ctx, d_bytes, nbytes, dsess, dxid, guess=None, device="cuda:0"):
    """Decode, growing the output capacities to what the library reports until everything fits."""
    caps = list(guess or (max(nbytes // 64, 16), max(nbytes // 128, 16), max(nbytes // 32, 16), 16))
    for _ in range(4):
        dc = DeviceColumns(dsess.n, *caps, device=device)
        rc = decode_dev(ctx, d_bytes, nbytes, dsess, dxid, dc)
        if rc == 0:
            return dc
        d = dc.struct
        caps = [max(caps[0], int(d.n_frames)), max(caps[1], int(d.n_stat), 16 if not d.n_stat else 0),
                max(caps[2], int(d.n_buf)), max(caps[3], int(d.n_acl))]
        if int(d.n_frames) > dc.struct.cap_frames:
            # side-table needs are unknown until the frame rows fit: start them at the frame count
            caps[1] = max(caps[1], caps[0])
            caps[2] = max(caps[2], 2 * caps[0])
            caps[3] = max(caps[3], caps[0] // 2)
        del dc
    raise ZKBatchError(_abi.EOVERFLOW, "could not size the output tables")


def xidtab_build_dev(ctx, dcols_req, nsess, dxid):
    rc = ctx._lib.zkb_xidtab_build_dev(ctx._h, C.byref(dcols_req.struct), nsess, C.byref(dxid.struct),
                                       dxid.ops.data_ptr())
    ctx._check(rc)


def session_stats_dev(ctx, dcols, nsess, flags=0, device="cuda:0"):
    """zkb_session_stats_dev over device-resident columns -> session.SessionStats (host copy)."""
    from .session import SessionStats
    mx = torch.empty(max(nsess, 1), dtype=torch.int64, device=device)
    nr = torch.empty(max(nsess, 1), dtype=torch.int32, device=device)
    nt = torch.empty(max(nsess, 1) * _abi.NOTIFY_SLOTS, dtype=torch.int32, device=device)
    st = _abi.SessionStats()
    st.max_zxid, st.n_replies, st.notifications = mx.data_ptr(), nr.data_ptr(), nt.data_ptr()
    ctx._check(ctx._lib.zkb_session_stats_dev(ctx._h, C.byref(dcols.struct), nsess, flags, C.byref(st)))
    return SessionStats(mx[:nsess].cpu().numpy().copy(), nr[:nsess].cpu().numpy().view(np.uint32).copy(),
                        nt[:nsess * _abi.NOTIFY_SLOTS].cpu().numpy().view(np.uint32).reshape(nsess, _abi.NOTIFY_SLOTS).copy())


class DeviceRequests:
    """A RequestBatch (encode.py) uploaded to HBM: arena and every descriptor column."""

    def __init__(self, batch, device="cuda:0"):
        self.n = batch.n
        self.t = {"arena": to_device(batch.arena, device, pad=16)}
        r = _abi.Requests()
        r.n, r.arena_len, r.n_elem, r.n_acl = batch.n, batch.arena.size, batch.n_elem, batch.n_acl
        r.arena = self.t["arena"].data_ptr()
        for name, dt in _abi.REQ_COLS + _abi.ELEM_COLS + _abi.REQ_ACL_COLS:
            self.t[name] = to_device(np.ascontiguousarray(batch.cols[name], dtype=_NP_DT[dt]), device)
            setattr(r, name, self.t[name].data_ptr())
        self.struct = r

    def nbytes(self):
        return sum(t.numel() for t in self.t.values())


def encode_dev(ctx, dreq, out_cap, device="cuda:0", out=None):
    """zkb_encode_dev into a device buffer of out_cap bytes.  Returns (out tensor, frame_off
    tensor [n+1] int64, status tensor [n] uint8, out_len)."""
    if out is None:
        out = (torch.empty(out_cap + 16, dtype=torch.uint8, device=device),
               torch.empty(dreq.n + 1, dtype=torch.int64, device=device),
               torch.empty(max(dreq.n, 1), dtype=torch.uint8, device=device))
    n = C.c_uint64()
    rc = ctx._lib.zkb_encode_dev(ctx._h, C.byref(dreq.struct), out[0].data_ptr(), out_cap, out[1].data_ptr(),
                                 out[2].data_ptr(), C.byref(n))
    if rc not in (0, _abi.EOVERFLOW):
        ctx._check(rc)
    return out[0], out[1], out[2], int(n.value), rc
