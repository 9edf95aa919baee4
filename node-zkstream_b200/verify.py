"""Order-sensitive digests of decoded columns that live in HBM, computed with integer tensor
arithmetic (no codec involved), for checking a whole multi-gigabyte result against the CPU
oracle without bringing the columns back:

    D(col) = sum over rows i of mix64(value_i zero-extended to 64 bits) * (i + 1)   (mod 2^64)

oracle/zkoracle.c (zko_digest) computes the same sums from its own decode of the same bytes;
bench.py compares the two before it times anything.  Same rows in the same places <=> equal
digests for every column (up to 64-bit hash collisions).
"""
import torch

from . import _abi

_M1 = 0xff51afd7ed558ccd - (1 << 64)
_M2 = 0xc4ceb9fe1a85ec53 - (1 << 64)

COLUMNS = ([n for n, _ in _abi.FRAME_COLS + _abi.STAT_COLS + _abi.BUF_COLS + _abi.ACL_COLS] +
           [n for n, _ in _abi.SESS_COLS])
_TABLE_OF = {}
for _cnt, _cols in (("n_frames", _abi.FRAME_COLS), ("n_stat", _abi.STAT_COLS), ("n_buf", _abi.BUF_COLS),
                    ("n_acl", _abi.ACL_COLS)):
    for _n, _dt in _cols:
        _TABLE_OF[_n] = _cnt


def _lsr(v, k):
    """logical shift right of int64 bit patterns"""
    return (v >> k) & ((1 << (64 - k)) - 1)


def mix64(v):
    v = v ^ _lsr(v, 33)
    v = v * _M1
    v = v ^ _lsr(v, 29)
    v = v * _M2
    v = v ^ _lsr(v, 32)
    return v


def digest_tensor(t, n, chunk=1 << 26):
    """Digest of the first n elements of an int32 / int64 tensor (any device), as a Python int in [0, 2^64)."""
    total = 0
    for lo in range(0, n, chunk):
        hi = min(n, lo + chunk)
        v = t[lo:hi]
        if v.dtype == torch.int32:
            v = v.to(torch.int64) & 0xFFFFFFFF
        elif v.dtype != torch.int64:
            raise TypeError("digest_tensor: int32 or int64 columns only, got %r" % (v.dtype,))
        w = torch.arange(lo + 1, hi + 1, dtype=torch.int64, device=v.device)
        total += int((mix64(v) * w).sum().item())
    return total & ((1 << 64) - 1)


def digest_device_columns(dc):
    """{column: digest} of a device.DeviceColumns after a decode (rows the call produced)."""
    d = dc.struct
    out = {}
    for name in COLUMNS:
        if name in _TABLE_OF:
            n = int(getattr(d, _TABLE_OF[name]))
        else:
            n = dc.nsess + 1 if name == "sess_first_frame" else dc.nsess
        out[name] = digest_tensor(dc.t[name], n)
    return out
