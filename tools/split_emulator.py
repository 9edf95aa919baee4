"""Development aid: a CPU emulation of the splitter's speculate/scan/fix loop
(csrc/zk_decode.cu k_split / k_scan_*), used to study how many iterations a corpus needs
and which granules stay dirty.  Not used by the library, the tests or the bench."""
import struct
import sys

import numpy as np

NONE = (1 << 64) - 1
INHERIT = NONE - 1
MAXP = 16 * 1024 * 1024


def be32(b, p):
    return struct.unpack_from(">i", b, p)[0]


def err_known(e):
    return (-8 <= e <= 0) or (-103 <= e <= -100) or e == -108 or (-115 <= e <= -110)


def op_known(v):
    return (0 <= v <= 14 and v != 10) or (100 <= v <= 102) or v in (-10, -11, -1)


def header_plausible(b, p, send, server):
    fld = 8 if server else 16
    if p + 4 + fld + 4 > send:
        return None
    l = be32(b, p)
    f = be32(b, p + fld)
    if l < fld or l > MAXP or p + 4 + l > send:
        return None
    if not (op_known(f) if server else err_known(f)):
        return None
    return p + 4 + l


def chain_plausible(b, q, send, server, hops):
    for _ in range(hops):
        if q == send or q + 4 > send:
            return True
        n = header_plausible(b, q, send, server)
        if n is None:
            # a frame that does not fit ends the chain legitimately only if its header is sane
            fld = 8 if server else 16
            if q + 4 + fld + 4 > send:
                return True
            l = be32(b, q)
            f = be32(b, q + fld)
            return fld <= l <= MAXP and q + 4 + l > send and (op_known(f) if server else err_known(f))
        q = n
    return True


def guess(b, gs, ge, send, server, hops=3):
    for p in range(gs, min(ge, send)):
        if b[p] > 1:
            continue
        n = header_plausible(b, p, send, server)
        if n is None:
            continue
        if chain_plausible(b, n, send, server, hops):
            return p
    return NONE


def walk(b, pos, send, ge):
    cnt = 0
    while True:
        if pos == send:
            return NONE, cnt
        if pos >= ge:
            return pos, cnt
        rem = send - pos
        if rem < 4:
            return NONE, cnt
        l = be32(b, pos)
        if l < 0 or l > MAXP or rem < 4 + l:
            return NONE, cnt
        cnt += 1
        pos += 4 + l


def granule(b, off, flags, g, G, nbytes, assumed_in, spec):
    gs = g * G
    last = gs + G >= nbytes
    ge = nbytes if last else gs + G
    sf = int(np.searchsorted(off, gs, side="left"))
    cont = off[sf] != gs
    assumed, pos, cnt, touched = NONE, NONE, 0, False
    if cont:
        s = sf - 1
        send = int(off[sf])
        assumed = guess(b, gs, ge, send, bool(flags[s] & 1)) if spec else assumed_in
        if gs <= assumed < ge:
            touched = True
            pos, c = walk(b, assumed, send, ge)
            cnt += c
    s = sf
    while s < len(off) - 1 and (off[s] < ge or last):
        touched = True
        pos, c = walk(b, int(off[s]), int(off[s + 1]), ge)
        cnt += c
        s += 1
    return assumed, (pos if touched else INHERIT), cnt


def run(b, off, flags, G, verbose=True):
    nbytes = len(b)
    n = (nbytes + G - 1) // G
    A = [granule(b, off, flags, g, G, nbytes, None, True) for g in range(n)]
    ver = [[A[g], A[g]] for g in range(n)]
    sel = [0] * n
    it = 0
    while True:
        carry, src = [NONE] * n, [0] * n
        cur, cs = INHERIT, 0
        for g in range(n):
            carry[g] = NONE if cur == INHERIT else cur
            src[g] = 0 if cur == INHERIT else cs
            o = ver[g][sel[g]][1]
            if o != INHERIT:
                cur, cs = o, g + 1
        dirty = [0] * n
        nd = 0
        for g in range(n):
            gs = g * G
            ge = min(gs + G, nbytes) if g < n - 1 else nbytes
            N = lambda x: x if gs <= x < ge else NONE
            nc = N(carry[g])
            n0, n1 = N(ver[g][0][0]), N(ver[g][1][0])
            old = sel[g]
            if (n1 if old else n0) == nc:
                pass
            elif n0 == nc:
                sel[g] = 0
            elif n1 == nc:
                sel[g] = 1
            else:
                dirty[g] = 1
            if dirty[g] or sel[g] != old:
                nd += 1
        if verbose:
            print("iter", it, "unstable", nd, "dirty", sum(dirty), "first dirty", [g for g in range(n) if dirty[g]][:12])
        if nd == 0:
            break
        for g in range(n):
            if not dirty[g]:
                continue
            r = granule(b, off, flags, g, G, nbytes, carry[g], False)
            ver[g][1] = r
            trusted = not (src[g] and dirty[src[g] - 1]) and not (carry[g] != NONE and carry[g] < g * G)
            sel[g] = 1 if trusted else 0
        it += 1
    return it


if __name__ == "__main__":
    import importlib
    sys.path.insert(0, ".")
    synth = importlib.import_module("node-zkstream_b200.synth")
    pc = synth.native_cfg2(20240607, 2, 20000)
    for name, data, off, fl in (("req", pc.req, pc.req_off, pc.req_flags), ("rsp", pc.rsp, pc.rsp_off, pc.rsp_flags)):
        print(name, data.size)
        run(data.tobytes(), off.astype(np.int64), fl, 4096)
