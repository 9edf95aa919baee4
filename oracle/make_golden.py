"""TEST INFRASTRUCTURE — (re)generates tests/golden/*.json by running the reference's
unmodified JavaScript (oracle/ref_js.py, needs /root/reference and oracle/_ref/js).

    python oracle/make_golden.py

ref_decode.json: sessions of wire bytes (every reader, every damaged-frame shape of
synth.fuzz_corpus, the reference's own captures and bad-length prefixes) with the packets /
errors ZKDecodeStream produced for them.  ref_encode.json: packets with the frames
ZKEncodeStream produced.  The tests compare the CPU oracle and the CUDA codec with these
ref_session_folds.json: client-side sessions with what ZKSession's own packet handler
(lib/zk-session.js) made of them.  The tests compare the CPU oracle and the CUDA codec with
these files, so the GPU box never needs the reference.
"""
import base64
import importlib
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import ref_js  # noqa: E402
from _canon import canon_events, xidmap_names  # noqa: E402

synth = importlib.import_module("node-zkstream_b200.synth")
GOLD = os.path.join(ROOT, "tests", "golden")

CAPTURE1 = ["AAAALQAAAAAAAAAAAAAAAAAAdTAAAAAAAAAAAAAAABAAAAAAAAAAAAAAAAAAAAAAAA==",
            "AAAAJQAAAAAAAHUwAVWjqFbbAAAAAAAQh19uvwgo25o9B6hUkSvqKQA=",
            "AAAADgAAAAEAAAAIAAAAAS8A",
            "AAAAKAAAAAEAAAAAAAAFFwAAAAAAAAACAAAACXpvb2tlZXBlcgAAAANmb28="]


def decode_corpora():
    out = []
    b = [base64.b64decode(x) for x in CAPTURE1]
    c = synth.Corpus()
    c.add(b[0] + b[2], server=True)
    c.add(b[1] + b[3], xidmap={1: 8})
    c.add(bytes.fromhex("40004000"))
    c.add(bytes.fromhex("000000000102"))
    c.add(bytes.fromhex("fffffffe0102"))
    out.append(("capture1+nasty", c))
    for seed in range(12):
        out.append(("fuzz%d" % seed, synth.fuzz_corpus(seed, nsessions=16, max_frames=24)))
    out.append(("cfg2", synth.cfg2_mixed(2, 40, seed=5)))
    out.append(("cfg3", synth.cfg3_tiny(4, 30, seed=6)))
    return out


ENCODE_PACKETS = [
    # (handshaking, isServer, packet) — the survey's known-answer list and more
    (True, False, dict(protocolVersion=0, lastZxidSeen=bytes(8), timeOut=30000, sessionId=bytes(8), passwd=bytes(8))),
    (True, True, dict(protocolVersion=0, timeOut=30000, sessionId=bytes.fromhex("0155a3a856db0000"),
                      passwd=bytes.fromhex("875f6ebf0828db9a3d07a854912bea29"))),
    (True, False, dict(protocolVersion=0, lastZxidSeen=bytes.fromhex("0517"), timeOut=4000,
                       sessionId=bytes.fromhex("0102030405"), passwd=b"")),
    (False, False, dict(xid=1, opcode="GET_CHILDREN", path="/", watch=False)),
    (False, False, dict(xid=-2, opcode="PING")),
    (False, False, dict(xid=11, opcode="CLOSE_SESSION")),
    (False, False, dict(xid=5, opcode="EXISTS", path="/foo", watch=True)),
    (False, False, dict(xid=6, opcode="GET_DATA", path="/foo", watch=False)),
    (False, False, dict(xid=7, opcode="GET_CHILDREN2", path="/", watch=False)),
    (False, False, dict(xid=8, opcode="DELETE", path="/foo", version=0)),
    (False, False, dict(xid=9, opcode="SYNC", path="/foo")),
    (False, False, dict(xid=10, opcode="GET_ACL", path="/foo")),
    (False, False, dict(xid=4, opcode="SET_DATA", path="/foo", data=b"hi", version=-1)),
    (False, False, dict(xid=0, opcode="CREATE", path="/foo", data=b"hi there",
                        acl=[dict(id=dict(scheme="world", id="anyone"), perms=["read", "write", "create", "delete", "admin"])],
                        flags=[])),
    (False, False, dict(xid=3, opcode="CREATE", path="/e", data=b"",
                        acl=[dict(id=dict(scheme="world", id="anyone"), perms=["READ", "WRITE", "CREATE", "DELETE", "ADMIN"])],
                        flags=["EPHEMERAL", "SEQUENTIAL"])),
    (False, False, dict(xid=12, opcode="CREATE", path="/two/acls", data=bytes(range(40)),
                        acl=[dict(id=dict(scheme="digest", id="u:p"), perms=["read"]),
                             dict(id=dict(scheme="ip", id=""), perms=[])],
                        flags=["SEQUENTIAL"])),
    (False, False, dict(xid=-8, opcode="SET_WATCHES", relZxid=bytes.fromhex("0000000000000517"),
                        events=dict(dataChanged=["/a"], createdOrDestroyed=[], childrenChanged=["/b", "/cc"]))),
    (False, False, dict(xid=13, opcode="SET_WATCHES", relZxid=bytes(8), events=dict())),
    (False, False, dict(xid=14, opcode="GET_DATA", path="", watch=True)),
    (False, False, dict(xid=15, opcode="SET_DATA", path="/café/中", data=bytes(300), version=7)),
    (False, False, dict(xid=16, opcode="SET_ACL", path="/x")),
    (False, False, dict(xid=17, opcode="MULTI")),
]


def main():
    if not ref_js.available():
        print("reference JS engine unavailable: run oracle/build_ref.sh where /root/reference exists")
        return 1
    os.makedirs(GOLD, exist_ok=True)
    dec = []
    for name, c in decode_corpora():
        sessions = [dict(chunks=[c.chunks[i]], isServer=bool(c.flags[i]), hs=c.hs[i],
                         xidMap=xidmap_names(c.xidmaps[i])) for i in range(len(c.chunks))]
        ref = ref_js.decode_sessions(sessions)
        for i, r in enumerate(ref):
            dec.append(dict(corpus=name, hex=c.chunks[i].hex(), isServer=bool(c.flags[i]), hs=c.hs[i],
                            xidMap={str(k): v for k, v in c.xidmaps[i].items()},
                            events=canon_events(r["events"]), used=r["used"]))
    with open(os.path.join(GOLD, "ref_decode.json"), "w") as f:
        json.dump(dict(generator="oracle/make_golden.py", reference="node-zkstream lib/zk-streams.js ZKDecodeStream "
                       "run under oracle/_ref/js", sessions=dec), f, separators=(",", ":"))
    enc = []
    for hs, server, pkt in ENCODE_PACKETS:
        r = ref_js.encode_sessions([dict(pkts=[pkt], isServer=server, hs=1 if hs else 0)])[0]
        enc.append(dict(handshaking=hs, isServer=server, pkt=_jsonable(pkt), result=r["frames"][0],
                        xidMap={k: (v[1].decode() if isinstance(v, tuple) and v[0] == "str" else None)
                                for k, v in r["xidMap"].items()}))
    with open(os.path.join(GOLD, "ref_encode.json"), "w") as f:
        json.dump(dict(generator="oracle/make_golden.py", reference="node-zkstream lib/zk-streams.js ZKEncodeStream "
                       "run under oracle/_ref/js", packets=enc), f, separators=(",", ":"))
    folds = []
    for name, c in session_corpora():
        sessions = [dict(chunks=[c.chunks[i]], hs=c.hs[i], xidMap=xidmap_names(c.xidmaps[i]))
                    for i in range(len(c.chunks)) if not c.flags[i]]
        keep = [i for i in range(len(c.chunks)) if not c.flags[i]]
        ref_all = ref_js.session_folds(sessions, until_error=False)
        ref_stop = ref_js.session_folds(sessions, until_error=True)
        for j, i in enumerate(keep):
            folds.append(dict(corpus=name, hex=c.chunks[i].hex(), hs=c.hs[i],
                              xidMap={str(k): v for k, v in c.xidmaps[i].items()},
                              all=ref_all[j], until_error=ref_stop[j]))
    with open(os.path.join(GOLD, "ref_session_folds.json"), "w") as f:
        json.dump(dict(generator="oracle/make_golden.py", reference="node-zkstream lib/zk-session.js: the 'packet' "
                       "handler of state_attached and processNotification, fed by lib/zk-streams.js ZKDecodeStream, "
                       "run under oracle/_ref/js (jsbn replaced by a signed big-endian stand-in)", sessions=folds),
                  f, separators=(",", ":"))
    print("wrote %d decode sessions, %d encode packets, %d session folds" % (len(dec), len(enc), len(folds)))
    return 0


def session_corpora():
    out = [("session%d" % seed, synth.session_corpus(seed)) for seed in range(6)]
    out.append(("cfg3", synth.cfg3_tiny(4, 60, seed=16)))
    out.append(("cfg2", synth.cfg2_mixed(2, 40, seed=15)))
    out.append(("fuzz", synth.fuzz_corpus(21, nsessions=24, max_frames=24)))
    return out


def _jsonable(v):
    if isinstance(v, (bytes, bytearray)):
        return {"$hex": bytes(v).hex()}
    if isinstance(v, dict):
        return {k: _jsonable(x) for k, x in v.items()}
    if isinstance(v, list):
        return [_jsonable(x) for x in v]
    return v


if __name__ == "__main__":
    sys.exit(main())
