"""Shared by the tests: one canonical, JSON-friendly form for packet objects, whether they
come from the reference's JavaScript (oracle/ref_js.py) or from Materialiser."""
import importlib

zk = importlib.import_module("node-zkstream_b200")
LongDate = zk.LongDate
consts = zk.consts


def canon(v):
    if isinstance(v, LongDate):
        return {"ld": v.ld_buffer.hex()}
    if isinstance(v, tuple):
        if v[0] == "str":
            return {"s": bytes(v[1]).hex()}
        if v[0] == "longdate":
            return {"ld": bytes(v[1]).hex()}
        if v[0] == "undefined":
            return None
    if isinstance(v, (bytes, bytearray)):
        return {"b": bytes(v).hex()}
    if isinstance(v, str):
        return {"s": v.encode("utf-8").hex()}
    if isinstance(v, dict):
        return {k: canon(x) for k, x in v.items()}
    if isinstance(v, list):
        return [canon(x) for x in v]
    return v


def canon_events(events):
    """[('pkt', dict) | ('error', ZKProtocolError)] or the reference harness' dicts -> canonical list."""
    out = []
    for e in events:
        if isinstance(e, dict):
            if "pkt" in e:
                out.append({"pkt": canon(e["pkt"]), "keys": list(e["pkt"].keys())})
            else:
                out.append({"error": {"code": e["error"]["code"], "message": e["error"]["message"]}})
        else:
            kind, item = e
            if kind == "pkt":
                out.append({"pkt": canon(item), "keys": list(item.keys())})
            else:
                out.append({"error": {"code": item.code, "message": item.message}})
    return out


def materialised_events(cols, data, session):
    return canon_events(list(zk.Materialiser(cols, data, decode_strings="tagged").events(session)))


def xidmap_names(m):
    return {k: consts.OP_CODE_LOOKUP[v] for k, v in m.items()}


_decanon_guard = True
