"""ctypes mirror of include/zkbatch.h (struct layouts and constants)."""
import ctypes as C

OK, EINVAL, ENOMEM, ENODEV, ECUDA, EOVERFLOW, ERANGE, EINTERNAL = 0, -22, -12, -19, -5, -75, -34, -71
MAX_PACKET = 16 * 1024 * 1024
OP_NONE = -128
ST_OK, ST_BAD_DECODE = 0, 1
KIND_RESPONSE, KIND_REQUEST, KIND_CONNECT_RESPONSE, KIND_CONNECT_REQUEST = 0, 1, 2, 3
CAUSE_RANGE, CAUSE_ASSERT, CAUSE_NO_MATCH, CAUSE_UNSUPPORTED = 1, 2, 3, 4
SESS_SERVER = 1
NO_OFFSET = 0xFFFFFFFFFFFFFFFF
ENC_CONNECT_REQUEST, ENC_CONNECT_RESPONSE = -100, -101
ENC_OK, ENC_UNSUPPORTED = 0, 1
NOALLOC_MAX_XIDS = 1 << 20
NOTIFY_SLOTS, NOTIFY_IGNORED, NOTIFY_BAD_TYPE = 6, 0, 5
STATS_UNTIL_ERROR = 1

FIELD_NAMES = [
    "NONE", "XID", "ZXID", "ERR", "OPCODE", "PATH", "DATA", "WATCH", "VERSION", "FLAGS",
    "COUNT", "ELEM", "ACL_PERMS", "ACL_SCHEME", "ACL_ID", "STAT", "NTYPE", "NSTATE",
    "PROTOCOL_VERSION", "LAST_ZXID", "TIMEOUT", "SESSION_ID", "PASSWD", "REL_ZXID",
]

P = C.c_void_p


class Sessions(C.Structure):
    _fields_ = [("nsessions", C.c_uint32), ("off", P), ("flags", P), ("handshake_frames", P)]


class XidTab(C.Structure):
    _fields_ = [("base", P), ("count", P), ("off", P), ("ops", P), ("nops", C.c_uint64)]


# (name, numpy dtype) in header order, grouped by table
FRAME_COLS = [("frame_off", "u8"), ("frame_len", "i4"), ("session", "u4"), ("meta", "u4"),
              ("xid", "i4"), ("zxid", "i8"), ("err", "i4"), ("a0", "i4"), ("a1", "i4")]
STAT_COLS = [("stat_frame", "u4"), ("czxid", "i8"), ("mzxid", "i8"), ("ctime", "i8"), ("mtime", "i8"),
             ("version", "i4"), ("cversion", "i4"), ("aversion", "i4"), ("ephemeralOwner", "i8"),
             ("dataLength", "i4"), ("numChildren", "i4"), ("pzxid", "i8")]
BUF_COLS = [("buf_frame", "u4"), ("buf_off", "u8"), ("buf_len", "i4")]
ACL_COLS = [("acl_frame", "u4"), ("acl_perms", "i4")]
SESS_COLS = [("sess_consumed", "u8"), ("sess_bad_length_at", "u8"), ("sess_first_frame", "u4")]


class Decoded(C.Structure):
    _fields_ = ([(n, C.c_uint64) for n in ("cap_frames", "cap_stat", "cap_buf", "cap_acl",
                                            "n_frames", "n_stat", "n_buf", "n_acl")] +
                [(n, P) for n, _ in FRAME_COLS + STAT_COLS + BUF_COLS + ACL_COLS + SESS_COLS] +
                [("buf_rel", P)])      # compact replay results only


REQ_COLS = [("xid", "i4"), ("op", "i1"), ("path_off", "u8"), ("path_len", "i4"), ("data_off", "u8"),
            ("data_len", "i4"), ("a0", "i4"), ("a1", "i4"), ("zxid", "i8"), ("zxid2", "i8"),
            ("vec_first", "u4"), ("vec_count", "i4"), ("n0", "i4"), ("n1", "i4")]
ELEM_COLS = [("elem_off", "u8"), ("elem_len", "i4")]
REQ_ACL_COLS = [("acl_perms", "i4"), ("acl_scheme_off", "u8"), ("acl_scheme_len", "i4"),
                ("acl_id_off", "u8"), ("acl_id_len", "i4")]


class Requests(C.Structure):
    _fields_ = [("n", C.c_uint64), ("xid", P), ("op", P), ("arena", P), ("arena_len", C.c_uint64),
                ("path_off", P), ("path_len", P), ("data_off", P), ("data_len", P),
                ("a0", P), ("a1", P), ("zxid", P), ("zxid2", P), ("vec_first", P), ("vec_count", P),
                ("n0", P), ("n1", P),
                ("n_elem", C.c_uint64), ("elem_off", P), ("elem_len", P),
                ("n_acl", C.c_uint64), ("acl_perms", P), ("acl_scheme_off", P), ("acl_scheme_len", P),
                ("acl_id_off", P), ("acl_id_len", P)]


class SessionStats(C.Structure):
    _fields_ = [("max_zxid", P), ("n_replies", P), ("notifications", P)]


SLICE_CB = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_uint32, C.c_uint32, C.POINTER(Decoded), C.POINTER(Decoded))

# every symbol include/zkbatch.h declares: (name, restype, argtypes)
SYMBOLS = [
    ("zkb_ctx_create", C.c_int, [C.c_int, C.POINTER(P)]),
    ("zkb_ctx_destroy", None, [P]),
    ("zkb_last_error", C.c_char_p, [P]),
    ("zkb_kernel_launches", C.c_uint64, [P]),
    ("zkb_split_fixups", C.c_uint64, [P]),
    ("zkb_row_mispredicts", C.c_uint64, [P]),
    ("zkb_last_slow_frames", C.c_uint64, [P]),
    ("zkb_ctx_set_timing", C.c_int, [P, C.c_int]),
    ("zkb_last_phase_ms", C.c_int, [P, C.POINTER(C.c_float * 3)]),
    ("zkb_ctx_set_replay_compact", C.c_int, [P, C.c_int]),
    ("zkb_ctx_stream", P, [P]),
    ("zkb_ctx_sync", C.c_int, [P]),
    ("zkb_decode_scratch_bytes", C.c_uint64, [C.c_uint64, C.c_uint32]),
    ("zkb_decode_dev", C.c_int, [P, P, C.c_uint64, C.POINTER(Sessions), C.POINTER(XidTab), C.POINTER(Decoded)]),
    ("zkb_decode_host", C.c_int, [P, P, C.c_uint64, C.POINTER(Sessions), C.POINTER(XidTab), C.POINTER(Decoded)]),
    ("zkb_replay_host", C.c_int, [P, P, C.c_uint64, C.POINTER(Sessions), P, C.c_uint64, C.POINTER(Sessions),
                                  P, P, C.c_uint64, SLICE_CB, P]),
    ("zkb_xidtab_build_host", C.c_int, [P, C.POINTER(Decoded), C.c_uint32, C.POINTER(XidTab), P]),
    ("zkb_xidtab_build_dev", C.c_int, [P, C.POINTER(Decoded), C.c_uint32, C.POINTER(XidTab), P]),
    ("zkb_xidtab_build_noalloc", C.c_int, [C.POINTER(Decoded), C.c_uint32, C.POINTER(XidTab), P, P,
                                           C.c_char_p, C.c_size_t]),
    ("zkb_session_stats_dev", C.c_int, [P, C.POINTER(Decoded), C.c_uint32, C.c_uint32, C.POINTER(SessionStats)]),
    ("zkb_session_stats_host", C.c_int, [P, C.POINTER(Decoded), C.c_uint32, C.c_uint32, C.POINTER(SessionStats)]),
    ("zkb_encode_dev", C.c_int, [P, C.POINTER(Requests), P, C.c_uint64, P, P, C.POINTER(C.c_uint64)]),
    ("zkb_encode_host", C.c_int, [P, C.POINTER(Requests), C.POINTER(P), C.POINTER(P), C.POINTER(P),
                                   C.POINTER(C.c_uint64)]),
]
