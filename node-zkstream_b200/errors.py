"""Error types of the reference (lib/errors.js:19-54), same names, codes and messages."""


class ZKProtocolError(Exception):
    """lib/errors.js:19-26: code + 'code: msg' message; name 'ZKProtocolError'."""

    def __init__(self, code, msg):
        self.code = code
        self.name = "ZKProtocolError"
        self.message = code + ": " + msg
        Exception.__init__(self, self.message)


class ZKError(Exception):
    """lib/errors.js:47-54."""

    def __init__(self, code, msg):
        self.code = code
        self.name = "ZKError"
        self.message = code + ": " + msg
        Exception.__init__(self, self.message)


class AssertionError_(AssertionError):
    """What assert-plus throws in the reference for a bad argument (synchronous throw)."""
    name = "AssertionError"
