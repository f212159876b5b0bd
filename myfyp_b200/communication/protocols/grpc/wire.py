"""Hand-written protobuf wire codec for the reference's ``node.proto``.

``grpc_tools``/``protoc`` are not available on the target image, and the schema is five tiny messages, so the
encoder/decoder is written out by hand.  Field numbers and types are the reference's
(communication/protocols/grpc/proto/node.proto:26-54), which keeps mixed deployments interoperable:

    RootMessage { string source=1; optional int32 round=2; string cmd=3; oneof { Message message=4; Weights weights=5; } }
    Message     { int32 ttl=1; int64 hash=2; repeated string args=3; [repeated string seen=4;] }

``seen`` is an extension of this framework: the addresses already known to have (or to be about to get) a flooded
message; a receiver forwards only to direct neighbours not listed.  In a full mesh that removes every redundant
forward (the reference sends (n-1)(n-2) of them per message); a reference peer ignores the
unknown field and simply floods as before, so mixed deployments still work.
    Weights     { bytes weights=1; repeated string contributors=2; int32 num_samples=3; }
    HandShakeRequest { string addr=1; }      ResponseMessage { optional string error=1; }
"""

from __future__ import annotations

from typing import Any, Dict, Optional, Tuple

_MASK64 = (1 << 64) - 1


def _varint(v: int) -> bytes:
    v &= _MASK64  # negative int32/int64 are sign-extended to 64 bits on the wire
    out = bytearray()
    while True:
        b = v & 0x7F
        v >>= 7
        if v:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _read_varint(buf: memoryview, pos: int) -> Tuple[int, int]:
    shift = result = 0
    while True:
        b = buf[pos]
        pos += 1
        result |= (b & 0x7F) << shift
        if not b & 0x80:
            return result, pos
        shift += 7
        if shift > 70:
            raise ValueError("varint too long")


def _signed(v: int, bits: int = 64) -> int:
    v &= _MASK64
    if v >= 1 << 63:
        v -= 1 << 64
    if bits == 32:
        v = ((v + (1 << 31)) % (1 << 32)) - (1 << 31)
    return v


def _key(field: int, wt: int) -> bytes:
    return _varint((field << 3) | wt)


def _ld(field: int, payload: bytes) -> bytes:
    return _key(field, 2) + _varint(len(payload)) + payload


def _fields(data: bytes):
    buf = memoryview(data)
    pos, n = 0, len(buf)
    while pos < n:
        key, pos = _read_varint(buf, pos)
        field, wt = key >> 3, key & 7
        if wt == 0:
            v, pos = _read_varint(buf, pos)
            yield field, wt, v
        elif wt == 2:
            ln, pos = _read_varint(buf, pos)
            yield field, wt, bytes(buf[pos : pos + ln])
            pos += ln
        elif wt == 1:
            yield field, wt, bytes(buf[pos : pos + 8])
            pos += 8
        elif wt == 5:
            yield field, wt, bytes(buf[pos : pos + 4])
            pos += 4
        else:
            raise ValueError(f"unsupported wire type {wt}")


# ---------------------------------------------------------------------------------- messages
def encode_root(msg: Dict[str, Any]) -> bytes:
    """``msg`` is the dict form used by the in-memory protocol (``"weight" in msg`` marks a weights message)."""
    out = bytearray()
    out += _ld(1, msg["source"].encode())
    rnd = msg.get("round")
    out += _key(2, 0) + _varint(-1 if rnd is None else int(rnd))
    out += _ld(3, msg["cmd"].encode())
    if "weight" in msg:
        w = bytearray()
        w += _ld(1, bytes(msg["weights"]))
        for c in msg.get("contributors", []):
            w += _ld(2, c.encode())
        w += _key(3, 0) + _varint(int(msg["weight"]))
        out += _ld(5, bytes(w))
    else:
        m = bytearray()
        m += _key(1, 0) + _varint(int(msg["ttl"]))
        m += _key(2, 0) + _varint(int(msg["hash"]))
        for a in msg.get("args", []):
            m += _ld(3, str(a).encode())
        for a in msg.get("seen") or []:  # extension field 4 (unknown to the reference's parser, which skips it)
            m += _ld(4, str(a).encode())
        out += _ld(4, bytes(m))
    return bytes(out)


def decode_root(data: bytes) -> Dict[str, Any]:
    msg: Dict[str, Any] = {"source": "", "round": -1, "cmd": ""}
    for f, wt, v in _fields(data):
        if f == 1:
            msg["source"] = v.decode()
        elif f == 2:
            msg["round"] = _signed(v, 32)
        elif f == 3:
            msg["cmd"] = v.decode()
        elif f == 4:
            msg.update(ttl=0, hash=0, args=[], seen=[])
            for f2, _, v2 in _fields(v):
                if f2 == 1:
                    msg["ttl"] = _signed(v2, 32)
                elif f2 == 2:
                    msg["hash"] = _signed(v2, 64)
                elif f2 == 3:
                    msg["args"].append(v2.decode())
                elif f2 == 4:
                    msg["seen"].append(v2.decode())
        elif f == 5:
            msg.update(weights=b"", contributors=[], weight=0)
            for f2, _, v2 in _fields(v):
                if f2 == 1:
                    msg["weights"] = v2
                elif f2 == 2:
                    msg["contributors"].append(v2.decode())
                elif f2 == 3:
                    msg["weight"] = _signed(v2, 32)
    return msg


def encode_handshake(addr: str) -> bytes:
    return _ld(1, addr.encode())


def decode_handshake(data: bytes) -> str:
    for f, _, v in _fields(data):
        if f == 1:
            return v.decode()
    return ""


def encode_response(error: Optional[str] = None) -> bytes:
    return b"" if error is None else _ld(1, error.encode())


def decode_response(data: bytes) -> str:
    """Error text, or "" when the optional field is absent (grpc treats a ``None`` from a deserializer as a failure)."""
    for f, _, v in _fields(data):
        if f == 1:
            return v.decode()
    return ""


def encode_empty(_: Any = None) -> bytes:
    return b""


def decode_empty(_: bytes) -> bytes:
    return b""
