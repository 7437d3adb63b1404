"""The wider ``pyspark.sql.functions`` surface: scalar functions over Arrow arrays.

The reference is a thin layer over Spark, so its users call whatever ``pyspark.sql.functions`` has
(reference ``examples/data_process.py:10-68`` and the DLRM notebook use a dozen of them; user code uses
many more).  :mod:`raydp_b200.sql.column` evaluates the handful its own examples need inline; every
other name lands here.  An implementation receives the already evaluated arguments -- Arrow arrays for
columns, ``pa.Scalar`` for literals, plain Python for non-column parameters -- plus the partition's row
count and :class:`~raydp_b200.sql.column.EvalContext`, and returns an Arrow array (or scalar).

Arrow compute kernels do the work where one exists; the long tail (soundex, levenshtein, conv, ...)
runs row by row in Python, which is what the JVM does too, just slower.
"""
from __future__ import annotations

import base64 as _b64
import datetime as _dt
import hashlib
import json
import math
import re
import zlib
from typing import Any, Callable, Dict, List, Optional

import numpy as np
import pyarrow as pa
import pyarrow.compute as pc

SCALAR: Dict[str, Callable] = {}


def scalar(*names: str):
    def deco(fn):
        for name in names:
            SCALAR[name] = fn
        return fn
    return deco


# ------------------------------------------------------------------------------------------ helpers
def _arrayish(v) -> bool:
    return isinstance(v, (pa.Array, pa.ChunkedArray))


def _arr(v, n: int) -> pa.Array:
    """Any argument -> a contiguous Arrow array of length ``n``."""
    if isinstance(v, pa.ChunkedArray):
        return v.combine_chunks()
    if isinstance(v, pa.Array):
        return v
    if isinstance(v, pa.Scalar):
        if not v.is_valid and pa.types.is_null(v.type):
            return pa.nulls(n)
        return pa.array([v.as_py()] * n, type=v.type)
    return pa.array([v] * n)


def _py(v):
    """A literal argument -> its Python value."""
    if isinstance(v, pa.Scalar):
        return v.as_py()
    if _arrayish(v):
        raise TypeError("this parameter must be a literal, not a column")
    return v


def _f64(v):
    return pc.cast(v, pa.float64())


def _rows(v, n: int) -> list:
    return _arr(v, n).to_pylist()


def _map_rows(fn: Callable, n: int, *args, type: Optional[pa.DataType] = None) -> pa.Array:  # noqa: A002
    """Row-wise Python fallback: ``fn(*row)`` with nulls passed through as null."""
    cols = [_rows(a, n) for a in args]
    out = [None if any(x is None for x in row) else fn(*row) for row in zip(*cols)]
    return pa.array(out, type=type)


def _string(v, n: int) -> pa.Array:
    a = _arr(v, n)
    return a if pa.types.is_string(a.type) else pc.cast(a, pa.string())


# ------------------------------------------------------------------------------------------ math
for _name in ("acos", "asin", "atan", "cosh", "sinh", "tan", "tanh", "acosh", "asinh", "atanh"):
    SCALAR[_name] = (lambda k: lambda args, n, ctx: getattr(pc, k)(_f64(args[0])))(_name)


@scalar("atan2")
def _atan2(args, n, ctx):
    return pc.atan2(_f64(args[0]), _f64(args[1]))


@scalar("cbrt")
def _cbrt(args, n, ctx):
    return pa.array(np.cbrt(np.asarray(_f64(_arr(args[0], n)).to_numpy(zero_copy_only=False), dtype=np.float64)),
                    mask=np.asarray(pc.is_null(_arr(args[0], n))))


@scalar("degrees")
def _degrees(args, n, ctx):
    return pc.multiply(_f64(args[0]), 180.0 / math.pi)


@scalar("radians")
def _radians(args, n, ctx):
    return pc.multiply(_f64(args[0]), math.pi / 180.0)


@scalar("expm1")
def _expm1(args, n, ctx):
    a = _arr(args[0], n)
    return pa.array(np.expm1(_f64(a).to_numpy(zero_copy_only=False)), mask=np.asarray(pc.is_null(a)))


@scalar("hypot")
def _hypot(args, n, ctx):
    x, y = _f64(args[0]), _f64(args[1])
    return pc.sqrt(pc.add(pc.multiply(x, x), pc.multiply(y, y)))


@scalar("rint")
def _rint(args, n, ctx):
    return pc.round(_f64(args[0]), round_mode="half_to_even")


@scalar("bround")
def _bround(args, n, ctx):
    return pc.round(args[0], ndigits=int(_py(args[1])) if len(args) > 1 else 0, round_mode="half_to_even")


@scalar("signum", "sign")
def _signum(args, n, ctx):
    return _f64(pc.sign(args[0]))


@scalar("factorial")
def _factorial(args, n, ctx):
    return _map_rows(lambda v: math.factorial(v) if 0 <= v <= 20 else None, n, args[0], type=pa.int64())


@scalar("pmod")
def _pmod(args, n, ctx):
    a = _arr(args[0], n).to_numpy(zero_copy_only=False)
    b = _arr(args[1], n).to_numpy(zero_copy_only=False)
    mask = np.asarray(pc.or_(pc.is_null(_arr(args[0], n)), pc.is_null(_arr(args[1], n))))
    with np.errstate(all="ignore"):
        return pa.array(np.mod(np.where(mask, 1, a), np.where(mask, 1, b)), mask=mask)


@scalar("bin")
def _bin(args, n, ctx):
    return _map_rows(lambda v: format(v & 0xFFFFFFFFFFFFFFFF, "b") if v < 0 else format(v, "b"), n, args[0],
                     type=pa.string())


@scalar("hex")
def _hex(args, n, ctx):
    def one(v):
        if isinstance(v, (bytes, bytearray)):
            return v.hex().upper()
        if isinstance(v, str):
            return v.encode().hex().upper()
        return format(v & 0xFFFFFFFFFFFFFFFF, "X")
    return _map_rows(one, n, args[0], type=pa.string())


@scalar("unhex")
def _unhex(args, n, ctx):
    def one(v):
        try:
            return bytes.fromhex(v if len(v) % 2 == 0 else "0" + v)
        except ValueError:
            return None
    return _map_rows(one, n, args[0], type=pa.binary())


@scalar("conv")
def _conv(args, n, ctx):
    src, dst = int(_py(args[1])), int(_py(args[2]))
    digits = "0123456789ABCDEFGHIJKLMNOPQRSTUVWXYZ"

    def one(v):
        try:
            value = int(str(v), src)
        except ValueError:
            return None
        if value == 0:
            return "0"
        out, value = [], abs(value)
        while value:
            value, r = divmod(value, abs(dst))
            out.append(digits[r])
        return "".join(reversed(out))
    return _map_rows(one, n, args[0], type=pa.string())


@scalar("shiftleft")
def _shiftleft(args, n, ctx):
    return pc.shift_left(args[0], pa.scalar(int(_py(args[1])), _arr(args[0], n).type))


@scalar("shiftright")
def _shiftright(args, n, ctx):
    return pc.shift_right(args[0], pa.scalar(int(_py(args[1])), _arr(args[0], n).type))


@scalar("bitwise_not")
def _bitwise_not(args, n, ctx):
    return pc.bit_wise_not(args[0])


@scalar("bitwise_and")
def _bitwise_and(args, n, ctx):
    return pc.bit_wise_and(args[0], args[1])


@scalar("bitwise_or")
def _bitwise_or(args, n, ctx):
    return pc.bit_wise_or(args[0], args[1])


@scalar("bitwise_xor")
def _bitwise_xor(args, n, ctx):
    return pc.bit_wise_xor(args[0], args[1])


# ------------------------------------------------------------------------------------------ nulls
@scalar("isnan")
def _isnan(args, n, ctx):
    a = _arr(args[0], n)
    if not pa.types.is_floating(a.type):
        return pa.array(np.zeros(n, dtype=bool))
    return pc.fill_null(pc.is_nan(a), False)


@scalar("nanvl")
def _nanvl(args, n, ctx):
    a, b = _f64(_arr(args[0], n)), _f64(_arr(args[1], n))
    return pc.if_else(pc.fill_null(pc.is_nan(a), False), b, a)


@scalar("nvl2")
def _nvl2(args, n, ctx):
    return pc.if_else(pc.is_valid(_arr(args[0], n)), _arr(args[1], n), _arr(args[2], n))


@scalar("nullif")
def _nullif(args, n, ctx):
    a, b = _arr(args[0], n), _arr(args[1], n)
    same = pc.fill_null(pc.equal(a, b), False)
    return pc.if_else(same, pa.nulls(n, a.type), a)


@scalar("eq_null_safe", "equal_null")
def _eq_null_safe(args, n, ctx):
    a, b = _arr(args[0], n), _arr(args[1], n)
    both_null = pc.and_(pc.is_null(a), pc.is_null(b))
    return pc.or_(both_null, pc.fill_null(pc.equal(a, b), False))


@scalar("typeof")
def _typeof(args, n, ctx):
    from raydp_b200.sql import types as T
    return pa.array([T.from_arrow_type(_arr(args[0], n).type).simpleString()] * n)


# ------------------------------------------------------------------------------------------ strings
@scalar("concat_ws")
def _concat_ws(args, n, ctx):
    sep = _py(args[0])
    cols = [_rows(a, n) for a in args[1:]]
    out = []
    for row in zip(*cols) if cols else [()] * n:
        parts: List[str] = []
        for v in row:
            if v is None:
                continue                      # concat_ws skips nulls
            parts.extend(str(x) for x in v if x is not None) if isinstance(v, list) else parts.append(str(v))
        out.append(sep.join(parts))
    return pa.array(out, type=pa.string())


@scalar("substring", "substr")
def _substring(args, n, ctx):
    pos, length = int(_py(args[1])), (int(_py(args[2])) if len(args) > 2 and args[2] is not None else None)
    s = _string(args[0], n)
    if pos > 0:
        start = pos - 1
        return pc.utf8_slice_codeunits(s, start, None if length is None else start + length)
    if pos == 0:
        return pc.utf8_slice_codeunits(s, 0, length)
    stop = None if length is None or pos + length >= 0 else pos + length
    return pc.utf8_slice_codeunits(s, pos, stop)


@scalar("left")
def _left(args, n, ctx):
    return pc.utf8_slice_codeunits(_string(args[0], n), 0, max(0, int(_py(args[1]))))


@scalar("right")
def _right(args, n, ctx):
    k = int(_py(args[1]))
    return pc.utf8_slice_codeunits(_string(args[0], n), -k, None) if k > 0 else pa.array([""] * n)


@scalar("substring_index")
def _substring_index(args, n, ctx):
    delim, count = _py(args[1]), int(_py(args[2]))

    def one(s):
        if count == 0 or not delim:
            return ""
        parts = s.split(delim)
        return delim.join(parts[:count]) if count > 0 else delim.join(parts[count:])
    return _map_rows(one, n, args[0], type=pa.string())


def _make_trim(kind: str):
    def impl(args, n, ctx):
        s = _string(args[0], n)
        if len(args) > 1 and args[1] is not None:
            return getattr(pc, f"utf8_{kind}")(s, characters=_py(args[1]))
        return getattr(pc, f"utf8_{kind}")(s, characters=" ")
    return impl


for _name in ("trim", "ltrim", "rtrim"):
    SCALAR[_name] = _make_trim(_name)


def _pad(kind: str):
    def impl(args, n, ctx):
        width, pad = int(_py(args[1])), (_py(args[2]) if len(args) > 2 else " ")

        def one(s):
            if len(s) >= width or not pad:
                return s[:width]
            fill = (pad * width)[: width - len(s)]
            return fill + s if kind == "lpad" else s + fill
        return _map_rows(one, n, args[0], type=pa.string())
    return impl


SCALAR["lpad"] = _pad("lpad")
SCALAR["rpad"] = _pad("rpad")


@scalar("repeat")
def _repeat(args, n, ctx):
    return pc.binary_repeat(_string(args[0], n), max(0, int(_py(args[1]))))


@scalar("reverse")
def _reverse(args, n, ctx):
    a = _arr(args[0], n)
    if pa.types.is_list(a.type) or pa.types.is_large_list(a.type):
        return pa.array([None if v is None else v[::-1] for v in a.to_pylist()], type=a.type)
    return pc.utf8_reverse(_string(a, n))


@scalar("initcap")
def _initcap(args, n, ctx):
    return _map_rows(lambda s: " ".join(w[:1].upper() + w[1:].lower() for w in s.split(" ")), n, args[0],
                     type=pa.string())


@scalar("instr")
def _instr(args, n, ctx):
    return pc.cast(pc.add(pc.find_substring(_string(args[0], n), _py(args[1])), 1), pa.int32())


@scalar("locate")
def _locate(args, n, ctx):        # locate(substr, str, pos=1)
    sub, pos = _py(args[0]), (int(_py(args[2])) if len(args) > 2 else 1)
    return _map_rows(lambda s: s.find(sub, max(pos - 1, 0)) + 1 if pos >= 1 else 0, n, args[1], type=pa.int32())


@scalar("translate")
def _translate(args, n, ctx):
    src, dst = _py(args[1]), _py(args[2])
    table = {ord(c): (dst[i] if i < len(dst) else None) for i, c in enumerate(src)}
    return _map_rows(lambda s: s.translate(table), n, args[0], type=pa.string())


def _java_regex(pattern: str) -> str:
    """Java and Python regex dialects agree on what SQL users write; possessive/atomic forms do not occur."""
    return pattern


@scalar("regexp_extract")
def _regexp_extract(args, n, ctx):
    rx, idx = re.compile(_java_regex(_py(args[1]))), (int(_py(args[2])) if len(args) > 2 else 1)

    def one(s):
        m = rx.search(s)
        return "" if m is None else (m.group(idx) or "")
    return _map_rows(one, n, args[0], type=pa.string())


@scalar("regexp_substr")
def _regexp_substr(args, n, ctx):
    rx = re.compile(_py(args[1]))

    def one(s):
        m = rx.search(s)
        return None if m is None else m.group(0)
    return _map_rows(one, n, args[0], type=pa.string())


@scalar("regexp_replace")
def _regexp_replace(args, n, ctx):
    replacement = re.sub(r"\$(\d+)", r"\\\1", _py(args[2]))        # $1 -> \1
    return pc.replace_substring_regex(_string(args[0], n), pattern=_py(args[1]), replacement=replacement)


@scalar("regexp_count")
def _regexp_count(args, n, ctx):
    return pc.cast(pc.count_substring_regex(_string(args[0], n), pattern=_py(args[1])), pa.int32())


@scalar("rlike", "regexp", "regexp_like")
def _rlike(args, n, ctx):
    return pc.match_substring_regex(_string(args[0], n), pattern=_py(args[1]))


@scalar("like")
def _like(args, n, ctx):
    return pc.match_like(_string(args[0], n), pattern=_py(args[1]))


@scalar("ilike")
def _ilike(args, n, ctx):
    return pc.match_like(_string(args[0], n), pattern=_py(args[1]), ignore_case=True)


@scalar("startswith")
def _startswith(args, n, ctx):
    if _arrayish(args[1]):
        return _map_rows(lambda s, p: s.startswith(p), n, args[0], args[1], type=pa.bool_())
    return pc.starts_with(_string(args[0], n), pattern=_py(args[1]))


@scalar("endswith")
def _endswith(args, n, ctx):
    if _arrayish(args[1]):
        return _map_rows(lambda s, p: s.endswith(p), n, args[0], args[1], type=pa.bool_())
    return pc.ends_with(_string(args[0], n), pattern=_py(args[1]))


@scalar("contains")
def _contains(args, n, ctx):
    if _arrayish(args[1]):
        return _map_rows(lambda s, p: p in s, n, args[0], args[1], type=pa.bool_())
    return pc.match_substring(_string(args[0], n), pattern=_py(args[1]))


@scalar("split")
def _split(args, n, ctx):
    limit = int(_py(args[2])) if len(args) > 2 else -1
    return pc.split_pattern_regex(_string(args[0], n), pattern=_py(args[1]),
                                  max_splits=None if limit <= 0 else limit - 1)


@scalar("split_part")
def _split_part(args, n, ctx):
    delim, part = _py(args[1]), int(_py(args[2]))

    def one(s):
        parts = s.split(delim)
        idx = part - 1 if part > 0 else part
        return parts[idx] if -len(parts) <= idx < len(parts) else ""
    return _map_rows(one, n, args[0], type=pa.string())


@scalar("format_number")
def _format_number(args, n, ctx):
    d = int(_py(args[1]))
    return _map_rows(lambda v: f"{v:,.{d}f}", n, args[0], type=pa.string())


@scalar("format_string")
def _format_string(args, n, ctx):
    fmt = _py(args[0])
    cols = [_rows(a, n) for a in args[1:]]
    return pa.array([fmt % tuple(row) for row in zip(*cols)] if cols else [fmt] * n, type=pa.string())


@scalar("levenshtein")
def _levenshtein(args, n, ctx):
    def dist(a, b):
        prev = list(range(len(b) + 1))
        for i, ca in enumerate(a, 1):
            cur = [i]
            for j, cb in enumerate(b, 1):
                cur.append(min(prev[j] + 1, cur[j - 1] + 1, prev[j - 1] + (ca != cb)))
            prev = cur
        return prev[-1]
    return _map_rows(dist, n, args[0], args[1], type=pa.int32())


@scalar("soundex")
def _soundex(args, n, ctx):
    codes = {c: d for d, letters in {"1": "BFPV", "2": "CGJKQSXZ", "3": "DT", "4": "L", "5": "MN", "6": "R"}.items()
             for c in letters}

    def one(s):
        if not s or not s[0].isalpha():
            return s
        up = s.upper()
        out, last = up[0], codes.get(up[0], "")
        for ch in up[1:]:
            code = codes.get(ch, "")
            if code and code != last:
                out += code
            if ch not in "HW":
                last = code
        return (out + "000")[:4]
    return _map_rows(one, n, args[0], type=pa.string())


@scalar("ascii")
def _ascii(args, n, ctx):
    return _map_rows(lambda s: ord(s[0]) if s else 0, n, args[0], type=pa.int32())


@scalar("char_length", "character_length")
def _char_length(args, n, ctx):
    return pc.utf8_length(_string(args[0], n))


@scalar("octet_length")
def _octet_length(args, n, ctx):
    return pc.binary_length(_arr(args[0], n))


@scalar("bit_length")
def _bit_length(args, n, ctx):
    return pc.multiply(pc.binary_length(_arr(args[0], n)), 8)


@scalar("base64")
def _base64(args, n, ctx):
    return _map_rows(lambda v: _b64.b64encode(v.encode() if isinstance(v, str) else v).decode(), n, args[0],
                     type=pa.string())


@scalar("unbase64")
def _unbase64(args, n, ctx):
    return _map_rows(lambda v: _b64.b64decode(v), n, args[0], type=pa.binary())


@scalar("encode")
def _encode(args, n, ctx):
    charset = _py(args[1])
    return _map_rows(lambda s: s.encode(charset), n, args[0], type=pa.binary())


@scalar("decode")
def _decode(args, n, ctx):
    charset = _py(args[1])
    return _map_rows(lambda b: b.decode(charset), n, args[0], type=pa.string())


def _bytes_of(v) -> bytes:
    return v if isinstance(v, (bytes, bytearray)) else str(v).encode()


@scalar("md5")
def _md5(args, n, ctx):
    return _map_rows(lambda v: hashlib.md5(_bytes_of(v)).hexdigest(), n, args[0], type=pa.string())


@scalar("sha1", "sha")
def _sha1(args, n, ctx):
    return _map_rows(lambda v: hashlib.sha1(_bytes_of(v)).hexdigest(), n, args[0], type=pa.string())


@scalar("sha2")
def _sha2(args, n, ctx):
    bits = int(_py(args[1])) or 256
    if bits not in (224, 256, 384, 512):
        return pa.nulls(n, pa.string())
    return _map_rows(lambda v: hashlib.new(f"sha{bits}", _bytes_of(v)).hexdigest(), n, args[0], type=pa.string())


@scalar("crc32")
def _crc32(args, n, ctx):
    return _map_rows(lambda v: zlib.crc32(_bytes_of(v)) & 0xFFFFFFFF, n, args[0], type=pa.int64())


@scalar("xxhash64")
def _xxhash64(args, n, ctx):
    """A stable 64-bit hash of the argument columns (not bit-compatible with Spark's xxHash64 seed 42)."""
    acc = np.full(n, 42, dtype=np.uint64)
    for a in args:
        rows = _rows(a, n)
        v = np.array([int.from_bytes(hashlib.blake2b(repr(x).encode(), digest_size=8).digest(), "little")
                      for x in rows], dtype=np.uint64)
        acc = (acc ^ v) * np.uint64(0x9E3779B97F4A7C15)
        acc ^= acc >> np.uint64(32)
    return pa.array(acc.view(np.int64))


@scalar("overlay")
def _overlay(args, n, ctx):
    pos = int(_py(args[2]))
    length = int(_py(args[3])) if len(args) > 3 and _py(args[3]) is not None else -1

    def one(s, r):
        ln = len(r) if length < 0 else length
        return s[: pos - 1] + r + s[pos - 1 + ln:]
    return _map_rows(one, n, args[0], args[1], type=pa.string())


@scalar("sentences")
def _sentences(args, n, ctx):
    return _map_rows(lambda s: [re.findall(r"[\w']+", part) for part in re.split(r"[.!?]+", s) if part.strip()],
                     n, args[0], type=pa.list_(pa.list_(pa.string())))


# ------------------------------------------------------------------------------------------ dates
def _as_ts(v, n: int, ctx) -> pa.Array:
    from raydp_b200.sql.column import _ts
    a = _arr(v, n)
    if pa.types.is_string(a.type) or pa.types.is_large_string(a.type):
        full = _ts(a, ctx)
        day = pc.strptime(pc.utf8_trim_whitespace(a), format="%Y-%m-%d", unit="us", error_is_null=True)
        return pc.coalesce(full, day)
    return _ts(a, ctx)


def _strftime(ts: pa.Array, fmt: str) -> pa.Array:
    """Arrow prints ``%S`` with the unit's fraction digits: pick the unit the pattern asks for."""
    if ".%f" in fmt or "%f" in fmt:
        return pc.strftime(pc.cast(ts, pa.timestamp("ms"), safe=False), format=fmt.replace(".%f", "").replace("%f", ""))
    return pc.strftime(pc.cast(ts, pa.timestamp("s"), safe=False), format=fmt)


def _as_date(v, n: int, ctx) -> pa.Array:
    a = _arr(v, n)
    if pa.types.is_date32(a.type):
        return a
    if pa.types.is_string(a.type) or pa.types.is_large_string(a.type):
        head = pc.utf8_slice_codeunits(pc.utf8_trim_whitespace(a), 0, 10)
        return pc.cast(pc.strptime(head, format="%Y-%m-%d", unit="s", error_is_null=True), pa.date32())
    return pc.cast(_as_ts(a, n, ctx), pa.date32())


_JAVA_TO_STRFTIME = [("yyyy", "%Y"), ("yy", "%y"), ("MMMM", "%B"), ("MMM", "%b"), ("MM", "%m"), ("M", "%-m"),
                     ("dd", "%d"), ("d", "%-d"), ("HH", "%H"), ("H", "%-H"), ("hh", "%I"), ("h", "%-I"),
                     ("mm", "%M"), ("m", "%-M"), ("ss", "%S"), ("s", "%-S"), ("SSS", "%f"), ("a", "%p"),
                     ("EEEE", "%A"), ("EEE", "%a"), ("E", "%a"), ("DDD", "%j"), ("D", "%-j"), ("XXX", "%z"),
                     ("Z", "%z"), ("z", "%Z")]


def java_datetime_pattern(pattern: str) -> str:
    """``yyyy-MM-dd HH:mm:ss`` (Java ``DateTimeFormatter`` letters) -> ``%Y-%m-%d %H:%M:%S``."""
    out, i = [], 0
    while i < len(pattern):
        ch = pattern[i]
        if ch == "'":                                           # quoted literal
            j = pattern.find("'", i + 1)
            j = len(pattern) if j < 0 else j
            out.append(pattern[i + 1:j].replace("%", "%%"))
            i = j + 1
            continue
        j = i
        while j < len(pattern) and pattern[j] == ch:
            j += 1
        run = pattern[i:j]
        if ch.isalpha():
            for java, c_fmt in _JAVA_TO_STRFTIME:
                if run == java or (java[0] == ch and len(java) == 1 and len(run) == 1):
                    out.append(c_fmt)
                    break
            else:
                best = next((c for jv, c in _JAVA_TO_STRFTIME if jv[0] == ch), run)
                out.append(best)
        else:
            out.append(run.replace("%", "%%"))
        i = j
    return "".join(out)


@scalar("current_date", "curdate")
def _current_date(args, n, ctx):
    return pa.array([_dt.date.today()] * n, type=pa.date32())


@scalar("current_timestamp", "now", "localtimestamp")
def _current_timestamp(args, n, ctx):
    return pa.array([_dt.datetime.utcnow()] * n, type=pa.timestamp("us"))


@scalar("to_date")
def _to_date(args, n, ctx):
    fmt = _py(args[1]) if len(args) > 1 and args[1] is not None else None
    if fmt is not None:
        a = _string(args[0], n)
        return pc.cast(pc.strptime(a, format=java_datetime_pattern(fmt), unit="s", error_is_null=True), pa.date32())
    return _as_date(args[0], n, ctx)


@scalar("to_timestamp_fmt")
def _to_timestamp_fmt(args, n, ctx):
    return pc.strptime(_string(args[0], n), format=java_datetime_pattern(_py(args[1])), unit="us",
                       error_is_null=True)


@scalar("date_add")
def _date_add(args, n, ctx):
    d = pc.cast(_as_date(args[0], n, ctx), pa.int32())
    return pc.cast(pc.add(d, pc.cast(_arr(args[1], n), pa.int32())), pa.date32())


@scalar("date_sub")
def _date_sub(args, n, ctx):
    d = pc.cast(_as_date(args[0], n, ctx), pa.int32())
    return pc.cast(pc.subtract(d, pc.cast(_arr(args[1], n), pa.int32())), pa.date32())


@scalar("datediff", "date_diff")
def _datediff(args, n, ctx):
    end, start = (pc.cast(_as_date(a, n, ctx), pa.int32()) for a in args[:2])
    return pc.subtract(end, start)


def _add_months_py(d: _dt.date, k: int) -> _dt.date:
    import calendar
    m = d.month - 1 + k
    y, m = d.year + m // 12, m % 12 + 1
    return _dt.date(y, m, min(d.day, calendar.monthrange(y, m)[1]))


@scalar("add_months")
def _add_months(args, n, ctx):
    return _map_rows(_add_months_py, n, _as_date(args[0], n, ctx), args[1], type=pa.date32())


@scalar("months_between")
def _months_between(args, n, ctx):
    import calendar
    round_off = bool(_py(args[2])) if len(args) > 2 else True

    def one(a, b):
        last_a = a.day == calendar.monthrange(a.year, a.month)[1]
        last_b = b.day == calendar.monthrange(b.year, b.month)[1]
        months = (a.year - b.year) * 12 + (a.month - b.month)
        if a.day == b.day or (last_a and last_b):
            return float(months)
        sec_a = a.day * 86400 + a.hour * 3600 + a.minute * 60 + a.second
        sec_b = b.day * 86400 + b.hour * 3600 + b.minute * 60 + b.second
        out = months + (sec_a - sec_b) / (31 * 86400.0)
        return round(out, 8) if round_off else out
    return _map_rows(one, n, _as_ts(args[0], n, ctx), _as_ts(args[1], n, ctx), type=pa.float64())


@scalar("last_day")
def _last_day(args, n, ctx):
    import calendar
    return _map_rows(lambda d: d.replace(day=calendar.monthrange(d.year, d.month)[1]), n,
                     _as_date(args[0], n, ctx), type=pa.date32())


@scalar("next_day")
def _next_day(args, n, ctx):
    names = ["mo", "tu", "we", "th", "fr", "sa", "su"]
    want = names.index(str(_py(args[1])).lower()[:2])
    return _map_rows(lambda d: d + _dt.timedelta(days=((want - d.weekday() - 1) % 7) + 1), n,
                     _as_date(args[0], n, ctx), type=pa.date32())


@scalar("date_format")
def _date_format(args, n, ctx):
    return _strftime(_as_ts(args[0], n, ctx), java_datetime_pattern(_py(args[1])))


_TRUNC_UNITS = {"year": "year", "yyyy": "year", "yy": "year", "quarter": "quarter", "month": "month", "mon": "month",
                "mm": "month", "week": "week", "day": "day", "dd": "day", "hour": "hour", "minute": "minute",
                "second": "second", "millisecond": "millisecond", "microsecond": "microsecond"}


@scalar("date_trunc")
def _date_trunc(args, n, ctx):        # date_trunc(format, timestamp)
    unit = _TRUNC_UNITS.get(str(_py(args[0])).lower())
    if unit is None:
        return pa.nulls(n, pa.timestamp("us"))
    return pc.floor_temporal(_as_ts(args[1], n, ctx), unit=unit, week_starts_monday=True)


@scalar("trunc")
def _trunc(args, n, ctx):             # trunc(date, format)
    unit = _TRUNC_UNITS.get(str(_py(args[1])).lower())
    if unit not in ("year", "quarter", "month", "week"):
        return pa.nulls(n, pa.date32())
    return pc.cast(pc.floor_temporal(pc.cast(_as_date(args[0], n, ctx), pa.timestamp("us")), unit=unit,
                                     week_starts_monday=True), pa.date32())


@scalar("from_unixtime")
def _from_unixtime(args, n, ctx):
    fmt = _py(args[1]) if len(args) > 1 and args[1] is not None else "yyyy-MM-dd HH:mm:ss"
    ts = pc.cast(pc.multiply(pc.cast(_arr(args[0], n), pa.int64()), 1000000), pa.timestamp("us"))
    return _strftime(ts, java_datetime_pattern(fmt))


@scalar("timestamp_seconds")
def _timestamp_seconds(args, n, ctx):
    micros = pc.cast(pc.multiply(_f64(_arr(args[0], n)), 1e6), pa.int64())
    return pc.cast(micros, pa.timestamp("us"))


@scalar("make_date")
def _make_date(args, n, ctx):
    def one(y, m, d):
        try:
            return _dt.date(y, m, d)
        except ValueError:
            return None
    return _map_rows(one, n, args[0], args[1], args[2], type=pa.date32())


def _shift_tz(args, n, ctx, to_utc: bool):
    ts, tz = _as_ts(args[0], n, ctx), _py(args[1])
    if to_utc:          # the wall clock is in ``tz``: find the instant
        return pc.cast(pc.assume_timezone(ts, tz, ambiguous="earliest", nonexistent="earliest"), pa.timestamp("us"))
    return pc.local_timestamp(pc.cast(pc.assume_timezone(ts, "UTC"), pa.timestamp("us", tz)))


@scalar("to_utc_timestamp")
def _to_utc_timestamp(args, n, ctx):
    return _shift_tz(args, n, ctx, True)


@scalar("from_utc_timestamp")
def _from_utc_timestamp(args, n, ctx):
    return _shift_tz(args, n, ctx, False)


@scalar("window_start")
def _window_start(args, n, ctx):
    """Start of the tumbling window of ``duration`` microseconds a timestamp falls into (``F.window``)."""
    width = int(_py(args[1]))
    micros = pc.cast(_as_ts(args[0], n, ctx), pa.int64()).to_numpy(zero_copy_only=False)
    mask = np.asarray(pc.is_null(_arr(args[0], n)))
    start = (np.where(mask, 0, micros).astype(np.int64) // width) * width + (int(_py(args[2])) if len(args) > 2 else 0)
    return pc.cast(pa.array(start, mask=mask), pa.timestamp("us"))


# ------------------------------------------------------------------------------------------ arrays
def _is_list(t: pa.DataType) -> bool:
    return pa.types.is_list(t) or pa.types.is_large_list(t) or pa.types.is_fixed_size_list(t)


def _list_rows(fn: Callable, n: int, lists, *others, type=None) -> pa.Array:  # noqa: A002
    a = _arr(lists, n)
    return _map_rows(fn, n, a, *others, type=type if type is not None else (a.type if _is_list(a.type) else None))


def _dedup(values: list) -> list:
    seen, out = [], []
    for v in values:
        key = ("__none__",) if v is None else v
        if key not in seen:
            seen.append(key)
            out.append(v)
    return out


@scalar("size", "cardinality", "array_size")
def _size(args, n, ctx):
    a = _arr(args[0], n)
    if pa.types.is_map(a.type):
        return pa.array([-1 if v is None else len(v) for v in a.to_pylist()], type=pa.int32())
    return pc.cast(pc.fill_null(pc.list_value_length(a), -1), pa.int32())


@scalar("array_contains")
def _array_contains(args, n, ctx):
    return _map_rows(lambda lst, v: v in lst, n, args[0], args[1], type=pa.bool_())


@scalar("array_distinct")
def _array_distinct(args, n, ctx):
    return _list_rows(_dedup, n, args[0])


@scalar("array_join")
def _array_join(args, n, ctx):
    sep = _py(args[1])
    null_rep = _py(args[2]) if len(args) > 2 else None

    def one(lst):
        parts = [(null_rep if v is None else str(v)) for v in lst]
        return sep.join(p for p in parts if p is not None)
    return _map_rows(one, n, args[0], type=pa.string())


def _elem_type(a: pa.Array) -> pa.DataType:
    return a.type.value_type if _is_list(a.type) else pa.null()


@scalar("array_max")
def _array_max(args, n, ctx):
    a = _arr(args[0], n)
    return _map_rows(lambda lst: max((v for v in lst if v is not None), default=None), n, a, type=_elem_type(a))


@scalar("array_min")
def _array_min(args, n, ctx):
    a = _arr(args[0], n)
    return _map_rows(lambda lst: min((v for v in lst if v is not None), default=None), n, a, type=_elem_type(a))


def _sorted_nulls(lst: list, asc: bool, nulls_first: bool) -> list:
    vals = sorted((v for v in lst if v is not None), reverse=not asc)
    nulls = [None] * (len(lst) - len(vals))
    return nulls + vals if nulls_first else vals + nulls


@scalar("array_sort")
def _array_sort(args, n, ctx):
    return _list_rows(lambda lst: _sorted_nulls(lst, True, False), n, args[0])


@scalar("sort_array")
def _sort_array(args, n, ctx):
    asc = bool(_py(args[1])) if len(args) > 1 else True
    return _list_rows(lambda lst: _sorted_nulls(lst, asc, asc), n, args[0])


@scalar("element_at", "try_element_at")
def _element_at(args, n, ctx):
    a = _arr(args[0], n)
    if pa.types.is_map(a.type):
        return _map_rows(lambda m, k: dict(m).get(k), n, a, args[1], type=a.type.item_type)

    def one(lst, i):
        if i == 0:
            raise ValueError("SQL array indices start at 1")
        i = i - 1 if i > 0 else i
        return lst[i] if -len(lst) <= i < len(lst) else None
    return _map_rows(one, n, a, args[1], type=_elem_type(a))


@scalar("get_item")
def _get_item(args, n, ctx):
    """``col[key]`` / ``col.getItem(key)``: 0-based for arrays, key lookup for maps, field for structs."""
    a, key = _arr(args[0], n), args[1]
    if pa.types.is_struct(a.type):
        return pc.struct_field(a, _py(key))
    if pa.types.is_map(a.type):
        return _map_rows(lambda m, k: dict(m).get(k), n, a, key, type=a.type.item_type)
    return _map_rows(lambda lst, i: lst[i] if 0 <= i < len(lst) else None, n, a, key, type=_elem_type(a))


@scalar("get_field")
def _get_field(args, n, ctx):
    return pc.struct_field(_arr(args[0], n), _py(args[1]))


@scalar("flatten")
def _flatten(args, n, ctx):
    a = _arr(args[0], n)
    return _map_rows(lambda lst: None if any(x is None for x in lst) else [v for sub in lst for v in sub], n, a,
                     type=a.type.value_type if _is_list(a.type) else None)


@scalar("slice")
def _slice(args, n, ctx):
    def one(lst, start, length):
        if start == 0:
            raise ValueError("SQL array indices start at 1")
        begin = start - 1 if start > 0 else max(len(lst) + start, 0)
        return lst[begin: begin + length]
    return _list_rows(one, n, args[0], args[1], args[2])


@scalar("array_position")
def _array_position(args, n, ctx):
    return _map_rows(lambda lst, v: lst.index(v) + 1 if v in lst else 0, n, args[0], args[1], type=pa.int64())


@scalar("array_remove")
def _array_remove(args, n, ctx):
    return _list_rows(lambda lst, v: [x for x in lst if x != v], n, args[0], args[1])


@scalar("array_union")
def _array_union(args, n, ctx):
    return _list_rows(lambda a, b: _dedup(a + b), n, args[0], args[1])


@scalar("array_intersect")
def _array_intersect(args, n, ctx):
    return _list_rows(lambda a, b: _dedup([x for x in a if x in b]), n, args[0], args[1])


@scalar("array_except")
def _array_except(args, n, ctx):
    return _list_rows(lambda a, b: _dedup([x for x in a if x not in b]), n, args[0], args[1])


@scalar("arrays_overlap")
def _arrays_overlap(args, n, ctx):
    return _map_rows(lambda a, b: any(x in b for x in a if x is not None), n, args[0], args[1], type=pa.bool_())


@scalar("array_append")
def _array_append(args, n, ctx):
    a = _arr(args[0], n)
    vals = _rows(args[1], n)
    return pa.array([None if lst is None else lst + [v] for lst, v in zip(a.to_pylist(), vals)], type=a.type)


@scalar("array_prepend")
def _array_prepend(args, n, ctx):
    a = _arr(args[0], n)
    vals = _rows(args[1], n)
    return pa.array([None if lst is None else [v] + lst for lst, v in zip(a.to_pylist(), vals)], type=a.type)


@scalar("array_insert")
def _array_insert(args, n, ctx):
    a = _arr(args[0], n)
    pos, vals = _rows(args[1], n), _rows(args[2], n)
    out = []
    for lst, p, v in zip(a.to_pylist(), pos, vals):
        if lst is None or p is None:
            out.append(None)
            continue
        idx = p - 1 if p > 0 else len(lst) + p + 1
        padded = lst + [None] * max(0, idx - len(lst))
        out.append(padded[:max(idx, 0)] + [v] + padded[max(idx, 0):])
    return pa.array(out, type=a.type)


@scalar("array_compact")
def _array_compact(args, n, ctx):
    return _list_rows(lambda lst: [x for x in lst if x is not None], n, args[0])


@scalar("array_repeat")
def _array_repeat(args, n, ctx):
    vals, counts = _rows(args[0], n), _rows(args[1], n)
    return pa.array([None if c is None else [v] * max(c, 0) for v, c in zip(vals, counts)])


@scalar("arrays_zip")
def _arrays_zip(args, n, ctx):
    names = [str(i) for i in range(len(args))]
    cols = [_rows(a, n) for a in args]
    out = []
    for row in zip(*cols):
        if any(v is None for v in row):
            out.append(None)
            continue
        width = max(len(v) for v in row)
        out.append([{nm: (v[i] if i < len(v) else None) for nm, v in zip(names, row)} for i in range(width)])
    fields = [pa.field(nm, _elem_type(_arr(a, n))) for nm, a in zip(names, args)]
    return pa.array(out, type=pa.list_(pa.struct(fields)))


@scalar("sequence")
def _sequence(args, n, ctx):
    def one(start, stop, step=None):
        step = step if step is not None else (1 if stop >= start else -1)
        return list(range(start, stop + (1 if step > 0 else -1), step))
    return _map_rows(one, n, *args, type=pa.list_(pa.int64()))


@scalar("shuffle")
def _shuffle(args, n, ctx):
    rng = np.random.default_rng([0x5eed, ctx.partition_index])
    return _list_rows(lambda lst: [lst[i] for i in rng.permutation(len(lst))], n, args[0])


# ------------------------------------------------------------------------------------------ structs, maps
@scalar("struct")
def _struct(args, n, ctx):
    names, values = list(args[0]), [_arr(a, n) for a in args[1:]]
    return pa.StructArray.from_arrays(values, names=names)


@scalar("create_map")
def _create_map(args, n, ctx):
    cols = [_rows(a, n) for a in args]
    keys, items = cols[0::2], cols[1::2]
    out = [[(k[i], v[i]) for k, v in zip(keys, items)] for i in range(n)]
    key_t, item_t = _arr(args[0], n).type, _arr(args[1], n).type
    return pa.array(out, type=pa.map_(key_t, item_t))


@scalar("map_from_arrays")
def _map_from_arrays(args, n, ctx):
    k, v = _arr(args[0], n), _arr(args[1], n)
    return _map_rows(lambda ks, vs: list(zip(ks, vs)), n, k, v, type=pa.map_(_elem_type(k), _elem_type(v)))


@scalar("map_from_entries")
def _map_from_entries(args, n, ctx):
    a = _arr(args[0], n)
    st = a.type.value_type
    return _map_rows(lambda ents: [tuple(e.values()) for e in ents], n, a,
                     type=pa.map_(st.field(0).type, st.field(1).type))


@scalar("map_keys")
def _map_keys(args, n, ctx):
    a = _arr(args[0], n)
    return _map_rows(lambda m: [k for k, _ in m], n, a, type=pa.list_(a.type.key_type))


@scalar("map_values")
def _map_values(args, n, ctx):
    a = _arr(args[0], n)
    return _map_rows(lambda m: [v for _, v in m], n, a, type=pa.list_(a.type.item_type))


@scalar("map_entries")
def _map_entries(args, n, ctx):
    a = _arr(args[0], n)
    st = pa.struct([pa.field("key", a.type.key_type), pa.field("value", a.type.item_type)])
    return _map_rows(lambda m: [{"key": k, "value": v} for k, v in m], n, a, type=pa.list_(st))


@scalar("map_concat")
def _map_concat(args, n, ctx):
    a = _arr(args[0], n)
    return _map_rows(lambda *ms: list({k: v for m in ms for k, v in m}.items()), n, *args, type=a.type)


@scalar("map_contains_key")
def _map_contains_key(args, n, ctx):
    return _map_rows(lambda m, k: any(kk == k for kk, _ in m), n, args[0], args[1], type=pa.bool_())


# ------------------------------------------------------------------------------------------ JSON
def _json_path(doc, path: str):
    if not path.startswith("$"):
        return None
    cur = doc
    for name, index in re.findall(r"\.([^.\[]+)|\[(\d+|\*)\]", path[1:]):
        if cur is None:
            return None
        if name:
            cur = cur.get(name) if isinstance(cur, dict) else None
        elif index == "*":
            continue
        else:
            i = int(index)
            cur = cur[i] if isinstance(cur, list) and i < len(cur) else None
    return cur


@scalar("get_json_object")
def _get_json_object(args, n, ctx):
    path = _py(args[1])

    def one(s):
        try:
            v = _json_path(json.loads(s), path)
        except ValueError:
            return None
        if v is None:
            return None
        return v if isinstance(v, str) else json.dumps(v, separators=(",", ":"))
    return _map_rows(one, n, args[0], type=pa.string())


@scalar("json_tuple_field")
def _json_tuple_field(args, n, ctx):
    return _get_json_object([args[0], "$." + str(_py(args[1]))], n, ctx)


def _coerce_json(value, t: pa.DataType):
    if value is None:
        return None
    try:
        if pa.types.is_struct(t):
            return {f.name: _coerce_json(value.get(f.name), f.type) for f in t} if isinstance(value, dict) else None
        if _is_list(t):
            return [_coerce_json(v, t.value_type) for v in value] if isinstance(value, list) else None
        if pa.types.is_map(t):
            return [(k, _coerce_json(v, t.item_type)) for k, v in value.items()] if isinstance(value, dict) else None
        if pa.types.is_integer(t):
            return int(value)
        if pa.types.is_floating(t):
            return float(value)
        if pa.types.is_boolean(t):
            return bool(value)
        if pa.types.is_string(t):
            return value if isinstance(value, str) else json.dumps(value, separators=(",", ":"))
        if pa.types.is_timestamp(t):
            return _dt.datetime.fromisoformat(str(value).replace("Z", ""))
        if pa.types.is_date(t):
            return _dt.date.fromisoformat(str(value)[:10])
    except (TypeError, ValueError):
        return None
    return value


@scalar("from_json")
def _from_json(args, n, ctx):
    target = args[1]            # an Arrow type prepared by functions.from_json

    def one(s):
        try:
            return _coerce_json(json.loads(s), target)
        except ValueError:
            return None
    return _map_rows(one, n, args[0], type=target)


def _jsonable(v):
    if isinstance(v, dict):
        return {k: _jsonable(x) for k, x in v.items() if x is not None}
    if isinstance(v, list):
        if v and all(isinstance(x, tuple) and len(x) == 2 for x in v):
            return {str(k): _jsonable(x) for k, x in v}
        return [_jsonable(x) for x in v]
    if isinstance(v, _dt.datetime):
        return v.isoformat(timespec="milliseconds")
    if isinstance(v, _dt.date):
        return v.isoformat()
    if isinstance(v, (bytes, bytearray)):
        return _b64.b64encode(v).decode()
    return v


@scalar("to_json")
def _to_json(args, n, ctx):
    return _map_rows(lambda v: json.dumps(_jsonable(v), separators=(",", ":")), n, args[0], type=pa.string())


@scalar("schema_of_json")
def _schema_of_json(args, n, ctx):
    from raydp_b200.sql import types as T
    sample = pa.array([json.loads(_py(args[0]))])
    return pa.array([T.from_arrow_type(sample.type).simpleString()] * n)


# ------------------------------------------------------------------------------------------ higher-order
def _list_layout(a: pa.Array):
    """``(values, parent_row_of_each_value, position_in_its_list, lengths, null_mask)`` of a list column."""
    n = len(a)
    null_mask = np.asarray(pc.is_null(a))
    lengths = np.asarray(pc.fill_null(pc.list_value_length(a), 0)).astype(np.int64)
    values = pc.list_flatten(a)
    parent = np.repeat(np.arange(n, dtype=np.int64), lengths)
    starts = np.concatenate([[0], np.cumsum(lengths)[:-1]]) if n else np.zeros(0, dtype=np.int64)
    pos = np.arange(len(parent), dtype=np.int64) - np.repeat(starts, lengths)
    return values, parent, pos, lengths, null_mask


def _assemble(values: pa.Array, lengths: np.ndarray, null_mask: np.ndarray) -> pa.Array:
    offsets = np.concatenate([[0], np.cumsum(lengths)]).astype(np.int32)
    return pa.ListArray.from_arrays(pa.array(offsets), values, mask=pa.array(null_mask))


def _inner_table(table: pa.Table, parent: np.ndarray, variables: Dict[str, pa.Array]) -> pa.Table:
    inner = table.take(pa.array(parent)) if table.num_columns else pa.table({})
    for name, arr in variables.items():
        if name in inner.column_names:
            inner = inner.drop_columns([name])
        inner = inner.append_column(name, arr)
    return inner


def eval_higher_order(name: str, arrays: list, body, variables: List[str], extra, table: pa.Table, ctx):
    """``transform / filter / exists / forall / zip_with / aggregate / map_filter / transform_values ...``:
    the lambda body is an ordinary expression over placeholder columns, evaluated once over the flattened
    elements (outer columns are repeated per element, so the body may use them)."""
    from raydp_b200.sql.column import _full
    n = table.num_rows
    a = _arr(arrays[0], n)
    if name == "aggregate":
        merge, finish = body, extra.get("finish")
        acc_var, x_var = variables
        values, parent, pos, lengths, null_mask = _list_layout(a)
        starts = np.concatenate([[0], np.cumsum(lengths)[:-1]]) if n else np.zeros(0, dtype=np.int64)
        acc = _arr(_full(extra["initial"].evaluate(table, ctx), n), n)
        for j in range(int(lengths.max()) if n else 0):
            live = lengths > j
            idx = np.where(live, starts + j, 0)
            x = values.take(pa.array(idx)) if len(values) else pa.nulls(n, values.type)
            x = pc.if_else(pa.array(live), x, pa.nulls(n, values.type))
            inner = _inner_table(table, np.arange(n), {acc_var: acc, x_var: x})
            new_acc = _arr(_full(merge.evaluate(inner, ctx), n), n)
            if new_acc.type != acc.type:
                acc = pc.cast(acc, new_acc.type)
            acc = pc.if_else(pa.array(live), new_acc, acc)
        if finish is not None:
            fin_var = extra["finish_var"]
            acc = _arr(_full(finish.evaluate(_inner_table(table, np.arange(n), {fin_var: acc}), ctx), n), n)
        return pc.if_else(pa.array(null_mask), pa.nulls(n, acc.type), acc)
    if name == "zip_with":
        b = _arr(arrays[1], n)
        la, lb = a.to_pylist(), b.to_pylist()
        null_mask = np.array([x is None or y is None for x, y in zip(la, lb)], dtype=bool)
        lengths = np.array([0 if m else max(len(x), len(y)) for x, y, m in zip(la, lb, null_mask)], dtype=np.int64)
        xs = [v for x, ln, m in zip(la, lengths, null_mask) if not m for v in x + [None] * (ln - len(x))]
        ys = [v for y, ln, m in zip(lb, lengths, null_mask) if not m for v in y + [None] * (ln - len(y))]
        parent = np.repeat(np.arange(n, dtype=np.int64), lengths)
        inner = _inner_table(table, parent, {variables[0]: pa.array(xs, type=_elem_type(a)),
                                             variables[1]: pa.array(ys, type=_elem_type(b))})
        out = _arr(_full(body.evaluate(inner, ctx), len(parent)), len(parent))
        return _assemble(out, lengths, null_mask)
    if pa.types.is_map(a.type):
        return _eval_map_lambda(name, a, body, variables, table, ctx)
    values, parent, pos, lengths, null_mask = _list_layout(a)
    bound = {variables[0]: values}
    if len(variables) > 1:
        bound[variables[1]] = pa.array(pos.astype(np.int32))
    inner = _inner_table(table, parent, bound)
    m = len(parent)
    out = _arr(_full(body.evaluate(inner, ctx), m), m)
    if name == "transform":
        return _assemble(out, lengths, null_mask)
    truth = np.asarray(pc.fill_null(out, False)).astype(bool)
    if name == "filter":
        return _assemble(values.filter(pa.array(truth)), np.bincount(parent[truth], minlength=n), null_mask)
    if name == "exists":
        return pa.array(np.bincount(parent[truth], minlength=n) > 0, mask=null_mask)
    if name == "forall":
        return pa.array(np.bincount(parent[~truth], minlength=n) == 0, mask=null_mask)
    raise NotImplementedError(f"higher-order function {name!r}")


def _eval_map_lambda(name: str, a: pa.Array, body, variables: List[str], table: pa.Table, ctx):
    from raydp_b200.sql.column import _full
    n = len(a)
    null_mask = np.asarray(pc.is_null(a))
    rows = a.to_pylist()
    lengths = np.array([0 if r is None else len(r) for r in rows], dtype=np.int64)
    keys = pa.array([k for r in rows if r is not None for k, _ in r], type=a.type.key_type)
    items = pa.array([v for r in rows if r is not None for _, v in r], type=a.type.item_type)
    parent = np.repeat(np.arange(n, dtype=np.int64), lengths)
    inner = _inner_table(table, parent, {variables[0]: keys, variables[1]: items})
    m = len(parent)
    out = _arr(_full(body.evaluate(inner, ctx), m), m)
    if name == "map_filter":
        truth = np.asarray(pc.fill_null(out, False)).astype(bool)
        keys, items, lengths = keys.filter(pa.array(truth)), items.filter(pa.array(truth)), \
            np.bincount(parent[truth], minlength=n)
    elif name == "transform_values":
        items = out
    elif name == "transform_keys":
        keys = out
    else:
        raise NotImplementedError(f"higher-order function {name!r} over a map")
    offsets = np.concatenate([[0], np.cumsum(lengths)]).astype(np.int32)
    result = pa.MapArray.from_arrays(pa.array(offsets), keys, items)
    if null_mask.any():
        return pc.if_else(pa.array(null_mask), pa.nulls(n, result.type), result)
    return result


@scalar("with_field")
def _with_field(args, n, ctx):
    a, name, value = _arr(args[0], n), args[1], _arr(args[2], n)
    names = [f.name for f in a.type]
    cols = [pc.struct_field(a, nm) for nm in names]
    if name in names:
        cols[names.index(name)] = value
    else:
        names.append(name)
        cols.append(value)
    return pa.StructArray.from_arrays(cols, names=names, mask=pc.is_null(a))


@scalar("drop_fields")
def _drop_fields(args, n, ctx):
    a, drop = _arr(args[0], n), set(args[1])
    names = [f.name for f in a.type if f.name not in drop]
    return pa.StructArray.from_arrays([pc.struct_field(a, nm) for nm in names], names=names, mask=pc.is_null(a))


@scalar("add_interval")
def _add_interval(args, n, ctx):
    """``date_or_timestamp + INTERVAL``: months move the calendar month (clamped to its last day), days and
    the sub-day part are exact.  A date stays a date unless the interval has a time part."""
    months, days, micros = int(_py(args[1])), int(_py(args[2])), int(_py(args[3]))
    a = _arr(args[0], n)
    as_date = pa.types.is_date(a.type)
    if as_date and micros == 0:
        out = a
        if months:
            out = _map_rows(_add_months_py, n, out, pa.scalar(months), type=pa.date32())
        if days:
            out = pc.cast(pc.add(pc.cast(out, pa.int32()), pa.scalar(days, pa.int32())), pa.date32())
        return out
    ts = _as_ts(a, n, ctx)
    if months:
        def shift(t):
            d = _add_months_py(t.date(), months)
            return t.replace(year=d.year, month=d.month, day=d.day)
        ts = _map_rows(shift, n, ts, type=pa.timestamp("us"))
    delta = days * 86400 * 10 ** 6 + micros
    if delta:
        ts = pc.cast(pc.add(pc.cast(ts, pa.int64()), delta), pa.timestamp("us"))
    return ts


@scalar("sliding_windows")
def _sliding_windows(args, n, ctx):
    """Every window ``[start, start + width)`` on the ``slide`` grid that contains the timestamp, as
    ``array<struct<start, end>>`` (earliest first); ``F.window`` explodes it."""
    width, slide = int(_py(args[1])), int(_py(args[2]))
    offset = int(_py(args[3])) if len(args) > 3 else 0
    micros = pc.cast(_as_ts(args[0], n, ctx), pa.int64()).to_numpy(zero_copy_only=False)
    mask = np.asarray(pc.is_null(_arr(args[0], n)))
    t = np.where(mask, 0, micros).astype(np.int64)
    last = ((t - offset) // slide) * slide + offset                    # latest window start <= t
    per_row = -(-width // slide)                                          # ceil(width / slide) candidates
    starts = last[:, None] - np.arange(per_row - 1, -1, -1, dtype=np.int64)[None, :] * slide
    inside = (starts + width > t[:, None]) & ~mask[:, None]
    lengths = inside.sum(axis=1)
    flat = starts[inside]
    offsets = np.concatenate([[0], np.cumsum(lengths)]).astype(np.int32)
    structs = pa.StructArray.from_arrays([pc.cast(pa.array(flat), pa.timestamp("us")),
                                          pc.cast(pa.array(flat + width), pa.timestamp("us"))], names=["start", "end"])
    return pa.ListArray.from_arrays(pa.array(offsets), structs, mask=pa.array(mask))


# ------------------------------------------------------------------------------------------ second batch
@scalar("cot")
def _cot(args, n, ctx):
    return pc.divide(1.0, pc.tan(_f64(args[0])))


@scalar("csc")
def _csc(args, n, ctx):
    return pc.divide(1.0, pc.sin(_f64(args[0])))


@scalar("sec")
def _sec(args, n, ctx):
    return pc.divide(1.0, pc.cos(_f64(args[0])))


@scalar("bit_count")
def _bit_count(args, n, ctx):
    return _map_rows(lambda v: bin(int(v) & 0xFFFFFFFFFFFFFFFF).count("1"), n, args[0], type=pa.int32())


@scalar("getbit", "bit_get")
def _getbit(args, n, ctx):
    return _map_rows(lambda v, p: (int(v) >> int(p)) & 1, n, args[0], args[1], type=pa.int8())


@scalar("shiftrightunsigned")
def _shiftrightunsigned(args, n, ctx):
    k = int(_py(args[1]))
    return _map_rows(lambda v: (int(v) & 0xFFFFFFFFFFFFFFFF) >> k, n, args[0], type=pa.int64())


@scalar("char", "chr")
def _char(args, n, ctx):
    return _map_rows(lambda v: chr(int(v) % 256), n, args[0], type=pa.string())


@scalar("elt")
def _elt(args, n, ctx):
    cols = [_rows(a, n) for a in args]
    out = []
    for row in zip(*cols):
        i = row[0]
        out.append(row[i] if i is not None and 1 <= i < len(row) else None)
    return pa.array(out)


@scalar("find_in_set")
def _find_in_set(args, n, ctx):
    return _map_rows(lambda s, lst: (lst.split(",").index(s) + 1) if s in lst.split(",") and "," not in s else 0, n,
                     args[0], args[1], type=pa.int32())


@scalar("position")
def _position(args, n, ctx):
    start = int(_py(args[2])) if len(args) > 2 else 1
    return _map_rows(lambda sub, s: s.find(sub, max(start - 1, 0)) + 1, n, args[0], args[1], type=pa.int32())


@scalar("replace")
def _replace_fn(args, n, ctx):
    repl = _py(args[2]) if len(args) > 2 else ""
    return pc.replace_substring(_string(args[0], n), pattern=_py(args[1]), replacement=repl)


@scalar("mask")
def _mask(args, n, ctx):
    up, low, dig, other = [(_py(args[i]) if len(args) > i else d) for i, d in ((1, "X"), (2, "x"), (3, "n"), (4, None))]

    def one(s):
        out = []
        for ch in s:
            if ch.isupper():
                out.append(up if up is not None else ch)
            elif ch.islower():
                out.append(low if low is not None else ch)
            elif ch.isdigit():
                out.append(dig if dig is not None else ch)
            else:
                out.append(other if other is not None else ch)
        return "".join(out)
    return _map_rows(one, n, args[0], type=pa.string())


@scalar("regexp_extract_all")
def _regexp_extract_all(args, n, ctx):
    rx, idx = re.compile(_py(args[1])), (int(_py(args[2])) if len(args) > 2 else 1)
    return _map_rows(lambda s: [m.group(idx) if rx.groups >= idx else m.group(0) for m in rx.finditer(s)], n, args[0],
                     type=pa.list_(pa.string()))


@scalar("regexp_instr")
def _regexp_instr(args, n, ctx):
    rx = re.compile(_py(args[1]))

    def one(s):
        m = rx.search(s)
        return 0 if m is None else m.start() + 1
    return _map_rows(one, n, args[0], type=pa.int32())


@scalar("str_to_map")
def _str_to_map(args, n, ctx):
    pair_delim = _py(args[1]) if len(args) > 1 else ","
    kv_delim = _py(args[2]) if len(args) > 2 else ":"

    def one(s):
        out = []
        for part in s.split(pair_delim):
            k, sep, v = part.partition(kv_delim)
            out.append((k, v if sep else None))
        return out
    return _map_rows(one, n, args[0], type=pa.map_(pa.string(), pa.string()))


@scalar("json_array_length")
def _json_array_length(args, n, ctx):
    def one(s):
        try:
            v = json.loads(s)
        except ValueError:
            return None
        return len(v) if isinstance(v, list) else None
    return _map_rows(one, n, args[0], type=pa.int32())


@scalar("json_object_keys")
def _json_object_keys(args, n, ctx):
    def one(s):
        try:
            v = json.loads(s)
        except ValueError:
            return None
        return list(v) if isinstance(v, dict) else None
    return _map_rows(one, n, args[0], type=pa.list_(pa.string()))


@scalar("to_csv")
def _to_csv(args, n, ctx):
    return _map_rows(lambda row: ",".join("" if v is None else str(v) for v in row.values()), n, args[0],
                     type=pa.string())


@scalar("uuid")
def _uuid(args, n, ctx):
    import uuid as _uuid_mod
    return pa.array([str(_uuid_mod.uuid4()) for _ in range(n)], type=pa.string())


@scalar("raise_error")
def _raise_error(args, n, ctx):
    if n:
        raise RuntimeError(str(_py(args[0]) if not _arrayish(args[0]) else _rows(args[0], n)[0]))
    return pa.nulls(0)


@scalar("width_bucket")
def _width_bucket(args, n, ctx):
    def one(v, lo, hi, k):
        if k <= 0 or lo == hi:
            return None
        if lo < hi:
            return 0 if v < lo else (k + 1 if v >= hi else int((v - lo) / (hi - lo) * k) + 1)
        return 0 if v > lo else (k + 1 if v <= hi else int((lo - v) / (lo - hi) * k) + 1)
    return _map_rows(one, n, args[0], args[1], args[2], args[3], type=pa.int64())


@scalar("date_from_unix_date")
def _date_from_unix_date(args, n, ctx):
    return pc.cast(pc.cast(_arr(args[0], n), pa.int32()), pa.date32())


@scalar("unix_date")
def _unix_date(args, n, ctx):
    return pc.cast(_as_date(args[0], n, ctx), pa.int32())


def _unix_scaled(divisor: int):
    def impl(args, n, ctx):
        micros = pc.cast(_as_ts(args[0], n, ctx), pa.int64())
        return micros if divisor == 1 else pc.cast(pc.floor(pc.divide(pc.cast(micros, pa.float64()), float(divisor))),
                                                   pa.int64())
    return impl


SCALAR["unix_micros"] = _unix_scaled(1)
SCALAR["unix_millis"] = _unix_scaled(1000)
SCALAR["unix_seconds"] = _unix_scaled(10 ** 6)


@scalar("timestamp_millis")
def _timestamp_millis(args, n, ctx):
    return pc.cast(pc.multiply(pc.cast(_arr(args[0], n), pa.int64()), 1000), pa.timestamp("us"))


@scalar("timestamp_micros")
def _timestamp_micros(args, n, ctx):
    return pc.cast(pc.cast(_arr(args[0], n), pa.int64()), pa.timestamp("us"))


@scalar("make_timestamp")
def _make_timestamp(args, n, ctx):
    def one(y, mo, d, h, mi, s):
        try:
            whole = int(s)
            return _dt.datetime(y, mo, d, h, mi, whole, int(round((s - whole) * 10 ** 6)))
        except ValueError:
            return None
    return _map_rows(one, n, *args[:6], type=pa.timestamp("us"))


@scalar("weekday")
def _weekday(args, n, ctx):          # Monday = 0
    return pc.cast(pc.day_of_week(_as_ts(args[0], n, ctx), count_from_zero=True, week_start=1), pa.int32())


@scalar("date_part")
def _date_part(args, n, ctx):
    field = str(_py(args[0])).lower()
    ts = _as_ts(args[1], n, ctx)
    table = {"year": pc.year, "yr": pc.year, "years": pc.year, "month": pc.month, "mon": pc.month, "day": pc.day,
             "d": pc.day, "hour": pc.hour, "h": pc.hour, "minute": pc.minute, "min": pc.minute, "quarter": pc.quarter,
             "qtr": pc.quarter, "week": pc.iso_week, "w": pc.iso_week, "doy": pc.day_of_year, "dayofyear": pc.day_of_year}
    if field in table:
        return pc.cast(table[field](ts), pa.int32())
    if field in ("second", "sec", "s", "seconds"):
        return pc.add(pc.cast(pc.second(ts), pa.float64()), pc.divide(pc.cast(pc.microsecond(ts), pa.float64()), 1e6))
    if field in ("dow", "dayofweek"):
        return pc.cast(pc.day_of_week(ts, count_from_zero=False, week_start=7), pa.int32())
    raise ValueError(f"date_part: unknown field {field!r}")


@scalar("convert_timezone")
def _convert_timezone(args, n, ctx):        # convert_timezone(source_tz, target_tz, ts)
    src, dst, ts = _py(args[0]), _py(args[1]), _as_ts(args[2], n, ctx)
    instant = pc.assume_timezone(ts, src, ambiguous="earliest", nonexistent="earliest")
    return pc.local_timestamp(pc.cast(instant, pa.timestamp("us", dst)))


@scalar("to_number")
def _to_number(args, n, ctx):
    def one(s):
        cleaned = re.sub(r"[,$\s]", "", s)
        try:
            return float(cleaned)
        except ValueError:
            return None
    return _map_rows(one, n, args[0], type=pa.float64())


def _try_arith(name: str):
    def impl(args, n, ctx):
        a, b = _arr(args[0], n), _arr(args[1], n)
        if name == "divide":
            fa, fb = _f64(a), _f64(b)
            safe = pc.if_else(pc.fill_null(pc.equal(fb, 0.0), False), pa.nulls(n, pa.float64()), fb)
            return pc.divide(fa, safe)
        try:
            return getattr(pc, name + "_checked")(a, b)
        except pa.ArrowInvalid:                                            # overflow somewhere: row by row
            op = {"add": lambda x, y: x + y, "subtract": lambda x, y: x - y, "multiply": lambda x, y: x * y}[name]

            def one(x, y):
                v = op(x, y)
                return v if isinstance(v, float) or -2 ** 63 <= v < 2 ** 63 else None
            return _map_rows(one, n, a, b)
    return impl


for _name in ("add", "subtract", "multiply", "divide"):
    SCALAR["try_" + _name] = _try_arith(_name)
