"""JavaScript number semantics needed by the host marshal layer.

The reference applies these ECMA-262 built-ins to raw Kubernetes / Prometheus strings before any
arithmetic happens; the marshal layer must reproduce them exactly so that the integers handed to
the CUDA kernels are the integers the reference would have summed:

  parseInt(v, 10)   src/api/k8s.ts:176,296; OverviewPage.tsx:98-99,110; NodesPage.tsx:173-174;
                    NodeDetailSection.tsx:74,86
  parseFloat(s)     src/api/metrics.ts:130,137
  Math.round        OverviewPage.tsx:116; NodesPage.tsx:37; NodeDetailSection.tsx:89;
                    MetricsPage.tsx:51; metrics.ts:167
  toFixed(1)        metrics.ts:162
  String(number)    k8s.ts:300
"""
from __future__ import annotations

import math
import re
from decimal import ROUND_HALF_UP, Decimal

__all__ = [
    "parse_int",
    "parse_float",
    "truthy",
    "strict_equal",
    "math_round",
    "number_to_string",
    "to_fixed_1",
    "EXACT_INT_LIMIT",
]

# StrWhiteSpaceChar: WhiteSpace (TAB VT FF SP NBSP ZWNBSP + Unicode Zs) and LineTerminator.
_WS = "\\t\\n\\v\\f\\r \\u00a0\\u1680\\u2000-\\u200a\\u2028\\u2029\\u202f\\u205f\\u3000\\ufeff"
_INT_RE = re.compile("^[" + _WS + "]*([+-]?)([0-9]+)")
_FLOAT_RE = re.compile(
    "^[" + _WS + "]*([+-]?)(Infinity|(?:[0-9]+\\.?[0-9]*|\\.[0-9]+)(?:[eE][+-]?[0-9]+)?)"
)

# Integers are carried as int64 on the device; they equal the reference's float64 sums only while
# every partial sum stays below 2^53.
EXACT_INT_LIMIT = 1 << 53


def _to_js_string(v) -> str:
    if isinstance(v, str):
        return v
    if v is None:
        return "undefined"
    if isinstance(v, bool):
        return "true" if v else "false"
    if isinstance(v, (int, float)):
        return number_to_string(float(v))
    return str(v)


def parse_int(v) -> float:
    """parseInt(v, 10) as a float64 (NaN when no digits follow the optional sign)."""
    m = _INT_RE.match(_to_js_string(v))
    if not m:
        return math.nan
    mag = float(int(m.group(2)))
    return -mag if m.group(1) == "-" else mag


def parse_float(v) -> float:
    """parseFloat(v): longest StrDecimalLiteral prefix, else NaN ('+Inf' and 'NaN' give NaN)."""
    m = _FLOAT_RE.match(_to_js_string(v))
    if not m:
        return math.nan
    body = m.group(2)
    if body == "Infinity":
        val = math.inf
    else:
        val = float(body if not body.endswith(".") else body + "0") if "." in body or "e" in body.lower() else float(int(body))
    return -val if m.group(1) == "-" else val


def truthy(v) -> bool:
    """ToBoolean for JSON-borne values."""
    if v is None or v is False:
        return False
    if isinstance(v, str):
        return v != ""
    if isinstance(v, (int, float)) and not isinstance(v, bool):
        return not (v == 0 or v != v)
    return True


def strict_equal(a, b) -> bool:
    """a === b for two values JSON.parse produced (PodsPage.tsx:69 `req === lim`): strings by content, numbers by
    value (1 === 1.0; JSON has no NaN), booleans and null by identity of the value; two containers (objects, arrays)
    are distinct objects after a parse, so never equal."""
    if isinstance(a, str) or isinstance(b, str):
        return isinstance(a, str) and isinstance(b, str) and a == b
    if isinstance(a, bool) or isinstance(b, bool):
        return isinstance(a, bool) and isinstance(b, bool) and a == b
    if isinstance(a, (int, float)) and isinstance(b, (int, float)):
        return float(a) == float(b)
    return a is None and b is None


def math_round(x: float) -> float:
    """Math.round: ties go toward +Infinity."""
    if math.isnan(x) or math.isinf(x):
        return x
    lo = math.floor(x)
    return float(lo + 1) if x - lo >= 0.5 else float(lo)


def number_to_string(x: float) -> str:
    """Number::toString, radix 10."""
    if math.isnan(x):
        return "NaN"
    if x == 0:
        return "0"
    if math.isinf(x):
        return "-Infinity" if x < 0 else "Infinity"
    sign = "-" if x < 0 else ""
    sgn, digs, exp = Decimal(repr(abs(x))).as_tuple()
    s = "".join(str(d) for d in digs).lstrip("0")
    while s.endswith("0"):
        s = s[:-1]
        exp += 1
    k, n = len(s), len(s) + exp
    if k <= n <= 21:
        out = s + "0" * (n - k)
    elif 0 < n <= 21:
        out = s[:n] + "." + s[n:]
    elif -6 < n <= 0:
        out = "0." + "0" * (-n) + s
    else:
        e = n - 1
        out = (s if k == 1 else s[0] + "." + s[1:]) + "e" + ("+" if e >= 0 else "-") + str(abs(e))
    return sign + out


def to_fixed_1(x: float) -> str:
    """Number.prototype.toFixed(1)."""
    if math.isnan(x):
        return "NaN"
    if abs(x) >= 1e21:
        return number_to_string(x)
    q = Decimal(abs(x)).quantize(Decimal("0.1"), rounding=ROUND_HALF_UP)  # exact binary value, ties up
    return ("-" if x < 0 else "") + format(q, "f")
