"""ECMAScript numeric built-ins the reference's rollup path leans on — TEST INFRASTRUCTURE ONLY.

This file belongs to the oracle (see oracle/README.md): only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference leg may import it.  The product's own host code has a
separately written implementation (headlamp_intel_gpu_plugin_b200/jsnum.py); tests cross-check
the two.

The reference never implements these itself — they are ECMA-262 built-ins used at:
  parseInt(v, 10)   src/api/k8s.ts:176, :296; src/components/OverviewPage.tsx:98-99, :110;
                    NodesPage.tsx:173-174; NodeDetailSection.tsx:74, :86
  parseFloat(s)     src/api/metrics.ts:130, :137
  Math.round(x)     OverviewPage.tsx:116; NodesPage.tsx:37; NodeDetailSection.tsx:89;
                    MetricsPage.tsx:51; src/api/metrics.ts:167
  toFixed(1)        src/api/metrics.ts:162
  String(number)    src/api/k8s.ts:300
"""
from __future__ import annotations

import math
from fractions import Fraction

NaN = float("nan")

# ECMA-262 WhiteSpace + LineTerminator code points (StrWhiteSpaceChar).
_JS_WS = frozenset(
    "\t\n\v\f\r \u00a0\u1680\u2000\u2001\u2002\u2003\u2004\u2005\u2006\u2007\u2008\u2009\u200a"
    "\u2028\u2029\u202f\u205f\u3000\ufeff"
)


def _skip_ws(s: str) -> int:
    i = 0
    while i < len(s) and s[i] in _JS_WS:
        i += 1
    return i


def js_to_string_for_parse(v) -> str:
    """ToString() of the argument as parseInt/parseFloat apply it (strings pass through)."""
    if isinstance(v, str):
        return v
    if v is None:
        return "undefined"
    if v is True:
        return "true"
    if v is False:
        return "false"
    if isinstance(v, (int, float)):
        return js_number_to_string(float(v))
    return str(v)


def js_parse_int10(v) -> float:
    """parseInt(v, 10): leading whitespace, optional sign, longest run of decimal digits, else NaN."""
    s = js_to_string_for_parse(v)
    i = _skip_ws(s)
    sign = 1.0
    if i < len(s) and s[i] in "+-":
        if s[i] == "-":
            sign = -1.0
        i += 1
    j = i
    while j < len(s) and "0" <= s[j] <= "9":
        j += 1
    if j == i:
        return NaN
    mag = float(int(s[i:j]))  # correctly rounded to the nearest double
    return sign * mag  # "-0" -> -0.0, like JS


def js_parse_float(v) -> float:
    """parseFloat(v): longest prefix that is a StrDecimalLiteral, else NaN."""
    s = js_to_string_for_parse(v)
    i = _skip_ws(s)
    n = len(s)
    start = i
    if i < n and s[i] in "+-":
        i += 1
    if s.startswith("Infinity", i):
        return -math.inf if s[start] == "-" else math.inf
    d0 = i
    while i < n and s[i].isascii() and s[i].isdigit():
        i += 1
    int_digits = i - d0
    frac_digits = 0
    if i < n and s[i] == ".":
        k = i + 1
        while k < n and s[k].isascii() and s[k].isdigit():
            k += 1
        frac_digits = k - (i + 1)
        if int_digits or frac_digits:
            i = k
    if int_digits == 0 and frac_digits == 0:
        return NaN
    if i < n and s[i] in "eE":
        k = i + 1
        if k < n and s[k] in "+-":
            k += 1
        e0 = k
        while k < n and s[k].isascii() and s[k].isdigit():
            k += 1
        if k > e0:
            i = k
    lit = s[start:i]
    if lit.endswith("."):
        lit += "0"
    return float(lit)


def js_or_zero(x: float) -> float:
    """`x || 0` for a number: NaN, +0 and -0 all become +0."""
    if x != x or x == 0.0:
        return 0.0
    return x


def js_truthy(v) -> bool:
    """ToBoolean for the value kinds that occur in K8s JSON."""
    if v is None or v is False:
        return False
    if isinstance(v, str):
        return len(v) > 0
    if isinstance(v, (int, float)) and not isinstance(v, bool):
        return not (v == 0 or v != v)
    return True


def js_strict_equal(a, b) -> bool:
    """a === b for values out of JSON.parse (undefined and null both arrive here as None: the call sites only compare
    two truthy values).  Strings by content; numbers by value; booleans by value; objects and arrays by identity."""
    if isinstance(a, bool) or isinstance(b, bool):
        return isinstance(a, bool) and isinstance(b, bool) and a == b
    if isinstance(a, str) or isinstance(b, str):
        return isinstance(a, str) and isinstance(b, str) and a == b
    if isinstance(a, (int, float)) and isinstance(b, (int, float)):
        return float(a) == float(b)
    if a is None and b is None:
        return True
    return a is b


def js_math_round(x: float) -> float:
    """Math.round: nearest integer, ties toward +Infinity."""
    if x != x or math.isinf(x):
        return x
    f = math.floor(x)
    return float(f + 1) if (x - f) >= 0.5 else float(f)


def js_number_to_string(x: float) -> str:
    """Number::toString(x) (radix 10): shortest round-trip digits, JS exponent rules."""
    if x != x:
        return "NaN"
    if x == 0:
        return "0"
    if math.isinf(x):
        return "Infinity" if x > 0 else "-Infinity"
    if x < 0:
        return "-" + js_number_to_string(-x)
    from decimal import Decimal

    t = Decimal(repr(x)).as_tuple()  # repr() = shortest digits that round-trip, as in JS
    digits = "".join(map(str, t.digits)).lstrip("0") or "0"
    exp = t.exponent
    stripped = digits.rstrip("0")
    exp += len(digits) - len(stripped)
    digits = stripped
    k = len(digits)
    n = k + exp  # value = 0.digits * 10^n
    if k <= n <= 21:
        return digits + "0" * (n - k)
    if 0 < n <= 21:
        return digits[:n] + "." + digits[n:]
    if -6 < n <= 0:
        return "0." + "0" * (-n) + digits
    e = n - 1
    sign = "+" if e >= 0 else "-"
    if k == 1:
        return f"{digits}e{sign}{abs(e)}"
    return f"{digits[0]}.{digits[1:]}e{sign}{abs(e)}"


def js_to_fixed1(x: float) -> str:
    """Number.prototype.toFixed(1): exact-value rounding, ties pick the larger n."""
    if x != x:
        return "NaN"
    if abs(x) >= 1e21:
        return js_number_to_string(x)
    neg = x < 0 or (x == 0 and math.copysign(1.0, x) < 0)
    ax = Fraction(abs(x))
    n = math.floor(ax * 10 + Fraction(1, 2))
    if n == 0:
        neg = neg and x != 0  # (-0).toFixed(1) is "0.0"; tiny negatives keep the sign: "-0.0"
    body = f"{n // 10}.{n % 10}"
    return ("-" if neg and x < 0 else "") + body
