"""CPU, property-based: the product's JS number semantics (jsnum.py) against the oracle's independent restatement
(oracle/js_semantics.py) on arbitrary strings and doubles — parseInt, parseFloat, truthiness, Math.round,
String(number), toFixed(1)."""
import math

from hypothesis import given, settings
from hypothesis import strategies as st

from headlamp_intel_gpu_plugin_b200 import jsnum
from oracle import js_semantics as ref

numberish = st.text(alphabet="0123456789+-.eEInfityNa xX\t\n ﻿١٢", max_size=12)
anytext = st.one_of(numberish, st.text(max_size=8))
doubles = st.one_of(st.floats(allow_nan=True, allow_infinity=True), st.integers(-(2**60), 2**60).map(float),
                    st.sampled_from([0.5, 1.5, 2.5, -0.5, -1.5, 1e21, 1e-7, 123456789012345680000.0, 0.000001, 5e-324, 1.7976931348623157e308]))
values = st.one_of(anytext, doubles, st.integers(-(10**18), 10**18), st.none(), st.booleans(), st.just([]), st.just({}), st.just([1, 2]))


def same(a, b):
    if isinstance(a, float) and isinstance(b, float):
        return (math.isnan(a) and math.isnan(b)) or (a == b and math.copysign(1, a) == math.copysign(1, b))
    return a == b


@settings(max_examples=1500, deadline=None)
@given(v=values)
def test_parse_int_parse_float_truthy(v):
    assert same(jsnum.parse_int(v), ref.js_parse_int10(v)), v
    assert same(jsnum.parse_float(v), ref.js_parse_float(v)), v
    assert jsnum.truthy(v) == ref.js_truthy(v), v


@settings(max_examples=1500, deadline=None)
@given(x=doubles)
def test_round_tostring_tofixed(x):
    assert same(jsnum.math_round(x), ref.js_math_round(x)), x
    assert jsnum.number_to_string(x) == ref.js_number_to_string(x), x
    if math.isfinite(x) and abs(x) < 1e21:
        assert jsnum.to_fixed_1(x) == ref.js_to_fixed1(x), x
