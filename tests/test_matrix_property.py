"""CPU, property-based: native Prometheus matrix ingest (igpu_marshal_matrix_json) against the Python marshal
(metrics.marshalMatrix over json.loads) on arbitrary query_range-shaped responses: off-grid and absurd timestamps,
every spelling of a float, junk in every position."""
import json
import math

import numpy as np
from hypothesis import HealthCheck, given, settings
from hypothesis import strategies as st

from headlamp_intel_gpu_plugin_b200 import metrics, native

START, STEP, T = 1_700_000_000_000, 15_000, 12

on_grid = st.integers(-2, T + 2).map(lambda i: (START + i * STEP) / 1000)
timestamps = st.one_of(on_grid, on_grid.map(lambda t: t + 0.007), st.integers(0, 2 * 10**9), st.sampled_from([1.7e9, 1700000000.0005, 0, -5, 1e15]))
value_text = st.one_of(
    st.floats(allow_nan=False, allow_infinity=False, width=64).map(repr),
    st.floats(min_value=0, max_value=1e9).map(lambda x: f"{x:.3f}"),
    st.floats(min_value=0, max_value=1e9).map(lambda x: f"{x:e}"),
    st.integers(-(10**20), 10**20).map(str),
    st.sampled_from(["NaN", "+Inf", "-Inf", "Inf", "0", "-0", "1e400", "1e-400", ".5", "5.", "+1.25", "0x10", "", "abc", "1 "]),
)
sample = st.tuples(timestamps, value_text).map(list)
label_text = st.text(alphabet="abc:/._-é\"\\ ", max_size=8)
series = st.fixed_dictionaries({}, optional={
    "metric": st.dictionaries(st.sampled_from(["chip", "instance", "__name__", "x"]), st.one_of(label_text, st.integers(0, 3)), max_size=3),
    "values": st.lists(sample, max_size=8),
    "extra": st.just([1, {"a": "}"}]),
})
response = st.one_of(
    st.lists(series, max_size=4),
    st.fixed_dictionaries({"status": st.sampled_from(["success", "error"]),
                           "data": st.fixed_dictionaries({"resultType": st.just("matrix"), "result": st.lists(series, max_size=4)})}),
)


def py_float_ok(text):
    try:
        float(text)
        return True
    except ValueError:
        return False


@settings(max_examples=400, deadline=None, suppress_health_check=[HealthCheck.too_slow])
@given(resp=response, spaced=st.booleans(), threads=st.sampled_from([1, 2]))
def test_native_matrix_ingest_equals_python(resp, spaced, threads):
    text = json.dumps(resp, indent=1 if spaced else None, ensure_ascii=not spaced)
    want_v, want_l, want_d = metrics.marshalMatrix(json.loads(text), START, STEP, T)
    got_v, got_l, got_d = metrics.marshalMatrixJson(text, START, STEP, T, threads=threads)
    assert got_v.shape == want_v.shape and got_l == want_l, text
    assert np.all((got_v == want_v) | (np.isnan(got_v) & np.isnan(want_v))), text
    assert got_d == want_d, text
