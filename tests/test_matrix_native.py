"""CPU: native Prometheus query_range ingest (igpu_marshal_matrix_json, csrc/marshal.cpp) against the Python marshal
(metrics.marshalMatrix over json.loads).  No GPU needed: the function is host code inside libigpu_rollup.so."""
import json
import math

import numpy as np
import pytest

from headlamp_intel_gpu_plugin_b200 import metrics, native

START, STEP = 1_700_000_000_000, 15_000


def same(a, b):
    return a.shape == b.shape and np.all((a == b) | (np.isnan(a) & np.isnan(b)))


def fmt_ts(ms):
    s, r = divmod(ms, 1000)
    return f"{s}.{r:03d}" if r else str(s)


def make_response(rng, n_series, n_time, gap=0.05):
    result = []
    for s in range(n_series):
        vals = []
        c = float(rng.integers(0, 10**7))
        for i in range(n_time):
            c += float(rng.integers(0, 5000)) * 0.125
            if rng.random() < gap:
                continue
            style = rng.integers(0, 6)
            text = [repr(c), f"{c:.3f}", f"{c:e}", f"{c:.17g}", str(int(c)), f"{c:.1f}"][style]
            vals.append([float(fmt_ts(START + i * STEP)) if rng.random() < 0.5 else json.loads(fmt_ts(START + i * STEP)), text])
        metric = {"__name__": "node_hwmon_energy_joule_total", "chip": f"pci0000:00_0000:{s:02x}:00_0", "instance": f"10.0.{s // 250}.{s % 250}:9100"}
        if s % 7 == 3:
            del metric["instance"]
        result.append({"metric": metric, "values": vals})
    return {"status": "success", "data": {"resultType": "matrix", "result": result}}


@pytest.mark.parametrize("threads", [1, 3, 0])
def test_native_equals_python_marshal(threads):
    rng = np.random.default_rng(5)
    resp = make_response(rng, 57, 300)
    text = json.dumps(resp).encode()
    want_v, want_l, want_d = metrics.marshalMatrix(resp, START, STEP, 300)
    got_v, got_l, got_d = metrics.marshalMatrixJson(text, START, STEP, 300, threads=threads)
    assert same(got_v, want_v)
    assert got_l == want_l and got_d == want_d == 0
    assert np.isnan(want_v).mean() > 0.02


def test_value_spellings_round_correctly():
    from oracle.js_semantics import js_parse_float

    texts = ["0", "-0", "1", "1.5", "123456789012345", "1234567890123456789", "0.1", "1e22", "1e23", "1.7976931348623157e308",
             "5e-324", "2.2250738585072014e-308", "0.000001", "123.456e-2", "1E5", "+3.25", "NaN", "+Inf", "-Inf", "Inf",
             "9007199254740993", "0.30000000000000004", "4.35", "Infinity", "-Infinity", " 7.5", "12abc", "0x10", "5.", ".5", "1e", "1e+",
             ".", "", "1e-22", "1e-23", "179769313486231580793728971405303415079934132710037826936173778980444968292764750946649017977587207096330286416692887910946555547851940402630657488671505820681908902000708383676273854845817711531764475730270069855571366959622842914819860834936475292719074168444365510704342711559699508093042880177904174497791.9999"]
    vals = [[json.loads(fmt_ts(START + i * STEP)), t] for i, t in enumerate(texts)]
    resp = {"status": "success", "data": {"resultType": "matrix", "result": [{"metric": {"chip": "a"}, "values": vals}]}}
    got, labs, dropped = metrics.marshalMatrixJson(json.dumps(resp), START, STEP, len(texts), labels=("chip",))
    for i, t in enumerate(texts):
        want = js_parse_float(t)  # the reference reads a sample with parseFloat (metrics.ts:130-137): "+Inf" is NaN there
        assert (math.isnan(want) and math.isnan(got[0, i])) or (got[0, i] == want and math.copysign(1, got[0, i]) == math.copysign(1, want)), t
    assert labs == [("a",)] and dropped == 0


def test_off_grid_and_out_of_range_samples_are_dropped():
    vals = [[(START - STEP) / 1000, "1"], [START / 1000, "2"], [(START + 7) / 1000, "3"], [(START + STEP) / 1000, "4"],
            [(START + 5 * STEP) / 1000, "5"], [(START + 2 * STEP) / 1000, "6"]]
    resp = {"status": "success", "data": {"resultType": "matrix", "result": [{"metric": {}, "values": vals}]}}
    want = metrics.marshalMatrix(resp, START, STEP, 4)
    got = metrics.marshalMatrixJson(json.dumps(resp), START, STEP, 4)
    assert same(got[0], want[0]) and got[2] == want[2] == 3
    assert got[0][0].tolist()[:3] == [2.0, 4.0, 6.0] and math.isnan(got[0][0, 3])
    assert got[1] == [(None, None)]


def test_envelope_variants():
    series = [{"metric": {"chip": "c\\u00e9\"x", "instance": "i"}, "values": [[START / 1000, "1"]]}]
    bare = json.dumps(series)
    v, l, _ = metrics.marshalMatrixJson(bare, START, STEP, 2)
    assert v.shape == (1, 2) and v[0, 0] == 1.0 and l == [("c\\u00e9\"x", "i")]
    failed = json.dumps({"status": "error", "errorType": "bad_data", "error": "x", "data": {"result": series}})
    v, l, _ = metrics.marshalMatrixJson(failed, START, STEP, 2)
    assert v.shape == (0, 2) and l == []
    assert metrics.marshalMatrix(json.loads(failed), START, STEP, 2)[0].shape == (0, 2)
    reordered = json.dumps({"data": {"result": series, "resultType": "matrix"}, "status": "success"})
    v, l, _ = metrics.marshalMatrixJson(reordered, START, STEP, 2)
    assert v.shape == (1, 2)
    empty = json.dumps({"status": "success", "data": {"resultType": "matrix", "result": []}})
    assert metrics.marshalMatrixJson(empty, START, STEP, 2)[0].shape == (0, 2)
    spaced = json.dumps({"status": "success", "data": {"resultType": "matrix", "result": series}}, indent=2)
    assert metrics.marshalMatrixJson(spaced, START, STEP, 2)[0][0, 0] == 1.0


def test_errors():
    with pytest.raises(native.IgpuError) as e:
        metrics.marshalMatrixJson(b'{"status":"success","data":{"result":[{"values":[[1,"2"', START, STEP, 2)
    assert e.value.status == -1
    with pytest.raises(native.IgpuError):
        metrics.marshalMatrixJson(b"[]", START, 0, 2)
    with pytest.raises(native.IgpuError):
        metrics.marshalMatrixJson(b"[]", START, STEP, 0)


def test_large_response_multithreaded():
    rng = np.random.default_rng(9)
    resp = make_response(rng, 400, 240, gap=0.01)
    text = json.dumps(resp).encode()
    want = metrics.marshalMatrix(resp, START, STEP, 240)
    for threads in (1, 8):
        got = metrics.marshalMatrixJson(text, START, STEP, 240, threads=threads)
        assert same(got[0], want[0]) and got[1] == want[1]
