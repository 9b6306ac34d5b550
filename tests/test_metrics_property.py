"""CPU, property-based: the instant power join.  marshalMetrics (product) -> columnar oracle -> decoded chip rows must
equal the literal restatement of metrics.ts:118-149 and the sums of MetricsPage.tsx:330-332 on arbitrary Prometheus
result lists: colliding chip labels (last write wins), missing labels, parseFloat junk, absent instance mappings."""
import json

from hypothesis import HealthCheck, given, settings
from hypothesis import strategies as st

from headlamp_intel_gpu_plugin_b200 import metrics
from oracle import ref_objects as ro

from .parity import nan_eq

chip = st.sampled_from(["c0", "c1", "c2", "c3", "", "ché", 'c"q'])
inst = st.sampled_from(["i0", "i1", "i2", "", "10.0.0.1:9100"])
value_text = st.one_of(st.floats(min_value=-1e6, max_value=1e6).map(repr),
                       st.sampled_from(["0", "-0", "12.5", "NaN", "+Inf", "-Inf", "Infinity", "1e3", "7abc", "", "abc", " 4", "0x10", ".5", "5."]))
metric = st.fixed_dictionaries({}, optional={"chip": chip, "instance": inst, "nodename": st.sampled_from(["n0", "n1", ""]),
                                             "node": st.sampled_from(["m0", ""]), "chip_name": st.just("i915")})
result = st.fixed_dictionaries({"metric": metric, "value": st.tuples(st.just(0), value_text).map(list)})
results = st.lists(result, max_size=7)


@settings(max_examples=400, deadline=None, suppress_health_check=[HealthCheck.too_slow])
@given(ch=results, en=results, mx=results, un=results)
def test_marshal_metrics_plus_oracle_equals_object_join(oracle_lib, ch, en, mx, un):
    ch, en, mx, un = (json.loads(json.dumps(x)) for x in (ch, en, mx, un))
    p = metrics.marshalMetrics(ch, en, mx, un)
    out = oracle_lib.power_join_sum(p)
    rows = metrics._decode_chips(p, out)
    ref = ro.join_gpu_metrics(ch, en, mx, un)
    assert nan_eq(rows, ref), (ch, en, mx, un)
    fp = ro.fleet_power(ref)
    assert nan_eq(float(out["fleet"][0]), fp["totalWatts"]) and nan_eq(float(out["fleet"][1]), fp["maxTotalWatts"]), (ch, en, mx)
