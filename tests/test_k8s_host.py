"""CPU: the host-only pieces of the product's k8s.py surface (no arithmetic, no device): getGpuResources (k8s.ts:159-168)
against the restatement of the reference, and the strict-equality helper the marshal uses for `req === lim`."""
import random

from hypothesis import given, settings
from hypothesis import strategies as st

from headlamp_intel_gpu_plugin_b200 import jsnum, k8s
from oracle import js_semantics as js
from oracle import ref_objects as ro

from .dirty import QUANTITIES, dirty_fleet


def test_get_gpu_resources_known_cases():
    # NodesPage.tsx:80-81 hands it status.capacity / status.allocatable
    cap = {"cpu": "64", "gpu.intel.com/i915": "4", "gpu.intel.com/xe": None, "gpu.intel.com/custom": "", "memory": "1Gi", "gpu.intel.com": "x"}
    assert k8s.getGpuResources(cap) == {"gpu.intel.com/i915": "4", "gpu.intel.com/xe": None, "gpu.intel.com/custom": ""}
    assert list(k8s.getGpuResources(cap)) == ["gpu.intel.com/i915", "gpu.intel.com/xe", "gpu.intel.com/custom"]  # Object.entries order
    for nothing in (None, {}, "gpu.intel.com/i915", 3, []):
        assert k8s.getGpuResources(nothing) == {} == ro.get_gpu_resources(nothing)


def test_get_gpu_resources_matches_the_restatement_on_dirty_nodes():
    for seed in range(6):
        nodes, _ = dirty_fleet(40, 0, seed)
        for n in nodes:
            for field in ("capacity", "allocatable"):
                res = n.get("status", {}).get(field)
                assert k8s.getGpuResources(res) == ro.get_gpu_resources(res)
                assert list(k8s.getGpuResources(res)) == list(ro.get_gpu_resources(res))


json_scalar = st.one_of(st.none(), st.booleans(), st.integers(-5, 5), st.floats(allow_nan=False, allow_infinity=False, width=32),
                        st.sampled_from([q for q in QUANTITIES if isinstance(q, str)]), st.lists(st.integers(0, 1), max_size=1))


@settings(max_examples=400, deadline=None)
@given(a=json_scalar, b=json_scalar)
def test_strict_equal_product_vs_oracle(a, b):
    assert jsnum.strict_equal(a, b) == js.js_strict_equal(a, b)
    assert jsnum.strict_equal(a, a) == js.js_strict_equal(a, a) or isinstance(a, list)


def test_strict_equal_cases():
    rnd = random.Random(1)
    assert jsnum.strict_equal("1", "1") and not jsnum.strict_equal("1", "01") and not jsnum.strict_equal("1", 1)
    assert jsnum.strict_equal(2, 2.0) and not jsnum.strict_equal(True, 1) and jsnum.strict_equal(None, None)
    assert not jsnum.strict_equal([], []) and not jsnum.strict_equal({}, {}) and rnd is not None
