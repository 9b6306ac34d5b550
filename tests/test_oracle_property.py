"""CPU, property-based: the two oracle layers and the host marshal against each other on arbitrary Kubernetes-shaped junk.
marshalFleet (product, Python) -> columnar C++ oracle -> FleetRollup views must equal the literal per-object restatement
of the reference's helpers and page loops (oracle/ref_objects.py) for every reference-visible quantity."""
import json

from hypothesis import HealthCheck, given, settings
from hypothesis import strategies as st

from headlamp_intel_gpu_plugin_b200 import k8s

from .parity import check_against_objects
from .test_marshal_property import node, pod


@settings(max_examples=300, deadline=None, suppress_health_check=[HealthCheck.too_slow, HealthCheck.data_too_large])
@given(nodes=st.lists(node, max_size=5), pods=st.lists(pod, max_size=8))
def test_marshal_plus_columnar_oracle_equals_object_oracle(oracle_lib, nodes, pods):
    nodes, pods = json.loads(json.dumps(nodes)), json.loads(json.dumps(pods))  # what a List response can carry
    try:
        cols = k8s.marshalFleet(nodes, pods)
    except ValueError:
        return  # outside the exact-integer domain (sum of |values| >= 2^53): rejected by design
    out = oracle_lib.alloc_rollup(cols)
    check_against_objects(nodes, pods, cols, out, container_rows=oracle_lib.container_rows(cols))
