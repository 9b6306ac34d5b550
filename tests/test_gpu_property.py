"""GPU, property-based: arbitrary Kubernetes-shaped junk -> JSON text -> native marshal -> CUDA rollup (+ per-container rows)
-> views, against the literal per-object restatement of the reference (oracle/ref_objects.py).  The CPU twins of this test
(test_oracle_property.py, test_marshal_property.py) found two real divergences in round 1; this one puts the device path
under the same pressure."""
import json

import pytest
from hypothesis import HealthCheck, given, settings
from hypothesis import strategies as st

from headlamp_intel_gpu_plugin_b200 import k8s, native

from .parity import check_against_objects
from .test_marshal_property import node, pod

pytestmark = pytest.mark.gpu


@settings(max_examples=200, deadline=None, suppress_health_check=[HealthCheck.too_slow, HealthCheck.data_too_large, HealthCheck.function_scoped_fixture])
@given(nodes=st.lists(node, max_size=5), pods=st.lists(pod, max_size=10), grouped=st.booleans())
def test_json_text_to_cuda_equals_the_object_restatement(ctx, nodes, pods, grouped):
    nodes, pods = json.loads(json.dumps(nodes)), json.loads(json.dumps(pods))  # what a List response can carry
    try:
        r = k8s.rollupFleetJson(json.dumps(nodes), json.dumps(pods), ctx, group_by_node=grouped, container_rows=True)
    except native.IgpuError as e:
        assert e.status == -6, e  # outside a documented limit (resource names, containers, 2^53): rejected by design
        return
    check_against_objects(nodes, pods, r.cols, r.out, r.order, r.pos, container_rows=r.container_rows)
