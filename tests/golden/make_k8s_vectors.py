"""Writes k8s_test_vectors.json: the known-answer cases of the reference's own unit tests for the
rollup helpers, transcribed by hand (the reference is TypeScript and cannot run in this image).

Each vector cites the test it comes from.  The fixture builders below restate
makeNode / makeGpuNode / makeGpuPod (src/api/k8s.test.ts:33-80) as data constructors.
Run: python tests/golden/make_k8s_vectors.py
"""
import json
import os

I915 = "gpu.intel.com/i915"
XE = "gpu.intel.com/xe"
NFD = "intel.feature.node.kubernetes.io/gpu"
T = "src/api/k8s.test.ts"


def make_node(**ov):
    d = {"apiVersion": "v1", "kind": "Node", "metadata": {"name": "test-node"}, "status": {}}
    d.update(ov)
    return d


def make_gpu_node(t="discrete"):
    labels = {}
    if t == "discrete":
        labels["node-role.kubernetes.io/gpu"] = "true"
    if t == "integrated":
        labels["node-role.kubernetes.io/igpu"] = "true"
    if t == "generic":
        labels[NFD] = "true"
    return {
        "apiVersion": "v1",
        "kind": "Node",
        "metadata": {"name": "gpu-node", "labels": labels},
        "status": {
            "capacity": {I915: "2"},
            "allocatable": {I915: "2"},
            "conditions": [{"type": "Ready", "status": "True"}],
        },
    }


def make_gpu_pod(key=I915, amount="1"):
    return {
        "apiVersion": "v1",
        "kind": "Pod",
        "metadata": {"name": "gpu-pod", "namespace": "default"},
        "spec": {
            "nodeName": "gpu-node",
            "containers": [{"name": "workload", "resources": {"requests": {key: amount}, "limits": {key: amount}}}],
        },
        "status": {"phase": "Running"},
    }


def pod(name, containers, **extra):
    d = {"apiVersion": "v1", "kind": "Pod", "metadata": {"name": name}, "spec": {"containers": containers}}
    d.update(extra)
    return d


V = []


def add(fn, inp, exp, ref, note=""):
    V.append({"fn": fn, "input": inp, "expect": exp, "ref": ref, "note": note})


add("isIntelGpuNode", make_gpu_node("discrete"), True, T + ":87-90", "discrete GPU label")
add("isIntelGpuNode", make_gpu_node("integrated"), True, T + ":92-95", "integrated GPU label")
add("isIntelGpuNode", make_gpu_node("generic"), True, T + ":97-100", "generic NFD label")
add("isIntelGpuNode", make_node(status={"capacity": {I915: "1"}}), True, T + ":102-107", "gpu.intel.com/* in capacity")
add(
    "isIntelGpuNode",
    make_node(metadata={"name": "regular-node", "labels": {}}, status={"capacity": {"cpu": "8", "memory": "16Gi"}}),
    False,
    T + ":109-115",
    "no GPU labels or resources",
)
add("isIntelGpuNode", None, False, T + ":117-120", "null / undefined")
add(
    "filterIntelGpuNodes",
    [make_gpu_node("discrete"), make_node(metadata={"name": "regular"})],
    ["gpu-node"],
    T + ":128-134",
    "filters out non-GPU nodes (names of survivors)",
)
add("filterIntelGpuNodes", [], [], T + ":136-138", "empty array")
add("getNodeGpuType", make_gpu_node("discrete"), "discrete", T + ":146-148")
add("getNodeGpuType", make_gpu_node("integrated"), "integrated", T + ":150-152")
add("getNodeGpuType", make_gpu_node("generic"), "unknown", T + ":154-156", "NFD label alone is unknown")
add("getNodeGpuType", make_node(status={"capacity": {I915: "1"}}), "unknown", T + ":158-161", "no labels")
add("getNodeGpuCount", make_node(status={"capacity": {I915: "4"}}), 4, T + ":169-174")
add("getNodeGpuCount", make_node(status={"capacity": {XE: "2"}}), 2, T + ":176-181")
add("getNodeGpuCount", make_node(status={"capacity": {I915: "2", XE: "1"}}), 3, T + ":183-193", "i915 + xe")
add("getNodeGpuCount", make_node(status={"capacity": {"cpu": "8"}}), 0, T + ":195-198", "no GPU capacity")
add("isNodeReady", make_node(status={"conditions": [{"type": "Ready", "status": "True"}]}), True, T + ":206-211")
add("isNodeReady", make_node(status={"conditions": [{"type": "Ready", "status": "False"}]}), False, T + ":213-218")
add("isNodeReady", make_node(status={}), False, T + ":220-223", "no conditions")
add("isGpuRequestingPod", make_gpu_pod(I915), True, T + ":231-233")
add("isGpuRequestingPod", make_gpu_pod(XE), True, T + ":235-237")
add("isGpuRequestingPod", make_gpu_pod("gpu.intel.com/millicores", "500"), True, T + ":239-241", "millicores")
add(
    "isGpuRequestingPod",
    pod("no-gpu-pod", [{"name": "app", "resources": {"requests": {"cpu": "1", "memory": "1Gi"}}}]),
    False,
    T + ":243-260",
    "cpu/memory only",
)
add("isGpuRequestingPod", None, False, T + ":262-264", "null")
add(
    "filterGpuRequestingPods",
    [make_gpu_pod(), pod("regular", [{"name": "app"}])],
    ["gpu-pod"],
    T + ":272-283",
    "names of survivors",
)
add("getPodGpuRequests", make_gpu_pod(I915, "2"), {I915: "2"}, T + ":291-295")
add(
    "getPodGpuRequests",
    pod("regular", [{"name": "app", "resources": {"requests": {"cpu": "1"}}}]),
    {},
    T + ":297-305",
    "non-GPU pod gives {}",
)
add(
    "getPodGpuRequests",
    pod(
        "multi",
        [
            {"name": "a", "resources": {"requests": {I915: "1"}}},
            {"name": "b", "resources": {"requests": {I915: "2"}}},
        ],
    ),
    {I915: "3"},
    T + ":307-321",
    "1 + 2 across two containers",
)
add(
    "isGpuRequestingPod",
    {
        "kind": "Pod",
        "metadata": {"name": "limits-only-pod", "namespace": "default"},
        "spec": {"containers": [{"name": "app", "resources": {"limits": {I915: "1"}}}]},
        "status": {"phase": "Pending"},
    },
    True,
    "src/components/PodDetailSection.test.tsx:63-78,119-124",
    "limits-only pod is still a GPU pod",
)

if __name__ == "__main__":
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "k8s_test_vectors.json")
    with open(out, "w") as f:
        json.dump(
            {
                "source": "transcribed by hand from the reference test files cited per vector; inputs are the "
                "literal fixtures built by makeNode/makeGpuNode/makeGpuPod (src/api/k8s.test.ts:33-80)",
                "vectors": V,
            },
            f,
            indent=1,
        )
    print(len(V), "vectors ->", out)
