"""Writes pod_detail_vectors.json: the fixtures and assertions of the reference's PodDetailSection test
(src/components/PodDetailSection.test.tsx), transcribed by hand (the reference is TypeScript and cannot run here).

`gpuContainers` and `rendersNothing` are what the upstream cases assert (the "GPU Containers" row :131-137, the
empty render :84-92, "still renders" :119-124).  `rows` is the NameValueTable content as data — (container, key,
'request' | 'limit', raw value or null for the dash) — of which upstream pins only the request-row labels (:101-105,
:122-123); the rest follows PodDetailSection.tsx:57-83 line by line.
Run: python tests/golden/make_pod_detail_vectors.py
"""
import json
import os

I915 = "gpu.intel.com/i915"
T = "src/components/PodDetailSection.test.tsx"

non_gpu_pod = {  # :34-42
    "kind": "Pod",
    "metadata": {"name": "plain-pod", "namespace": "default"},
    "spec": {"containers": [{"name": "main", "resources": {"requests": {"cpu": "100m", "memory": "128Mi"}}}]},
    "status": {"phase": "Running"},
}
gpu_pod = {  # :45-61
    "kind": "Pod",
    "metadata": {"name": "gpu-workload", "namespace": "default"},
    "spec": {
        "nodeName": "gpu-node-1",
        "containers": [{"name": "trainer", "resources": {"requests": {I915: "1", "cpu": "2"}, "limits": {I915: "1", "cpu": "2"}}}],
    },
    "status": {"phase": "Running"},
}
limits_only_pod = {  # :64-78
    "kind": "Pod",
    "metadata": {"name": "limits-only-pod", "namespace": "default"},
    "spec": {"containers": [{"name": "app", "resources": {"limits": {I915: "1"}}}]},
    "status": {"phase": "Pending"},
}

vectors = [
    {"ref": f"{T}:84-87", "pod": non_gpu_pod, "rendersNothing": True, "gpuContainers": 0, "rows": []},
    {"ref": f"{T}:101-105,131-137", "pod": gpu_pod, "rendersNothing": False, "gpuContainers": 1,
     "rows": [["trainer", I915, "request", "1"]], "upstream_asserts": ["trainer → GPU (i915) request", "GPU Containers = 1"]},
    {"ref": f"{T}:119-124", "pod": limits_only_pod, "rendersNothing": False, "gpuContainers": 1,
     "rows": [["app", I915, "request", None], ["app", I915, "limit", "1"]], "upstream_asserts": ["app → GPU (i915) request"]},
]

if __name__ == "__main__":
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "pod_detail_vectors.json")
    with open(out, "w") as f:
        json.dump({"source": T, "vectors": vectors}, f, indent=1, ensure_ascii=False)
    print(f"wrote {len(vectors)} vectors to {out}")
