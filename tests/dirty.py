"""Adversarial K8s objects: parseInt edge strings, init containers, limits-only pods, odd phases,
unscheduled pods, NaN-poisoned capacities — the cases SURVEY.md §4 lists as untested upstream."""
import random

I915 = "gpu.intel.com/i915"
XE = "gpu.intel.com/xe"
MC = "gpu.intel.com/millicores"
PFX = "gpu.intel.com/"

QUANTITIES = ["1", "2", "8", "500m", "1Gi", "1e3", "0x10", "", " 42", "-0", "+7x", "abc", "  3", "-2", "007", "1.9", "٣", "9999999",
              None, 0, 3, 2.5, False, True, " 5", "﻿6", "1_000", "- 1", "+", "-"]
PHASES = ["Running", "Pending", "Succeeded", "Failed", "Unknown", "Other", None, "toString", "constructor", "running", "", 5,
          "__proto__", "hasOwnProperty"]
LABEL_VALUES = ["true", "True", "1", "false", "", None, "TRUE", " true"]


def dirty_fleet(n_nodes=40, n_pods=400, seed=7):
    rng = random.Random(seed)
    nodes = []
    for i in range(n_nodes):
        labels = {}
        for key in ("node-role.kubernetes.io/gpu", "node-role.kubernetes.io/igpu", "intel.feature.node.kubernetes.io/gpu"):
            if rng.random() < 0.35:
                labels[key] = rng.choice(LABEL_VALUES)
        cap, alloc = {"cpu": "8"}, {"cpu": "8"}
        for key in (I915, XE, MC, PFX + "tiles"):
            if rng.random() < 0.45:
                cap[key] = rng.choice(QUANTITIES)
            if rng.random() < 0.45:
                alloc[key] = rng.choice(QUANTITIES)
        status = {}
        if rng.random() < 0.9:
            status["capacity"] = cap
        if rng.random() < 0.9:
            status["allocatable"] = alloc
        r = rng.random()
        if r < 0.6:
            status["conditions"] = [{"type": "Ready", "status": rng.choice(["True", "False", "Unknown", "true"])}]
        elif r < 0.8:
            status["conditions"] = [{"type": "DiskPressure", "status": "True"}, {"type": "Ready", "status": "True"}]
        meta = {"name": f"n{i}"}
        if rng.random() < 0.85:
            meta["labels"] = labels
        nodes.append({"kind": "Node", "metadata": meta, "status": status})
    nodes.append({"kind": "Node", "metadata": {"name": "bare"}})
    pods = []
    for p in range(n_pods):
        def container(j):
            c = {"name": f"c{j}"}
            r = rng.random()
            if r < 0.1:
                return c
            res = {}
            for field in ("requests", "limits"):
                if rng.random() < 0.7:
                    m = {"cpu": "1"} if rng.random() < 0.5 else {}
                    for key in (I915, XE, MC, PFX + "memory.max", PFX + "tiles", PFX + "custom-" + str(rng.randrange(3))):
                        if rng.random() < 0.3:
                            m[key] = rng.choice(QUANTITIES)
                    res[field] = m
            c["resources"] = res
            return c
        spec = {}
        if rng.random() < 0.95:
            spec["containers"] = [container(j) for j in range(rng.randrange(0, 4))]
        if rng.random() < 0.3:
            spec["initContainers"] = [container(9)]
        r = rng.random()
        if r < 0.8:
            spec["nodeName"] = f"n{rng.randrange(n_nodes + 3)}"  # some point at unknown nodes
        elif r < 0.85:
            spec["nodeName"] = ""
        status = {}
        ph = rng.choice(PHASES + ["Running"] * 6)
        if ph is not None or rng.random() < 0.5:
            status["phase"] = ph
        pod = {"kind": "Pod", "metadata": {"name": f"p{p}"}, "spec": spec}
        if rng.random() < 0.95:
            pod["status"] = status
        pods.append(pod)
    pods.append({"kind": "Pod", "metadata": {"name": "nospec"}})
    return nodes, pods
