"""Host marshal throughput: native extractor (igpu_marshal_fleet_json) vs json.loads + Python marshal."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from headlamp_intel_gpu_plugin_b200 import k8s, synth

n_nodes, n_pods = int(sys.argv[1]) if len(sys.argv) > 1 else 2000, int(sys.argv[2]) if len(sys.argv) > 2 else 200_000
nodes, pods = synth.fleet_objects(n_nodes, n_pods)
nj, pj = json.dumps(nodes).encode(), json.dumps(pods).encode()
print(f"{n_nodes} nodes / {n_pods} pods: {len(nj) / 1e6:.1f} + {len(pj) / 1e6:.1f} MB of JSON, {os.cpu_count()} cores")
for threads in (1, 2, 4, 8, 16, 0):
    best = 1e9
    for _ in range(3):
        t = time.perf_counter(); c = k8s.marshalFleetJson(nj, pj, threads=threads); best = min(best, time.perf_counter() - t)
    print(f"native threads={threads or 'all'}: {best * 1e3:8.1f} ms  {n_pods / best / 1e6:6.2f} M pods/s  {(len(nj) + len(pj)) / best / 1e9:5.2f} GB/s")
t = time.perf_counter(); n2, p2 = json.loads(nj), json.loads(pj); t1 = time.perf_counter() - t
t = time.perf_counter(); c2 = k8s.marshalFleet(n2, p2); t2 = time.perf_counter() - t
print(f"python: json.loads {t1 * 1e3:.0f} ms + marshalFleet {t2 * 1e3:.0f} ms = {n_pods / (t1 + t2) / 1e6:.3f} M pods/s")
