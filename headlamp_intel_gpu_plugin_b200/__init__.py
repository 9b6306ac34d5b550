"""B200-native fleet rollup behind the exported API of headlamp-intel-gpu-plugin's src/api.

Sub-modules: k8s (allocation rollup + classification, mirror of src/api/k8s.ts), metrics (power
join / series statistics, mirror of src/api/metrics.ts), columns (columnar layout), native (ctypes
binding of libigpu_rollup.so), shard (node-range sharding for one rank per GPU), synth (synthetic
fleets).  The compute path is CUDA only; see DESIGN.md.
"""
__version__ = "0.1.0"
