"""CPU oracle for the fleet rollup — TEST INFRASTRUCTURE ONLY (see oracle/README.md).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may import
this package; nothing under headlamp_intel_gpu_plugin_b200/ does.
"""
