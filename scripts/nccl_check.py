"""Multi-rank correctness of the exchange step (run under torchrun on N GPUs): every rank rolls up its node-range shard,
the library all-reduces the fleet vectors / totals over its own NCCL communicator, and rank 0 compares them with the
CPU oracle over the whole fleet.  usage: torchrun --nproc-per-node N scripts/nccl_check.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
    del os.environ["NCCL_DEBUG"]  # (NCCL prints the banner at WARN too; only an unset variable keeps it quiet)
import numpy as np
import torch
import torch.distributed as dist

from headlamp_intel_gpu_plugin_b200 import columns as K
from headlamp_intel_gpu_plugin_b200 import native, shard, synth

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
ctx = native.Context(local)
uid = [ctx.comm_unique_id() if rank == 0 else None]
dist.broadcast_object_list(uid, src=0)
ctx.comm_init(uid[0], rank, world)

NODES, CHIPS, T, PODS = 40 * world, 4, 2880, 20_000 * world
lo, hi = shard.node_range(NODES, rank, world)
s = ctx.series_create((hi - lo) * CHIPS, T, synth.even_node_offsets(hi - lo, CHIPS))
s.generate(synth.DEFAULT_SEED, global_first=lo * CHIPS, nan_per_million=1000)
_, q, st, fleet = s.rollup(60, 0.95, fetch_windows=False)
full = synth.synth_fleet(NODES, PODS)
cols, _ = shard.shard_fleet(full, rank, world)  # node range + this rank's share of the pods without a node
out = ctx.alloc_rollup(cols)
ok = True
if rank == 0:
    from oracle import columnar as oracle

    x = oracle.generate_series(NODES * CHIPS, T, synth.DEFAULT_SEED, 0, 1000, threads=4)
    _, _, ofleet = oracle.quantile_by_node(x, synth.even_node_offsets(NODES, CHIPS), 0.95, threads=4)
    ok &= bool(np.allclose(fleet[:5], ofleet[:5], rtol=1e-12))
    want = oracle.alloc_rollup(full, threads=4)
    bad = np.nonzero(out.fleet_totals != want.fleet_totals)[0]  # all 96 slots, node-borne and pod-borne
    ok &= len(bad) == 0
    if len(bad):
        print(f"nccl_check: totals slots {bad.tolist()} differ: {out.fleet_totals[bad]} vs oracle {want.fleet_totals[bad]}")
    print(f"nccl_check world={world}: fleet vector {fleet[:5]} vs oracle {ofleet[:5]}; totals ok -> {'PASS' if ok else 'FAIL'}")
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
