"""Chunk-length sweep of the sequence-sharded step (run under torchrun, one rank per GPU):
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29520 tools/shard_sweep.py --T 250000
times the one-launch route (hmm_b200_genepred_shard_run_x) at every chunk length given and the four-launch route at its own
choice, for ONE sequence of T positions cut over the ranks.  --T 250000 on two ranks = the per-rank work of 1 Mbp on eight."""
import argparse
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from hmm_layer_b200 import ops  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--T", type=int, default=250_000)
    ap.add_argument("--chunks", type=str, default="0,14,16,20,24,28,32,36,40")
    ap.add_argument("--steps", type=int, default=100)
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    params = bench.gene_params_on(dev)
    tl = a.T // world
    rows = []
    ms, ll, launches, _ = bench.measure_c5(a.T, rank, world, dev, params, steps=a.steps, one_launch=False)
    rows.append(("four launches", 0, launches, ms, ll))
    for L in [int(v) for v in a.chunks.split(",")]:
        ops.set_tuning(L, 0)
        got = ops.genepred_shard_fused_chunk(1, 1, tl, world, dev)
        if got == 0:
            rows.append((f"one launch L={L}", 0, 0, float("nan"), float("nan")))
            continue
        ms, ll, launches, _ = bench.measure_c5(a.T, rank, world, dev, params, steps=a.steps, one_launch=True)
        rows.append((f"one launch L={L}", got, launches, ms, ll))
    ops.set_tuning(0, 0)
    if rank == 0:
        print(f"T={a.T} world={world} positions per rank={tl}")
        for name, got, launches, ms, ll in rows:
            tiles = -(-(-(-tl // got)) // 32) if got else 0
            print(f"{name:22s} chunk={got:2d} tiles={tiles:4d} launches={launches} {ms * 1e3:8.1f} us  loglik={ll:.1f}")
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
