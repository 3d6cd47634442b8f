"""Only the partitioned leg of bench.py (bench_partitioned.run_partitioned), for A/B runs:

    [B200_HALO_MODE=push|split|pull|direct] [B200_BENCH_MODES=split,push@2] [B200_BENCH_PROFILE=gpurun_out/x.txt] python -m torch.distributed.run \
        --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 \
        profiles/bench_partitioned_only.py [--scale 1] [--steps 5] [--tag name]

Prints rank 0's `partitioned` record as one JSON line (with the tag and the env switches).
"""
import argparse
import json
import os
import sys

os.environ.setdefault("TORCH_NCCL_HIGH_PRIORITY", "1")   # barriers are not queued behind SpMM CTAs
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scale", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--epochs", type=int, default=5)
    ap.add_argument("--tag", default="")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from bench_partitioned import run_partitioned
    from dgl_hack_b200 import kernel as K
    l0 = K.launch_count()
    rec = run_partitioned(rank, world, dev, scale=args.scale, steps=args.steps,
                          warmup=args.warmup, epochs=args.epochs)
    if rank == 0:
        rec["gpu_launches_rank0"] = K.launch_count() - l0
        rec["tag"] = args.tag
        rec["env"] = {k: v for k, v in os.environ.items() if k.startswith("B200_")}
        print(json.dumps(rec), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
