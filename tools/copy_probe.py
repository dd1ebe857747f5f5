#!/usr/bin/env python3
"""Copy-only probe of the host side of the end-to-end path: every rank (one per GPU, torchrun) moves what one decode
step of config 2 moves -- 15.7 MB of PCM device -> page-locked host, 1 MB of packets host -> device -- with NO kernels,
as fast as the copy engines go, and reports the aggregate rate.  If N ranks together reach no more than the decode
bench's end-to-end byte rate, the end-to-end scaling limit at N GPUs is the host's memory system, not the pipeline.
    python -m torch.distributed.run --nproc-per-node N tools/copy_probe.py [--mb 15.7] [--iters 300]"""
import argparse
import json
import os
import time

import torch
import torch.distributed as dist


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=float, default=15.73)
    ap.add_argument("--iters", type=int, default=300)
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group(backend="nccl", device_id=torch.device("cuda", local))
    n = int(args.mb * 1e6)
    d_out = [torch.zeros(n, dtype=torch.uint8, device="cuda") for _ in range(3)]
    h_out = [torch.zeros(n, dtype=torch.uint8).pin_memory() for _ in range(3)]
    h_in = torch.zeros(1000000, dtype=torch.uint8).pin_memory()
    d_in = torch.zeros(1000000, dtype=torch.uint8, device="cuda")
    s_out, s_in = torch.cuda.Stream(), torch.cuda.Stream()

    def run(iters):
        for k in range(iters):
            with torch.cuda.stream(s_in):
                d_in.copy_(h_in, non_blocking=True)
            with torch.cuda.stream(s_out):
                h_out[k % 3].copy_(d_out[k % 3], non_blocking=True)
        torch.cuda.synchronize()

    run(20)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    run(args.iters)
    dt = time.perf_counter() - t0
    tt = torch.tensor([dt], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    dt = float(tt.item())
    if rank == 0:
        gbs = world * args.iters * n / dt / 1e9
        print(json.dumps({"probe": "copy-only", "n_gpus": world, "d2h_mb_per_iter_per_gpu": args.mb, "iters": args.iters,
                          "ms_per_iter": 1e3 * dt / args.iters, "aggregate_d2h_gb_s": gbs,
                          "equivalent_config2_stream_s_per_s": world * 81.92 * args.iters / dt}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
