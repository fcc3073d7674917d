"""Whole exhaustive-matching job of BASELINE config C4 / C5 through the product's own sharded driver:
synthetic descriptor sets [n_images, kpts, d] bf16 generated on the device (each rank its shard), ONE NCCL
all-gather, pair blocks of 50 x 50 images (ExhaustivePairGenerator) dealt to the ranks by
pairing.shard_blocks, every block matched on the resident set (vc_match_pairs_resident), match lists
gathered on rank 0.  Reports job time (max over ranks), pairs/s, TFLOP/s and the fraction of the measured
sustained bf16 peak.

    python tools/run_exhaustive_job.py [n_images=1024] [kpts=8192] [d=768]
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/run_exhaustive_job.py 1024
"""
import json
import os
import sys
import time
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from bench import _peaks, make_descriptors_device  # noqa: E402
from vit_colmap_b200 import distributed as vd  # noqa: E402
from vit_colmap_b200.matching import FeatureMatchingOptions, Matcher, ResidentDescriptors  # noqa: E402


def main():
    n_images = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
    kpts = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
    d = int(sys.argv[3]) if len(sys.argv) > 3 else 768
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lo, hi = vd.shard_range(n_images, rank, world)
    # the C4 / C5 recipe (bench.make_descriptors_device) in chunks of 32 images, seeded per chunk so that a
    # rank generates exactly its own images
    shard = torch.empty((hi - lo, kpts, d), dtype=torch.bfloat16, device=dev)
    for s in range(lo, hi, 32):
        e = min(hi, s + 32)
        shard[s - lo:e - lo] = make_descriptors_device(torch, dev, e - s, kpts, d, seed=1000 + s)
    counts_local = torch.full((hi - lo,), kpts, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    # the exchange starts asynchronously; blocks whose images are all this rank's own are matched meanwhile
    t1 = time.perf_counter()
    exchange = vd.DescriptorExchange(shard, counts_local, n_images)
    matcher = Matcher(local)
    options = FeatureMatchingOptions()
    counts_host = np.full((n_images,), kpts, np.int32)
    res_local = ResidentDescriptors(shard, counts_host[lo:hi].copy(), hi - lo, 1.0, 0, 0.0, 0)
    fn_local = vd.gpu_match_fn(matcher, res_local, options)
    from vit_colmap_b200.pairing import ExhaustivePairGenerator, shard_blocks_local_first
    ids = list(range(1, n_images + 1))
    blocks = list(ExhaustivePairGenerator(ids, 50).blocks())
    owners = vd.block_owners(blocks, {i: k for k, i in enumerate(ids)}, n_images, world)
    own, other = shard_blocks_local_first([len(b) for b in blocks], owners, world)
    n_pairs = n_rows = n_local_pairs = 0
    overlap = os.environ.get("VC_JOB_NO_OVERLAP") != "1"   # =1: wait for the exchange first (comparison runs)
    if not overlap:
        exchange.wait()
        torch.cuda.synchronize()
    for b in own[rank]:
        idx = np.array([(a - 1 - lo, c - 1 - lo) for a, c in blocks[b]], dtype=np.int32)
        for m in fn_local(idx):
            n_rows += len(m)
        n_pairs += len(idx)
    n_local_pairs = n_pairs
    t_wait0 = time.perf_counter()
    desc, counts = exchange.wait()
    torch.cuda.synchronize()
    t_gather = time.perf_counter() - t_wait0     # what is left of the exchange after the local blocks
    resident = ResidentDescriptors(desc, counts_host, n_images, 1.0, 0, 0.0, 0)
    fn = vd.gpu_match_fn(matcher, resident, options)
    for b in other[rank]:
        idx = np.array([(a - 1, c - 1) for a, c in blocks[b]], dtype=np.int32)
        for m in fn(idx):
            n_rows += len(m)
        n_pairs += len(idx)
    torch.cuda.synchronize()
    t_match = time.perf_counter() - t1
    stats = torch.tensor([t_match, t_gather, float(n_pairs), float(n_rows)], dtype=torch.float64, device=dev)
    mx = stats.clone()
    if world > 1:
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        dist.all_reduce(stats, op=dist.ReduceOp.SUM)
    if rank == 0:
        total_pairs = int(stats[2].item())
        assert total_pairs == n_images * (n_images - 1) // 2, total_pairs
        secs = float(mx[0].item())
        tf = 2.0 * kpts * kpts * d * total_pairs / secs / 1e12
        peak, _, src = _peaks()
        print(json.dumps({"job": f"exhaustive matching of {n_images} images x {kpts} kpts x d={d}", "n_gpus": world, "exchange_overlapped": overlap,
                          "pairs": total_pairs, "match_rows": int(stats[3].item()), "seconds": secs,
                          "pairs_per_s": total_pairs / secs, "tflops_per_gpu": tf / world,
                          "frac_of_sustained_bf16_per_gpu": tf / world / peak, "peak_source": src,
                          "exchange_wait_seconds_after_local_blocks": float(mx[1].item()),
                          "slowest_rank_pairs": int(mx[2].item()), "includes": "all-gather (overlapped with the rank-local pair blocks), pair lists up, packed match lists down to host"}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
