"""world_size-2 gloo tests (CPU) of the multi-GPU host logic: image sharding, PCA broadcast,
descriptor all-gather, pair-block sharding and the gather to the single database writer.  The
matcher itself is replaced by the CPU oracle here (the checker), the collectives are real."""
import os
import sys
from pathlib import Path

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parent.parent


def _make_descriptors(n_images, K, d, seed=0):
    rng = np.random.default_rng(seed)
    desc = np.zeros((n_images, K, d), np.float32)
    counts = rng.integers(K // 2, K + 1, n_images).astype(np.int32)
    prev = None
    for j in range(n_images):
        x = rng.standard_normal((K, d)).astype(np.float32)
        if prev is not None:
            m = K // 2
            x[rng.permutation(K)[:m]] = prev[rng.permutation(K)[:m]] + 0.02 * rng.standard_normal((m, d)).astype(np.float32)
        x /= np.linalg.norm(x, axis=1, keepdims=True)
        desc[j], prev = x, x
    return desc, counts


def _worker(rank, world, port, tmp, n_images, block_size, overlap=False):
    sys.path.insert(0, str(ROOT))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import torch.distributed as dist
    from oracle import matching as om
    from vit_colmap_b200 import distributed as vd
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        K, d = 48, 32
        full, counts = _make_descriptors(n_images, K, d)
        lo, hi = vd.shard_range(n_images, rank, world)
        # PCA projection: owner of image 0 broadcasts
        proj = torch.arange(12, dtype=torch.float32).reshape(4, 3) if rank == 0 else None
        proj = vd.broadcast_projection(proj, 4, 3, torch.device("cpu"))
        assert proj.tolist() == torch.arange(12, dtype=torch.float32).reshape(4, 3).tolist()
        # one all-gather makes every shard resident everywhere
        g_desc, g_cnt = vd.all_gather_descriptors(torch.from_numpy(full[lo:hi]), torch.from_numpy(counts[lo:hi]), n_images)
        assert g_desc.shape == (n_images, K, d) and np.array_equal(g_desc.numpy(), full)
        assert np.array_equal(g_cnt.numpy(), counts)
        ids = list(range(1, n_images + 1))

        def match_fn(idx):
            return [om.match_c(full[a, :counts[a]], full[b, :counts[b]], 0.8, 0.7, True) for a, b in idx]

        seen = []
        if not overlap:
            res = vd.match_exhaustive_sharded(ids, lambda idx: (seen.extend(map(tuple, idx.tolist())), match_fn(idx))[1],
                                              block_size=block_size, min_num_matches=3)
        else:
            # the exchange is started asynchronously; blocks whose images are all local are matched before
            # wait(), against the local shard only, the rest afterwards against the gathered set
            ex = vd.DescriptorExchange(torch.from_numpy(full[lo:hi]), torch.from_numpy(counts[lo:hi]), n_images)
            state = {"waited": False, "local_pairs": 0}

            def match_local(idx):
                assert not state["waited"] and idx.min() >= lo and idx.max() < hi   # own images only
                state["local_pairs"] += len(idx)
                seen.extend(map(tuple, idx.tolist()))
                return [om.match_c(full[lo:hi][a - lo, :counts[a]], full[lo:hi][b - lo, :counts[b]], 0.8, 0.7, True)
                        for a, b in idx]

            def wait():
                d_all, c_all = ex.wait()
                assert np.array_equal(d_all.numpy(), full) and np.array_equal(c_all.numpy(), counts)
                state["waited"] = True

            def match_rest(idx):
                assert state["waited"]
                seen.extend(map(tuple, idx.tolist()))
                return match_fn(idx)

            res = vd.match_exhaustive_sharded(ids, match_rest, block_size=block_size, min_num_matches=3,
                                              local_first=(match_local, wait))
            assert state["waited"]
            np.save(os.path.join(tmp, f"local_{rank}.npy"), np.array([state["local_pairs"]]))
        np.save(os.path.join(tmp, f"seen_{rank}.npy"), np.array(seen, dtype=np.int64).reshape(-1, 2))
        if rank == 0:
            assert res is not None and len(res) == n_images * (n_images - 1) // 2
            for (a, b), m in res.items():
                ref = om.match_c(full[a - 1, :counts[a - 1]], full[b - 1, :counts[b - 1]], 0.8, 0.7, True)
                if len(ref) < 3:
                    ref = np.zeros((0, 2), np.uint32)
                assert m.tolist() == ref.tolist()
            np.save(os.path.join(tmp, "ok.npy"), np.array([len(res)]))
        else:
            assert res is None
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_images,block_size", [(7, 3), (5, 50)])
def test_world_size_2_gloo(tmp_path, n_images, block_size):
    port = 29500 + (os.getpid() % 2000) + n_images
    mp.spawn(_worker, args=(2, port, str(tmp_path), n_images, block_size), nprocs=2, join=True)
    assert (tmp_path / "ok.npy").exists()
    a, b = np.load(tmp_path / "seen_0.npy"), np.load(tmp_path / "seen_1.npy")
    pairs = {tuple(x) for x in a.tolist()} | {tuple(x) for x in b.tolist()}
    assert len(pairs) == len(a) + len(b) == n_images * (n_images - 1) // 2   # disjoint, complete
    if block_size < n_images:
        assert len(a) > 0 and len(b) > 0                                      # both ranks got blocks


@pytest.mark.parametrize("n_images,block_size", [(12, 3), (9, 2)])
def test_world_size_2_gloo_overlapped_exchange(tmp_path, n_images, block_size):
    """the descriptor all-gather overlaps the pair blocks that need no remote data.  Same
    results as the plain sharded run; every pair exactly once; both ranks match local blocks before waiting."""
    port = 31500 + (os.getpid() % 2000) + n_images
    mp.spawn(_worker, args=(2, port, str(tmp_path), n_images, block_size, True), nprocs=2, join=True)
    assert (tmp_path / "ok.npy").exists()
    a, b = np.load(tmp_path / "seen_0.npy"), np.load(tmp_path / "seen_1.npy")
    pairs = {tuple(x) for x in a.tolist()} | {tuple(x) for x in b.tolist()}
    assert len(pairs) == len(a) + len(b) == n_images * (n_images - 1) // 2
    assert int(np.load(tmp_path / "local_0.npy")[0]) > 0 and int(np.load(tmp_path / "local_1.npy")[0]) > 0


def test_shard_blocks_local_first_is_balanced_and_complete():
    from vit_colmap_b200.distributed import block_owners
    from vit_colmap_b200.pairing import ExhaustivePairGenerator, shard_blocks_local_first
    n, world, bs = 1024, 8, 50
    ids = list(range(1, n + 1))
    blocks = list(ExhaustivePairGenerator(ids, bs).blocks())
    owners = block_owners(blocks, {i: k for k, i in enumerate(ids)}, n, world)
    sizes = [len(b) for b in blocks]
    local, other = shard_blocks_local_first(sizes, owners, world)
    flat = sorted(b for r in range(world) for b in local[r] + other[r])
    assert flat == list(range(len(blocks)))
    for r in range(world):
        assert local[r] and all(owners[b] == r for b in local[r])        # every rank has local work to overlap with
        assert sum(sizes[b] for b in local[r]) >= 1000   # >= 75 ms of matching at C5 sizes: more than the 54 ms exchange
    loads = [sum(sizes[b] for b in local[r] + other[r]) for r in range(world)]
    assert max(loads) - min(loads) <= max(sizes) and sum(loads) == n * (n - 1) // 2


def test_shard_range_partitions():
    from vit_colmap_b200.distributed import shard_range
    for n in (0, 1, 7, 8, 1024):
        for world in (1, 2, 3, 8):
            parts = [shard_range(n, r, world) for r in range(world)]
            assert parts[0][0] == 0 and parts[-1][1] == n
            assert all(parts[i][1] == parts[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in parts]
            assert max(sizes) - min(sizes) <= 1
