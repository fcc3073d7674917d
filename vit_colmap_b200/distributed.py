"""Multi-GPU orchestration of the hot path: one process per GPU (torchrun), `torch.distributed` for
the plumbing (NCCL on the GPUs, gloo in the CPU tests).

The reference's only parallelism on this path is independent image pairs handed to per-GPU worker
threads, with descriptors re-uploaded from the host per pair
(COLMAP/controllers/feature_matching_utils.cc:249-283, SIFTGPU/SiftMatchCU.cpp:99-112).  Here:
  * extraction shards by image (contiguous ranges of the sorted file list, vit_extractor.py:684-687);
    the PCA projection fitted on the first image (vit_extractor.py:601-639) is broadcast from its
    owner so the result does not depend on the sharding;
  * ONE all-gather makes every rank's descriptors resident everywhere;
  * matching shards by pair block (pairing.shard_blocks, collective-free, same table on every rank);
  * match lists are gathered to rank 0, the single database writer (as in the reference,
    feature_matching_utils.cc:480-500).
The compute calls are injected (`match_fn`), so the host logic is testable without a GPU.
"""
from __future__ import annotations

from typing import Callable, Dict, List, Optional, Sequence, Tuple

import numpy as np

from .pairing import ExhaustivePairGenerator, shard_blocks, shard_blocks_local_first


def _dist():
    import torch.distributed as dist
    return dist


def world_info() -> Tuple[int, int]:
    dist = _dist()
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_range(n: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous split of n sorted images; the first ranks take the remainder."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def broadcast_projection(projection, channels: int, dim: int, device, src: int = 0):
    """PCA projection [C, d] f32 from the rank that owns image 0 to everyone."""
    import torch
    dist = _dist()
    rank, world = world_info()
    if world == 1:
        return projection
    if rank != src or projection is None:
        projection = torch.empty((channels, dim), dtype=torch.float32, device=device)
    projection = projection.to(device=device, dtype=torch.float32).contiguous()
    dist.broadcast(projection, src=src)
    return projection


def all_gather_descriptors(desc_local, counts_local, n_total: int):
    """desc_local [n_local, K, d], counts_local int32 [n_local] (shard_range order) ->
    (desc [n_total, K, d], counts [n_total]) on every rank.  One all-gather for the payload, one
    for the counts; shards are padded to the largest shard so the collective is regular."""
    import torch
    dist = _dist()
    rank, world = world_info()
    if world == 1:
        return desc_local, counts_local
    n_max = -(-n_total // world)
    K, d = desc_local.shape[1], desc_local.shape[2]
    pad_d = torch.zeros((n_max, K, d), dtype=desc_local.dtype, device=desc_local.device)
    pad_c = torch.zeros((n_max,), dtype=torch.int32, device=desc_local.device)
    pad_d[: desc_local.shape[0]] = desc_local
    pad_c[: counts_local.shape[0]] = counts_local
    out_d = torch.empty((world * n_max, K, d), dtype=desc_local.dtype, device=desc_local.device)
    out_c = torch.empty((world * n_max,), dtype=torch.int32, device=desc_local.device)
    dist.all_gather_into_tensor(out_d, pad_d)
    dist.all_gather_into_tensor(out_c, pad_c)
    keep = []
    for r in range(world):
        lo, hi = shard_range(n_total, r, world)
        keep.extend(range(r * n_max, r * n_max + (hi - lo)))
    idx = torch.tensor(keep, dtype=torch.long, device=desc_local.device)
    return out_d.index_select(0, idx).contiguous(), out_c.index_select(0, idx).contiguous()


class DescriptorExchange:
    """The one all-gather of the path, started asynchronously: ``start`` enqueues the collective(s) on the
    communication stream and returns at once; ``wait`` makes the current stream wait for them and returns
    (desc [n_total, K, d], counts [n_total]).  Between the two the caller matches the pair blocks whose
    images are all its own (pairing.shard_blocks_local_first), so the exchange costs no job time."""

    def __init__(self, desc_local, counts_local, n_total: int):
        import torch
        dist = _dist()
        self.rank, self.world = world_info()
        self.n_total = int(n_total)
        self.local = (desc_local, counts_local)
        self.works = []
        if self.world == 1:
            return
        n_max = -(-self.n_total // self.world)
        K, d = desc_local.shape[1], desc_local.shape[2]
        if desc_local.shape[0] == n_max:
            pad_d, pad_c = desc_local.contiguous(), counts_local.contiguous()
        else:  # shards are padded to the largest one so that the collective is regular
            pad_d = torch.zeros((n_max, K, d), dtype=desc_local.dtype, device=desc_local.device)
            pad_c = torch.zeros((n_max,), dtype=torch.int32, device=desc_local.device)
            pad_d[: desc_local.shape[0]] = desc_local
            pad_c[: counts_local.shape[0]] = counts_local
        self.out_d = torch.empty((self.world * n_max, K, d), dtype=desc_local.dtype, device=desc_local.device)
        self.out_c = torch.empty((self.world * n_max,), dtype=torch.int32, device=desc_local.device)
        self._keep = (pad_d, pad_c)
        self.n_max = n_max
        self.works = [dist.all_gather_into_tensor(self.out_d, pad_d, async_op=True),
                      dist.all_gather_into_tensor(self.out_c, pad_c, async_op=True)]

    def wait(self):
        import torch
        if self.world == 1:
            return self.local
        for w in self.works:
            w.wait()
        if self.n_total == self.world * self.n_max:  # equal shards: the gathered buffer is the result
            return self.out_d, self.out_c
        keep = []
        for r in range(self.world):
            lo, hi = shard_range(self.n_total, r, self.world)
            keep.extend(range(r * self.n_max, r * self.n_max + (hi - lo)))
        idx = torch.tensor(keep, dtype=torch.long, device=self.out_d.device)
        return self.out_d.index_select(0, idx).contiguous(), self.out_c.index_select(0, idx).contiguous()


def block_owners(blocks: Sequence[Sequence[Tuple[int, int]]], index_of: Dict[int, int], n_total: int,
                 world: int) -> List[int]:
    """Rank that holds every image of a pair block under shard_range, or -1."""
    bounds = [shard_range(n_total, r, world) for r in range(world)]

    def owner_of(i: int) -> int:
        for r, (lo, hi) in enumerate(bounds):
            if lo <= i < hi:
                return r
        return -1

    owners = []
    for blk in blocks:
        ranks = {owner_of(index_of[a]) for pr in blk for a in pr}
        owners.append(ranks.pop() if len(ranks) == 1 else -1)
    return owners


def gather_to_writer(local: dict, dst: int = 0) -> Optional[dict]:
    """Merge per-rank {key: ndarray} dictionaries on the writer rank (None elsewhere)."""
    dist = _dist()
    rank, world = world_info()
    if world == 1:
        return local
    bucket = [None] * world if rank == dst else None
    dist.gather_object(local, bucket, dst=dst)
    if rank != dst:
        return None
    merged: dict = {}
    for part in bucket:
        merged.update(part)
    return merged


MatchFn = Callable[[np.ndarray], List[np.ndarray]]


def match_exhaustive_sharded(image_ids: Sequence[int], match_fn: Callable[[np.ndarray], List[np.ndarray]],
                             block_size: int = 50, min_num_matches: int = 15,
                             skip: Optional[Callable[[int, int], bool]] = None,
                             local_first: Optional[Tuple[Callable[[np.ndarray], List[np.ndarray]], Callable[[], None]]] = None
                             ) -> Optional[Dict[Tuple[int, int], np.ndarray]]:
    """Exhaustive matching of the images `image_ids` (database order) sharded by pair block.

    `match_fn(index_pairs int32 [n,2]) -> list of [m,2] uint32` matches pairs given as indices into
    `image_ids` against the descriptors resident on this rank (GPU: Matcher.match_pairs on the
    all-gathered tensor).  Blocks follow ExhaustivePairGenerator (pairing.cc:202-236); pairs with
    fewer than `min_num_matches` matches are blanked (feature_matching_utils.cc:485-488).
    `local_first = (match_fn_local, wait)`: the descriptor exchange is still in flight.  Blocks whose
    images all sit in this rank's shard_range are matched first with `match_fn_local` (same index space,
    only this rank's images are valid), then `wait()` is called once and the other blocks follow.
    Returns {(id1, id2): matches} on rank 0, None elsewhere."""
    rank, world = world_info()
    ids = [int(i) for i in image_ids]
    index_of = {i: k for k, i in enumerate(ids)}
    blocks = list(ExhaustivePairGenerator(ids, block_size).blocks())
    sizes = [len(b) for b in blocks]
    if local_first is None:
        phases = [(shard_blocks(sizes, world)[rank], match_fn)]
    else:
        own, other = shard_blocks_local_first(sizes, block_owners(blocks, index_of, len(ids), world), world)
        phases = [(own[rank], local_first[0]), (other[rank], match_fn)]
    local: Dict[Tuple[int, int], np.ndarray] = {}
    for k, (mine, fn) in enumerate(phases):
        if k == 1:
            local_first[1]()
        for b in mine:
            pairs = [p for p in blocks[b] if not (skip and skip(*p))]
            if not pairs:
                continue
            idx = np.array([(index_of[a], index_of[c]) for a, c in pairs], dtype=np.int32)
            for pair, m in zip(pairs, fn(idx)):
                m = np.asarray(m, dtype=np.uint32).reshape(-1, 2)
                local[pair] = m if m.shape[0] >= min_num_matches else np.zeros((0, 2), np.uint32)
    return gather_to_writer(local, 0)


def gpu_match_fn(matcher, resident, options, pairs_per_call: int = 1024) -> MatchFn:
    """match_fn over descriptors resident in HBM (`Matcher.upload`) with the fused sm_100a matcher:
    per call only the pair list goes up and the packed match lists come back."""

    def fn(idx: np.ndarray) -> List[np.ndarray]:
        return matcher.match_resident(resident, idx, options.max_ratio, options.max_distance,
                                      options.cross_check, pairs_per_call)
    return fn


def all_gather_objects(obj) -> list:
    dist = _dist()
    rank, world = world_info()
    if world == 1:
        return [obj]
    bucket = [None] * world
    dist.all_gather_object(bucket, obj)
    return bucket


def broadcast_object(obj, src: int = 0):
    dist = _dist()
    rank, world = world_info()
    if world == 1:
        return obj
    box = [obj if rank == src else None]
    dist.broadcast_object_list(box, src=src)
    return box[0]


def extract_and_match(image_dir, db_path, extractor, camera_model: str = "SIMPLE_PINHOLE", camera_params=None,
                      matching_options=None, block_size: int = 50, matcher=None):
    """Whole hot path on `world` GPUs: every rank extracts its contiguous shard of the sorted image
    list with `extractor` (a ViTExtractor bound to this rank's GPU), descriptors are all-gathered
    once, pair blocks are matched where they were dealt, rank 0 writes every database row.

    Matching runs on the uint8 descriptors -- the bytes COLMAP would read back from the database --
    converted exactly to bf16 on the device (inv_norm = 1/512^2), so the result equals the
    single-process `extract()` + `match_exhaustive()` pipeline.  Unreadable files are reported and
    skipped exactly as `extract()` does (no image row, vit_extractor.py:733-737); image ids are the ones
    the single writer hands out (the database may already hold images).  Returns the matches dict on
    rank 0."""
    import cv2
    import torch
    from .database import ColmapDatabase, make_camera
    from .features.base_extractor import default_camera_params, list_images
    from .matching import FeatureMatchingOptions, Matcher
    rank, world = world_info()
    options = matching_options or FeatureMatchingOptions()
    files = list_images(image_dir)
    if not files:
        raise ValueError(f"No images found in {image_dir}")
    n = len(files)
    lo, hi = shard_range(n, rank, world)
    dev = extractor.device
    K, d = extractor.num_keypoints, min(extractor.descriptor_dim, extractor.model.embed_dim)
    needs_pca = extractor.model.embed_dim > extractor.descriptor_dim

    def load(i):
        """(resized image, original size, resized size) or None for a file extract() would skip."""
        img = cv2.imread(str(files[i]))
        if img is None:
            print(f"[{i + 1}/{n}] {files[i].name}: ⚠ Warning: Failed to read image, skipping")
            return None
        try:
            return extractor._resize(img)
        except Exception as e:  # noqa: BLE001  (vit_extractor.py:757-762: report and go on)
            print(f"  ✗ Error during feature extraction: {e}")
            return None

    def run_batch(items):
        return extractor._extract_batch([it[0] for it in items], items[0][1], items[0][2])[0]

    if needs_pca:  # fitted on the first readable image by rank 0, then shared (vit_extractor.py:601-639)
        if rank == 0 and extractor.descriptor_projection is None:
            for i in range(n):
                it = load(i)
                if it is not None:
                    run_batch([it])
                    break
        extractor.descriptor_projection = broadcast_projection(extractor.descriptor_projection,
                                                               extractor.model.embed_dim, d, dev)
    feats: Dict[int, Tuple[np.ndarray, np.ndarray]] = {}
    sizes: Dict[int, Tuple[int, int]] = {}
    pending: list = []

    def flush():
        if pending:
            for (i, _), f in zip(pending, run_batch([it for _, it in pending])):
                feats[i] = f
            pending.clear()

    for i in range(lo, hi):
        it = load(i)
        if it is None:
            continue
        sizes[i] = it[1]
        if pending and (pending[0][1][1:] != it[1:] or len(pending) == extractor.batch_size):
            flush()  # a batch holds equally sized images
        pending.append((i, it))
    flush()
    # uint8 descriptors of this shard, padded to K rows -> one all-gather (skipped files keep a
    # zero-row slot so that the image axis is the file index on every rank)
    desc_local = torch.zeros((hi - lo, K, d), dtype=torch.uint8)
    counts_local = torch.zeros((hi - lo,), dtype=torch.int32)
    for i, (_, du8) in feats.items():
        m = min(len(du8), K)
        desc_local[i - lo, :m] = torch.from_numpy(du8[:m])
        counts_local[i - lo] = m
    desc_u8, counts = all_gather_descriptors(desc_local.to(dev), counts_local.to(dev), n)
    matcher = matcher or Matcher(extractor.device_index)
    resident = matcher.upload(desc_u8, counts.cpu().numpy())

    # rank 0 is the single writer: camera + image rows first, so that matching uses the ids it got
    size_of = dict(kv for part in all_gather_objects(sorted(sizes.items())) for kv in part)
    readable = sorted(size_of)
    db, file_ids = None, None
    if rank == 0:
        if not readable:
            raise ValueError(f"Failed to read any image in {image_dir}")
        db = ColmapDatabase(str(db_path))
        w, h = size_of[readable[0]]
        params = camera_params if camera_params is not None else default_camera_params(camera_model, w, h)
        cam = db.db.write_camera(make_camera(camera_model, w, h, params))
        file_ids = [(i, db.add_image(files[i].name, camera_id=cam)) for i in readable]
    file_ids = broadcast_object(file_ids, 0)
    file_of_pos = np.array([i for i, _ in file_ids], dtype=np.int32)
    inner = gpu_match_fn(matcher, resident, options)
    matches = match_exhaustive_sharded([image_id for _, image_id in file_ids],
                                       lambda idx: inner(file_of_pos[idx]), block_size, options.min_num_matches)
    all_feats = gather_to_writer(feats, 0)
    if rank != 0:
        return None
    for i, image_id in file_ids:
        kps, du8 = all_feats[i]
        if len(kps):
            db.add_keypoints(image_id, kps)
            db.add_descriptors(image_id, du8)
    for (a, b), m in matches.items():
        db.add_matches(a, b, m)
    db.close()
    return matches
