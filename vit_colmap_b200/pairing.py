"""Exhaustive pair generation and pair-block sharding (host logic).

Mirrors ``ExhaustivePairGenerator`` (COLMAP/feature/pairing.cc:171-236, block_size default 50 from
COLMAP/feature/pairing.h:43-45) and ``ImagePairToPairId`` (COLMAP/util/types.h:188-208).  The
multi-GPU split follows the reference's only parallelism on this path -- independent image pairs
handed to per-GPU workers (COLMAP/controllers/feature_matching_utils.cc:249-283) -- as a static,
collective-free assignment of pair blocks to ranks.
"""
from __future__ import annotations

from typing import Iterator, List, Sequence, Tuple

import numpy as np

K_MAX_NUM_IMAGES = 2147483647  # COLMAP/util/types.h:100-101


def image_pair_to_pair_id(image_id1: int, image_id2: int) -> int:
    """COLMAP/util/types.h:200-208."""
    if image_id1 > image_id2:
        image_id1, image_id2 = image_id2, image_id1
    return K_MAX_NUM_IMAGES * int(image_id1) + int(image_id2)


def pair_id_to_image_pair(pair_id: int) -> Tuple[int, int]:
    """COLMAP/util/types.h:210-218."""
    image_id2 = pair_id % K_MAX_NUM_IMAGES
    image_id1 = (pair_id - image_id2) // K_MAX_NUM_IMAGES
    return int(image_id1), int(image_id2)


class ExhaustivePairGenerator:
    """Same iteration contract as the reference class: ``next()`` returns the pairs of one
    block_size x block_size image block, ``has_finished()`` tells when all blocks are done."""

    def __init__(self, image_ids: Sequence[int], block_size: int = 50):
        if block_size <= 0:
            raise ValueError("block_size must be positive")  # ExhaustivePairingOptions::Check
        self.image_ids = [int(i) for i in image_ids]
        self.block_size = int(block_size)
        self.num_blocks = -(-len(self.image_ids) // self.block_size)
        self.reset()

    def reset(self) -> None:  # pairing.cc:193-196
        self._start1 = 0
        self._start2 = 0

    def has_finished(self) -> bool:  # pairing.cc:198-200
        return self._start1 >= len(self.image_ids)

    def next(self) -> List[Tuple[int, int]]:  # pairing.cc:202-236
        if self.has_finished():
            return []
        n, bs = len(self.image_ids), self.block_size
        s1, s2 = self._start1, self._start2
        i1 = np.arange(s1, min(n, s1 + bs))[:, None]
        i2 = np.arange(s2, min(n, s2 + bs))[None, :]
        b1, b2 = i1 % bs, i2 % bs
        keep = ((i1 > i2) & (b1 <= b2)) | ((i1 < i2) & (b1 < b2))  # pairing.cc:223-224
        r, c = np.nonzero(keep)  # row-major == idx1 outer, idx2 inner
        ids = self.image_ids
        pairs = [(ids[s1 + a], ids[s2 + b]) for a, b in zip(r.tolist(), c.tolist())]
        self._start2 += bs  # pairing.cc:230-234
        if self._start2 >= n:
            self._start2 = 0
            self._start1 += bs
        return pairs

    def all_pairs(self) -> List[Tuple[int, int]]:
        self.reset()
        out: List[Tuple[int, int]] = []
        while not self.has_finished():
            out.extend(self.next())
        return out

    def blocks(self) -> Iterator[List[Tuple[int, int]]]:
        self.reset()
        while not self.has_finished():
            yield self.next()


def shard_blocks(block_sizes: Sequence[int], world_size: int) -> List[List[int]]:
    """Static, deterministic assignment of pair blocks to ranks: longest-processing-time greedy on
    the number of pairs per block (diagonal blocks hold about half as many pairs).  No collective:
    every rank computes the same table."""
    order = sorted(range(len(block_sizes)), key=lambda b: (-int(block_sizes[b]), b))
    loads = [0] * world_size
    assign: List[List[int]] = [[] for _ in range(world_size)]
    for b in order:
        r = min(range(world_size), key=lambda k: (loads[k], k))
        assign[r].append(b)
        loads[r] += int(block_sizes[b])
    for a in assign:
        a.sort()
    return assign


def shard_blocks_local_first(block_sizes: Sequence[int], block_owner: Sequence[int],
                             world_size: int) -> Tuple[List[List[int]], List[List[int]]]:
    """As shard_blocks, but a block whose images all live on one rank (block_owner[b] = that rank, -1
    otherwise) goes to its owner: those pairs need no remote descriptors, so the owner matches them while
    the all-gather is still in flight.  The remaining blocks are dealt by the same
    longest-processing-time rule on top of the local loads.  Returns (local blocks per rank, other blocks
    per rank); no collective -- every rank computes the same tables."""
    loads = [0] * world_size
    local: List[List[int]] = [[] for _ in range(world_size)]
    rest = []
    for b, size in enumerate(block_sizes):
        r = int(block_owner[b])
        if 0 <= r < world_size:
            local[r].append(b)
            loads[r] += int(size)
        else:
            rest.append(b)
    other: List[List[int]] = [[] for _ in range(world_size)]
    for b in sorted(rest, key=lambda b: (-int(block_sizes[b]), b)):
        r = min(range(world_size), key=lambda k: (loads[k], k))
        other[r].append(b)
        loads[r] += int(block_sizes[b])
    for a in other:
        a.sort()
    return local, other


def shard_pairs(pairs: Sequence[Tuple[int, int]], rank: int, world_size: int) -> List[Tuple[int, int]]:
    """Contiguous split of a flat pair list (used when there are fewer blocks than ranks)."""
    n = len(pairs)
    lo = n * rank // world_size
    hi = n * (rank + 1) // world_size
    return list(pairs[lo:hi])
