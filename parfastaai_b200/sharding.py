"""Host side of the multi-GPU row-block sharding (include/pfaai_gpu.h, pfaai_gpu_set_shard).

Every GPU computes only the blocks of 384 target genomes it owns.  With the replicated build every GPU holds the whole index and
there is no data-path collective; with the partitioned / routed builds every GPU sorts the postings and
builds the column dictionaries of its own row blocks only and the dictionaries are all-gathered over NVLink -- the one exchange
step of the path.  This module restates the ownership rule and the slot layout of that exchange (so that hosts can plan buffers
before any GPU work) and assembles the shards' compact matrices into the reference's dense output (src/main.cpp:143-154),
mirroring the upper triangle in all-vs-all mode.  (With pfaai_gpu_set_global_matrix the shards store straight into one shared
G x G matrix, mirror images included, and no assembly is needed.)
"""
from __future__ import annotations

import numpy as np

BLOCK = 384          # == TILE_ROWS of csrc/common.cuh (24 warps x 16 rows)
ALL_VS_ALL, QUERY_SUBSET, QUERY_DB = 0, 1, 2


def owner(block: int, n_shards: int) -> int:
    """Folded assignment: blocks 0..n-1 go to shards 0..n-1, blocks n..2n-1 to shards n-1..0, and so on."""
    f = block % (2 * n_shards)
    return f if f < n_shards else 2 * n_shards - 1 - f


def owned_indices(extent: int, shard: int, n_shards: int) -> np.ndarray:
    """Global indices (along the sharded axis) stored by `shard`, in storage order."""
    nb = (extent + BLOCK - 1) // BLOCK
    parts = [np.arange(b * BLOCK, min((b + 1) * BLOCK, extent), dtype=np.int32) for b in range(nb) if owner(b, n_shards) == shard]
    return np.concatenate(parts) if parts else np.zeros(0, np.int32)


def shard_shape(mode: int, n_genomes: int, n_qry: int, shard: int, n_shards: int):
    """Shape of the compact matrix of one shard (matches pfaai_gpu_matrix_shape after a compute)."""
    if mode == ALL_VS_ALL:
        return owned_indices(n_genomes, shard, n_shards).shape[0], n_genomes
    n_tgt = n_genomes if mode == QUERY_SUBSET else n_genomes - n_qry
    return n_qry, owned_indices(n_tgt, shard, n_shards).shape[0]


def pair_count(mode: int, n_genomes: int, n_qry: int, shard: int, n_shards: int, qry_ids=None) -> int:
    """Genome pairs (BASELINE metric unit) a shard computes: all-vs-all counts cells with row < column."""
    if mode == ALL_VS_ALL:
        rows = owned_indices(n_genomes, shard, n_shards).astype(np.int64)
        return int((n_genomes - 1 - rows).sum())
    r, c = shard_shape(mode, n_genomes, n_qry, shard, n_shards)
    return r * c


# ---- partitioned / routed build (include/pfaai_gpu.h, pfaai_gpu_build_partitioned): where a column block's dictionary lives ----
COLS = 64            # == COLS_PER_TILE: genomes per column block
RB_CBS = BLOCK // COLS   # column blocks per row block


def slots_per_shard(n_genomes: int, n_shards: int) -> int:
    """Dictionary slots in one shard's chunk of the exchange buffers: the largest number of row blocks any shard owns x 6."""
    nb = (n_genomes + BLOCK - 1) // BLOCK
    per = [0] * n_shards
    for b in range(nb):
        per[owner(b, n_shards)] += 1
    return max(per) * RB_CBS


def slot_of_column_block(J: int, n_genomes: int, n_shards: int) -> int:
    """Slot (index into the all-gathered stage-range / column-T tables) of column block J: the shard that owns J's row block
    built it, as the k-th column block of its own list (ascending), into chunk `owner` of every exchange buffer."""
    I = J // RB_CBS
    o = owner(I, n_shards)
    before = sum(1 for b in range(I) if owner(b, n_shards) == o)
    return o * slots_per_shard(n_genomes, n_shards) + before * RB_CBS + (J - I * RB_CBS)


def owned_column_blocks(n_genomes: int, shard: int, n_shards: int):
    ncb = (n_genomes + COLS - 1) // COLS
    return [J for J in range(ncb) if owner(J // RB_CBS, n_shards) == shard]


def assemble(mode: int, parts, layouts, out: np.ndarray) -> np.ndarray:
    """Scatter the shards' compact matrices into the dense output `out` (zero-initialised by the caller or not: every
    cell is written).  parts[i] is shard i's matrix, layouts[i] its global indices along the sharded axis."""
    if mode == ALL_VS_ALL:
        for part, ids in zip(parts, layouts):
            out[ids, :] = part                    # upper triangle of the owned rows (lower part of `part` is zero)
        iu = np.triu_indices(out.shape[0], 1)
        out[(iu[1], iu[0])] = out[iu]             # M[j][i] = M[i][j]
        np.fill_diagonal(out, 0.0)
        # drop whatever the zero-filled lower parts of the shards wrote below the diagonal: done by the mirror above
    else:
        for part, ids in zip(parts, layouts):
            out[:, ids] = part
    return out
