"""World-size-2 (gloo, CPU) test of the N>1 host path: ownership rule, per-rank compact parts, gather to rank 0 and the
mirroring assembly.  No GPU: each rank cuts "its" part out of the oracle's matrix exactly as a sharded
context would deliver it (upper triangle of the owned rows), so everything around the kernels is exercised."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _worker(rank, world, port, N, q):
    import torch
    import torch.distributed as dist
    from oracle import pfaai_oracle as orc
    from parfastaai_b200 import sharding, synth
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    P = 2
    off, vals = synth.generate(N, P, seed=5)
    T = synth.t_matrix(N, P, off)
    ids = np.arange(N, dtype=np.int32)
    full = orc.direct_aji(P, N, off, vals, T, ids, ids)["aji"]
    mine = sharding.owned_indices(N, rank, world)
    part = np.triu(full, 1)[mine, :]                       # what pfaai_gpu_download_matrix returns on this rank
    assert part.shape == sharding.shard_shape(sharding.ALL_VS_ALL, N, 0, rank, world)
    pairs = torch.tensor([sharding.pair_count(sharding.ALL_VS_ALL, N, 0, rank, world)], dtype=torch.int64)
    dist.all_reduce(pairs)
    assert int(pairs.item()) == N * (N - 1) // 2
    # gather the compact parts on rank 0 (sizes differ per rank)
    sizes = [sharding.shard_shape(sharding.ALL_VS_ALL, N, 0, r, world) for r in range(world)]
    if rank == 0:
        parts = [part]
        for r in range(1, world):
            buf = torch.empty(sizes[r], dtype=torch.float64)
            dist.recv(buf, src=r)
            parts.append(buf.numpy())
        lay = [sharding.owned_indices(N, r, world) for r in range(world)]
        out = sharding.assemble(sharding.ALL_VS_ALL, parts, lay, np.empty((N, N)))
        q.put(bool(np.array_equal(out, full)))
    else:
        dist.send(torch.from_numpy(np.ascontiguousarray(part)), dst=0)
    dist.barrier()
    dist.destroy_process_group()


def _exchange_worker(rank, world, port, G, q):
    """The in-place chunk all-gather of the partitioned build, on CPU: every rank fills its chunk of a slot table with the ids of
    the column blocks it built (position k of its own list), all_gather_into_tensor-style exchange, then every column block must
    be found at slot_of_column_block() on every rank."""
    import torch
    import torch.distributed as dist
    from parfastaai_b200 import sharding
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    cap = sharding.slots_per_shard(G, world)
    mine = sharding.owned_column_blocks(G, rank, world)
    chunk = torch.full((cap,), -1, dtype=torch.int64)
    chunk[:len(mine)] = torch.tensor(mine, dtype=torch.int64)
    table = [torch.empty(cap, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(table, chunk)
    table = torch.cat(table).numpy()
    ncb = (G + sharding.COLS - 1) // sharding.COLS
    ok = all(table[sharding.slot_of_column_block(J, G, world)] == J for J in range(ncb))
    ok = ok and int((table >= 0).sum()) == ncb
    q.put(bool(ok))
    dist.barrier()
    dist.destroy_process_group()


def test_partitioned_slot_layout_after_the_chunk_exchange():
    import torch.multiprocessing as mp
    from parfastaai_b200 import sharding
    # the layout rule itself: slots are a bijection of the column blocks, inside the owner's chunk, for ragged sizes too
    for n in (2, 3, 4, 8):
        for G in (900, 1000, 6000, 50000):
            ncb = (G + sharding.COLS - 1) // sharding.COLS
            cap = sharding.slots_per_shard(G, n)
            slots = [sharding.slot_of_column_block(J, G, n) for J in range(ncb)]
            assert len(set(slots)) == ncb
            assert all(s // cap == sharding.owner(J // sharding.RB_CBS, n) for J, s in enumerate(slots))
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + os.getpid() % 2000
    procs = [ctx.Process(target=_exchange_worker, args=(r, 2, port, 5000, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=240)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True and q.get(timeout=5) is True


def test_ownership_rule():
    from parfastaai_b200 import sharding
    for n in (1, 2, 3, 4, 8):
        for extent in (1, 383, 384, 385, 5000, 50000):
            got = np.sort(np.concatenate([sharding.owned_indices(extent, s, n) for s in range(n)]))
            assert np.array_equal(got, np.arange(extent))
    # folded assignment: 8 shards, 50,000 genomes -> the triangular work is balanced to a few percent
    w = [sharding.pair_count(0, 50000, 0, s, 8) for s in range(8)]
    assert sum(w) == 50000 * 49999 // 2 and max(w) / min(w) < 1.05


def test_two_rank_gather_and_mirror():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, 900, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=240)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True
