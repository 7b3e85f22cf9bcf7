"""World-size-2 gloo tests (CPU) of the multi-GPU host logic: utterance sharding of the
feature path (no collective) and the row-sharded score matrix with its one broadcast and
one small reduce.  The CUDA block computation is replaced by the oracle here -- only the
partition / exchange logic is under test; the GPU tests cover the kernels."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import scoring_oracle as so
from speaker_verification_b200 import batch, scoring, synth


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _oracle_block(E_block, T, ti_local, tj):
    S = so.cosine_scores(E_block.numpy(), T.numpy()).astype(np.float32)
    return torch.from_numpy(S[ti_local.numpy(), tj.numpy()]), None


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # ---- feature path: every rank derives the same partition without talking
        lengths = synth.cfg3_lengths(999, seed=3)
        shards = batch.shard_indices(lengths, world)
        mine = shards[rank]
        counts = [torch.zeros(2, dtype=torch.int64) for _ in range(world)]
        dist.all_gather(counts, torch.tensor([len(mine), int(lengths[mine].sum())]))
        ids = [None] * world
        dist.all_gather_object(ids, mine.tolist())
        if rank == 0:
            allc = torch.stack(counts).numpy()
            assert allc[:, 0].sum() == len(lengths)
            assert sorted(sum(ids, [])) == list(range(len(lengths)))          # disjoint cover
            assert allc[:, 1].max() - allc[:, 1].min() <= lengths.max()       # balanced by samples
        # ---- score path: rows sharded, T broadcast from rank 0, trial scores reduced to rank 0
        m, n, d = 203, 157, 64
        E, T = so.synth_embeddings(m, n, d, seed=8)
        ti, tj, lab = so.synth_trials(m, n, 3000, seed=8)
        bounds = scoring.row_partition(m, world)
        r0, r1 = bounds[rank], bounds[rank + 1]
        T_local = torch.from_numpy(T.copy()) if rank == 0 else torch.zeros(n, d)  # only rank 0 has T
        eer, auc, full = scoring.score_trials_sharded(
            torch.from_numpy(E[r0:r1].copy()), r0, m, T_local, torch.from_numpy(ti), torch.from_numpy(tj),
            torch.from_numpy(lab), block_scores=_oracle_block)
        assert torch.equal(T_local, torch.from_numpy(T))                     # broadcast happened
        if rank == 0:
            ref = so.cosine_scores(E, T).astype(np.float32)[ti, tj]
            assert np.array_equal(full.numpy(), ref)                          # reduce == exact gather
            np.save(os.path.join(out_dir, "ok.npy"), np.array([1]))
        else:
            assert full is None
    finally:
        dist.destroy_process_group()


def test_two_rank_partition_and_score_exchange(tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert os.path.exists(tmp_path / "ok.npy")


def test_row_partition_sizes():
    b = scoring.row_partition(4874, 8)
    sizes = np.diff(b)
    assert b[0] == 0 and b[-1] == 4874 and set(sizes.tolist()) == {609, 610}
    assert scoring.row_partition(5, 8)[-1] == 5
