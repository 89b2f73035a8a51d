"""CPU, world_size 2 over gloo: the host-side logic of the two sharded workloads.
  * relocalisation kNN: contiguous row shards, per-shard top-2 keys, all_gather, merge == one sequential strict-< scan
  * frame-parallel replay: contiguous frame chunks per rank cover every frame exactly once, results gather in order
The per-shard scan itself runs on the oracle here (no GPU in this container); on the GPU box the same exchange runs over
NCCL with the CUDA scan (tests/test_knn_gpu.py, bench.py)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from lpslam_b200.synth import synth_descriptors

WORLD = 2


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _pack(best_idx, best_dist, second_dist, second_idx):
    from lpslam_b200 import knn
    k = np.full((len(best_idx), 2), knn.EMPTY_KEY, np.uint64)
    ok = best_idx >= 0
    k[ok, 0] = (best_dist[ok].astype(np.uint64) << np.uint64(40)) | best_idx[ok].astype(np.uint64)
    ok2 = second_idx >= 0
    k[ok2, 1] = (second_dist[ok2].astype(np.uint64) << np.uint64(40)) | second_idx[ok2].astype(np.uint64)
    return k


def _shard_top2(q, db, base):
    """per-shard (best, second) as packed keys; numpy statement of the shard scan (ties -> lowest global index)"""
    d = np.unpackbits(q[:, None, :] ^ db[None, :, :], axis=2).sum(axis=2).astype(np.int64)  # (nq, n)
    key = (d << 40) | (np.arange(len(db), dtype=np.int64) + base)[None, :]
    key.sort(axis=1)
    out = np.full((len(q), 2), -1, np.int64)
    out[:, : min(2, key.shape[1])] = key[:, :2]
    return out.view(np.uint64)


def _worker(rank, port, n_db, n_q, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=WORLD)
    from lpslam_b200 import knn, replay
    db, q, _ = synth_descriptors(n_db, n_q, seed=6)
    db[n_db - 3] = db[1]  # duplicate rows in different shards: the lower global index must win
    lo, hi = knn.shard_range(n_db, rank, WORLD)
    mine = torch.from_numpy(_shard_top2(q, db[lo:hi], lo).view(np.int64).copy())
    gathered = [torch.empty_like(mine) for _ in range(WORLD)]
    dist.all_gather(gathered, mine)
    merged = knn.merge_keys_host(np.stack([g.numpy().view(np.uint64) for g in gathered]))
    # frame-parallel replay partition
    frames = list(range(37))
    flo, fhi = replay.frame_range(len(frames), rank, WORLD)
    local = [f * f for f in frames[flo:fhi]]  # stand-in for per-frame results
    allres = [None] * WORLD
    dist.all_gather_object(allres, (flo, fhi, local))
    if rank == 0:
        ret["merged"] = merged
        ret["frames"] = allres
    dist.barrier()
    dist.destroy_process_group()


def test_world2_knn_merge_and_replay_partition(oracle):
    n_db, n_q = 4001, 50
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(_free_port(), n_db, n_q, ret), nprocs=WORLD, join=True)
    from lpslam_b200 import knn
    db, q, _ = synth_descriptors(n_db, n_q, seed=6)
    db[n_db - 3] = db[1]
    bi, bd, sd = knn.unpack_keys_host(ret["merged"])
    obi, obd, osd = oracle.knn_top2(q, db)
    assert np.array_equal(bi, obi) and np.array_equal(bd, obd) and np.array_equal(sd, osd)
    covered = []
    for flo, fhi, local in ret["frames"]:
        assert local == [f * f for f in range(flo, fhi)]
        covered += list(range(flo, fhi))
    assert covered == list(range(37))


def test_shard_ranges_cover():
    from lpslam_b200 import knn, replay
    for n in (0, 1, 7, 8, 10_000_000, 4096):
        for world in (1, 2, 3, 4, 8):
            prev = 0
            for r in range(world):
                lo, hi = knn.shard_range(n, r, world)
                assert lo == prev and hi >= lo
                prev = hi
                assert replay.frame_range(n, r, world) == (lo, hi)
            assert prev == n
    assert knn.shard_range(10_000_000, 7, 8) == (8_750_000, 10_000_000)
