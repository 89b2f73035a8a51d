"""Relocalisation / loop-query kNN: brute-force Hamming top-2 of query descriptors against a device-resident keyframe
descriptor database, optionally sharded over the ranks of a torch.distributed job (one shard per GPU, per-shard top-2
keys all-gathered over NCCL and merged on the device). Generalises the scan of match/robust.cc:252-274."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import check, lib, register

register("lpslam_knn_create", C.c_int, [C.POINTER(C.c_void_p), C.c_int, C.c_uint64, C.c_uint64])
register("lpslam_knn_destroy", None, [C.c_void_p])
register("lpslam_knn_add", C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64])
register("lpslam_knn_add_dev", C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64])
register("lpslam_knn_size", C.c_int, [C.c_void_p, C.POINTER(C.c_uint64)])
register("lpslam_knn_query", C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p])
register("lpslam_knn_query_dev", C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p])
register("lpslam_knn_merge_dev", C.c_int, [C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p])
register("lpslam_knn_merge_peer_dev", C.c_int, [C.POINTER(C.c_void_p), C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p])
register("lpslam_knn_sharded_query_dev", C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p,
                                                   C.c_void_p])
register("lpslam_knn_unpack", C.c_int, [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p])

EMPTY_KEY = 0xFFFFFFFFFFFFFFFF
IDX_MASK = (1 << 40) - 1


def shard_range(n_rows, rank, world):
    """contiguous row range of `rank` (the first n_rows % world ranks get one extra row)"""
    base, rem = divmod(n_rows, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def merge_keys_host(keys):
    """CPU statement of the shard merge (host logic, used by the gloo tests): keys (n_shards, nq, 2) uint64 -> (nq, 2)"""
    k = np.asarray(keys, np.uint64)
    flat = np.sort(k.transpose(1, 0, 2).reshape(k.shape[1], -1), axis=1)
    return flat[:, :2].copy()


def unpack_keys_host(keys):
    k = np.asarray(keys, np.uint64)
    empty = k == np.uint64(EMPTY_KEY)
    idx = np.where(empty[:, 0], -1, (k[:, 0] & np.uint64(IDX_MASK)).astype(np.int64))
    bd = np.where(empty[:, 0], 256, (k[:, 0] >> np.uint64(40)).astype(np.int64)).astype(np.uint32)
    sd = np.where(empty[:, 1], 256, (k[:, 1] >> np.uint64(40)).astype(np.int64)).astype(np.uint32)
    return idx, bd, sd


class KnnDatabase:
    """One shard of the keyframe-descriptor database on one GPU."""

    def __init__(self, capacity_rows, device=0, index_base=0):
        h = C.c_void_p()
        self._h = None
        check(lib().lpslam_knn_create(C.byref(h), device, int(capacity_rows), int(index_base)))
        self._h = h
        self.device = device

    def close(self):
        if getattr(self, "_h", None) and lib is not None:  # `lib` is already gone at interpreter shutdown
            lib().lpslam_knn_destroy(self._h)
            self._h = None

    __del__ = close

    def add(self, descriptors):
        d = np.ascontiguousarray(descriptors, np.uint8).reshape(-1, 32)
        check(lib().lpslam_knn_add(self._h, d.ctypes.data, len(d)))

    def add_dev(self, ptr, n_rows):
        check(lib().lpslam_knn_add_dev(self._h, ptr, int(n_rows)))

    def __len__(self):
        n = C.c_uint64()
        check(lib().lpslam_knn_size(self._h, C.byref(n)))
        return int(n.value)

    def query(self, queries):
        q = np.ascontiguousarray(queries, np.uint8).reshape(-1, 32)
        bi = np.empty(len(q), np.int64)
        bd = np.empty(len(q), np.uint32)
        sd = np.empty(len(q), np.uint32)
        check(lib().lpslam_knn_query(self._h, q.ctypes.data, len(q), bi.ctypes.data, bd.ctypes.data, sd.ctypes.data))
        return bi, bd, sd

    def query_dev(self, q_ptr, nq, keys_ptr, stream=None):
        check(lib().lpslam_knn_query_dev(self._h, q_ptr, nq, keys_ptr, stream))


def sharded_query(db, q_dev, nq, stream_ptr=None):
    """All ranks call this with the same device-resident queries (torch uint8 tensor (nq, 32)): per-shard top-2 on the
    local shard, all_gather of the (nq, 2) uint64 keys over NCCL, device merge. Returns a torch int64 tensor (nq, 2) of
    merged keys on every rank."""
    import torch
    import torch.distributed as dist
    if stream_ptr is None:  # the scan, the all_gather and the merge must be ordered on ONE stream: torch's current one
        stream_ptr = torch.cuda.current_stream(q_dev.device).cuda_stream
    keys = torch.empty((nq, 2), dtype=torch.int64, device=q_dev.device)
    db.query_dev(q_dev.data_ptr(), nq, keys.data_ptr(), stream_ptr)
    world = dist.get_world_size() if dist.is_initialized() else 1
    if world == 1:
        return keys
    gathered = torch.empty((world, nq, 2), dtype=torch.int64, device=q_dev.device)
    dist.all_gather_into_tensor(gathered, keys)
    merged = torch.empty_like(keys)
    check(lib().lpslam_knn_merge_dev(gathered.data_ptr(), world, nq, merged.data_ptr(), stream_ptr))
    return merged
