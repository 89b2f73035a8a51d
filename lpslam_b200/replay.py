"""Frame-parallel offline replay (BASELINE config 4): frames are independent (the extractor keeps no cross-frame state,
feature/orb_extractor.cc:85-88 aside), so a replay of N frames is split into contiguous per-rank ranges — lpslam's
ReplayEngine reads its log in contiguous chunks as well (Manager/ReplayEngine.h:53) — and every rank runs its range in
batches through its own extractor context. No data-path collective; results stay with the rank that produced them
(gathered on the host only when the caller asks).

PipelinedReplay is the host loop: ONE thread per GPU keeps two batches in flight through lpslam_orb_extract_batch_submit /
lpslam_stereo_frames_submit and their _collect calls, so the host -> device copy of batch n + 1 and the read-back of batch n - 1
run beside the kernels of batch n (the role of ReplayEngine's reader thread, Manager/ReplayEngine.h:40-53)."""
import ctypes as C

import numpy as np

from . import knn
from ._lib import KP_DTYPE, check, lib, register

register("lpslam_orb_extract_batch_submit", C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.c_uint32, C.c_int, C.c_int, C.c_size_t,
                                                      C.POINTER(C.c_uint64)])
register("lpslam_orb_extract_batch_collect", C.c_int, [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p])


def frame_range(n_frames, rank, world):
    """contiguous frame range of `rank`; the first n_frames % world ranks get one extra frame"""
    return knn.shard_range(n_frames, rank, world)


def replay(extractor, frames, batch=None):
    """Extract every frame of `frames` (sequence of equally sized 8UC1 images) in batches of `batch` (default: the
    context's max_batch). Yields (keypoints, descriptors) per frame, in order. Blocking calls, one batch at a time."""
    batch = batch or extractor.max_batch
    for i in range(0, len(frames), batch):
        for res in extractor.extract_batch(frames[i:i + batch]):
            yield res


def replay_rank(extractor, frames, rank, world, batch=None):
    """this rank's share of a replay: (first_frame_index, list of (keypoints, descriptors))"""
    lo, hi = frame_range(len(frames), rank, world)
    return lo, list(replay(extractor, frames[lo:hi], batch))


def pinned_array(shape, dtype=np.uint8):
    """page-locked host array (cudaHostAlloc through the C ABI); the array keeps its allocation alive"""
    dtype = np.dtype(dtype)
    n = int(np.prod(shape)) * dtype.itemsize
    p = C.c_void_p()
    check(lib().lpslam_cuda_host_alloc(C.byref(p), max(n, 1)))
    buf = (C.c_uint8 * max(n, 1)).from_address(p.value)
    arr = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
    _keep[arr.ctypes.data] = (p, buf)
    return arr


_keep = {}


class PipelinedReplay:
    """Two batches in flight on one extractor context. stereo = (focal_x_baseline, true_baseline) runs data::frame's stereo
    constructor on consecutive (left, right) frames; None extracts every frame on its own."""

    def __init__(self, extractor, height, width, batch_frames, stereo=None, cap_per_frame=4096):
        assert batch_frames <= extractor.max_batch and (stereo is None or batch_frames % 2 == 0)
        self.e, self.h, self.w, self.bf, self.stereo = extractor, height, width, batch_frames, stereo
        self.cap = batch_frames * cap_per_frame
        self.out = [dict(kp=pinned_array(self.cap, KP_DTYPE), desc=pinned_array((self.cap, 32)), xr=pinned_array(self.cap, np.float32),
                         dep=pinned_array(self.cap, np.float32), offs=np.zeros(batch_frames + 1, np.uint32)) for _ in range(2)]

    def _submit(self, frames, lo, n):
        img = self.h * self.w
        base = frames.ctypes.data + lo * img
        t = C.c_uint64(0)
        if self.stereo:
            pl = (C.c_void_p * (n // 2))(*[base + 2 * p * img for p in range(n // 2)])
            pr = (C.c_void_p * (n // 2))(*[base + (2 * p + 1) * img for p in range(n // 2)])
            check(lib().lpslam_stereo_frames_submit(self.e._h, pl, pr, n // 2, self.w, self.h, self.w, self.stereo[0], self.stereo[1],
                                                    C.byref(t)))
        else:
            ptrs = (C.c_void_p * n)(*[base + f * img for f in range(n)])
            check(lib().lpslam_orb_extract_batch_submit(self.e._h, ptrs, n, self.w, self.h, self.w, C.byref(t)))
        return t.value

    def _collect(self, ticket, o):
        if self.stereo:
            check(lib().lpslam_stereo_frames_collect(self.e._h, ticket, o["kp"].ctypes.data, o["desc"].ctypes.data, self.cap,
                                                     o["offs"].ctypes.data, o["xr"].ctypes.data, o["dep"].ctypes.data))
        else:
            check(lib().lpslam_orb_extract_batch_collect(self.e._h, ticket, o["kp"].ctypes.data, o["desc"].ctypes.data, self.cap,
                                                         o["offs"].ctypes.data))

    def run(self, frames, lo=0, hi=None, sink=None, cycles=1):
        """frames: contiguous (N, H, W) uint8 host array (page-locked for full copy speed). Processes frames[lo:hi] `cycles` times over;
        sink(first_frame, n_frames, out) sees every finished batch (out = dict kp / desc / xr / dep / offs, valid until the next but
        one batch). Returns (frames, keypoints)."""
        from . import match  # noqa: F401  (registers the stereo entry points)
        assert frames.flags.c_contiguous and frames.dtype == np.uint8 and frames.shape[1:] == (self.h, self.w)
        hi = len(frames) if hi is None else hi
        pending = []
        n_kp = 0
        seq = 0
        for f0 in list(range(lo, hi, self.bf)) * cycles + [None]:
            if f0 is not None:
                n = min(self.bf, hi - f0)
                if self.stereo:
                    n -= n % 2
                if n > 0:
                    pending.append((self._submit(frames, f0, n), f0, n, self.out[seq & 1]))
                    seq += 1
            if len(pending) == 2 or (f0 is None and pending):
                while pending and (len(pending) == 2 or f0 is None):
                    ticket, first, n, o = pending.pop(0)
                    self._collect(ticket, o)
                    n_kp += int(o["offs"][n])
                    if sink:
                        sink(first, n, o)
        return (hi - lo) * cycles, n_kp
