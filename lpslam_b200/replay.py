"""Frame-parallel offline replay (BASELINE config 4): frames are independent (the extractor keeps no cross-frame state,
feature/orb_extractor.cc:85-88 aside), so a replay of N frames is split into contiguous per-rank ranges — lpslam's
ReplayEngine reads its log in contiguous chunks as well (Manager/ReplayEngine.h:53) — and every rank runs its range in
batches through its own extractor context. No data-path collective; results stay with the rank that produced them
(gathered on the host only when the caller asks)."""
from . import knn


def frame_range(n_frames, rank, world):
    """contiguous frame range of `rank`; the first n_frames % world ranks get one extra frame"""
    return knn.shard_range(n_frames, rank, world)


def replay(extractor, frames, batch=None):
    """Extract every frame of `frames` (sequence of equally sized 8UC1 images) in batches of `batch` (default: the
    context's max_batch). Yields (keypoints, descriptors) per frame, in order."""
    batch = batch or extractor.max_batch
    for i in range(0, len(frames), batch):
        for res in extractor.extract_batch(frames[i:i + batch]):
            yield res


def replay_rank(extractor, frames, rank, world, batch=None):
    """this rank's share of a replay: (first_frame_index, list of (keypoints, descriptors))"""
    lo, hi = frame_range(len(frames), rank, world)
    return lo, list(replay(extractor, frames[lo:hi], batch))
