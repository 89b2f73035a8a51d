"""Descriptor (de)serialisation of the reference's map files, so that the relocalisation kNN database (knn.py) can be filled
straight from a saved map.

 * convert_descriptors_to_json / convert_json_to_descriptors: data/common.cc:71-95. A descriptor row is stored as 8 numbers,
   the row's 32 bytes read as native (little-endian) uint32 words.
 * load_message_pack / save_message_pack: the container of io/map_database_io.cc:23-91 (nlohmann::json::to_msgpack of
   {"cameras", "keyframes", "landmarks", "frame_next_id", "keyframe_next_id", "landmark_next_id"}; keyframes and landmarks are
   objects keyed by the decimal id, data/map_database.cc:382-416). Only the fields the descriptor path needs are interpreted
   ("n_keypts", "descs", "lm_ids" of data/keyframe.cc:113-138); everything else passes through untouched.
 * landmark_observations: the association step of data/map_database.cc:358-380 over those fields: landmark -> the
   descriptors of the keypoints that observe it. The reference keeps a landmark's observations in a std::map keyed by
   keyframe POINTER (data/landmark.h), so their order - and with it the first-minimum tie-break of compute_descriptor - depends
   on the allocator there; here observations are ordered by keyframe id, then keypoint index.
 * fill_knn_from_map: keyframe keypoint descriptors or landmark representative descriptors (computed on the device by
   lpslam_landmark_descriptors, the batched data::landmark::compute_descriptor that map_database::from_json runs per landmark,
   data/map_database.cc:252-263) appended to a KnnDatabase shard.

The file format is host-side plumbing; the compute (median-distance selection, kNN scan) stays in liblpslam_cuda."""
import numpy as np


def convert_descriptors_to_json(descriptors):
    """data/common.cc:71-83: (n, 32) uint8 -> n lists of 8 uint32 numbers"""
    d = np.ascontiguousarray(descriptors, np.uint8).reshape(-1, 32)
    return [[int(v) for v in row] for row in d.view("<u4")]


def convert_json_to_descriptors(json_descriptors):
    """data/common.cc:85-95: n lists of 8 numbers -> (n, 32) uint8. Raises like json::get<uint32_t> would not: numbers are
    checked to fit 32 bits and rows to hold at least 8 entries (json::at throws on a short row)."""
    n = len(json_descriptors)
    words = np.empty((n, 8), "<u4")
    for i, row in enumerate(json_descriptors):
        if len(row) < 8:
            raise IndexError(f"descriptor {i} has {len(row)} words, 8 needed")
        for j in range(8):
            v = int(row[j])
            if not 0 <= v <= 0xFFFFFFFF:
                raise ValueError(f"descriptor {i} word {j} does not fit uint32")
            words[i, j] = v
    return words.view(np.uint8).reshape(n, 32)


def load_message_pack(path):
    """io/map_database_io.cc:50-91 up to the parse: returns the map as nested dicts / lists"""
    import msgpack
    with open(path, "rb") as f:
        return msgpack.unpackb(f.read(), raw=False, strict_map_key=False)


def save_message_pack(path, cameras, keyframes, landmarks, frame_next_id=0, keyframe_next_id=0, landmark_next_id=0):
    """io/map_database_io.cc:22-48"""
    import msgpack
    doc = {"cameras": cameras, "keyframes": keyframes, "landmarks": landmarks, "frame_next_id": int(frame_next_id),
           "keyframe_next_id": int(keyframe_next_id), "landmark_next_id": int(landmark_next_id)}
    with open(path, "wb") as f:
        f.write(msgpack.packb(doc, use_bin_type=True))


def _keyframes_in_id_order(doc):
    return sorted(((int(k), v) for k, v in doc["keyframes"].items()), key=lambda kv: kv[0])


def keyframe_descriptors(doc):
    """[(keyframe id, (n_keypts, 32) uint8 descriptors, (n_keypts,) int64 landmark ids)] in keyframe-id order"""
    out = []
    for kid, kf in _keyframes_in_id_order(doc):
        descs = convert_json_to_descriptors(kf["descs"])
        lm_ids = np.asarray(kf["lm_ids"], np.int64)
        n = int(kf["n_keypts"])
        if len(descs) != n or len(lm_ids) != n:  # the asserts of data/map_database.cc:297-306 / :362
            raise ValueError(f"keyframe {kid}: n_keypts {n}, {len(descs)} descriptors, {len(lm_ids)} landmark ids")
        out.append((kid, descs, lm_ids))
    return out


def landmark_observations(doc):
    """data/map_database.cc:358-380: returns (landmark ids ascending, obs_descs (total, 32) uint8, offsets (n + 1,) uint32);
    ids missing from "landmarks" are skipped like the reference does (with its warning), landmarks nobody observes get an
    empty range, a landmark referenced twice from one keyframe keeps its first keypoint only."""
    known = sorted(int(k) for k in doc["landmarks"].keys())
    slot = {lm: i for i, lm in enumerate(known)}
    per_lm = [[] for _ in known]
    for _, descs, lm_ids in keyframe_descriptors(doc):
        seen = set()  # landmark::add_observation (data/landmark.cc:45-50) ignores a keyframe that already observes the landmark:
        for idx in np.nonzero(lm_ids >= 0)[0]:  # of several keypoints of one keyframe carrying the same id, the first one counts
            lm = int(lm_ids[idx])
            s = slot.get(lm)
            if s is not None and lm not in seen:
                seen.add(lm)
                per_lm[s].append(descs[idx])
    offsets = np.zeros(len(known) + 1, np.uint32)
    offsets[1:] = np.cumsum([len(p) for p in per_lm])
    flat = [d for p in per_lm for d in p]
    obs = np.stack(flat).astype(np.uint8) if flat else np.zeros((0, 32), np.uint8)
    return np.asarray(known, np.int64), obs, offsets


def fill_knn_from_map(db, doc_or_path, what="keypoints"):
    """Appends the map's descriptors to `db` (a knn.KnnDatabase shard). what = "keypoints": every keyframe's descriptor rows
    in keyframe-id order; returns (keyframe id, keypoint index) per appended row. what = "landmarks": one representative
    descriptor per observed landmark, chosen on the device; returns the landmark id per appended row."""
    doc = load_message_pack(doc_or_path) if isinstance(doc_or_path, (str, bytes)) else doc_or_path
    if what == "keypoints":
        rows, tags = [], []
        for kid, descs, _ in keyframe_descriptors(doc):
            rows.append(descs)
            tags.append(np.stack([np.full(len(descs), kid, np.int64), np.arange(len(descs), dtype=np.int64)], axis=1))
        if rows:
            db.add(np.concatenate(rows))
        return np.concatenate(tags) if tags else np.zeros((0, 2), np.int64)
    if what == "landmarks":
        from .match import compute_landmark_descriptors  # device kernel; fails loudly without the CUDA library
        ids, obs, offsets = landmark_observations(doc)
        seen = np.nonzero(offsets[1:] > offsets[:-1])[0]
        if len(seen) == 0:
            return np.zeros(0, np.int64)
        _, reps = compute_landmark_descriptors(obs, offsets)
        db.add(reps[seen])
        return ids[seen]
    raise ValueError("what must be 'keypoints' or 'landmarks'")
