"""Descriptor (de)serialisation of the reference's map files (data/common.cc:71-95, io/map_database_io.cc:23-91,
data/map_database.cc:358-380) and the kNN database filled from a map."""
import numpy as np
import pytest

from lpslam_b200 import mapio


def synthetic_map(seed, n_keyfrms=5, n_kp=60, n_lm=40):
    """a map in the reference's schema: keyframes keyed by id with n_keypts / descs / lm_ids (plus fields that only pass
    through), landmarks keyed by id; some lm_ids point at landmarks that are not in the file"""
    rng = np.random.default_rng(seed)
    protos = rng.integers(0, 256, (n_lm + 3, 32), dtype=np.uint8)
    keyfrms, truth = {}, {}
    for k in range(n_keyfrms):
        kid = 3 * k + 1
        lm_ids = np.where(rng.random(n_kp) < 0.6, rng.integers(0, n_lm + 3, n_kp), -1)
        descs = rng.integers(0, 256, (n_kp, 32), dtype=np.uint8)
        has = lm_ids >= 0
        descs[has] = protos[lm_ids[has]] ^ np.packbits(rng.random((int(has.sum()), 256)) < 0.03, axis=1, bitorder="little")
        keyfrms[str(kid)] = {"n_keypts": n_kp, "descs": mapio.convert_descriptors_to_json(descs), "lm_ids": [int(v) for v in lm_ids],
                             "ts": 0.1 * k, "cam": "synthetic", "depth_thr": 40.0}
        truth[kid] = (descs, lm_ids)
    landmarks = {str(i): {"pos_w": [0.0, 0.0, 1.0], "ref_keyfrm": 1, "1st_keyfrm": 1, "n_vis": 1, "n_fnd": 1} for i in range(n_lm)}
    return keyfrms, landmarks, truth


def test_descriptor_json_known_answers():
    # data/common.cc:71-95: a row's 32 bytes read as eight native uint32 words
    d = np.zeros((2, 32), np.uint8)
    d[0, 0], d[0, 1], d[0, 3], d[0, 4], d[0, 31] = 0x01, 0x02, 0x80, 0xFF, 0x7F
    d[1] = np.arange(32)
    j = mapio.convert_descriptors_to_json(d)
    assert j[0] == [0x80000201, 0xFF, 0, 0, 0, 0, 0, 0x7F000000]
    assert j[1][0] == 0x03020100 and j[1][7] == 0x1F1E1D1C
    assert np.array_equal(mapio.convert_json_to_descriptors(j), d)
    rng = np.random.default_rng(3)
    r = rng.integers(0, 256, (500, 32), dtype=np.uint8)
    assert np.array_equal(mapio.convert_json_to_descriptors(mapio.convert_descriptors_to_json(r)), r)
    assert mapio.convert_json_to_descriptors([]).shape == (0, 32)
    with pytest.raises(IndexError):
        mapio.convert_json_to_descriptors([[1, 2, 3]])
    with pytest.raises(ValueError):
        mapio.convert_json_to_descriptors([[1 << 32, 0, 0, 0, 0, 0, 0, 0]])


def test_message_pack_round_trip_and_association(tmp_path):
    keyfrms, landmarks, truth = synthetic_map(1)
    path = str(tmp_path / "map.msg")
    mapio.save_message_pack(path, {"synthetic": {"model_type": "perspective"}}, keyfrms, landmarks, 7, 14, 40)
    doc = mapio.load_message_pack(path)
    assert doc["frame_next_id"] == 7 and doc["keyframe_next_id"] == 14 and doc["landmark_next_id"] == 40
    assert set(doc["keyframes"]) == set(keyfrms) and doc["keyframes"]["1"]["cam"] == "synthetic"
    kfs = mapio.keyframe_descriptors(doc)
    assert [k for k, _, _ in kfs] == sorted(truth)
    for kid, descs, lm_ids in kfs:
        assert np.array_equal(descs, truth[kid][0]) and np.array_equal(lm_ids, truth[kid][1])
    ids, obs, offsets = mapio.landmark_observations(doc)
    assert list(ids) == list(range(40)) and offsets[-1] == len(obs)
    for s, lm in enumerate(ids):  # observations in keyframe-id order, one per keyframe (the first keypoint carrying the id,
        # data/landmark.cc:45-50); ids 40..42 are not in the file
        want = [truth[kid][0][i] for kid in sorted(truth) for i in np.nonzero(truth[kid][1] == lm)[0][:1]]
        got = obs[offsets[s]:offsets[s + 1]]
        assert len(got) == len(want) and all(np.array_equal(a, b) for a, b in zip(got, want))
    doc["keyframes"]["1"]["n_keypts"] += 1
    with pytest.raises(ValueError):
        mapio.keyframe_descriptors(doc)


def test_landmark_referenced_twice_from_one_keyframe_keeps_its_first_keypoint():
    """landmark::add_observation (data/landmark.cc:45-50) returns early when the keyframe already observes the landmark"""
    rng = np.random.default_rng(3)
    d0, d1 = rng.integers(0, 256, (4, 32), dtype=np.uint8), rng.integers(0, 256, (3, 32), dtype=np.uint8)
    doc = {"landmarks": {"5": {}, "9": {}},
           "keyframes": {"0": {"n_keypts": 4, "descs": mapio.convert_descriptors_to_json(d0), "lm_ids": [5, 9, 5, -1]},
                         "1": {"n_keypts": 3, "descs": mapio.convert_descriptors_to_json(d1), "lm_ids": [9, 9, 5]}}}
    ids, obs, offsets = mapio.landmark_observations(doc)
    assert list(ids) == [5, 9] and list(offsets) == [0, 2, 4]
    assert np.array_equal(obs, np.stack([d0[0], d1[2], d0[1], d1[0]]))


@pytest.mark.gpu
def test_knn_filled_from_map(oracle, tmp_path):
    from lpslam_b200 import knn
    keyfrms, landmarks, truth = synthetic_map(2, n_keyfrms=8, n_kp=300, n_lm=150)
    path = str(tmp_path / "map.msg")
    mapio.save_message_pack(path, {}, keyfrms, landmarks)
    rng = np.random.default_rng(9)
    # keypoint database: rows in keyframe-id order, tagged (keyframe, keypoint)
    db = knn.KnnDatabase(8 * 300)
    tags = mapio.fill_knn_from_map(db, path, "keypoints")
    all_rows = np.concatenate([truth[k][0] for k in sorted(truth)])
    assert len(db) == len(all_rows) == len(tags)
    q = all_rows[rng.integers(0, len(all_rows), 64)] ^ np.packbits(rng.random((64, 256)) < 0.05, axis=1, bitorder="little")
    bi, bd, sd = db.query(q)
    obi, obd, osd = oracle.knn_top2(q, all_rows)
    assert np.array_equal(bi, obi) and np.array_equal(bd, obd) and np.array_equal(sd, osd)
    assert tags[bi[0]][0] in truth
    # landmark database: one representative per observed landmark = the oracle's compute_descriptor over its observations
    ldb = knn.KnnDatabase(150)
    lm_ids = mapio.fill_knn_from_map(ldb, path, "landmarks")
    ids, obs, offsets = mapio.landmark_observations(mapio.load_message_pack(path))
    reps = []
    for s, lm in enumerate(ids):
        if offsets[s + 1] > offsets[s]:
            o = obs[offsets[s]:offsets[s + 1]]
            reps.append(o[oracle.landmark_descriptor(o)])
    reps = np.stack(reps)
    assert len(ldb) == len(reps) == len(lm_ids)
    bi, bd, sd = ldb.query(q)
    obi, obd, osd = oracle.knn_top2(q, reps)
    assert np.array_equal(bi, obi) and np.array_equal(bd, obd) and np.array_equal(sd, osd)
