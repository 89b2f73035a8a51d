"""GPU parity of the stereo row matcher (match/stereo.cc:18-249) against the oracle: bit-exact x_right / depth."""
import os

import numpy as np
import pytest

from lpslam_b200.synth import synth_stereo_pair

pytestmark = pytest.mark.gpu

FX_B, B = 700.0 * 0.12, 0.12


def oracle_pair(oracle, left, right, n):
    ol, orr = oracle.extractor(max_kpts=n), oracle.extractor(max_kpts=n)
    kl, dl = ol.extract(left)
    kr, dr = orr.extract(right)
    pl = [ol.level(l) for l in range(8)]
    pr = [orr.level(l) for l in range(8)]
    xr, dep = oracle.match_stereo(pl, pr, kl, kr, dl, dr, ol.table(0), ol.table(1), FX_B, B)
    return kl, dl, kr, dr, pl, pr, ol.table(0), ol.table(1), xr, dep


@pytest.mark.parametrize("h,w,n", [(480, 640, 1000), (720, 1280, 2000)])
def test_standalone_matcher(oracle, h, w, n):
    from lpslam_b200 import match
    left, right = synth_stereo_pair(h, w, 2000 + h)
    kl, dl, kr, dr, pl, pr, sf, isf, xr, dep = oracle_pair(oracle, left, right, n)
    gxr, gdep = match.stereo(pl, pr, kl, kr, dl, dr, sf, isf, FX_B, B).compute()
    assert (xr >= 0).sum() > n // 10, "synthetic pair should produce stereo matches"
    assert np.array_equal(gxr, xr, equal_nan=True)
    assert np.array_equal(gdep, dep, equal_nan=True)


def test_fused_frames(oracle):
    from lpslam_b200 import match, orb
    pairs = [synth_stereo_pair(360, 640, 2100 + k) for k in range(3)]
    e = orb.orb_extractor(max_num_keypts=800, max_batch=6)
    res = match.stereo_frames(e, [p[0] for p in pairs], [p[1] for p in pairs], FX_B, B)
    for (left, right), r in zip(pairs, res):
        kl, dl, kr, dr, pl, pr, sf, isf, xr, dep = oracle_pair(oracle, left, right, 800)
        assert np.array_equal(r["keypts_left"]["x"], kl["x"]) and np.array_equal(r["keypts_right"]["x"], kr["x"])
        assert (r["descs_left"] == dl).all() and (r["descs_right"] == dr).all()
        assert np.array_equal(r["stereo_x_right"], xr, equal_nan=True)
        assert np.array_equal(r["depths"], dep, equal_nan=True)


def test_matcher_on_contexts_waits_for_enqueued_extraction(oracle):
    """lpslam_match_stereo_ctx reads the pyramids two contexts hold. Behind a submit (which returns before the device has rebuilt
    them) it must see the new pyramids, not those of the frame before: the matcher orders itself behind both contexts' streams."""
    import ctypes as C
    from lpslam_b200 import match, orb, replay  # noqa: F401  (replay registers submit / collect)
    from lpslam_b200._lib import KP_DTYPE, check, lib
    h, w, n = 480, 640, 1000
    left, right = synth_stereo_pair(h, w, 2480)
    other_l, other_r = synth_stereo_pair(h, w, 77)
    kl, dl, kr, dr, pl, pr, sf, isf, xr, dep = oracle_pair(oracle, left, right, n)
    el, er = orb.orb_extractor(max_num_keypts=n), orb.orb_extractor(max_num_keypts=n)
    # blocking extraction, then the matcher on what the contexts hold
    gkl, gdl = el.extract(left)
    gkr, gdr = er.extract(right)
    assert np.array_equal(gkl["x"], kl["x"]) and (gdl == dl).all() and (gdr == dr).all()
    gxr, gdep = match.stereo_on_contexts(el, er, gkl, gkr, gdl, gdr, FX_B, B)
    assert np.array_equal(gxr, xr, equal_nan=True) and np.array_equal(gdep, dep, equal_nan=True)
    for _ in range(3):
        # the contexts hold another pair ...
        el.extract(other_l)
        er.extract(other_r)
        # ... and get the first pair back through submit, with the matcher called right behind it
        tickets = []
        for e, img in ((el, left), (er, right)):
            t = C.c_uint64(0)
            ptrs = (C.c_void_p * 1)(img.ctypes.data)
            check(lib().lpslam_orb_extract_batch_submit(e._h, ptrs, 1, w, h, w, C.byref(t)))
            tickets.append(t)
        gxr, gdep = match.stereo_on_contexts(el, er, gkl, gkr, gdl, gdr, FX_B, B)
        assert np.array_equal(gxr, xr, equal_nan=True) and np.array_equal(gdep, dep, equal_nan=True)
        for e, t, want in ((el, tickets[0], gkl), (er, tickets[1], gkr)):
            kp, desc, offs = np.empty(8 * n + 256, KP_DTYPE), np.empty((8 * n + 256, 32), np.uint8), np.zeros(2, np.uint32)
            check(lib().lpslam_orb_extract_batch_collect(e._h, t, kp.ctypes.data, desc.ctypes.data, len(kp), offs.ctypes.data))
            assert offs[1] == len(want) and np.array_equal(kp[:offs[1]]["x"], want["x"])


def test_empty_right(oracle):
    from lpslam_b200 import match
    left, right = synth_stereo_pair(240, 320, 5)
    o = oracle.extractor(max_kpts=300)
    kl, dl = o.extract(left)
    pl = [o.level(l) for l in range(8)]
    xr, dep = match.stereo(pl, pl, kl, kl[:0], dl, dl[:0], o.table(0), o.table(1), FX_B, B).compute()
    assert (xr == -1).all() and (dep == -1).all()


def test_stereo_frames_submit_collect():
    """lpslam_stereo_frames_submit / _collect: two batches in flight on one context give the results of the blocking call;
    a third submit before a collect is refused, tickets are collected in order."""
    from lpslam_b200 import _lib, match, orb
    from lpslam_b200.synth import synth_stereo_pair
    e = orb.orb_extractor(max_num_keypts=600, max_batch=4)
    batches = [[synth_stereo_pair(240, 320, 300 + 2 * b + k) for k in range(2)] for b in range(4)]
    want = [match.stereo_frames(e, [p[0] for p in bt], [p[1] for p in bt], 84.0, 0.12) for bt in batches]
    tickets = []
    got = []
    for b, bt in enumerate(batches):
        tickets.append(match.stereo_frames_submit(e, [p[0] for p in bt], [p[1] for p in bt], 84.0, 0.12))
        if len(tickets) == 2:
            with pytest.raises(_lib.LpslamCudaError):  # both slots taken
                match.stereo_frames_submit(e, [p[0] for p in bt], [p[1] for p in bt], 84.0, 0.12)
            t, keep = tickets.pop(0)
            got.append(match.stereo_frames_collect(e, t, keep))
    while tickets:
        t, keep = tickets.pop(0)
        got.append(match.stereo_frames_collect(e, t, keep))
    assert len(got) == len(want)
    for g, w in zip(got, want):
        for a, b in zip(g, w):
            for k in a:
                assert np.array_equal(a[k], b[k], equal_nan=True) if a[k].dtype.kind == "f" else a[k].tobytes() == b[k].tobytes(), k
    # the blocking entry point still works on the same context afterwards
    again = match.stereo_frames(e, [p[0] for p in batches[0]], [p[1] for p in batches[0]], 84.0, 0.12)
    assert again[0]["keypts_left"].tobytes() == want[0][0]["keypts_left"].tobytes()


def test_single_pair_calls_replay_a_graph():
    """lpslam_stereo_frames with one pair per call (data/frame.cc:55-93's pattern): the first call of a setting runs launch by launch,
    the second is captured into a CUDA graph, later ones replay it. Every call must give what a fresh context gives for the same
    pair; changing a parameter or the stereo constants in between drops the graph."""
    from lpslam_b200 import match, orb
    pairs = [synth_stereo_pair(360, 640, 7000 + k) for k in range(6)]
    e = orb.orb_extractor(max_num_keypts=700, max_batch=2)

    def fresh(pair, n, fxb):
        f = orb.orb_extractor(max_num_keypts=n, max_batch=2)
        return match.stereo_frames(f, [pair[0]], [pair[1]], fxb, B)[0]

    def same(a, b):
        return all(a[k].tobytes() == b[k].tobytes() for k in a)

    def graph_launches(x):
        return int(x._get(100))  # LPSLAM_ORB_DEBUG_GRAPH_LAUNCHES

    graphs_on = os.environ.get("LPSLAM_CUDA_GRAPH", "1")[:1] != "0"
    for k, pair in enumerate(pairs):  # calls 0, 1: plain / capture, 2..5: graph replays on new images
        got = match.stereo_frames(e, [pair[0]], [pair[1]], FX_B, B)[0]
        assert same(got, fresh(pair, 700, FX_B)), k
        # the capture call launches the new graph once, every later call replays it: a silent fall-back to plain launches fails here
        assert graph_launches(e) == (k if graphs_on else 0), k
    e.set_max_num_keypoints(500)  # kernel arguments change: the graph is dropped and rebuilt
    for k, pair in enumerate(pairs[:4]):
        assert same(match.stereo_frames(e, [pair[0]], [pair[1]], FX_B, B)[0], fresh(pair, 500, FX_B)), k
    for k, pair in enumerate(pairs[:4]):  # other stereo constants
        assert same(match.stereo_frames(e, [pair[0]], [pair[1]], 2 * FX_B, B)[0], fresh(pair, 500, 2 * FX_B)), k
    # a two-pair call in between runs without the graph (and drops it: the next one-pair calls build a new one)
    two = orb.orb_extractor(max_num_keypts=500, max_batch=4)
    want = match.stereo_frames(two, [pairs[0][0], pairs[1][0]], [pairs[0][1], pairs[1][1]], 2 * FX_B, B)
    e4 = orb.orb_extractor(max_num_keypts=500, max_batch=4)
    for _ in range(3):
        match.stereo_frames(e4, [pairs[2][0]], [pairs[2][1]], 2 * FX_B, B)
    before = graph_launches(e4)
    assert before == (2 if graphs_on else 0)
    got = match.stereo_frames(e4, [pairs[0][0], pairs[1][0]], [pairs[0][1], pairs[1][1]], 2 * FX_B, B)
    assert all(same(a, b) for a, b in zip(got, want))
    assert graph_launches(e4) == before
    assert same(match.stereo_frames(e4, [pairs[2][0]], [pairs[2][1]], 2 * FX_B, B)[0], fresh(pairs[2], 500, 2 * FX_B))
