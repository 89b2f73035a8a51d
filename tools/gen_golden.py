#!/usr/bin/env python
"""Generates tests/golden/*.npz — outputs of the REFERENCE's own sources (oracle/_ref: feature/orb_extractor.cc,
orb_extractor_node.cc, orb_params.cc, match/stereo.cc compiled unmodified from /root/reference against the cv:: shim)
on seeded inputs. Runs only in the build container (needs /root/reference and, for the real-image case, cv2); the
fixtures it writes are committed and are what the GPU box checks against.

  python tools/gen_golden.py            # rewrites every fixture

Each fixture stores the input recipe (or the pixels, for the real-image crop), the extractor parameters, and
  kp / desc                       final keypoints (cv::KeyPoint fields) and descriptors, reference order
  pyr_crc / blur_crc              CRC32 of every pyramid level / blurred level (-1 where the reference skipped the blur)
  cand_n / cand_crc               FAST candidates handed to the quadtree per level: count and CRC32 of (x, y, score) int32
  sel_l{l}                        quadtree selection per level, (x, y, score) int32 in border-relative coordinates
  stereo_x_right / depths         match::stereo::compute outputs (stereo fixtures)
"""
import os
import sys
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from lpslam_b200.synth import noise_image, synth_image, synth_stereo_pair  # noqa: E402
from oracle import pyoracle  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
FX_B, B = 700.0 * 0.12, 0.12


def crc(a):
    return zlib.crc32(np.ascontiguousarray(a).tobytes()) & 0xFFFFFFFF


def make_input(spec, store):
    kind = spec["kind"]
    if kind == "synth":
        return synth_image(spec["h"], spec["w"], spec["seed"])
    if kind == "noise":
        return noise_image(spec["h"], spec["w"], spec["seed"])
    if kind == "flat":
        return np.full((spec["h"], spec["w"]), 77, np.uint8)
    if kind == "toy":  # test/openvslam/feature/orb_extractor.cc:24-49: white canvas, black rectangle with a corner at (300, 300)
        img = np.full((spec["h"], spec["w"]), 255, np.uint8)
        img[300:, 300:] = 0
        return img
    if kind == "pixels":
        return store["image"]
    raise ValueError(kind)


def _extractor(o, params):
    return o.extractor(max_kpts=params["max_kpts"], scale_factor=params["scale_factor"], num_levels=params["num_levels"],
                       ini_fast_thr=params["ini_fast_thr"], min_fast_thr=params["min_fast_thr"], mask_rects=params["mask_rects"])


def run_extract(o, img, params, mask=None):
    """kp / desc / pyramid / selection come from the reference build `o`. The FAST candidate list and the blurred level
    are locals of the reference's functions (orb_extractor.cc:135-139,352-369) that its build cannot expose, so those two
    come from the port — after the port has been checked here against the reference on everything they share."""
    e = _extractor(o, params)
    kp, desc = e.extract(img, mask)
    pe = _extractor(pyoracle.port(), params)
    pkp, pdesc = pe.extract(img, mask)
    assert pkp.tobytes() == kp.tobytes() and pdesc.tobytes() == desc.tobytes(), "port and reference disagree"
    L = params["num_levels"]
    out = {"kp": kp, "desc": desc, "pyr_crc": np.zeros(L, np.int64), "blur_crc": np.full(L, -1, np.int64),
           "cand_n": np.zeros(L, np.int64), "cand_crc": np.zeros(L, np.int64)}
    for l in range(L):
        out["pyr_crc"][l] = crc(e.level(l))
        assert np.array_equal(e.level(l), pe.level(l))
        c = pe.candidates(l)
        out["cand_n"][l], out["cand_crc"][l] = len(c), crc(c.astype(np.int32))
        s = e.selected(l)
        assert np.array_equal(s, pe.selected(l)), "port and reference quadtree selections disagree"
        out[f"sel_l{l}"] = s.astype(np.int32)
        if len(s):
            out["blur_crc"][l] = crc(pe.level(l, blurred=True))
    return e, out


def default_params(**kw):
    p = dict(max_kpts=1000, scale_factor=1.2, num_levels=8, ini_fast_thr=20, min_fast_thr=7, mask_rects=[])
    p.update(kw)
    return p


def pack_params(p):
    return {"p_max_kpts": p["max_kpts"], "p_scale_factor": np.float32(p["scale_factor"]), "p_num_levels": p["num_levels"],
            "p_ini_fast_thr": p["ini_fast_thr"], "p_min_fast_thr": p["min_fast_thr"],
            "p_mask_rects": np.asarray(p["mask_rects"], np.float32).reshape(-1, 4)}


def main():
    pyoracle.build()
    o = pyoracle.ref()
    if o is None:
        raise SystemExit("oracle/_ref is not built: this script needs /root/reference")
    os.makedirs(OUT, exist_ok=True)
    cases = [
        ("c1_synth_640x480", dict(kind="synth", h=480, w=640, seed=1000), default_params(), None),
        ("c2_synth_1280x720", dict(kind="synth", h=720, w=1280, seed=2000), default_params(max_kpts=2000), None),
        ("noise_320x240", dict(kind="noise", h=240, w=320, seed=3), default_params(max_kpts=500), None),
        ("portrait_240x320", dict(kind="synth", h=320, w=240, seed=11), default_params(max_kpts=400), None),
        ("odd_517x301", dict(kind="synth", h=301, w=517, seed=12), default_params(max_kpts=500), None),
        ("flat_320x240", dict(kind="flat", h=240, w=320, seed=0), default_params(max_kpts=500), None),
        ("toy_corner_600x600", dict(kind="toy", h=600, w=600, seed=0), default_params(), None),
        ("levels3_640x480", dict(kind="synth", h=480, w=640, seed=31), default_params(max_kpts=1200, num_levels=3), None),
        ("scale2_640x480", dict(kind="synth", h=480, w=640, seed=31), default_params(max_kpts=500, num_levels=4, scale_factor=2.0), None),
        ("tiny_quota_640x480", dict(kind="synth", h=480, w=640, seed=31), default_params(max_kpts=20), None),
        ("rectmask_640x480", dict(kind="synth", h=480, w=640, seed=21),
         default_params(mask_rects=[[0.2, 0.5, 0.1, 0.6], [0.7, 0.9, 0.7, 0.95]]), None),
        ("imgmask_640x480", dict(kind="synth", h=480, w=640, seed=21), default_params(), "box"),
        # BASELINE configs[2]: 4K, 4000 keypoints
        ("c3_synth_3840x2160", dict(kind="synth", h=2160, w=3840, seed=3000), default_params(max_kpts=4000), None),
        # the parameter set of the reference's own test/openvslam/feature/orb_params.cc:5-58 (3000 keypoints, scale 1.3, 12 levels,
        # thresholds 25 / 9, its two mask rectangles), on an image large enough for level 11 (1280 / 1.3^11 = 71 px)
        ("params_ref_test_1280x960", dict(kind="synth", h=960, w=1280, seed=41),
         default_params(max_kpts=3000, scale_factor=1.3, num_levels=12, ini_fast_thr=25, min_fast_thr=9,
                        mask_rects=[[0.2, 0.5, 0.3, 0.8], [0.28, 0.59, 0.1, 0.2]]), None),
    ]
    only = [a for a in sys.argv[1:] if not a.startswith("-")]
    if only:
        cases = [c for c in cases if c[0] in only]
    for name, spec, params, mask_kind in cases:
        store = {}
        img = make_input(spec, store)
        mask = None
        if mask_kind == "box":
            mask = np.full(img.shape, 255, np.uint8)
            mask[100:300, 200:500] = 0
        _, out = run_extract(o, img, params, mask)
        out.update(pack_params(params))
        out.update({"in_kind": spec["kind"], "in_h": spec["h"], "in_w": spec["w"], "in_seed": spec["seed"],
                    "in_mask": "" if mask_kind is None else mask_kind})
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
        print(f"{name}: {len(out['kp'])} keypoints, candidates {out['cand_n'].tolist()}")

    if only:
        return
    # the reference's own real test image (test/data/equirectangular_image_001.jpg, decoded as its tests do,
    # test/openvslam/feature/orb_extractor.cc:83), cropped so the fixture stays small; pixels are stored.
    try:
        import cv2
        src = "/root/reference/external/openvslam/test/data/equirectangular_image_001.jpg"
        full = cv2.imread(src, cv2.IMREAD_GRAYSCALE)
        img = np.ascontiguousarray(full[full.shape[0] // 2 - 120: full.shape[0] // 2 + 120, 400:720])
        params = default_params(max_kpts=600)
        _, out = run_extract(o, img, params)
        out.update(pack_params(params))
        out.update({"in_kind": "pixels", "in_h": img.shape[0], "in_w": img.shape[1], "in_seed": 0, "in_mask": "", "image": img})
        np.savez_compressed(os.path.join(OUT, "real_equirect_crop_320x240.npz"), **out)
        print(f"real crop: {len(out['kp'])} keypoints")
    except ImportError:
        print("cv2 missing: real-image fixture not regenerated")

    # stereo: two extractions + match::stereo::compute (data/frame.cc:55-93)
    for name, h, w, seed, n in [("stereo_640x360", 360, 640, 2100, 800), ("stereo_1280x720", 720, 1280, 2720, 2000)]:
        left, right = synth_stereo_pair(h, w, seed)
        el, er = o.extractor(max_kpts=n), o.extractor(max_kpts=n)
        kl, dl = el.extract(left)
        kr, dr = er.extract(right)
        xr, dep = o.match_stereo([el.level(l) for l in range(8)], [er.level(l) for l in range(8)], kl, kr, dl, dr,
                                 el.table(0), el.table(1), FX_B, B)
        np.savez_compressed(os.path.join(OUT, name + ".npz"), in_h=h, in_w=w, in_seed=seed, p_max_kpts=n,
                            fx_b=np.float32(FX_B), baseline=np.float32(B), kp_l_xy=np.stack([kl["x"], kl["y"]], 1),
                            kp_r_xy=np.stack([kr["x"], kr["y"]], 1), desc_l_crc=crc(dl), desc_r_crc=crc(dr),
                            stereo_x_right=xr, depths=dep)
        print(f"{name}: {len(kl)}/{len(kr)} keypoints, {(xr >= 0).sum()} stereo matches")


if __name__ == "__main__":
    main()
