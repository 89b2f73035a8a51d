"""Loader for tests/golden/*.npz (written by tools/gen_golden.py / tools/gen_golden_match.py from the reference's own sources,
oracle/_ref). Provenance of the fields: keypoints, descriptors, pyramid levels, quadtree selections, stereo and matcher outputs come
from the reference build; `cand_crc` (FAST candidates before the quadtree) and `blur_crc` (blurred levels) are locals of the
reference's extract() that its interface does not expose (oracle/ref_glue.cc: ref_level_candidates / blurred ref_level_copy return
an error), so those two come from the PORT, whose FAST and Gaussian primitives are pinned bit-for-bit against cv2 and whose final
keypoints / descriptors, which depend on both, are asserted equal to the reference build on every fixture."""
import glob
import os
import zlib

import numpy as np

from lpslam_b200.synth import noise_image, synth_image, synth_stereo_pair

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
EXTRACT_FIXTURES = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "*.npz"))
                          if not os.path.basename(p).startswith(("stereo_", "match_")))
MATCH_FIXTURES = sorted(os.path.basename(p)[6:-4] for p in glob.glob(os.path.join(GOLDEN, "match_*.npz")))
STEREO_FIXTURES = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "stereo_*.npz")))


def crc(a):
    return zlib.crc32(np.ascontiguousarray(a).tobytes()) & 0xFFFFFFFF


def load(name):
    return dict(np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False))


def case_input(g):
    """(image, mask or None, extractor kwargs) of an extraction fixture"""
    kind, h, w, seed = str(g["in_kind"]), int(g["in_h"]), int(g["in_w"]), int(g["in_seed"])
    if kind == "synth":
        img = synth_image(h, w, seed)
    elif kind == "noise":
        img = noise_image(h, w, seed)
    elif kind == "flat":
        img = np.full((h, w), 77, np.uint8)
    elif kind == "toy":
        img = np.full((h, w), 255, np.uint8)
        img[300:, 300:] = 0
    elif kind == "pixels":
        img = g["image"]
    else:
        raise ValueError(kind)
    mask = None
    if str(g["in_mask"]) == "box":
        mask = np.full(img.shape, 255, np.uint8)
        mask[100:300, 200:500] = 0
    params = dict(max_kpts=int(g["p_max_kpts"]), scale_factor=float(g["p_scale_factor"]), num_levels=int(g["p_num_levels"]),
                  ini_fast_thr=int(g["p_ini_fast_thr"]), min_fast_thr=int(g["p_min_fast_thr"]),
                  mask_rects=[list(map(float, r)) for r in g["p_mask_rects"]])
    return img, mask, params


def stereo_input(g):
    return synth_stereo_pair(int(g["in_h"]), int(g["in_w"]), int(g["in_seed"])), int(g["p_max_kpts"])
