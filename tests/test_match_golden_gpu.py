"""Matcher parity against committed outputs of the REFERENCE's own sources (tests/golden/match_*.npz, written by
tools/gen_golden_match.py from oracle/_ref/libmatch_ref.so): C1 / C2 / C3-sized brute force with and without the orientation check,
clustered descriptors, frame-to-frame projection (mono, stereo, temporaries), local-map search, initializer area matcher.
The CPU test holds the port against the same fixtures, so the fixtures stay meaningful where /root/reference does not exist."""
import numpy as np
import pytest

import golden_util as gu
import match_cases


class CudaMatcher:
    """lpslam_b200.match behind the method names the case runners use"""

    def match_bruteforce(self, d1, d2, valid, a1, a2, ratio, chk):
        from lpslam_b200 import match
        pairs, out = match.robust(ratio, chk).brute_force_match(d1, d2, valid, a1, a2)
        return out, len(pairs)

    def match_projection(self, cam, Tc, Tl, has, outl, pos, ld, loct, lang, kp, cd, xr, tk, sf, margin, chk, obs):
        from lpslam_b200 import match
        return match.projection(0.6, chk).match_current_and_last_frames(self._cam(cam), Tc, Tl, has, outl, pos, ld, loct, lang, kp, cd, xr,
                                                                        tk, sf, margin, obs)

    def match_landmarks(self, cam, act, rp, lxr, lvl, ld, kp, cd, xr, tk, sf, margin, ratio, obs):
        from lpslam_b200 import match
        return match.projection(ratio, True).match_frame_and_landmarks(self._cam(cam), act, rp, lxr, lvl, ld, kp, cd, xr, tk, sf, margin, obs)

    def match_area(self, cam, k1, d1, prev, k2, d2, margin, ratio, chk):
        from lpslam_b200 import match
        return match.area(ratio, chk).match_in_consistent_area(self._cam(cam), k1, d1, k2, d2, prev, margin)

    @staticmethod
    def _cam(c):
        from lpslam_b200 import _lib
        cam = _lib.Camera()
        for f, _ in cam._fields_:
            setattr(cam, f, getattr(c, f))
        return cam


def _check(name, got):
    want = gu.load("match_" + name)
    assert set(want) == set(got)
    for k in want:
        assert np.array_equal(np.asarray(got[k]), want[k]), (name, k)


@pytest.mark.parametrize("name", gu.MATCH_FIXTURES)
def test_port_equals_reference_golden(oracle, name):
    _check(name, match_cases.CASES[name](oracle._port, oracle._port))


@pytest.mark.gpu
@pytest.mark.parametrize("name", gu.MATCH_FIXTURES)
def test_cuda_equals_reference_golden(oracle, name):
    _check(name, match_cases.CASES[name](CudaMatcher(), oracle._port))
