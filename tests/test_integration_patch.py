"""CPU: cpp/patches/lpslam_use_cuda_orb.patch — the four hunks a maintainer applies to the reference tree (INTEGRATION.md section 2:
the `use_cuda` read site in tracking_module.cc:77-84 that lpslam's OpenVSLAMTrackerBase.cpp:209 already feeds, the source swap in
feature/ and match/ CMakeLists.txt, the option in the top-level CMakeLists.txt) — applies cleanly to a copy of the reference files.
Runs where /root/reference exists (this container); the GPU box has no reference tree."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PATCH = os.path.join(ROOT, "cpp", "patches", "lpslam_use_cuda_orb.patch")
REF = "/root/reference"
FILES = ["CMakeLists.txt", "external/openvslam/src/openvslam/tracking_module.cc",
         "external/openvslam/src/openvslam/feature/CMakeLists.txt", "external/openvslam/src/openvslam/match/CMakeLists.txt"]


def test_patch_applies_to_the_reference(tmp_path):
    if not os.path.isdir(REF) or shutil.which("patch") is None:
        pytest.skip("needs /root/reference and patch(1)")
    touched = [l[6:].strip() for l in open(PATCH) if l.startswith("+++ b/")]
    assert sorted(touched) == sorted(FILES)
    for f in FILES:
        os.makedirs(os.path.dirname(tmp_path / f), exist_ok=True)
        shutil.copy(os.path.join(REF, f), tmp_path / f)
    for extra in ([], ["--dry-run"])[::-1]:
        r = subprocess.run(["patch", "-p1", "--fuzz=0", "-i", PATCH] + extra, cwd=tmp_path, capture_output=True, text=True)
        assert r.returncode == 0, r.stdout + r.stderr
    tm = open(tmp_path / FILES[1]).read()
    read_site = tm.index('cfg->yaml_node_["use_cuda"].as<bool>(false)')
    assert tm.index("get_orb_params(cfg->yaml_node_[\"Feature\"])") < read_site < tm.index("extractor_left_ = new feature::orb_extractor(orb_params);")
    assert "#include <stdexcept>" in tm and "std::runtime_error(\"use_cuda is set" in tm  # a CPU build refuses use_cuda: no silent fallback
    feat = open(tmp_path / FILES[2]).read()
    head, _, tail = feat.partition("if(USE_CUDA_ORB)")
    assert "orb_extractor.cc" not in head and "orb_params.cc" in head  # orb_params.cc stays the reference's in both builds
    assert "lpslam_cuda_dropin" in tail and "orb_extractor_node.cc" in tail.split("else()")[1]
    mat = open(tmp_path / FILES[3]).read()
    assert "stereo.cc" not in mat.split("if(NOT USE_CUDA_ORB)")[0] and "robust.cc" in mat.split("if(NOT USE_CUDA_ORB)")[0]
    top = open(tmp_path / FILES[0]).read()
    assert top.index("add_subdirectory(external/lpslam_cuda)") < top.index("add_subdirectory(external/openvslam)")
    # the names the hunks rely on exist in this repository's own CMakeLists.txt
    ours = open(os.path.join(ROOT, "CMakeLists.txt")).read()
    for name in ("lpslam_cuda_dropin", "LPSLAM_BUILD_DROPIN", "OPENVSLAM_SRC", "project(lpslam_cuda"):
        assert name in ours, name
