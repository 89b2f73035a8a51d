// Test harness for the C++ drop-in layer (cpp/openvslam/...): reads a raw stereo pair, runs the drop-in
// openvslam::feature::orb_extractor twice and the drop-in match::stereo::compute, and dumps the results for
// tests/test_dropin_gpu.py. Built by cpp/Makefile against the oracle's minimal cv:: shim because the OpenCV SDK is not
// installed in this image; with a real OpenCV the same sources compile unchanged.
//   usage: dropin_smoke <w> <h> <max_kpts> <left.raw> <right.raw> <out.bin> <fx_b> <baseline>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include <opencv2/core.hpp>

#include "openvslam/feature/orb_extractor.h"
#include "openvslam/match/stereo.h"

static std::vector<uint8_t> slurp(const char* path, size_t n) {
    std::vector<uint8_t> v(n);
    FILE* f = std::fopen(path, "rb");
    if (!f || std::fread(v.data(), 1, n, f) != n) {
        std::fprintf(stderr, "cannot read %s\n", path);
        std::exit(2);
    }
    std::fclose(f);
    return v;
}

int main(int argc, char** argv) {
    if (argc != 9) return 2;
    const int w = std::atoi(argv[1]), h = std::atoi(argv[2]);
    const unsigned max_kpts = static_cast<unsigned>(std::atoi(argv[3]));
    std::vector<uint8_t> lbuf = slurp(argv[4], static_cast<size_t>(w) * h), rbuf = slurp(argv[5], static_cast<size_t>(w) * h);
    const float fx_b = static_cast<float>(std::atof(argv[7])), baseline = static_cast<float>(std::atof(argv[8]));
    cv::Mat left(h, w, CV_8UC1, lbuf.data(), w), right(h, w, CV_8UC1, rbuf.data(), w);

    using openvslam::feature::orb_extractor;
    orb_extractor* ext_l = new orb_extractor(max_kpts, 2 * max_kpts, 1.2f, 8, 20, 7);  // tracking_module.cc:78-84 style ownership
    orb_extractor* ext_r = new orb_extractor(max_kpts, 2 * max_kpts, 1.2f, 8, 20, 7);
    std::vector<cv::KeyPoint> kl, kr;
    cv::Mat dl, dr;
    ext_l->extract(left, cv::Mat(), kl, dl);
    ext_r->extract(right, cv::Mat(), kr, dr);
    std::vector<float> xr, depth;
    // match::stereo keeps references (stereo.h:68-88): the tables must outlive it, as data::frame's members do (frame.cc:78)
    const std::vector<float> scale_factors = ext_l->get_scale_factors(), inv_scale_factors = ext_l->get_inv_scale_factors();
    openvslam::match::stereo matcher(ext_l->image_pyramid_, ext_r->image_pyramid_, kl, kr, dl, dr, scale_factors, inv_scale_factors,
                                     fx_b, baseline);
    matcher.compute(xr, depth);

    FILE* f = std::fopen(argv[6], "wb");
    if (!f) return 3;
    const uint32_t nl = static_cast<uint32_t>(kl.size()), nr = static_cast<uint32_t>(kr.size());
    std::fwrite(&nl, 4, 1, f);
    std::fwrite(&nr, 4, 1, f);
    std::fwrite(kl.data(), sizeof(cv::KeyPoint), nl, f);
    for (uint32_t i = 0; i < nl; ++i) std::fwrite(dl.ptr(static_cast<int>(i)), 1, 32, f);
    std::fwrite(kr.data(), sizeof(cv::KeyPoint), nr, f);
    for (uint32_t i = 0; i < nr; ++i) std::fwrite(dr.ptr(static_cast<int>(i)), 1, 32, f);
    std::fwrite(xr.data(), 4, nl, f);
    std::fwrite(depth.data(), 4, nl, f);
    // pyramid level 3 of the left eye, to check the image_pyramid_ side effect
    const cv::Mat& p3 = ext_l->image_pyramid_.at(3);
    const int32_t pw = p3.cols, ph = p3.rows;
    std::fwrite(&pw, 4, 1, f);
    std::fwrite(&ph, 4, 1, f);
    for (int y = 0; y < ph; ++y) std::fwrite(p3.ptr(y), 1, static_cast<size_t>(pw), f);
    std::fclose(f);
    delete ext_l;
    delete ext_r;
    std::printf("%u %u\n", nl, nr);
    return 0;
}
