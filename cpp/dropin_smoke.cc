// Test harness for the C++ drop-in layer (cpp/openvslam/...): reads a raw stereo pair, runs the drop-in
// openvslam::feature::orb_extractor twice and the drop-in match::stereo::compute, and dumps the results for
// tests/test_dropin_gpu.py. Built by cpp/Makefile against the oracle's minimal cv:: shim because the OpenCV SDK is not
// installed in this image; with a real OpenCV the same sources compile unchanged.
//   usage: dropin_smoke <w> <h> <max_kpts> <left.raw> <right.raw> <out.bin> <fx_b> <baseline>
//          dropin_smoke <w> <h> <max_kpts> <left.raw> <right.raw> latency:<reps> <fx_b> <baseline>
//              -> one JSON line: median / p99 wall time of what data::frame's stereo constructor does per frame
//                 (frame.cc:55-93: two extract calls on two threads, then match::stereo::compute)
#include <algorithm>
#include <chrono>
#include <cstring>
#include <thread>
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
    if (std::strncmp(argv[6], "latency:", 8) == 0) {
        const int reps = std::atoi(argv[6] + 8);
        const std::vector<float> sf = ext_l->get_scale_factors(), isf = ext_l->get_inv_scale_factors();
        std::vector<double> ms, ms_extract;
        for (int i = 0; i < reps + 10; ++i) {
            const auto t0 = std::chrono::steady_clock::now();
            std::thread tl([&] { ext_l->extract(left, cv::Mat(), kl, dl); });   // frame.cc:60-63
            std::thread tr([&] { ext_r->extract(right, cv::Mat(), kr, dr); });
            tl.join();
            tr.join();
            const auto t1 = std::chrono::steady_clock::now();
            std::vector<float> xr_, depth_;
            openvslam::match::stereo m(ext_l->image_pyramid_, ext_r->image_pyramid_, kl, kr, dl, dr, sf, isf, fx_b, baseline);
            m.compute(xr_, depth_);
            if (i >= 10) {
                ms.push_back(std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
                ms_extract.push_back(std::chrono::duration<double, std::milli>(t1 - t0).count());
            }
        }
        std::sort(ms.begin(), ms.end());
        std::sort(ms_extract.begin(), ms_extract.end());
        std::printf("{\"call\": \"drop-in orb_extractor::extract x 2 (two threads) + match::stereo::compute, image_pyramid_ filled\", "
                    "\"calls\": %d, \"median_ms\": %.4f, \"p99_ms\": %.4f, \"extract_median_ms\": %.4f, \"keypoints_left\": %zu}\n",
                    reps, ms[ms.size() / 2], ms[static_cast<size_t>(0.99 * (ms.size() - 1))], ms_extract[ms_extract.size() / 2], kl.size());
        delete ext_l;
        delete ext_r;
        return 0;
    }
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
    // a copy of the extractor (the reference's class is copyable) gives the same keypoints; its pyramid levels 1.. follow
    {
        orb_extractor copy(*ext_l);
        std::vector<cv::KeyPoint> kc;
        cv::Mat dc;
        copy.extract(left, cv::Mat(), kc, dc);
        int32_t same = kc.size() == kl.size();
        for (size_t i = 0; same && i < kc.size(); ++i)
            same = kc[i].pt.x == kl[i].pt.x && kc[i].pt.y == kl[i].pt.y && kc[i].octave == kl[i].octave
                   && std::memcmp(dc.ptr(static_cast<int>(i)), dl.ptr(static_cast<int>(i)), 32) == 0;
        std::fwrite(&same, 4, 1, f);
        for (size_t level = 1; level < copy.image_pyramid_.size(); ++level) {
            const cv::Mat& m = copy.image_pyramid_.at(level);
            const int32_t mw = m.cols, mh = m.rows;
            std::fwrite(&mw, 4, 1, f);
            std::fwrite(&mh, 4, 1, f);
            for (int y = 0; y < mh; ++y) std::fwrite(m.ptr(y), 1, static_cast<size_t>(mw), f);
        }
        // an empty input image returns at once and leaves the caller's keypoints and descriptors as they were (orb_extractor.cc:73-75)
        const size_t n_before = kc.size();
        const int rows_before = dc.rows;
        const uint8_t* data_before = dc.data;
        copy.extract(cv::Mat(), cv::Mat(), kc, dc);
        const int32_t untouched = kc.size() == n_before && n_before > 0 && dc.rows == rows_before && dc.data == data_before;
        std::fwrite(&untouched, 4, 1, f);
    }
    std::fclose(f);
    delete ext_l;
    delete ext_r;
    std::printf("%u %u\n", nl, nr);
    return 0;
}
