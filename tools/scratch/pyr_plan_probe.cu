// host-only probe: does the TMA pyramid accept a geometry? (prints the rejecting check with LPS_PYR_DEBUG=1)
#include "../../lpslam_b200/csrc/common.cuh"
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
using namespace lps;
int main(int argc, char** argv) {
    const int W = argc > 1 ? atoi(argv[1]) : 1280, H = argc > 2 ? atoi(argv[2]) : 720, L = 8;
    LevelGeom lv[kMaxLevels];
    memset(lv, 0, sizeof lv);
    float sc = 1.f;
    for (int l = 0; l < L; ++l) {
        lv[l].w = l ? (int)std::round(W * 1.0 / (double)sc) : W;
        lv[l].h = l ? (int)std::round(H * 1.0 / (double)sc) : H;
        lv[l].pitch = (int)align_up(lv[l].w + 32, 64);
        lv[l].frame_stride = align_up((size_t)lv[l].pitch * (lv[l].h + 8), 256);
        lv[l].img = (uint8_t*)0x10000000 + 4 * lv[l].pitch + 16;
        sc *= 1.2f;
    }
    setenv("LPS_PYR_DEBUG", "1", 1);
    PyramidPlan* p = pyramid_plan_create(lv, L, 64);
    printf("%dx%d plan %s\n", W, H, p ? "accepted" : "rejected (no message above = CUDA side)");
    return 0;
}
