// liblpslam_cuda internals shared by all translation units (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <cstring>
#include <stdint.h>
#include <stddef.h>

#include <string>
#include <vector>

#include "../../include/lpslam_cuda.h"

namespace lps {

// ---- constants fixed by the reference (feature/orb_extractor.h:128-133, orb_extractor.cc:285-286) ----
constexpr int kPatch = 31;
constexpr int kHalfPatch = 15;
constexpr int kBorder = 19;
constexpr int kCell = 64;
constexpr int kOverlap = 6;
constexpr int kCellSlots = 1024;  // a 64x64 cell holds at most 32x32 strict 8-neighbourhood maxima
constexpr int kMaxLevels = 16;
// Level images live inside a reflected apron (BORDER_REFLECT_101, written by k_pad_borders): kApron rows above and below,
// kPadLeft bytes left of pixel 0 and at least kPadLeft bytes right of the last pixel of every row. LevelGeom::img / blur
// point at pixel (0, 0); [-kPadLeft, pitch - kPadLeft) of every row from -kApron to h + kApron - 1 is readable.
constexpr int kApron = 4;
constexpr int kPadLeft = 16;

// ---- error plumbing ----
void set_error(const std::string& msg);
int cuda_fail(cudaError_t e, const char* what, const char* file, int line);

#define LPS_CUDA(call)                                                       \
    do {                                                                     \
        cudaError_t e__ = (call);                                            \
        if (e__ != cudaSuccess) return lps::cuda_fail(e__, #call, __FILE__, __LINE__); \
    } while (0)

#define LPS_ARG(cond, msg)               \
    do {                                 \
        if (!(cond)) {                   \
            lps::set_error(msg);         \
            return LPSLAM_ERR_ARG;       \
        }                                \
    } while (0)

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// per-thread, grow-only device scratch for the short entry points (scratch.cu); scratch_free must follow every scratch_alloc
cudaError_t scratch_alloc(void** out, size_t bytes);
void scratch_free(void* p);

// One short call (a matcher, frame post-processing, ...) = one copy in, kernels, one copy out, all on a stream of the calling
// thread: `dev` is a device arena and `host` a page-locked mirror of it with the same offsets. The caller packs its inputs with
// put() / zero(), sends the input prefix with upload(), launches on `st`, and fetches the output suffix with download()
// (which also waits for the stream). Both buffers and the stream belong to the host thread and only grow (scratch.cu).
struct CallStage {
    uint8_t* dev = nullptr;
    uint8_t* host = nullptr;
    cudaStream_t st = nullptr;
    cudaStream_t side = nullptr;  // second stream of the thread, for a copy that runs beside kernels already queued on st
    cudaEvent_t joined = nullptr;
    cudaError_t open(size_t bytes);
    void put(size_t off, const void* src, size_t bytes) {
        if (bytes) memcpy(host + off, src, bytes);
    }
    void zero(size_t off, size_t bytes) { memset(host + off, 0, bytes); }
    cudaError_t upload(size_t begin, size_t end) {
        return end > begin ? cudaMemcpyAsync(dev + begin, host + begin, end - begin, cudaMemcpyHostToDevice, st) : cudaSuccess;
    }
    // a second input range copied on the side stream: what is queued on st so far does not wait for it, what follows does
    cudaError_t upload_beside(size_t begin, size_t end) {
        if (end <= begin) return cudaSuccess;
        cudaError_t e = cudaMemcpyAsync(dev + begin, host + begin, end - begin, cudaMemcpyHostToDevice, side);
        if (e == cudaSuccess) e = cudaEventRecord(joined, side);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(st, joined, 0);
        return e;
    }
    cudaError_t download(size_t begin, size_t end) {
        if (end > begin) {
            const cudaError_t e = cudaMemcpyAsync(host + begin, dev + begin, end - begin, cudaMemcpyDeviceToHost, st);
            if (e != cudaSuccess) return e;
        }
        return cudaStreamSynchronize(st);
    }
    template <class T>
    const T* out(size_t off) const {
        return reinterpret_cast<const T*>(host + off);
    }
    CallStage() = default;
    CallStage(const CallStage&) = delete;
    CallStage& operator=(const CallStage&) = delete;
    ~CallStage();

private:
    bool owned_ = false;  // true when the thread's buffers were busy (nested call) and this stage allocated its own
};

// bump planner for one-shot device scratch (offsets aligned to 256 bytes)
struct ArenaBump {
    size_t off = 0;
    size_t take(size_t bytes) {
        const size_t o = off;
        off = align_up(off + (bytes ? bytes : 1), 256);
        return o;
    }
};

// One pyramid level of a batch: frame f lives at base + f * frame_stride (row pitch `pitch`).
struct LevelGeom {
    int w, h, pitch;
    size_t frame_stride;
    uint8_t* img;   // un-blurred level (level 0 of the batch is a copy of / alias to the input)
    uint8_t* blur;  // Gaussian-blurred level
    // resize tables (device), valid for level >= 1: per destination column / row
    const int* xofs;     // source column
    const short2* xcoef; // (c0, c1), 11-bit fixed point
    const int* yofs;
    const short2* ycoef;
    // schedule of the column-walk resize, indexed by SOURCE row r = 0 .. src.h (r = src.h stands for the clamped row
    // src.h - 1): x = destination row completed by the source rows (r - 1, r), or -1; y = its row coefficients b0 | b1 << 16
    const int2* srctab;
    // per 2-pixel window of the column-walk resize (window i covers destination columns 2 i - 2 kApron, + 1; apron and
    // overhang columns included): two uint4 = (source byte offset of the window's first word, PRMT selector of pixel 0,
    // of pixel 1, coefficient pair of pixel 0 | coefficient pair of pixel 1, live flag, -, -)
    const uint4* wintab;
    int walk_ok;         // the column-walk kernel applies (down-scaling by less than 3 in x, monotonic rows)
    int mode;            // 0 = general bilinear, 1 = exact 2x2 decimation (OpenCV routes it to INTER_AREA), 2 = same size copy
    // FAST cell grid of the level
    int cell_begin, cell_cols, cell_rows;  // first cell in the per-frame cell table, grid extent (only cells the reference visits)
    // quadtree
    unsigned quota;
    int nx, ny;         // initial node grid
    double dx, dy;      // initial node extents
    int depth;          // digits (2 bits each) in the path key
    int rbits;          // bits of the initial-node index in the key
    // integer-only statement of the initial-node assignment (orb_extractor.cc:532-539): node column of every border-relative
    // x (qix[x] = unsigned(float(x) / dx)), node row of every y, and the node boundaries int(dx * i), int(dy * i)
    const uint16_t* qix;
    const uint16_t* qiy;
    const int* qbx;     // nx + 1 entries
    const int* qby;     // ny + 1 entries
    // path key of a candidate = qxk[x] + qyk[y] (initial node index above 2 bits per depth, host-built)
    const uint32_t* qxk;
    const uint32_t* qyk;
    int cand_begin;     // first candidate slot of the level in the per-frame candidate scratch
    int cand_cap;       // cell_count * kCellSlots
    int node_begin, node_cap;
    int sel_begin, sel_cap;
    float scale, inv_scale, kp_size;
};

struct CellDesc {
    short level;
    short x0, y0;  // top-left of the 70x70 FAST window in level coordinates (>= 19)
    short cw, ch;  // window size after clipping to the border (<= 70)
    short j, i;    // cell column / row (candidate coordinates are shifted by 64*j, 64*i)
};

// Kernel parameter block for one extraction launch sequence (passed by value / from constant memory)
struct FrameParams {
    int num_levels;
    int batch;
    LevelGeom lv[kMaxLevels];
};

// ---- launchers (defined in the kernel translation units) ----
void launch_resize(const LevelGeom& src, const LevelGeom& dst, int batch, cudaStream_t st);
void launch_copy_level0(const uint8_t* src, size_t src_pitch, size_t src_frame_stride, const LevelGeom& dst, int batch,
                        cudaStream_t st);
// apron of the levels in `level_mask` (bit l); the column-walk resize writes the apron of its own output
void launch_pad_borders(const LevelGeom* lv_dev_table, const LevelGeom* lv_host, int num_levels, uint32_t level_mask, int batch,
                        cudaStream_t st);
bool resize_writes_apron(const LevelGeom& src, const LevelGeom& dst);
void launch_blur(const LevelGeom* lv_dev_table, const LevelGeom* lv_host, int num_levels, int batch, cudaStream_t st);
// OpenCV's INTER_LINEAR 8-bit coefficient tables of one axis (ctx.cu): source offset and (c0, c1) per destination index
void make_axis_tab(int sn, int dn, bool clamp_reset, std::vector<int>& ofs, std::vector<short2>& coef);
// TMA-staged pyramid (pyramid_tma.cu): levels 1 .. L-1 of a batch in ONE launch, level l+1 of a frame starts when level l of
// that frame is complete (device counters). nullptr when the geometry is outside what the kernel covers (the per-level
// k_resize_walk chain stays the general path).
struct PyramidPlan;
PyramidPlan* pyramid_plan_create(const LevelGeom* lv_host, int num_levels, int max_batch);
void pyramid_plan_destroy(PyramidPlan* p);
int launch_pyramid_tma(PyramidPlan* p, const LevelGeom* lv_dev_table, int batch, cudaStream_t st);
// TMA-staged blur (blur_tma.cu): tensor maps over the level images of one arena, unit tables per batch size
struct BlurPlan;
BlurPlan* blur_plan_create(const LevelGeom* lv_host, int num_levels, int max_batch);
void blur_plan_destroy(BlurPlan* p);
int launch_blur_tma(BlurPlan* p, const LevelGeom* lv_dev_table, int batch, cudaStream_t st);

struct FastArgs {
    const LevelGeom* lv;      // device table
    const CellDesc* cells;    // device table
    int num_cells;            // per frame
    int batch;
    int frame0;               // first frame of this launch (sub-batches of a step run on their own streams)
    uint32_t ini_thr, min_thr;
    uint32_t* cell_counts;    // [batch][num_cells]
    uint32_t* cell_slots;     // [batch][num_cells][kCellSlots] packed lx | ly<<8 | score<<16
    const uint8_t* mask;      // level-0 mask or nullptr (shared by the batch)
    size_t mask_pitch;
};
void launch_fast(const FastArgs& a, cudaStream_t st);

struct TreeArgs {
    const LevelGeom* lv;  // device table
    int num_levels, batch, num_cells;
    int frame0;  // first frame of this launch
    const CellDesc* cells;
    const uint32_t* cell_counts;
    const uint32_t* cell_slots;
    // per-frame scratch (strides in elements)
    uint32_t* cand_xy;  size_t cand_stride;   // [batch][cand_total]  x | y<<16 (border-relative)
    uint8_t* cand_sc;                          // [batch][cand_total]
    uint32_t* keys[2];                         // sort ping-pong, same stride
    uint32_t* vals[2];
    uint32_t* cell_base; size_t cell_stride;   // [batch][num_cells]
    int* node[2][3];     size_t node_stride;   // lo, hi, depth|flags — ping-pong
    int* tmp[8];                               // per-node temporaries, node_stride
    uint32_t* cand_count;                      // [batch][levels] candidates per level
    uint32_t* sel_xys;   size_t sel_stride;    // [batch][sel_total] selected: x | y<<16, score separately
    uint8_t* sel_sc;
    uint32_t* sel_count;                       // [batch][levels]
};
void launch_quadtree(const TreeArgs& a, const LevelGeom* lv_host, cudaStream_t st);

struct DescribeArgs {
    const LevelGeom* lv;
    int num_levels, batch;
    const uint32_t* sel_xys; size_t sel_stride;
    const uint8_t* sel_sc;
    const uint32_t* sel_count;   // [batch][levels]
    uint32_t* level_off;         // [batch*levels + 1] packed offsets (written by the offsets kernel)
    uint32_t* frame_off;         // [batch + 1]
    lpslam_keypoint* kp;         // packed output
    uint8_t* desc;
    uint32_t out_cap;
    uint32_t* overflow;          // set to needed count when out_cap is too small
};
void launch_offsets(const DescribeArgs& a, cudaStream_t st);

// match::stereo on device-resident extraction results: frame 2p = left eye, frame 2p + 1 = right eye of pair p
struct StereoArgs {
    const LevelGeom* lv;
    int num_levels, n_pairs;
    const lpslam_keypoint* kp;   // packed keypoints of all 2 * n_pairs frames
    const uint8_t* desc;
    const uint32_t* frame_off;   // [2 * n_pairs + 1]
    uint32_t max_rows;           // rows in the packed arrays that may be touched
    uint2* rec;                  // per packed row: (min_r | max_r << 16, octave) of the keypoint as a RIGHT candidate
    float* x_right;              // per packed row (left-eye rows are meaningful), -1 = no match
    float* depth;
    int* corr;                   // best SAD as int, -1 = no match
    uint4* perm;                 // right-eye keypoints of a pair in row order: (row interval, octave, x bits, packed row)
    uint32_t* row_start;         // [n_pairs][rows0 + 1] first perm slot (relative to the right eye's first row) of every image row
    float focal_x_baseline, true_baseline;
    int rows0;                   // level-0 image rows
};
void launch_stereo(const StereoArgs& a, cudaStream_t st);
void launch_describe(const DescribeArgs& a, cudaStream_t st);

}  // namespace lps
