// rct_api.cu -- C ABI (include/rct.h) over the kernels in rct_kernels.cuh.
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -fmad=false
//        -Xcompiler -fPIC -shared   (see csrc/Makefile, __graft_entry__.build()).
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <atomic>
#include <condition_variable>
#include <mutex>
#include <thread>
#include <new>
#include <string>
#include <vector>
#if defined(__SSE2__)
#include <emmintrin.h>
#endif

#include "../../include/rct.h"
#include "rct_kernels.cuh"
#include "rct_jpeg.cuh"
#include "rct_jpeg_dec.cuh"
#include "rct_png.h"

using namespace rct;

namespace {

thread_local std::string t_err;
thread_local long long t_launches = 0;

int fail(int code, const char *fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    t_err = buf;
    return code;
}

#define CK(call)                                                                                       \
    do {                                                                                               \
        cudaError_t e_ = (call);                                                                       \
        if (e_ != cudaSuccess)                                                                         \
            return fail(RCT_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, \
                        __LINE__);                                                                     \
    } while (0)

#define LAUNCH(kernel, grid, block, stream, ...)                                     \
    do {                                                                             \
        kernel<<<grid, block, 0, stream>>>(__VA_ARGS__);                             \
        ++t_launches;                                                                \
        cudaError_t e_ = cudaGetLastError();                                         \
        if (e_ != cudaSuccess)                                                       \
            return fail(RCT_ERR_CUDA, "launch %s failed: %s", #kernel, cudaGetErrorString(e_)); \
    } while (0)

// ---- matrices / tables ---------------------------------------------------------------------
std::mutex g_mu;
double g_mat[36] = {
    // RGB_TO_LMS (LST:12-14)
    0x1.863f141205bc0p-2, 0x1.2816f0068db8cp-1, 0x1.495182a9930bep-5, 0x1.92d77318fc505p-3, 0x1.72e48e8a71de7p-1,
    0x1.404ea4a8c154dp-4, 0x1.8adab9f559b3dp-6, 0x1.07c84b5dcc63fp-3, 0x1.b05532617c1bep-1,
    // LMS_TO_LAB (LST:15-17)
    0x1.279a74590331dp-1, 0x1.279a74590331dp-1, 0x1.279a74590331dp-1, 0x1.a20bd700c2c3fp-2, 0x1.a20bd700c2c3fp-2,
    -0x1.a20bd700c2c3fp-1, 0x1.6a09e667f3bccp-1, -0x1.6a09e667f3bccp-1, 0x0.0p+0,
    // LAB_TO_LMS = inv(LMS_TO_LAB) (LST:18), as numpy 2.3.5 / OpenBLAS 0.3.30 computes it
    0x1.279a74590331cp-1, 0x1.a20bd700c2c3dp-2, 0x1.6a09e667f3bcdp-1, 0x1.279a74590331cp-1, 0x1.a20bd700c2c3dp-2,
    -0x1.6a09e667f3bcdp-1, 0x1.279a74590331ap-1, -0x1.a20bd700c2c3dp-1, 0x1.06f92bd2a7fb7p-53,
    // LMS_TO_RGB = inv(RGB_TO_LMS) (LST:19)
    0x1.1dfeafe21594bp+2, -0x1.cb59bb64ad92ap+1, 0x1.e9e644ae9f486p-4, -0x1.383f59520175cp+0, 0x1.31090654d17b8p+1,
    -0x1.4d11041a8966dp-3, 0x1.df4d2b5faca4cp-5, -0x1.0b582575423a0p-2, 0x1.34a6855f0509ap+0};
unsigned g_mat_version = 1;
unsigned g_dev_version[64] = {0};
int g_sm_count[64] = {0};

int ensure_device(int *dev_out)
{
    int dev = 0;
    CK(cudaGetDevice(&dev));
    if (dev_out) *dev_out = dev;
    if (dev < 0 || dev >= 64) return fail(RCT_ERR_ARG, "device index %d out of range", dev);
    std::lock_guard<std::mutex> lk(g_mu);
    if (g_dev_version[dev] == g_mat_version) return RCT_OK;
    MathTabs *h = new (std::nothrow) MathTabs;
    if (!h) return fail(RCT_ERR_NOMEM, "out of host memory");
    memcpy(h->log_inv, RCT_LOG_INV, sizeof h->log_inv);
    memcpy(h->log_t10, RCT_LOG_T10, sizeof h->log_t10);
    memcpy(h->exp2_t, RCT_EXP2_T, sizeof h->exp2_t);
    for (int k = 0; k < 256; ++k) h->unit[k] = (double)((float)k / 255.0f);  // LST:71 float32 divide
    cudaError_t e1 = cudaMemcpyToSymbol(g_tabs, h, sizeof(MathTabs));
    delete h;
    if (e1 != cudaSuccess) return fail(RCT_ERR_CUDA, "table upload failed: %s", cudaGetErrorString(e1));
    CK(cudaMemcpyToSymbol(c_mat, g_mat, sizeof g_mat));
    float ktm[9];
    for (int i = 0; i < 9; ++i) ktm[i] = (float)(g_mat[i] / 255.0);
    CK(cudaMemcpyToSymbol(c_ktm, ktm, sizeof ktm));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, dev));
    g_sm_count[dev] = prop.multiProcessorCount;
    g_dev_version[dev] = g_mat_version;
    return RCT_OK;
}

inline size_t align_up(size_t v, size_t a = 256) { return (v + a - 1) / a * a; }

// Switches to `device` for the scope and restores the caller's current device afterwards
// (a plan of another GPU must not change the calling thread's device: multi-GPU processes).
struct DeviceGuard {
    int prev = -1;
    bool switched = false;
    explicit DeviceGuard(int device)
    {
        if (cudaGetDevice(&prev) == cudaSuccess && prev != device) switched = cudaSetDevice(device) == cudaSuccess;
    }
    ~DeviceGuard()
    {
        if (switched) cudaSetDevice(prev);
    }
};

constexpr size_t kXchgBytes = 64 << 10;
struct Layout {
    size_t cap, maxparts;
    size_t off_ctrl, off_jobs, off_status, off_xtab, off_ytab;
    size_t off_hist, hist_stride_u32, off_lut, lut_stride_u32;
    size_t off_region, region_bytes;  // [slots][2] blocks of per-region arrays
    size_t r_col, r_cnt, r_keyA, r_recP, r_pof, r_pstart, r_ppix, r_items;
    size_t off_cmap, cmap_stride_u32;  // [slots] chunk maps of k_front_hist
    size_t off_zero, zero_bytes, z_fcnt, z_pfill;  // [slots][2] blocks zeroed before every sort
    size_t off_dbg_m, off_dbg_j, dbg_stride;
    size_t off_xchg;  // exchange buffer of the row-sharded binned path (kXchgBytes)
    size_t off_kt;    // [slots] key tables (rct_keytab.cuh)
    size_t total;
};

// per-job sort arrays that must start at zero (fine counts, part fill counters)
void zero_zone(size_t maxparts, size_t &z_fcnt, size_t &z_pfill, size_t &bytes)
{
    size_t z = 0;
    z_fcnt = z; z = align_up(z + 4 * (size_t)(kFineMax + 1));
    z_pfill = z; z = align_up(z + 4 * maxparts * (size_t)kFillStride);
    bytes = z;
}

Layout make_layout(int H, int W, int slots, unsigned flags)
{
    Layout L{};
    size_t npix = (size_t)H * (size_t)W;
    L.cap = npix < (size_t)kCube ? npix : (size_t)kCube;
    L.maxparts = L.cap / kPartT + 2;
    size_t o = 0;
    L.off_ctrl = o; o = align_up(o + sizeof(FrameCtrl) * slots);
    L.off_jobs = o; o = align_up(o + sizeof(Job) * 2 * slots);
    L.off_status = o; o = align_up(o + sizeof(int) * slots);
    L.off_xtab = o; o = align_up(o + sizeof(int) * (size_t)W);
    L.off_ytab = o; o = align_up(o + sizeof(int) * (size_t)H);
    L.hist_stride_u32 = 2 * (size_t)kCube;
    L.off_hist = o; o = align_up(o + 4 * L.hist_stride_u32 * slots);
    L.lut_stride_u32 = kCube;
    L.off_lut = o; o = align_up(o + 4 * L.lut_stride_u32 * slots);
    size_t r = 0;
    L.r_col = r; r = align_up(r + 4 * L.cap);
    L.r_cnt = r; r = align_up(r + 4 * L.cap);
    L.r_keyA = r; r = align_up(r + 8 * L.cap);
    L.r_recP = r; r = align_up(r + 16 * L.cap);
    L.r_pof = r; r = align_up(r + 2 * (size_t)kFineMax);
    L.r_pstart = r; r = align_up(r + 4 * (L.maxparts + 1));
    L.r_ppix = r; r = align_up(r + 4 * (L.maxparts + 1));
    // front-end items (4 B per region pixel) die before the sort records are born: share the memory
    if (4 * npix <= 16 * L.cap) L.r_items = L.r_recP;
    else { L.r_items = r; r = align_up(r + 4 * npix); }
    L.region_bytes = r;
    L.off_region = o; o += L.region_bytes * 2 * slots;
    L.cmap_stride_u32 = 2 * (size_t)kNB + 2 * npix / kChunk + 2;
    L.off_cmap = o; o = align_up(o + 4 * L.cmap_stride_u32 * slots);
    zero_zone(L.maxparts, L.z_fcnt, L.z_pfill, L.zero_bytes);
    L.off_zero = o; o += L.zero_bytes * 2 * slots;
    L.dbg_stride = L.cap;
    if (flags & RCT_FLAG_DEBUG) {
        L.off_dbg_m = o; o = align_up(o + 8 * L.cap * slots);
        L.off_dbg_j = o; o = align_up(o + 8 * L.cap * slots);
    }
    L.off_xchg = o; o = align_up(o + kXchgBytes);
    L.off_kt = o; o = align_up(o + kKtBytes * (size_t)slots);
    L.total = o;
    return L;
}

}  // namespace

constexpr int kStageSetsMax = 4;
struct rct_plan {
    int device, H, W, sh, sw, slots;
    unsigned flags;
    bool resample, owns_ws;
    bool rs_tma_ok;  // geometry admits k_apply_warp_rs (TMA-staged NEAREST resample)
    uint8_t *ws;
    Layout L;
    FrameCtrl *ctrl;
    Job *jobs;
    int *status_slots;
    int *xtab, *ytab;
    unsigned *hist, *lut;
    double *dbg_m;
    long long *dbg_j;
    int sm;
    int aggregate;
    int apply_tma;    // RCT_OPT_APPLY: 1 = bulk-async staged apply kernel where the geometry allows
    int match_interleave;  // RCT_OPT_MATCH
    int front;        // RCT_OPT_FRONTEND: 1 = colour-partitioned shared-memory histograms, 0 = colour cube
    bool cube_dirty;  // the partition front end needs (and leaves) an all-zero colour cube
    unsigned part_cap;  // entries a part may have for the shared-memory sort (RCT_OPT_PART_CAP)
    // lazily allocated helpers
    int *status_frames;
    int status_frames_cap;
    uint8_t *stage;  // device staging of the host entry point
    size_t stage_bytes;
    uint8_t *hstage;  // pinned host staging for callers that hand over pageable arrays (rct_transfer_host)
    size_t hstage_bytes;
    cudaStream_t own_stream;            // compute stream of the host entry point
    cudaStream_t h2d_stream, d2h_stream; // copy streams of the host entry point (overlap with compute)
    cudaEvent_t ev_in[kStageSetsMax], ev_comp[kStageSetsMax], ev_out[kStageSetsMax];
    // The workspace (ctrl, lists, LUT) is shared by every call on this plan: each entry point makes its
    // stream wait for the previous call's last kernel (ev_last) and records it again when it has enqueued
    // its own work, so calls from different streams (or the host entry point's own stream) never overlap.
    cudaEvent_t ev_last;
    bool ev_last_valid;
    int lpf;         // labels per frame of the call in progress (rct_transfer_labels_async), else 1
    int lut_mode;    // RCT_OPT_LUT: 1 = uint16 run indices + palette, 0 = uint32 RGB per colour
    uint8_t *files_ws;   // decode / encode workspaces and the JPEG output of rct_transfer_files_host
    size_t files_ws_bytes;
    int keytab;      // RCT_OPT_KEYTAB: 1 = key table in front of the colour LUT (rct_keytab.cuh)
    bool kt_live;    // the tables of the launch group in flight were built (exact path, interleaved match)
    int kt_mode;     // mode of the call in progress: the key table is built only where the apply pass can use it
    int stage_sets;  // RCT_OPT_STAGE_SETS: staging sets of the host entry point (2..kStageSetsMax)
    // RCT_OPT_TIMING: five events per launch group (memset | k_hist | solve | k_apply)
    int timing;
    std::vector<cudaEvent_t> ev_busy, ev_free;
    double phase_ms[4];
    long long phase_groups;
    // RCT_OPT_TIMING == 2: one event after every kernel of rct_transfer_async (ktiming ids)
    std::vector<cudaEvent_t> kev;
    std::vector<int> kev_id;
    double kernel_ms[RCT_TK_COUNT];
    long long kernel_groups;
};

extern "C" {

int rct_version(void) { return 100; }
const char *rct_last_error(void) { return t_err.c_str(); }
long long rct_launch_count(int reset)
{
    long long v = t_launches;
    if (reset) t_launches = 0;
    return v;
}

int rct_set_matrices(const double *m36)
{
    if (!m36) return fail(RCT_ERR_ARG, "null matrices");
    std::lock_guard<std::mutex> lk(g_mu);
    memcpy(g_mat, m36, sizeof g_mat);
    ++g_mat_version;
    return RCT_OK;
}
int rct_get_matrices(double *m36)
{
    if (!m36) return fail(RCT_ERR_ARG, "null matrices");
    std::lock_guard<std::mutex> lk(g_mu);
    memcpy(m36, g_mat, sizeof g_mat);
    return RCT_OK;
}

int rct_resample_table(int in_size, int out_size, int32_t *tab)
{
    if (in_size <= 0 || out_size <= 0 || !tab) return fail(RCT_ERR_ARG, "bad resample sizes");
    // Pillow's nearest filter: accumulated double source coordinate
    double scale = (double)in_size / (double)out_size;
    double pos = 0.5 * scale;
    for (int i = 0; i < out_size; ++i, pos += scale) {
        int src = (int)pos;
        tab[i] = src < in_size ? src : in_size - 1;
    }
    return RCT_OK;
}

size_t rct_plan_workspace_bytes(int H, int W, int slots, unsigned flags)
{
    if (H <= 0 || W <= 0 || slots <= 0) return 0;
    return make_layout(H, W, slots, flags).total;
}

int rct_plan_create(int device, int H, int W, int sh, int sw, int slots, unsigned flags, void *workspace,
                    size_t workspace_bytes, rct_plan **out)
{
    if (!out) return fail(RCT_ERR_ARG, "null plan pointer");
    *out = nullptr;
    if (H <= 0 || W <= 0 || sh <= 0 || sw <= 0 || slots <= 0 || slots > 4096)
        return fail(RCT_ERR_ARG, "bad geometry H=%d W=%d sh=%d sw=%d slots=%d", H, W, sh, sw, slots);
    if ((size_t)H * (size_t)W >= ((size_t)1 << 32)) return fail(RCT_ERR_ARG, "image too large (>= 2^32 pixels)");
    DeviceGuard guard(device);
    {
        int cur = -1;
        CK(cudaGetDevice(&cur));
        if (cur != device) return fail(RCT_ERR_CUDA, "cannot switch to device %d", device);
    }
    int rc = ensure_device(nullptr);
    if (rc) return rc;
    rct_plan *p = new (std::nothrow) rct_plan();
    if (!p) return fail(RCT_ERR_NOMEM, "out of host memory");
    p->device = device; p->H = H; p->W = W; p->sh = sh; p->sw = sw; p->slots = slots; p->flags = flags;
    p->resample = (sh != H) || (sw != W);
    p->L = make_layout(H, W, slots, flags);
    p->sm = g_sm_count[device] > 0 ? g_sm_count[device] : 148;
    p->aggregate = 0;
    p->front = 1;
    p->match_interleave = 1;
    p->apply_tma = 1;
    p->cube_dirty = true;
    p->part_cap = kPartCap;
    p->stage_sets = 4;
    p->lut_mode = 0;
    p->keytab = 1;
    p->lpf = 1;
    if (workspace) {
        if (workspace_bytes < p->L.total) {
            delete p;
            return fail(RCT_ERR_ARG, "workspace too small: %zu < %zu", workspace_bytes, p->L.total);
        }
        if (((uintptr_t)workspace & 255) != 0) {
            delete p;
            return fail(RCT_ERR_ARG, "workspace must be 256-byte aligned");
        }
        p->ws = (uint8_t *)workspace;
        p->owns_ws = false;
    } else {
        cudaError_t e = cudaMalloc((void **)&p->ws, p->L.total);
        if (e != cudaSuccess) {
            size_t need = p->L.total;
            delete p;
            return fail(RCT_ERR_NOMEM, "cudaMalloc(%zu) failed: %s", need, cudaGetErrorString(e));
        }
        p->owns_ws = true;
    }
    const Layout &L = p->L;
    p->ctrl = (FrameCtrl *)(p->ws + L.off_ctrl);
    p->jobs = (Job *)(p->ws + L.off_jobs);
    p->status_slots = (int *)(p->ws + L.off_status);
    p->xtab = (int *)(p->ws + L.off_xtab);
    p->ytab = (int *)(p->ws + L.off_ytab);
    p->hist = (unsigned *)(p->ws + L.off_hist);
    p->lut = (unsigned *)(p->ws + L.off_lut);
    p->dbg_m = (flags & RCT_FLAG_DEBUG) ? (double *)(p->ws + L.off_dbg_m) : nullptr;
    p->dbg_j = (flags & RCT_FLAG_DEBUG) ? (long long *)(p->ws + L.off_dbg_j) : nullptr;
    // job descriptors
    Job *hj = new (std::nothrow) Job[2 * slots];
    int32_t *tab = new (std::nothrow) int32_t[(size_t)(W > H ? W : H)];
    if (!hj || !tab) {
        delete[] hj; delete[] tab;
        rct_plan_destroy(p);
        return fail(RCT_ERR_NOMEM, "out of host memory");
    }
    for (int s = 0; s < slots; ++s)
        for (int r = 0; r < 2; ++r) {
            uint8_t *b = p->ws + L.off_region + L.region_bytes * (size_t)(2 * s + r);
            uint8_t *z = p->ws + L.off_zero + L.zero_bytes * (size_t)(2 * s + r);
            Job &j = hj[2 * s + r];
            j.rc = &p->ctrl[s].r[r];
            j.payA = (unsigned *)(b + L.r_col);
            j.cntA = (unsigned *)(b + L.r_cnt);
            j.keyA = (double *)(b + L.r_keyA);
            j.recP = (ulonglong2 *)(b + L.r_recP);
            j.pof = (unsigned short *)(b + L.r_pof);
            j.pstart = (unsigned *)(b + L.r_pstart);
            j.ppix = (unsigned *)(b + L.r_ppix);
            j.fcnt = (unsigned *)(z + L.z_fcnt);
            j.pfill = (unsigned *)(z + L.z_pfill);
            j.cap = (unsigned)L.cap;
            j.maxparts = (unsigned)L.maxparts;
        }
    cudaError_t e = cudaMemcpy(p->jobs, hj, sizeof(Job) * 2 * slots, cudaMemcpyHostToDevice);
    delete[] hj;
    if (e == cudaSuccess && p->resample) {
        rct_resample_table(sw, W, tab);
        e = cudaMemcpy(p->xtab, tab, sizeof(int) * (size_t)W, cudaMemcpyHostToDevice);
        // k_apply_warp_rs stages, per 256-pixel tile, the source spans of its (at most two) row
        // segments: every pair must fit kRsCap bytes (3 bytes per source pixel, each span widened
        // to 16-byte boundaries), the stylised frames must be whole 16-byte units, W >= 256
        bool ok = W >= 256 && ((size_t)sh * (size_t)sw * 3) % 16 == 0;
        auto seg_bytes = [&](int xa, int npx) -> size_t {
            if (npx <= 0) return 0;
            size_t span = (size_t)(tab[xa + npx - 1] - tab[xa] + 1);
            return (3 * span + 15 + 15 + 15) & ~(size_t)15;
        };
        for (int x0 = 0; ok && x0 < W; ++x0) {
            int na = W - x0 < 256 ? W - x0 : 256;
            if (seg_bytes(x0, na) + seg_bytes(0, 256 - na) > (size_t)kRsCap) ok = false;
        }
        p->rs_tma_ok = ok;
        if (e == cudaSuccess) {
            rct_resample_table(sh, H, tab);
            e = cudaMemcpy(p->ytab, tab, sizeof(int) * (size_t)H, cudaMemcpyHostToDevice);
        }
    }
    delete[] tab;
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&p->ev_last, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&p->own_stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&p->h2d_stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&p->d2h_stream, cudaStreamNonBlocking);
    for (int b = 0; b < kStageSetsMax && e == cudaSuccess; ++b) {
        e = cudaEventCreateWithFlags(&p->ev_in[b], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&p->ev_comp[b], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&p->ev_out[b], cudaEventDisableTiming);
    }
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
        rct_plan_destroy(p);
        return fail(RCT_ERR_CUDA, "plan setup failed: %s", cudaGetErrorString(e));
    }
    *out = p;
    return RCT_OK;
}

int rct_plan_destroy(rct_plan *p)
{
    if (!p) return RCT_OK;
    DeviceGuard guard(p->device);
    if (p->ev_last) cudaEventDestroy(p->ev_last);
    if (p->own_stream) cudaStreamDestroy(p->own_stream);
    if (p->h2d_stream) cudaStreamDestroy(p->h2d_stream);
    if (p->d2h_stream) cudaStreamDestroy(p->d2h_stream);
    for (int b = 0; b < kStageSetsMax; ++b) {
        if (p->ev_in[b]) cudaEventDestroy(p->ev_in[b]);
        if (p->ev_comp[b]) cudaEventDestroy(p->ev_comp[b]);
        if (p->ev_out[b]) cudaEventDestroy(p->ev_out[b]);
    }
    for (cudaEvent_t e : p->ev_busy) cudaEventDestroy(e);
    for (cudaEvent_t e : p->ev_free) cudaEventDestroy(e);
    for (cudaEvent_t e : p->kev) cudaEventDestroy(e);
    if (p->owns_ws && p->ws) cudaFree(p->ws);
    if (p->status_frames) cudaFree(p->status_frames);
    if (p->stage) cudaFree(p->stage);
    if (p->hstage) cudaFreeHost(p->hstage);
    if (p->files_ws) cudaFree(p->files_ws);
    delete p;
    return RCT_OK;
}

int rct_plan_set_option(rct_plan *p, int option, int value)
{
    if (!p) return fail(RCT_ERR_ARG, "null plan");
    if (option == RCT_OPT_AGGREGATE) p->aggregate = value;
    else if (option == RCT_OPT_TIMING) p->timing = value;
    else if (option == RCT_OPT_FRONTEND) p->front = value ? 1 : 0;
    else if (option == RCT_OPT_MATCH) p->match_interleave = value ? 1 : 0;
    else if (option == RCT_OPT_APPLY) p->apply_tma = (value >= 0 && value <= 8) ? value : 1;
    else if (option == RCT_OPT_LUT) p->lut_mode = value ? 1 : 0;
    else if (option == RCT_OPT_KEYTAB) p->keytab = value ? 1 : 0;
    else if (option == RCT_OPT_STAGE_SETS) p->stage_sets = value < 2 ? 2 : (value > kStageSetsMax ? kStageSetsMax : value);
    else if (option == RCT_OPT_PART_CAP) p->part_cap = (value > 0 && value <= kPartCap) ? (unsigned)value : (unsigned)kPartCap;
    else return fail(RCT_ERR_ARG, "unknown option %d", option);
    return RCT_OK;
}

}  // extern "C"

namespace {

struct FrameSet {
    const uint8_t *content, *sty, *mask;
    uint8_t *out;
    size_t cs, ss, ms, os;  // strides in bytes
};

PixArgs pix_args(rct_plan *p, const FrameSet &f, int mode, int label, size_t p_begin, size_t p_end, int *status_out)
{
    PixArgs a{};
    a.content = f.content; a.sty = f.sty; a.mask = f.mask; a.out = f.out;
    a.content_stride = f.cs; a.sty_stride = f.ss; a.mask_stride = f.ms; a.out_stride = f.os;
    a.xtab = p->resample ? p->xtab : nullptr;
    a.ytab = p->resample ? p->ytab : nullptr;
    a.hist = p->hist; a.hist_stride = p->L.hist_stride_u32;
    a.lut = p->lut; a.lut_stride = p->L.lut_stride_u32; a.lut_mode = p->lut_mode;
    a.kt = p->kt_live ? p->ws + p->L.off_kt : nullptr;
    a.ctrl = p->ctrl; a.status_out = status_out;
    a.W = p->W; a.sw = p->sw; a.sh_dbg = p->sh;
    a.p_begin = p_begin; a.p_end = p_end;
    a.mode = mode; a.label = label; a.lpf = p->lpf > 0 ? p->lpf : 1;
    a.aggregate = p->aggregate;
    return a;
}

unsigned pix_blocks(rct_plan *p, size_t npix, int nslots)
{
    size_t groups = (npix + 15) / 16;
    size_t blocks = (groups + 255) / 256;
#ifndef RCT_PB
#define RCT_PB 8
#endif
    size_t maxb = (size_t)p->sm * RCT_PB;
    if (nslots > 1) maxb = (maxb + nslots - 1) / nslots < (size_t)p->sm ? (size_t)p->sm : (maxb + nslots - 1) / nslots;
    if (blocks > maxb) blocks = maxb;
    return (unsigned)(blocks ? blocks : 1);
}

unsigned entry_blocks(rct_plan *p, int nslots)
{
    size_t blocks = (p->L.cap + 255) / 256;
#ifndef RCT_EB
#define RCT_EB 8
#endif
    size_t maxb = (size_t)p->sm * RCT_EB;
    if (nslots > 1) maxb = maxb / nslots < 32 ? 32 : maxb / nslots;
    if (blocks > maxb) blocks = maxb;
    return (unsigned)(blocks ? blocks : 1);
}

// timing == 2: an event after each kernel; the interval since the previous event is that kernel's
int kmark(rct_plan *p, cudaStream_t st, int id)
{
    if (!p || p->timing != 2) return RCT_OK;
    cudaEvent_t e;
    if (!p->ev_free.empty()) {
        e = p->ev_free.back();
        p->ev_free.pop_back();
    } else {
        CK(cudaEventCreate(&e));
    }
    p->kev.push_back(e);
    p->kev_id.push_back(id);
    CK(cudaEventRecord(e, st));
    return RCT_OK;
}
#define KMARK(id)                               \
    do {                                        \
        int rc_ = kmark(p, st, id);             \
        if (rc_) return rc_;                    \
    } while (0)

int mark(rct_plan *p, cudaStream_t st)
{
    if (p->timing != 1) return RCT_OK;
    cudaEvent_t e;
    if (!p->ev_free.empty()) {
        e = p->ev_free.back();
        p->ev_free.pop_back();
    } else {
        CK(cudaEventCreate(&e));
    }
    p->ev_busy.push_back(e);
    CK(cudaEventRecord(e, st));
    return RCT_OK;
}

int phase_hist(rct_plan *p, const FrameSet &f, int nslots, int mode, int label, size_t p_begin, size_t p_end,
               cudaStream_t st)
{
    int rc;
    if ((rc = mark(p, st))) return rc;
    CK(cudaMemsetAsync(p->ctrl, 0, sizeof(FrameCtrl) * nslots, st));
    KMARK(RCT_TK_BEGIN);
    CK(cudaMemsetAsync(p->hist, 0, 4 * p->L.hist_stride_u32 * nslots, st));
    p->cube_dirty = true;
    KMARK(RCT_TK_MEMSET);
    if ((rc = mark(p, st))) return rc;
    PixArgs a = pix_args(p, f, mode, label, p_begin, p_end, nullptr);
    dim3 grid(pix_blocks(p, p_end - p_begin, nslots), nslots);
    if (p->resample) LAUNCH(k_hist<true>, grid, 256, st, a);
    else LAUNCH(k_hist<false>, grid, 256, st, a);
    KMARK(RCT_TK_HIST);
    return mark(p, st);
}

FrontArgs front_args(rct_plan *p)
{
    FrontArgs f{};
    const Layout &L = p->L;
    f.items0 = (unsigned *)(p->ws + L.off_region + L.r_items);
    f.items_stride = L.region_bytes / 4;
    f.cmap0 = (unsigned *)(p->ws + L.off_cmap);
    f.cmap_stride = L.cmap_stride_u32;
    f.cmap_cap = (unsigned)L.cmap_stride_u32;
    f.items_cap = (unsigned)((size_t)p->H * (size_t)p->W);
    return f;
}

// colour-partitioned front end: pixels -> (colour, count) lists, no colour cube (rct_front.cuh)
int phase_front(rct_plan *p, const FrameSet &f, int nslots, int mode, int label, size_t p_begin, size_t p_end,
                cudaStream_t st)
{
    int rc;
    if ((rc = mark(p, st))) return rc;
    CK(cudaMemsetAsync(p->ctrl, 0, sizeof(FrameCtrl) * nslots, st));
    KMARK(RCT_TK_BEGIN);
    if (p->cube_dirty) {  // split buckets merge through windows of the cube and leave it zero again
        CK(cudaMemsetAsync(p->hist, 0, 4 * p->L.hist_stride_u32 * p->slots, st));
        p->cube_dirty = false;
    }
    KMARK(RCT_TK_MEMSET);
    if ((rc = mark(p, st))) return rc;
    static thread_local int attr_dev = -1;
    int dev = 0;
    CK(cudaGetDevice(&dev));
    if (attr_dev != dev) {
        CK(cudaFuncSetAttribute(k_front_hist, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemFrontHist));
        attr_dev = dev;
    }
    PixArgs a = pix_args(p, f, mode, label, p_begin, p_end, nullptr);
    FrontArgs fa = front_args(p);
    const size_t npix = p_end - p_begin;
    dim3 gcount(pix_blocks(p, npix, nslots), nslots);
    size_t ntiles = ((npix + 15) / 16 + kPartThreads - 1) / kPartThreads;
#ifndef RCT_FPG
#define RCT_FPG 64
#endif
    size_t maxb = ((size_t)p->sm * RCT_FPG + nslots - 1) / nslots;  // two waves of 2 CTAs per SM over all slots
    if (ntiles < 1) ntiles = 1;
    dim3 gpart((unsigned)(ntiles < maxb ? ntiles : maxb), nslots);
    const bool two = mode != RCT_MODE_MASK;
#define FRONT_DISPATCH(K, grid, threads, ...)                                      \
    do {                                                                           \
        if (p->resample) {                                                         \
            if (two) LAUNCH((K<true, true>), grid, threads, st, __VA_ARGS__);      \
            else LAUNCH((K<true, false>), grid, threads, st, __VA_ARGS__);         \
        } else {                                                                   \
            if (two) LAUNCH((K<false, true>), grid, threads, st, __VA_ARGS__);     \
            else LAUNCH((K<false, false>), grid, threads, st, __VA_ARGS__);        \
        }                                                                          \
    } while (0)
    FRONT_DISPATCH(k_front_count, gcount, 256, a);
    KMARK(RCT_TK_FRONT_COUNT);
    LAUNCH(k_front_scan, dim3(nslots), kNB, st, p->ctrl, fa);
    KMARK(RCT_TK_FRONT_SCAN);
    FRONT_DISPATCH(k_front_part, gpart, kPartThreads, a, fa);
    KMARK(RCT_TK_FRONT_PART);
#undef FRONT_DISPATCH
    k_front_hist<<<dim3((unsigned)p->L.cmap_stride_u32, nslots), kFrontHistThreads, kSmemFrontHist, st>>>(
        p->ctrl, p->jobs, fa, p->hist, p->L.hist_stride_u32);
    ++t_launches;
    CK(cudaGetLastError());
    KMARK(RCT_TK_FRONT_HIST);
    return mark(p, st);
}

constexpr size_t kSmemFineHist = (size_t)(kFineMax + kCoarse + 1) * 4;
constexpr size_t kSmemPartSort = (size_t)kPartCap * 16 + ((size_t)kPartCap + kPartCap / 32 + 2) * 4;  // records + padded counters

int ensure_kernel_attrs()
{
    static thread_local int done_dev = -1;
    int dev = 0;
    CK(cudaGetDevice(&dev));
    if (done_dev == dev) return RCT_OK;
    CK(cudaFuncSetAttribute(k_fine_hist, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemFineHist));
    CK(cudaFuncSetAttribute(k_part_sort, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemPartSort));
    CK(cudaFuncSetAttribute(k_scatter, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemScatter));
    CK(cudaFuncSetAttribute(k_scatter_tiled, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemScatter2));
    CK(cudaFuncSetAttribute(k_scatter_runs, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemScatter3));
    done_dev = dev;
    return RCT_OK;
}

// exact sort of njobs key lists (keys + coarse histograms ready; zero zones cleared)
int sort_jobs(rct_plan *p, const Job *jobs, int njobs, unsigned eblocks, size_t cap, size_t maxparts, unsigned part_cap,
              cudaStream_t st)
{
    int rc = ensure_kernel_attrs();
    if (rc) return rc;
    unsigned hblocks = (unsigned)((cap + kFineHistChunk - 1) / kFineHistChunk);
    LAUNCH(k_plan_fine, dim3(1, njobs), 1024, st, jobs);
    KMARK(RCT_TK_PLAN_FINE);
    k_fine_hist<<<dim3(hblocks, njobs), RCT_FHTH, kSmemFineHist, st>>>(jobs);
    ++t_launches;
    CK(cudaGetLastError());
    KMARK(RCT_TK_FINE_HIST);
    LAUNCH(k_plan_parts, dim3(1, njobs), 1024, st, jobs);
    KMARK(RCT_TK_PLAN_PARTS);
    // few, long-lived CTAs per job: each stages the fine-bin tables (kSmemScatter bytes) once
#ifndef RCT_SB
#define RCT_SB 12
#endif
#ifndef RCT_SCATTER_DEFAULT
#define RCT_SCATTER_DEFAULT 0
#endif
    unsigned sblocks = (unsigned)((p ? p->sm : 148) * RCT_SB / (njobs > 0 ? njobs : 1));  // ~4 waves of 3 CTAs per SM
    if (sblocks < 16u) sblocks = 16u;
    // Three scatters, the same time on the worst case (8 x 4K U, ms; N in brackets): k_scatter 0.996 (0.576), one returning
    // global atomic + one 16-byte store per entry; k_scatter_tiled 1.079 (0.507), one global reservation per 16K-entry tile
    // and part; k_scatter_runs 0.999 (0.532), the tile sorted by part in shared memory and streamed out in runs.  Neither
    // the per-entry atomics nor the store granularity at the L1 is what bounds it.  The tiled kernels hold a whole SM
    // (1024 threads, 120-200 KB), which costs the overlap with the other launch group: the U step is 10.83 ms with
    // k_scatter_runs against 10.57 ms with k_scatter, so k_scatter stays the default (profiles/r02_summary.md).
    // RCT_SCATTER_TILED: 0 = k_scatter, 1 = k_scatter_tiled, 2 = k_scatter_runs (tile sorted by part in shared memory)
    static const int tiled = getenv("RCT_SCATTER_TILED") ? atoi(getenv("RCT_SCATTER_TILED")) : RCT_SCATTER_DEFAULT;
    const unsigned sm = (unsigned)(p ? p->sm : 148), nj = (unsigned)(njobs > 0 ? njobs : 1);
    if (tiled == 2 && maxparts <= (size_t)kScat3MaxParts) {
        // one resident CTA per SM; ~3 CTAs per SM in the grid so that jobs of different length balance
        unsigned tb = (unsigned)((cap + kScat3Tile - 1) / kScat3Tile);
        unsigned per_job = sm * 3u / nj;
        if (per_job < 1u) per_job = 1u;
        if (tb > per_job) tb = per_job;
        k_scatter_runs<<<dim3(tb ? tb : 1u, njobs), kScat3Threads, kSmemScatter3, st>>>(jobs);
    } else if (tiled == 1 && maxparts <= (size_t)kScat2MaxParts) {
        // one CTA per SM and job at most; tiles of 16 K entries
        unsigned tb = (unsigned)((cap + kScat2Tile - 1) / kScat2Tile);
        const unsigned per_job = sm * 2u / nj;
        if (tb > (per_job ? per_job : 1u)) tb = per_job ? per_job : 1u;
        k_scatter_tiled<<<dim3(tb ? tb : 1u, njobs), kScat2Threads, kSmemScatter2, st>>>(jobs);
    } else {
        k_scatter<<<dim3(sblocks, njobs), RCT_SCT, kSmemScatter, st>>>(jobs);
    }
    ++t_launches;
    CK(cudaGetLastError());
    KMARK(RCT_TK_SCATTER);
    k_part_sort<<<dim3((unsigned)maxparts, njobs), kSortThreads, kSmemPartSort, st>>>(jobs, part_cap);
    ++t_launches;
    CK(cudaGetLastError());
    KMARK(RCT_TK_PART_SORT);
    LAUNCH(k_part_scan, dim3(1, njobs), 1024, st, jobs);
    KMARK(RCT_TK_PART_SCAN);
    return RCT_OK;
}

// The key table pays for itself only in front of the bulk-async apply kernels: MASK mode, one label per frame, stylised
// image at the content's size (k_apply_kt) or resampled through k_apply_warp_rs.  Elsewhere (label geometries, the
// register-staged resample kernel) it is not built at all.
inline bool want_keytab(const rct_plan *p)
{
    if (!p->keytab || p->lut_mode || p->lpf != 1 || p->kt_mode != RCT_MODE_MASK) return false;
    return !p->resample || (p->apply_tma == 8 && p->rs_tma_ok);  // resampled: only the TMA-staged kernel takes it
}

int phase_solve(rct_plan *p, int nslots, cudaStream_t st, bool from_cube)
{
    unsigned eb = entry_blocks(p, nslots);
    CK(cudaMemsetAsync(p->ws + p->L.off_zero, 0, p->L.zero_bytes * 2 * (size_t)nslots, st));
    if (want_keytab(p))  // parameters, seeds and flags of the key tables (rct_keytab.cuh)
        CK(cudaMemset2DAsync(p->ws + p->L.off_kt, kKtBytes, 0, kKtZeroBytes, (size_t)nslots, st));
    KMARK(RCT_TK_MEMSET);
    if (from_cube) {
        LAUNCH(k_compact, dim3(kCube / 1024, 2, nslots), 256, st, p->hist, p->L.hist_stride_u32, p->ctrl, p->jobs);
        KMARK(RCT_TK_COMPACT);
    }
    LAUNCH(k_lab_moments, dim3(eb, 2, nslots), 256, st, p->ctrl, p->jobs);
    KMARK(RCT_TK_LAB_MOMENTS);
    LAUNCH(k_pca, dim3((nslots + 31) / 32), 32, st, p->ctrl, nslots);
    KMARK(RCT_TK_PCA);
#ifndef RCT_EBK
#define RCT_EBK 2
#endif
    LAUNCH(k_keys, dim3(eb * RCT_EBK, 2, nslots), 256, st, p->ctrl, p->jobs);
    KMARK(RCT_TK_KEYS);
    int rc = sort_jobs(p, p->jobs, 2 * nslots, eb, p->L.cap, p->L.maxparts, p->part_cap, st);
    if (rc) return rc;
    if (p->lut_mode)
        LAUNCH(k_match_pal, dim3((unsigned)p->L.maxparts, nslots), kMatchThreads, st, p->ctrl, p->jobs, p->lut,
               p->L.lut_stride_u32, p->dbg_m, p->dbg_j, p->L.dbg_stride, p->match_interleave);
    else
        LAUNCH(k_match, dim3((unsigned)p->L.maxparts, nslots), kMatchThreads, st, p->ctrl, p->jobs, p->lut,
               p->L.lut_stride_u32, p->dbg_m, p->dbg_j, p->L.dbg_stride, p->match_interleave,
               want_keytab(p) ? p->ws + p->L.off_kt : nullptr);
    if (p->lut_mode) {  // palette LUT: runs of equal output colour along the sorted target -> uint16 indices + palette
        LAUNCH(k_pal_count, dim3((unsigned)p->L.maxparts, nslots), kPalThreads, st, p->ctrl, p->jobs);
        LAUNCH(k_pal_scan, dim3(nslots), 1024, st, p->ctrl, p->jobs);
        LAUNCH(k_pal_write, dim3((unsigned)p->L.maxparts, nslots), kPalThreads, st, p->ctrl, p->jobs, p->lut, p->L.lut_stride_u32);
    }
    KMARK(RCT_TK_MATCH);
    p->kt_live = false;
    if (want_keytab(p)) {  // key table: run starts -> flagged cells + seeds -> prefix maximum -> cells
        uint8_t *kt0 = p->ws + p->L.off_kt;
        LAUNCH(k_kt_mark, dim3(8, nslots), 256, st, p->ctrl, kt0, p->match_interleave);
        LAUNCH(k_kt_fill, dim3(nslots), 1024, st, p->ctrl, kt0, p->match_interleave);
        KMARK(RCT_TK_KEYTAB);
        p->kt_live = true;
    }
    return mark(p, st);
}

int phase_apply(rct_plan *p, const FrameSet &f, int nslots, int mode, int label, size_t p_begin, size_t p_end,
                int *status_out, cudaStream_t st)
{
    PixArgs a = pix_args(p, f, mode, label, p_begin, p_end, status_out);
    // bulk-async staged pipeline (rct_apply.cuh) when the geometry allows 16-byte bulk copies
    uintptr_t al = (uintptr_t)f.content | (uintptr_t)f.sty | (uintptr_t)f.mask | (uintptr_t)f.out | (uintptr_t)p_begin;
    if (nslots > 1) al |= f.cs | f.ss | f.ms | f.os;
    const size_t npix16 = (p_end - p_begin) & ~(size_t)15;
    if (p->apply_tma && mode == RCT_MODE_MASK && !p->resample && f.mask && (al & 15) == 0 && npix16 > 0) {
        static thread_local int attr_dev = -1;
        int dev = 0;
        CK(cudaGetDevice(&dev));
        if (attr_dev != dev) {
            CK(cudaFuncSetAttribute(k_apply_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(ApSmem)));
            CK(cudaFuncSetAttribute((k_apply_warp<true, 8, 4>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(AwSmem<8>)));
            CK(cudaFuncSetAttribute((k_apply_warp<true, 8, 5>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(AwSmem<8>)));
            CK(cudaFuncSetAttribute((k_apply_warp<false, 8, 4>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(AwSmem<8>)));
            CK(cudaFuncSetAttribute((k_apply_warp<true, 16, 3>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(AwSmem<16>)));
            CK(cudaFuncSetAttribute((k_apply_warp<true, 8, 4, 3>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(AwSmem<8, 3>)));
            CK(cudaFuncSetAttribute((k_apply_warp<true, 8, 4, 2, true>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(AwSmem<8, 2, true>)));
            CK(cudaFuncSetAttribute((k_apply_warp<true, 8, 4, 2, false, 1>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(AwSmem<8>)));
            CK(cudaFuncSetAttribute(k_apply_kt, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(AwSmem<8>)));
            CK(cudaFuncSetAttribute((k_apply_warp<true, 16, 3>), cudaFuncAttributePreferredSharedMemoryCarveout, 100));
            CK(cudaFuncSetAttribute(k_apply_tma, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
            attr_dev = dev;
        }
        PixArgs b = a;
        b.p_end = p_begin + npix16;
        if (p->lut_mode) {  // palette LUT: the per-warp pipeline with the two-level lookup
            const size_t tiles = (npix16 + AwCfg<8>::kTile - 1) / AwCfg<8>::kTile, total = tiles * (size_t)nslots;
            const size_t need = (total + AwCfg<8>::kWarps - 1) / AwCfg<8>::kWarps, maxb = (size_t)p->sm * 4;
            k_apply_warp<true, 8, 4, 2, false, 1><<<(unsigned)(need < maxb ? need : maxb), AwCfg<8>::kThreads, sizeof(AwSmem<8>), st>>>(
                b, (unsigned)tiles, (unsigned)nslots);
        } else if (p->apply_tma == 3) {  // CTA-wide ring (kept for A/B)
            const size_t tiles = (npix16 + kApTile - 1) / kApTile, total = tiles * (size_t)nslots;
            const size_t maxb = (size_t)p->sm * 2;
            k_apply_tma<<<(unsigned)(total < maxb ? total : maxb), kApThreads, sizeof(ApSmem), st>>>(b, (unsigned)tiles,
                                                                                                       (unsigned)nslots);
        } else if (p->apply_tma == 4) {  // one pipeline per warp, 512-pixel tiles, bulk-async loads
            const size_t tiles = (npix16 + AwCfg<16>::kTile - 1) / AwCfg<16>::kTile, total = tiles * (size_t)nslots;
            const size_t need = (total + AwCfg<16>::kWarps - 1) / AwCfg<16>::kWarps, maxb = (size_t)p->sm * AwCfg<16>::kMinBlocks;
            k_apply_warp<true, 16, 3><<<(unsigned)(need < maxb ? need : maxb), AwCfg<16>::kThreads, sizeof(AwSmem<16>), st>>>(
                b, (unsigned)tiles, (unsigned)nslots);
        } else {  // one pipeline per warp, 256-pixel tiles: 1 = bulk-async (TMA 1-D) loads, 2 = cp.async (LDGSTS)
            const size_t tiles = (npix16 + AwCfg<8>::kTile - 1) / AwCfg<8>::kTile, total = tiles * (size_t)nslots;
            const size_t need = (total + AwCfg<8>::kWarps - 1) / AwCfg<8>::kWarps;
            const size_t maxb = (size_t)p->sm * (p->apply_tma == 5 ? 5 : 4);
            const unsigned grid = (unsigned)(need < maxb ? need : maxb);
            if (p->apply_tma == 2)
                k_apply_warp<false, 8, 4><<<grid, AwCfg<8>::kThreads, sizeof(AwSmem<8>), st>>>(b, (unsigned)tiles, (unsigned)nslots);
            else if (p->apply_tma == 7)
                k_apply_warp<true, 8, 4, 2, true><<<grid, AwCfg<8>::kThreads, sizeof(AwSmem<8, 2, true>), st>>>(b, (unsigned)tiles, (unsigned)nslots);
            else if (p->apply_tma == 6)
                k_apply_warp<true, 8, 4, 3><<<grid, AwCfg<8>::kThreads, sizeof(AwSmem<8, 3>), st>>>(b, (unsigned)tiles, (unsigned)nslots);
            else if (p->apply_tma == 5)
                k_apply_warp<true, 8, 5><<<grid, AwCfg<8>::kThreads, sizeof(AwSmem<8>), st>>>(b, (unsigned)tiles, (unsigned)nslots);
            else if (b.kt)
                k_apply_kt<<<grid, 256, sizeof(AwSmem<8>), st>>>(b, (unsigned)tiles, (unsigned)nslots);
            else
                k_apply_warp<true, 8, 4><<<grid, AwCfg<8>::kThreads, sizeof(AwSmem<8>), st>>>(b, (unsigned)tiles, (unsigned)nslots);
        }
        ++t_launches;
        CK(cudaGetLastError());
        if (npix16 < p_end - p_begin) {  // ragged tail of < 16 pixels per frame
            PixArgs c = a;
            c.p_begin = p_begin + npix16;
            if (p->lut_mode) LAUNCH((k_apply<false, 1>), dim3(1, nslots), 256, st, c);
            else LAUNCH(k_apply<false>, dim3(1, nslots), 256, st, c);
        }
        KMARK(RCT_TK_APPLY);
        return mark(p, st);
    }
    static const bool rs_disabled = getenv("RCT_DISABLE_RS") != nullptr;  // debugging aid
    if (!rs_disabled && !p->lut_mode && p->apply_tma == 8 && mode == RCT_MODE_MASK && p->resample && p->rs_tma_ok && f.mask && npix16 > 0 &&
        (((uintptr_t)f.content | (uintptr_t)f.sty | (uintptr_t)f.mask | (uintptr_t)f.out | (uintptr_t)p_begin |
          (nslots > 1 ? (f.cs | f.ss | f.ms | f.os) : 0)) & 15) == 0) {
        // the a-1 resample staged by bulk-async copies (k_apply_warp_rs): 256-pixel tiles, one pipeline per warp
        static thread_local int attr_dev_rs = -1;
        int dev = 0;
        CK(cudaGetDevice(&dev));
        if (attr_dev_rs != dev) {
            CK(cudaFuncSetAttribute(k_apply_warp_rs, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(RsSmem)));
            attr_dev_rs = dev;
        }
        PixArgs b = a;
        b.p_end = p_begin + npix16;
        const size_t tiles = (npix16 + 255) / 256, total = tiles * (size_t)nslots;
        const size_t need = (total + 7) / 8, maxb = (size_t)p->sm * 4;
        k_apply_warp_rs<<<(unsigned)(need < maxb ? need : maxb), 256, sizeof(RsSmem), st>>>(b, (unsigned)tiles, (unsigned)nslots);
        ++t_launches;
        CK(cudaGetLastError());
        if (npix16 < p_end - p_begin) {
            PixArgs c = a;
            c.p_begin = p_begin + npix16;
            LAUNCH(k_apply<true>, dim3(1, nslots), 256, st, c);
        }
        KMARK(RCT_TK_APPLY);
        return mark(p, st);
    }
    dim3 grid(pix_blocks(p, p_end - p_begin, nslots), nslots);
    if (p->lut_mode) {
        if (p->resample) LAUNCH((k_apply<true, 1>), grid, 256, st, a);
        else LAUNCH((k_apply<false, 1>), grid, 256, st, a);
    } else if (p->resample) LAUNCH(k_apply<true>, grid, 256, st, a);
    else LAUNCH(k_apply<false>, grid, 256, st, a);
    KMARK(RCT_TK_APPLY);
    return mark(p, st);
}

// ordering between successive users of a plan's workspace (see rct_plan::ev_last)
int plan_enter(rct_plan *p, cudaStream_t st)
{
    if (p->ev_last_valid) CK(cudaStreamWaitEvent(st, p->ev_last, 0));
    return RCT_OK;
}
int plan_leave(rct_plan *p, cudaStream_t st)
{
    CK(cudaEventRecord(p->ev_last, st));
    p->ev_last_valid = true;
    return RCT_OK;
}

// Every entry point that enqueues work on a plan's workspace runs inside one of these.  On the success path the entry
// point calls leave() itself (so that a failing event record is reported); on ANY other exit the destructor records the
// plan's ordering event behind the work already enqueued - a later call on the plan from another stream then waits for
// it instead of overlapping it in the shared workspace - and puts back the plan fields the failed call had changed.
struct PlanScope {
    rct_plan *p;
    cudaStream_t st;
    bool entered = false;
    int front_saved, kt_saved;
    cudaStream_t drain[2] = {nullptr, nullptr};  // host entry points: copy streams that touch the CALLER's host buffers
    PlanScope(rct_plan *plan, cudaStream_t stream) : p(plan), st(stream), front_saved(plan->front), kt_saved(plan->kt_mode) {}
    PlanScope(const PlanScope &) = delete;
    PlanScope &operator=(const PlanScope &) = delete;
    int enter()
    {
        int rc = plan_enter(p, st);
        entered = rc == RCT_OK;
        return rc;
    }
    int leave()
    {
        entered = false;
        return plan_leave(p, st);
    }
    ~PlanScope()
    {
        if (!entered) return;
        if (cudaEventRecord(p->ev_last, st) == cudaSuccess) p->ev_last_valid = true;
        for (cudaStream_t d : drain)  // the caller may release its host buffers as soon as the error is returned
            if (d) cudaStreamSynchronize(d);
        if (drain[0] || drain[1]) cudaStreamSynchronize(st);
        p->front = front_saved;
        p->kt_mode = kt_saved;
    }
};

int check_mode(int mode, const void *mask)
{
    if (mode < 0 || mode > 2) return fail(RCT_ERR_ARG, "bad mode %d", mode);
    if (mode != RCT_MODE_ALL && !mask) return fail(RCT_ERR_ARG, "mask required for mode %d", mode);
    return RCT_OK;
}

int ensure_status_frames(rct_plan *p, int n)
{
    if (n <= p->status_frames_cap) return RCT_OK;
    if (p->status_frames) CK(cudaFree(p->status_frames));
    p->status_frames = nullptr;
    p->status_frames_cap = 0;
    int cap = n < 64 ? 64 : n;
    CK(cudaMalloc((void **)&p->status_frames, sizeof(int) * (size_t)cap));
    p->status_frames_cap = cap;
    return RCT_OK;
}

}  // namespace

// ---- pageable callers of the host entry point ------------------------------------------------------------------
// np.array(Image.open(...)) (LST:200-201, 219-220) is pageable memory.  Handed straight to cudaMemcpy2DAsync the driver
// stages it through its own small pinned buffer on ONE thread, synchronously: ~11 GB/s, 1.1 GP/s end to end against
// 6.7 GP/s from pinned buffers.  So rct_transfer_host copies pageable buffers into / out of a pinned ring of the plan
// with a few host threads (below) while the DMA engines and the kernels work on the neighbouring groups.
namespace {
// Large one-way copies between the caller's arrays and the pinned ring: neither side is read again by this core, so
// the destination is written with non-temporal stores (no read-for-ownership of the destination lines: two bytes of
// memory traffic per byte copied instead of three, and the copier does not evict the other threads' data).
inline void copy_streaming(uint8_t *dst, const uint8_t *src, size_t n)
{
#if defined(__SSE2__)
    static const bool on = !(getenv("RCT_HOST_NT") && atoi(getenv("RCT_HOST_NT")) == 0);
    if (on && n >= 4096) {
        const size_t head = (16 - ((uintptr_t)dst & 15)) & 15;
        memcpy(dst, src, head);
        dst += head; src += head; n -= head;
        const size_t blocks = n / 64;
        for (size_t i = 0; i < blocks; ++i, src += 64, dst += 64) {
            const __m128i a = _mm_loadu_si128((const __m128i *)src), b = _mm_loadu_si128((const __m128i *)(src + 16)),
                          c = _mm_loadu_si128((const __m128i *)(src + 32)), d = _mm_loadu_si128((const __m128i *)(src + 48));
            _mm_stream_si128((__m128i *)dst, a);
            _mm_stream_si128((__m128i *)(dst + 16), b);
            _mm_stream_si128((__m128i *)(dst + 32), c);
            _mm_stream_si128((__m128i *)(dst + 48), d);
        }
        _mm_sfence();  // the DMA engine / the caller must see these stores once the batch is reported done
        n -= blocks * 64;
    }
#endif
    memcpy(dst, src, n);
}

class HostPool {
public:
    struct Job { uint8_t *dst; const uint8_t *src; size_t bytes; };
    static HostPool &get()
    {
        static HostPool pool;
        return pool;
    }
    // copies every job, split into chunks of kChunk bytes over the workers and the calling thread; returns when done
    void copy(const Job *jobs, int njobs)
    {
        std::lock_guard<std::mutex> one_call(call_mu_);  // one batch at a time: calls on different plans take turns
        size_t total = 0;
        for (int i = 0; i < njobs; ++i) total += (jobs[i].bytes + kChunk - 1) / kChunk;
        if (total == 0) return;
        if (total == 1 || workers_.empty()) {
            for (int i = 0; i < njobs; ++i) copy_streaming(jobs[i].dst, jobs[i].src, jobs[i].bytes);
            return;
        }
        {
            std::lock_guard<std::mutex> lk(mu_);
            jobs_ = jobs; njobs_ = njobs; total_ = total;
            next_.store(0); done_.store(0);
            ++epoch_;
        }
        cv_.notify_all();
        work();
        std::unique_lock<std::mutex> lk(mu_);
        cv_done_.wait(lk, [&] { return done_.load() == total_ && active_ == 0; });
        jobs_ = nullptr;
    }

private:
    static constexpr size_t kChunk = (size_t)1 << 20;
    HostPool()
    {
        int n = 0;
        if (const char *e = getenv("RCT_HOST_THREADS")) n = atoi(e);
        if (n <= 0) {
            unsigned hw = std::thread::hardware_concurrency();
            n = hw >= 16 ? 8 : (hw >= 4 ? (int)hw / 2 : 1);
        }
        if (n > 32) n = 32;
        for (int i = 0; i + 1 < n; ++i) {  // the calling thread is the n-th copier
            try {
                workers_.emplace_back([this] { loop(); });
            } catch (...) {
                break;  // fewer threads: slower, still correct
            }
        }
    }
    ~HostPool()
    {
        {
            std::lock_guard<std::mutex> lk(mu_);
            stop_ = true;
        }
        cv_.notify_all();
        for (auto &t : workers_) t.join();
    }
    void work()
    {
        for (;;) {
            size_t c = next_.fetch_add(1);
            if (c >= total_) break;
            size_t k = c;
            for (int i = 0; i < njobs_; ++i) {
                const size_t nc = (jobs_[i].bytes + kChunk - 1) / kChunk;
                if (k < nc) {
                    const size_t off = k * kChunk, len = jobs_[i].bytes - off < kChunk ? jobs_[i].bytes - off : kChunk;
                    copy_streaming(jobs_[i].dst + off, jobs_[i].src + off, len);
                    break;
                }
                k -= nc;
            }
            done_.fetch_add(1);
        }
    }
    void loop()
    {
        unsigned long long seen = 0;
        std::unique_lock<std::mutex> lk(mu_);
        for (;;) {
            cv_.wait(lk, [&] { return stop_ || epoch_ != seen; });
            if (stop_) return;
            seen = epoch_;
            if (!jobs_) continue;  // woke up after that batch was finished by the others: never touch a batch half set up
            ++active_;
            lk.unlock();
            work();
            lk.lock();
            --active_;
            if (done_.load() == total_ && active_ == 0) cv_done_.notify_all();
        }
    }
    std::vector<std::thread> workers_;
    std::mutex mu_, call_mu_;
    std::condition_variable cv_, cv_done_;
    const Job *jobs_ = nullptr;
    int njobs_ = 0, active_ = 0;
    size_t total_ = 0;
    std::atomic<size_t> next_{0}, done_{0};
    unsigned long long epoch_ = 0;
    bool stop_ = false;
};

// true when the driver knows nothing about this host pointer: plain malloc / numpy memory, not cudaHostAlloc'ed or registered
bool is_pageable(const void *ptr)
{
    if (!ptr) return false;
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, ptr) != cudaSuccess) {
        cudaGetLastError();  // older drivers report unregistered memory as an error: clear it
        return true;
    }
    return a.type == cudaMemoryTypeUnregistered;
}
}  // namespace

extern "C" {

int rct_transfer_async(rct_plan *p, const uint8_t *content, size_t content_stride, const uint8_t *stylised,
                       size_t stylised_stride, const uint8_t *mask, size_t mask_stride, uint8_t *out,
                       size_t out_stride, int n_frames, int mode, int label, void *stream)
{
    if (!p || !content || !stylised || !out || n_frames <= 0) return fail(RCT_ERR_ARG, "null/invalid argument");
    int rc = check_mode(mode, mask);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    rc = ensure_status_frames(p, n_frames);
    if (rc) return rc;
    PlanScope scope(p, st);
    if ((rc = scope.enter())) return rc;
    p->kt_mode = mode;
    const size_t npix = (size_t)p->H * (size_t)p->W;
    for (int f0 = 0; f0 < n_frames; f0 += p->slots) {
        int ns = n_frames - f0 < p->slots ? n_frames - f0 : p->slots;
        FrameSet f{content + (size_t)f0 * content_stride, stylised + (size_t)f0 * stylised_stride,
                   mask ? mask + (size_t)f0 * mask_stride : nullptr, out + (size_t)f0 * out_stride,
                   content_stride, stylised_stride, mask_stride, out_stride};
        if ((rc = p->front ? phase_front(p, f, ns, mode, label, 0, npix, st)
                           : phase_hist(p, f, ns, mode, label, 0, npix, st))) return rc;
        if ((rc = phase_solve(p, ns, st, !p->front))) return rc;
        if ((rc = phase_apply(p, f, ns, mode, label, 0, npix, p->status_frames + f0, st))) return rc;
    }
    return scope.leave();
}

int rct_plan_status(rct_plan *p, int *status_host, int n_frames, void *stream)
{
    if (!p || !status_host || n_frames <= 0 || n_frames > p->status_frames_cap)
        return fail(RCT_ERR_ARG, "bad status request");
    cudaStream_t st = (cudaStream_t)stream;
    CK(cudaMemcpyAsync(status_host, p->status_frames, sizeof(int) * (size_t)n_frames, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    return RCT_OK;
}

// BASELINE.json configs[4]: every label of a label map is an independent exact transfer (target =
// content pixels with that label, source = stylised pixels with that label).  One launch group carries
// slots / n_labels frames x n_labels labels as virtual slots (slot = frame * n_labels + label): the
// solve runs all labels of all frames of the group side by side instead of n_labels small passes.
// `out` must hold the values pixels of labels >= n_labels keep (normally a copy of content).
int rct_transfer_labels_async(rct_plan *p, const uint8_t *content, size_t content_stride, const uint8_t *stylised,
                              size_t stylised_stride, const uint8_t *labels, size_t labels_stride, uint8_t *out,
                              size_t out_stride, int n_frames, int n_labels, void *stream)
{
    if (!p || !content || !stylised || !labels || !out || n_frames <= 0) return fail(RCT_ERR_ARG, "null/invalid argument");
    if (n_labels <= 0 || n_labels > 256 || n_labels > p->slots)
        return fail(RCT_ERR_ARG, "n_labels %d needs a plan with at least as many slots (has %d)", n_labels, p->slots);
    cudaStream_t st = (cudaStream_t)stream;
    int rc = ensure_status_frames(p, n_frames * n_labels);
    if (rc) return rc;
    const size_t npix = (size_t)p->H * (size_t)p->W;
    const int fpg = p->slots / n_labels;  // frames per launch group
    PlanScope scope(p, st);
    if ((rc = scope.enter())) return rc;
    p->kt_mode = RCT_MODE_LABEL;
    p->lpf = n_labels;
    for (int f0 = 0; f0 < n_frames && rc == RCT_OK; f0 += fpg) {
        const int nf = n_frames - f0 < fpg ? n_frames - f0 : fpg, ns = nf * n_labels;
        FrameSet f{content + (size_t)f0 * content_stride, stylised + (size_t)f0 * stylised_stride,
                   labels + (size_t)f0 * labels_stride, out + (size_t)f0 * out_stride,
                   content_stride, stylised_stride, labels_stride, out_stride};
        rc = p->front ? phase_front(p, f, ns, RCT_MODE_LABEL, 0, 0, npix, st) : phase_hist(p, f, ns, RCT_MODE_LABEL, 0, 0, npix, st);
        if (rc == RCT_OK) rc = phase_solve(p, ns, st, !p->front);
        if (rc == RCT_OK) rc = phase_apply(p, f, ns, RCT_MODE_LABEL, 0, 0, npix, p->status_frames + (size_t)f0 * n_labels, st);
    }
    p->lpf = 1;
    if (rc == RCT_OK) rc = scope.leave();
    return rc;
}

int rct_transfer_host(rct_plan *p, const uint8_t *content_h, const uint8_t *stylised_h, const uint8_t *mask_h,
                      uint8_t *out_h, int n_frames, int mode, int label, int *status_host)
{
    if (!p || !content_h || !stylised_h || !out_h || n_frames <= 0) return fail(RCT_ERR_ARG, "null/invalid argument");
    int rc = check_mode(mode, mask_h);
    if (rc) return rc;
    DeviceGuard guard(p->device);
    const size_t npix = (size_t)p->H * (size_t)p->W, spix = (size_t)p->sh * (size_t)p->sw;
    const size_t cb = align_up(npix * 3), sb = align_up(spix * 3), mb = align_up(npix);
    // A ring of S staging sets (RCT_OPT_STAGE_SETS, default 4): group k is copied in on the H2D
    // stream, computed on the plan's stream and copied out on the D2H stream; a set's inputs are
    // free again once its compute is done, its output once the copy-out is done, so the H2D
    // stream (the bottleneck: 7 of the 10 bytes per pixel) never waits for a copy-out.
    // Frames are contiguous on the host; device staging keeps aligned per-slot strides.
    const int S = p->stage_sets;
    const size_t per_set = (cb + sb + mb + cb) * (size_t)p->slots;
    const size_t need = (size_t)S * per_set;
    if (p->stage_bytes < need) {
        if (p->stage) CK(cudaFree(p->stage));
        p->stage = nullptr;
        p->stage_bytes = 0;
        CK(cudaMalloc((void **)&p->stage, need));
        p->stage_bytes = need;
    }
    // Pageable host buffers (what numpy / PIL hand over) go through a pinned ring of the plan, copied by HostPool's
    // threads: group k is staged in while the GPU works on group k-1, group k-1's output is copied out to the caller
    // while the GPU works on group k.  Pinned or registered buffers keep the direct DMA path.  RCT_HOST_STAGING=0: off.
    static const bool staging_on = !(getenv("RCT_HOST_STAGING") && atoi(getenv("RCT_HOST_STAGING")) == 0);
    const bool pg_c = staging_on && is_pageable(content_h), pg_s = staging_on && is_pageable(stylised_h),
               pg_m = staging_on && is_pageable(mask_h), pg_o = staging_on && is_pageable(out_h);
    const bool pg_in = pg_c || pg_s || pg_m || (pg_o && mode == RCT_MODE_LABEL);
    const size_t hper_set = per_set + cb * (size_t)p->slots;  // inputs (device layout) + a separate output region
    if (pg_in || pg_o) {
        const size_t hneed = (size_t)S * hper_set;
        if (p->hstage_bytes < hneed) {
            if (p->hstage) CK(cudaFreeHost(p->hstage));
            p->hstage = nullptr;
            p->hstage_bytes = 0;
            if (cudaHostAlloc((void **)&p->hstage, hneed, cudaHostAllocDefault) != cudaSuccess) {
                cudaGetLastError();
                return fail(RCT_ERR_NOMEM, "cudaHostAlloc(%zu) for the pinned staging ring failed", hneed);
            }
            p->hstage_bytes = hneed;
        }
    }
    rc = ensure_status_frames(p, n_frames);
    if (rc) return rc;
    cudaStream_t st = p->own_stream, sin = p->h2d_stream, sout = p->d2h_stream;
    HostPool::Job jobs[4 * 64];
    // copies the finished output of group `kk` (frames g0 .. g0+gn) from its pinned set to the caller's array
    auto drain_out = [&](int kk, int g0, int gn) -> int {
        const int bb = kk % S;
        CK(cudaEventSynchronize(p->ev_out[bb]));
        const uint8_t *h_o = p->hstage + hper_set * bb + per_set;
        int nj = 0;
        for (int i = 0; i < gn; ++i) jobs[nj++] = {out_h + (size_t)(g0 + i) * npix * 3, h_o + cb * i, npix * 3};
        HostPool::get().copy(jobs, nj);
        return RCT_OK;
    };
    if (p->slots > 64 && (pg_in || pg_o)) return fail(RCT_ERR_ARG, "pageable host buffers: at most 64 slots per plan");
    PlanScope scope(p, st);
    scope.drain[0] = sin;
    scope.drain[1] = sout;
    if ((rc = scope.enter())) return rc;
    p->kt_mode = mode;
    int k = 0;
    for (int f0 = 0; f0 < n_frames; f0 += p->slots, ++k) {
        const int b = k % S;
        int ns = n_frames - f0 < p->slots ? n_frames - f0 : p->slots;
        uint8_t *d_c = p->stage + per_set * b, *d_s = d_c + cb * p->slots, *d_m = d_s + sb * p->slots,
                *d_o = d_m + mb * p->slots;
        // pinned set b (only touched when something is pageable, in which case hstage exists): same layout as the device set
        uint8_t *const h_base = (pg_in || pg_o) ? p->hstage + hper_set * b : p->stage + per_set * b;
        uint8_t *h_c = h_base, *h_s = h_c + cb * p->slots, *h_m = h_s + sb * p->slots, *h_i = h_m + mb * p->slots;
        uint8_t *h_o = (pg_in || pg_o) ? h_base + per_set : h_base;
        if (pg_in) {  // stage this group's pageable inputs into pinned set b (free once group k-S's copy-in is done)
            if (k >= S) CK(cudaEventSynchronize(p->ev_in[b]));
            int nj = 0;
            for (int i = 0; i < ns; ++i) {
                if (pg_c) jobs[nj++] = {h_c + cb * i, content_h + (size_t)(f0 + i) * npix * 3, npix * 3};
                if (pg_s) jobs[nj++] = {h_s + sb * i, stylised_h + (size_t)(f0 + i) * spix * 3, spix * 3};
                if (pg_m) jobs[nj++] = {h_m + mb * i, mask_h + (size_t)(f0 + i) * npix, npix};
                if (pg_o && mode == RCT_MODE_LABEL) jobs[nj++] = {h_i + cb * i, out_h + (size_t)(f0 + i) * npix * 3, npix * 3};
            }
            HostPool::get().copy(jobs, nj);
        }
        if (k >= S) {  // set b was last used by group k-S
            CK(cudaStreamWaitEvent(sin, p->ev_comp[b], 0));
            if (mode == RCT_MODE_LABEL) CK(cudaStreamWaitEvent(sin, p->ev_out[b], 0));
            else CK(cudaStreamWaitEvent(st, p->ev_out[b], 0));
        }
        // pinned staging has the device layout (per-slot strides): one 2-D copy either way
        if (pg_c) CK(cudaMemcpy2DAsync(d_c, cb, h_c, cb, npix * 3, ns, cudaMemcpyHostToDevice, sin));
        else CK(cudaMemcpy2DAsync(d_c, cb, content_h + (size_t)f0 * npix * 3, npix * 3, npix * 3, ns, cudaMemcpyHostToDevice, sin));
        if (pg_s) CK(cudaMemcpy2DAsync(d_s, sb, h_s, sb, spix * 3, ns, cudaMemcpyHostToDevice, sin));
        else CK(cudaMemcpy2DAsync(d_s, sb, stylised_h + (size_t)f0 * spix * 3, spix * 3, spix * 3, ns, cudaMemcpyHostToDevice, sin));
        if (mask_h) {
            if (pg_m) CK(cudaMemcpy2DAsync(d_m, mb, h_m, mb, npix, ns, cudaMemcpyHostToDevice, sin));
            else CK(cudaMemcpy2DAsync(d_m, mb, mask_h + (size_t)f0 * npix, npix, npix, ns, cudaMemcpyHostToDevice, sin));
        }
        if (mode == RCT_MODE_LABEL) {  // untouched pixels keep the caller's values
            if (pg_o) CK(cudaMemcpy2DAsync(d_o, cb, h_i, cb, npix * 3, ns, cudaMemcpyHostToDevice, sin));
            else CK(cudaMemcpy2DAsync(d_o, cb, out_h + (size_t)f0 * npix * 3, npix * 3, npix * 3, ns, cudaMemcpyHostToDevice, sin));
        }
        CK(cudaEventRecord(p->ev_in[b], sin));
        CK(cudaStreamWaitEvent(st, p->ev_in[b], 0));
        FrameSet f{d_c, d_s, mask_h ? d_m : nullptr, d_o, cb, sb, mb, cb};
        if ((rc = p->front ? phase_front(p, f, ns, mode, label, 0, npix, st)
                           : phase_hist(p, f, ns, mode, label, 0, npix, st))) return rc;
        if ((rc = phase_solve(p, ns, st, !p->front))) return rc;
        if ((rc = phase_apply(p, f, ns, mode, label, 0, npix, p->status_frames + f0, st))) return rc;
        CK(cudaEventRecord(p->ev_comp[b], st));
        CK(cudaStreamWaitEvent(sout, p->ev_comp[b], 0));
        if (pg_o) CK(cudaMemcpy2DAsync(h_o, cb, d_o, cb, npix * 3, ns, cudaMemcpyDeviceToHost, sout));
        else CK(cudaMemcpy2DAsync(out_h + (size_t)f0 * npix * 3, npix * 3, d_o, cb, npix * 3, ns, cudaMemcpyDeviceToHost, sout));
        CK(cudaEventRecord(p->ev_out[b], sout));
        // while the GPU works on this group, hand the previous group's output to the caller
        if (pg_o && k >= 1 && (rc = drain_out(k - 1, f0 - p->slots, p->slots))) return rc;
    }
    if (pg_o && k >= 1) {  // the last group
        const int g0 = (k - 1) * p->slots;
        if ((rc = drain_out(k - 1, g0, n_frames - g0))) return rc;
    }
    CK(cudaStreamSynchronize(sout));
    int worst = 0;
    int *tmp = status_host ? status_host : new (std::nothrow) int[n_frames];
    if (!tmp) return fail(RCT_ERR_NOMEM, "out of host memory");
    cudaError_t e = cudaMemcpyAsync(tmp, p->status_frames, sizeof(int) * (size_t)n_frames, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e == cudaSuccess)
        for (int i = 0; i < n_frames; ++i)
            if (tmp[i] > worst) worst = tmp[i];
    if (!status_host) delete[] tmp;
    if (e != cudaSuccess) return fail(RCT_ERR_CUDA, "host transfer failed: %s", cudaGetErrorString(e));
    if ((rc = scope.leave())) return rc;
    return worst;
}

// ---- row-sharded phases ---------------------------------------------------------------------
int rct_shard_hist_async(rct_plan *p, const uint8_t *content, const uint8_t *stylised, const uint8_t *mask, int mode,
                         int label, int row0, int rows, void *stream)
{
    if (!p || !content || !stylised || row0 < 0 || rows < 0 || row0 + rows > p->H)
        return fail(RCT_ERR_ARG, "bad shard arguments");
    int rc = check_mode(mode, mask);
    if (rc) return rc;
    FrameSet f{content, stylised, mask, nullptr, 0, 0, 0, 0};
    cudaStream_t st = (cudaStream_t)stream;
    const size_t pb = (size_t)row0 * p->W, pe = (size_t)(row0 + rows) * p->W;
    PlanScope scope(p, st);
    if ((rc = scope.enter())) return rc;
    p->kt_mode = mode;  // rct_shard_solve_async follows on the same plan
    if (!p->front) {
        rc = phase_hist(p, f, 1, mode, label, pb, pe, st);
        return rc ? rc : scope.leave();
    }
    // Colour-partitioned front end on the band (streaming passes, no global atomics), then the band's
    // (colour, count) lists are written into the zeroed colour cubes the ranks all-reduce: one plain
    // store per distinct colour instead of one atomic per pixel (k_hist).
    rc = phase_front(p, f, 1, mode, label, pb, pe, st);
    if (rc) return rc;
    LAUNCH(k_list_to_cube, (unsigned)p->sm * 8, 256, st, p->ctrl, p->jobs, 0, 0, p->hist);
    LAUNCH(k_list_to_cube, (unsigned)p->sm * 8, 256, st, p->ctrl, p->jobs, 0, 1, p->hist + kCube);
    p->cube_dirty = true;
    CK(cudaMemsetAsync(p->ctrl, 0, sizeof(FrameCtrl), st));  // the solve rebuilds the lists from the reduced cubes
    if ((rc = mark(p, st))) return rc;
    return scope.leave();
}
void *rct_plan_hist_ptr(rct_plan *p) { return p ? (void *)p->hist : nullptr; }
size_t rct_plan_hist_bytes(rct_plan *p) { return p ? 4 * p->L.hist_stride_u32 : 0; }
int rct_shard_solve_async(rct_plan *p, void *stream)
{
    if (!p) return fail(RCT_ERR_ARG, "null plan");
    PlanScope scope(p, (cudaStream_t)stream);
    int rc = scope.enter();
    if (rc == RCT_OK) rc = phase_solve(p, 1, (cudaStream_t)stream, true);
    return rc ? rc : scope.leave();
}
int rct_shard_apply_async(rct_plan *p, const uint8_t *content, const uint8_t *stylised, const uint8_t *mask,
                          uint8_t *out, int mode, int label, int row0, int rows, void *stream)
{
    if (!p || !content || !stylised || !out || row0 < 0 || rows < 0 || row0 + rows > p->H)
        return fail(RCT_ERR_ARG, "bad shard arguments");
    int rc = check_mode(mode, mask);
    if (rc) return rc;
    rc = ensure_status_frames(p, 1);
    if (rc) return rc;
    FrameSet f{content, stylised, mask, out, 0, 0, 0, 0};
    PlanScope scope(p, (cudaStream_t)stream);
    if ((rc = scope.enter())) return rc;
    rc = phase_apply(p, f, 1, mode, label, (size_t)row0 * p->W, (size_t)(row0 + rows) * p->W, p->status_frames,
                     (cudaStream_t)stream);
    return rc ? rc : scope.leave();
}


// ---- BINNED mode (approximation; rct_binned.cuh) ---------------------------------------------------
namespace {
int binned_solve(rct_plan *p, int nslots, int bins, cudaStream_t st, int first_phase, int last_phase)
{
    // phases: 0 moments | 1 pca + keys + key range | 2 histograms | 3 CDF build / inverse lookup -> LUT
    const unsigned eb = entry_blocks(p, nslots);
    if (first_phase <= 0 && last_phase >= 0) {
        LAUNCH(k_lab_moments, dim3(eb, 2, nslots), 256, st, p->ctrl, p->jobs);
        KMARK(RCT_TK_LAB_MOMENTS);
    }
    if (first_phase <= 1 && last_phase >= 1) {
        LAUNCH(k_pca, dim3((nslots + 31) / 32), 32, st, p->ctrl, nslots);
        KMARK(RCT_TK_PCA);
        LAUNCH(k_bin_keys, dim3(eb * RCT_EBK, 2, nslots), 256, st, p->ctrl, p->jobs);
        KMARK(RCT_TK_KEYS);
    }
    if (first_phase <= 2 && last_phase >= 2) {
        LAUNCH(k_bin_hist, dim3(eb, 2, nslots), 256, st, p->ctrl, p->jobs, bins);
        KMARK(RCT_TK_FINE_HIST);
    }
    if (first_phase <= 3 && last_phase >= 3) {
        p->kt_live = false;  // the key table belongs to the exact path
        LAUNCH(k_bin_match, dim3(eb * 2, nslots), 256, st, p->ctrl, p->jobs, p->lut, p->L.lut_stride_u32, bins, p->dbg_m,
               p->dbg_j, p->L.dbg_stride);
        KMARK(RCT_TK_MATCH);
    }
    return RCT_OK;
}
int check_bins(int bins)
{
    if (bins < 2 || bins > kBinMax) return fail(RCT_ERR_ARG, "bins %d outside 2..%d", bins, kBinMax);
    return RCT_OK;
}
}  // namespace

int rct_transfer_binned_async(rct_plan *p, const uint8_t *content, size_t content_stride, const uint8_t *stylised,
                              size_t stylised_stride, const uint8_t *mask, size_t mask_stride, uint8_t *out,
                              size_t out_stride, int n_frames, int mode, int label, int bins, void *stream)
{
    if (!p || !content || !stylised || !out || n_frames <= 0) return fail(RCT_ERR_ARG, "null/invalid argument");
    int rc = check_mode(mode, mask);
    if (rc == RCT_OK) rc = check_bins(bins);
    if (rc) return rc;
    if (p->lut_mode) return fail(RCT_ERR_ARG, "the binned mode writes the uint32 colour LUT: set RCT_OPT_LUT to 0");
    cudaStream_t st = (cudaStream_t)stream;
    rc = ensure_status_frames(p, n_frames);
    if (rc) return rc;
    PlanScope scope(p, st);
    if ((rc = scope.enter())) return rc;
    const size_t npix = (size_t)p->H * (size_t)p->W;
    const int front_saved = p->front;
    p->front = 1;  // the colour-partition front end produces the (colour, count) lists directly
    for (int f0 = 0; f0 < n_frames && rc == RCT_OK; f0 += p->slots) {
        int ns = n_frames - f0 < p->slots ? n_frames - f0 : p->slots;
        FrameSet f{content + (size_t)f0 * content_stride, stylised + (size_t)f0 * stylised_stride,
                   mask ? mask + (size_t)f0 * mask_stride : nullptr, out + (size_t)f0 * out_stride,
                   content_stride, stylised_stride, mask_stride, out_stride};
        rc = phase_front(p, f, ns, mode, label, 0, npix, st);
        if (rc == RCT_OK) rc = binned_solve(p, ns, bins, st, 0, 3);
        if (rc == RCT_OK) rc = mark(p, st);
        if (rc == RCT_OK) rc = phase_apply(p, f, ns, mode, label, 0, npix, p->status_frames + f0, st);
    }
    p->front = front_saved;
    return rc ? rc : scope.leave();
}

/* Row-sharded binned path (one image, rows [row0, row0+rows) on this rank; slot 0).  Between the phases the caller
 * all-reduces the exchange buffer (rct_plan_xchg_ptr):
 *   phase 0: band -> colour lists -> exact moments          -> xchg 0: int64[56]   SUM
 *   phase 1: PCA (global moments) -> keys -> key range      -> xchg 1: double[4]   MAX
 *   phase 2: histograms over the global range               -> xchg 2: uint32[2*bins] SUM
 *   phase 3: CDFs -> LUT of the band's colours -> apply the band's rows */
int rct_shard_binned_phase_async(rct_plan *p, int phase, const uint8_t *content, const uint8_t *stylised,
                                 const uint8_t *mask, uint8_t *out, int mode, int label, int row0, int rows, int bins,
                                 void *stream)
{
    if (!p || !content || !stylised || row0 < 0 || rows < 0 || row0 + rows > p->H || phase < 0 || phase > 3)
        return fail(RCT_ERR_ARG, "bad shard arguments");
    int rc = check_mode(mode, mask);
    if (rc == RCT_OK) rc = check_bins(bins);
    if (rc) return rc;
    if (p->lut_mode) return fail(RCT_ERR_ARG, "the binned mode writes the uint32 colour LUT: set RCT_OPT_LUT to 0");
    cudaStream_t st = (cudaStream_t)stream;
    void *xchg = p->ws + p->L.off_xchg;
    const size_t pb = (size_t)row0 * p->W, pe = (size_t)(row0 + rows) * p->W;
    PlanScope scope(p, st);
    if ((rc = scope.enter())) return rc;
    if (phase == 0) {
        FrameSet f{content, stylised, mask, nullptr, 0, 0, 0, 0};
        const int front_saved = p->front;
        p->front = 1;
        rc = phase_front(p, f, 1, mode, label, pb, pe, st);
        p->front = front_saved;
        if (rc == RCT_OK) rc = binned_solve(p, 1, bins, st, 0, 0);
        if (rc) return rc;
        LAUNCH(k_xchg_pack, 1, 64, st, p->ctrl, xchg, 0, bins);
    } else if (phase == 1) {
        LAUNCH(k_xchg_unpack, 1, 64, st, p->ctrl, xchg, 0, bins);
        if ((rc = binned_solve(p, 1, bins, st, 1, 1))) return rc;
        LAUNCH(k_xchg_pack, 1, 64, st, p->ctrl, xchg, 1, bins);
    } else if (phase == 2) {
        LAUNCH(k_xchg_unpack, 1, 64, st, p->ctrl, xchg, 1, bins);
        if ((rc = binned_solve(p, 1, bins, st, 2, 2))) return rc;
        LAUNCH(k_xchg_pack, (2 * bins + 255) / 256, 256, st, p->ctrl, xchg, 2, bins);
    } else {
        if (!out) return fail(RCT_ERR_ARG, "null output");
        LAUNCH(k_xchg_unpack, (2 * bins + 255) / 256, 256, st, p->ctrl, xchg, 2, bins);
        if ((rc = binned_solve(p, 1, bins, st, 3, 3))) return rc;
        if ((rc = ensure_status_frames(p, 1))) return rc;
        FrameSet f{content, stylised, mask, out, 0, 0, 0, 0};
        if ((rc = phase_apply(p, f, 1, mode, label, pb, pe, p->status_frames, st))) return rc;
    }
    return scope.leave();
}
void *rct_plan_xchg_ptr(rct_plan *p) { return p ? (void *)(p->ws + p->L.off_xchg) : nullptr; }
size_t rct_plan_xchg_bytes(int which, int bins)
{
    if (which == 0) return 56 * sizeof(long long);
    if (which == 1) return 4 * sizeof(double);
    if (which == 2 && bins >= 2 && bins <= kBinMax) return 2 * (size_t)bins * sizeof(unsigned);
    return 0;
}

/* binned introspection (parity tests): the two B-bin histograms, the key ranges (lo_t hi_t lo_s hi_s) and -- with
 * RCT_FLAG_DEBUG -- per entry of slot's target list its colour, CDF index pair b*65536+b' and matched value */
int rct_plan_debug_binned(rct_plan *p, int slot, int bins, uint32_t *hist_t, uint32_t *hist_s, double *range4,
                          uint32_t *colour, int64_t *code, double *matched)
{
    if (!p || slot < 0 || slot >= p->slots) return fail(RCT_ERR_ARG, "bad slot");
    int rc = check_bins(bins);
    if (rc) return rc;
    DeviceGuard guard(p->device);
    CK(cudaDeviceSynchronize());
    FrameCtrl *h = new (std::nothrow) FrameCtrl;
    if (!h) return fail(RCT_ERR_NOMEM, "out of host memory");
    cudaError_t e = cudaMemcpy(h, p->ctrl + slot, sizeof(FrameCtrl), cudaMemcpyDeviceToHost);
    size_t d = 0;
    if (e == cudaSuccess) {
        if (hist_t) memcpy(hist_t, h->r[0].coarse, 4 * (size_t)bins);
        if (hist_s) memcpy(hist_s, h->r[1].coarse, 4 * (size_t)bins);
        if (range4)
            for (int r = 0; r < 2; ++r) {
                unsigned long long mx = h->r[r].lab_max[0], mn = ~h->r[r].lab_min[0];
                unsigned long long umx = (mx & 0x8000000000000000ULL) ? (mx & 0x7fffffffffffffffULL) : ~mx;
                unsigned long long umn = (mn & 0x8000000000000000ULL) ? (mn & 0x7fffffffffffffffULL) : ~mn;
                memcpy(&range4[2 * r], &umn, 8);
                memcpy(&range4[2 * r + 1], &umx, 8);
            }
        d = h->status == 0 ? h->r[0].d : 0;
    }
    delete h;
    if (e != cudaSuccess) return fail(RCT_ERR_CUDA, "debug copy failed: %s", cudaGetErrorString(e));
    if (d && (colour || code || matched)) {
        if (!(p->flags & RCT_FLAG_DEBUG)) return fail(RCT_ERR_ARG, "plan was not created with RCT_FLAG_DEBUG");
        const Layout &L = p->L;
        uint8_t *b = p->ws + L.off_region + L.region_bytes * (size_t)(2 * slot);
        if (colour) CK(cudaMemcpy(colour, b + L.r_col, 4 * d, cudaMemcpyDeviceToHost));
        if (code) CK(cudaMemcpy(code, p->dbg_j + (size_t)slot * L.dbg_stride, 8 * d, cudaMemcpyDeviceToHost));
        if (matched) CK(cudaMemcpy(matched, p->dbg_m + (size_t)slot * L.dbg_stride, 8 * d, cudaMemcpyDeviceToHost));
    }
    return RCT_OK;
}

// ---- stand-alone functions -----------------------------------------------------------------
static unsigned flat_blocks(size_t n, int dev)
{
    size_t b = (n + 255) / 256, maxb = (size_t)(g_sm_count[dev] > 0 ? g_sm_count[dev] : 148) * 8;
    if (b > maxb) b = maxb;
    return (unsigned)(b ? b : 1);
}

int rct_rgb_to_lab(const uint8_t *rgb, double *lab, size_t n, void *stream)
{
    if (n == 0) return RCT_OK;
    if (!rgb || !lab) return fail(RCT_ERR_ARG, "null pointer");
    int dev, rc = ensure_device(&dev);
    if (rc) return rc;
    LAUNCH(k_rgb_to_lab, flat_blocks(n, dev), 256, (cudaStream_t)stream, rgb, lab, n);
    return RCT_OK;
}

int rct_lab_to_rgb(const double *lab, uint8_t *rgb, double *rgb_float, size_t n, void *stream)
{
    if (n == 0) return RCT_OK;
    if (!rgb || !lab) return fail(RCT_ERR_ARG, "null pointer");
    int dev, rc = ensure_device(&dev);
    if (rc) return rc;
    LAUNCH(k_lab_to_rgb, flat_blocks(n, dev), 256, (cudaStream_t)stream, lab, rgb, rgb_float, n);
    return RCT_OK;
}

size_t rct_pca_scratch_bytes(void) { return align_up(sizeof(RegionCtrl)) + 256; }

int rct_apply_pca(const double *lab, size_t n, double *proj, double *mean3, double *comp3, void *scratch, void *stream)
{
    if (n == 0) return fail(RCT_ERR_ARG, "empty input");
    if (!lab || !proj || !mean3 || !comp3 || !scratch) return fail(RCT_ERR_ARG, "null pointer");
    int dev, rc = ensure_device(&dev);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    RegionCtrl *rcd = (RegionCtrl *)scratch;
    int *bad = (int *)((uint8_t *)scratch + align_up(sizeof(RegionCtrl)));
    CK(cudaMemsetAsync(scratch, 0, rct_pca_scratch_bytes(), st));
    unsigned b = flat_blocks(n, dev);
    if (b > 1024) b = 1024;
    LAUNCH(k_pca_moments, b, 256, st, lab, n, rcd, bad);
    LAUNCH(k_pca_solve, 1, 32, st, rcd, mean3, comp3);
    LAUNCH(k_pca_project, flat_blocks(n, dev), 256, st, lab, n, rcd, proj);
    return RCT_OK;
}

int rct_pca_inverse(const double *matched, size_t n, const double *mean3, const double *comp3, double *lab, void *stream)
{
    if (n == 0) return RCT_OK;
    if (!matched || !mean3 || !comp3 || !lab) return fail(RCT_ERR_ARG, "null pointer");
    int dev, rc = ensure_device(&dev);
    if (rc) return rc;
    LAUNCH(k_pca_inverse, flat_blocks(n, dev), 256, (cudaStream_t)stream, matched, n, mean3, comp3, lab);
    return RCT_OK;
}

namespace {
struct MatchLayout {
    size_t off_rc[2], off_zero[2], zero_bytes[2], z_fcnt[2], z_pfill[2], off_jobs, off_arr[2], total;
    size_t maxparts[2];
};
MatchLayout match_layout(size_t nt, size_t ns)
{
    MatchLayout M{};
    size_t n[2] = {nt ? nt : 1, ns ? ns : 1};
    size_t o = 0;
    for (int r = 0; r < 2; ++r) { M.off_rc[r] = o; o = align_up(o + sizeof(RegionCtrl)); }
    for (int r = 0; r < 2; ++r) {  // everything up to off_jobs is zeroed per call
        M.maxparts[r] = n[r] / kPartT + 2;
        zero_zone(M.maxparts[r], M.z_fcnt[r], M.z_pfill[r], M.zero_bytes[r]);
        M.off_zero[r] = o; o += M.zero_bytes[r];
    }
    M.off_jobs = o; o = align_up(o + 2 * sizeof(Job));
    for (int r = 0; r < 2; ++r) {
        size_t a = 0;
        a = align_up(a + 16 * n[r]);                   // recP
        a = align_up(a + 2 * (size_t)kFineMax);        // pof
        a = align_up(a + 4 * (M.maxparts[r] + 1));     // pstart
        a = align_up(a + 4 * (M.maxparts[r] + 1));     // ppix
        M.off_arr[r] = o;
        o += a;
    }
    M.total = o;
    return M;
}
}  // namespace

size_t rct_match_scratch_bytes(size_t nt, size_t ns) { return match_layout(nt, ns).total; }

int rct_match_cdf(const double *target_proj, size_t nt, const double *source_proj, size_t ns, double *matched,
                  int64_t *j_index, void *scratch, void *stream)
{
    if (nt == 0 || ns == 0) return fail(RCT_ERR_ARG, "empty input");
    // part ids are 16-bit: at most 65535 parts of >= kPartT entries
    if (nt >= ((size_t)1 << 26) || ns >= ((size_t)1 << 26)) return fail(RCT_ERR_ARG, "input too long (>= 2^26)");
    if (!target_proj || !source_proj || !matched || !scratch) return fail(RCT_ERR_ARG, "null pointer");
    int dev, rc = ensure_device(&dev);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    MatchLayout M = match_layout(nt, ns);
    uint8_t *ws = (uint8_t *)scratch;
    Job hj[2];
    const double *keys[2] = {target_proj, source_proj};
    size_t n[2] = {nt, ns};
    for (int r = 0; r < 2; ++r) {
        uint8_t *b = ws + M.off_arr[r];
        size_t a = 0;
        Job &j = hj[r];
        j.rc = (RegionCtrl *)(ws + M.off_rc[r]);
        j.keyA = keys[r]; j.payA = nullptr; j.cntA = nullptr;
        j.recP = (ulonglong2 *)(b + a); a = align_up(a + 16 * n[r]);
        j.pof = (unsigned short *)(b + a); a = align_up(a + 2 * (size_t)kFineMax);
        j.pstart = (unsigned *)(b + a); a = align_up(a + 4 * (M.maxparts[r] + 1));
        j.ppix = (unsigned *)(b + a);
        j.fcnt = (unsigned *)(ws + M.off_zero[r] + M.z_fcnt[r]);
        j.pfill = (unsigned *)(ws + M.off_zero[r] + M.z_pfill[r]);
        j.cap = (unsigned)n[r]; j.maxparts = (unsigned)M.maxparts[r];
    }
    CK(cudaMemsetAsync(ws, 0, M.off_jobs, st));
    CK(cudaMemcpyAsync(ws + M.off_jobs, hj, sizeof hj, cudaMemcpyHostToDevice, st));
    const Job *jobs = (const Job *)(ws + M.off_jobs);
    size_t nmax = nt > ns ? nt : ns;
    unsigned eb = flat_blocks(nmax, dev);
    LAUNCH(k_key_range, flat_blocks(nt, dev), 256, st, target_proj, nt, hj[0].rc);
    LAUNCH(k_key_range, flat_blocks(ns, dev), 256, st, source_proj, ns, hj[1].rc);
    LAUNCH(k_key_range_finish, 1, 32, st, hj[0].rc, hj[1].rc, (unsigned long long)nt, (unsigned long long)ns);
    LAUNCH(k_coarse_hist, dim3(eb, 2), 256, st, jobs);
    size_t mp = M.maxparts[0] > M.maxparts[1] ? M.maxparts[0] : M.maxparts[1];
    rc = sort_jobs(nullptr, jobs, 2, eb, nmax, mp, kPartCap, st);
    if (rc) return rc;
    LAUNCH(k_match_generic, (unsigned)M.maxparts[0], kMatchThreads, st, jobs, matched, (long long *)j_index);
    return RCT_OK;
}

// ---- match_cdf with built-in scratch: for callers that must not allocate device memory -------------
// Some hosts of this library are not allowed to call cudaMalloc at the point where they need a CDF match: code running
// while a CUDA graph is being built or under a memory pool whose budget is fixed at start-up, plug-ins of an
// application that owns the device heap and hands out no allocator, and latency-critical loops in which an implicit
// device synchronisation (cudaMalloc / cudaFree can cause one) is not acceptable.  For them the work arrays of the
// sort (rct_match_scratch_bytes) live in the library image itself: one array per GPU that loads librct.so, present
// from module load until the process exits.  Bounded on purpose: RCT_MATCH_STATIC_MAX entries on either side.
namespace {
constexpr size_t kStaticMatchMax = RCT_MATCH_STATIC_MAX;
constexpr size_t kStaticMatchBytes = (size_t)34 << 20;  // >= match_layout(max, max).total = 33 961 728; checked at set-up
}  // namespace
static __device__ __align__(256) unsigned char g_match_static[kStaticMatchBytes];
namespace {
// g_match_static above is, in host code, only a handle of the symbol: it is NOT a device address and must never be
// given to a kernel or a cudaMemset.  The address a kernel can use differs per GPU and comes from
// cudaGetSymbolAddress with that GPU current; it is looked up once per device and kept here (guarded by g_mu).
void *g_match_static_ptr[64] = {nullptr};

int match_static_scratch(int dev, void **out)
{
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_match_static_ptr[dev]) {
        if (match_layout(kStaticMatchMax, kStaticMatchMax).total > kStaticMatchBytes)
            return fail(RCT_ERR_ARG, "built-in match scratch (%zu bytes) is smaller than the layout of %zu entries needs",
                        kStaticMatchBytes, kStaticMatchMax);
        void *ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&ptr, g_match_static);
        if (e != cudaSuccess || !ptr)
            return fail(RCT_ERR_CUDA, "cudaGetSymbolAddress(built-in match scratch) failed on device %d: %s", dev,
                        cudaGetErrorString(e));
        g_match_static_ptr[dev] = ptr;
    }
    *out = g_match_static_ptr[dev];
    return RCT_OK;
}
}  // namespace

int rct_match_cdf_static(const double *target_proj, size_t nt, const double *source_proj, size_t ns, double *matched,
                         int64_t *j_index, void *stream)
{
    if (nt == 0 || ns == 0) return fail(RCT_ERR_ARG, "empty input");
    if (nt > kStaticMatchMax || ns > kStaticMatchMax)
        return fail(RCT_ERR_ARG,
                    "rct_match_cdf_static: %zu target / %zu source entries exceed the built-in scratch "
                    "(RCT_MATCH_STATIC_MAX = %zu per side); use rct_match_cdf with caller-provided scratch",
                    nt, ns, kStaticMatchMax);
    if (!target_proj || !source_proj || !matched) return fail(RCT_ERR_ARG, "null pointer");
    int dev, rc = ensure_device(&dev);  // first call on a device: uploads the constant tables (no device allocation)
    if (rc) return rc;
    void *scratch = nullptr;
    rc = match_static_scratch(dev, &scratch);  // first call on a device: resolves the array's device address
    if (rc) return rc;
    if (match_layout(nt, ns).total > kStaticMatchBytes) return fail(RCT_ERR_ARG, "layout exceeds the built-in scratch");
    // The array still holds what the previous call left there; rct_match_cdf zeroes the counters it needs at zero
    // (cudaMemsetAsync of the head of the layout, on `stream`) before the first kernel and overwrites the rest.
    return rct_match_cdf(target_proj, nt, source_proj, ns, matched, j_index, scratch, stream);
}

// ---- adjacent rows: AdaIN mask-blend tail, DeepLab post-processing ------------------
int rct_adain_mask_blend(const float *content, const float *output, const float *mask, float *stylized,
                         uint8_t *rgb8, int n, int H, int W, int oh, int ow, int mh, int mw, size_t mask_stride,
                         void *stream)
{
    if (n <= 0 || H <= 0 || W <= 0 || oh <= 0 || ow <= 0 || mh <= 0 || mw <= 0) return fail(RCT_ERR_ARG, "bad geometry");
    if (!content || !output || !mask || !stylized) return fail(RCT_ERR_ARG, "null pointer");
    int dev, rc = ensure_device(&dev);
    if (rc) return rc;
    BlendArgs a{content, output, mask, stylized, rgb8, n, H, W, oh, ow, mh, mw, mask_stride};
    const bool vec4 = W % 4 == 0 && ((((uintptr_t)content) | ((uintptr_t)stylized)) & 15) == 0 && (((uintptr_t)rgb8) & 3) == 0;
    if (vec4) LAUNCH(k_mask_blend4, flat_blocks((size_t)n * H * W / 4, dev), 256, (cudaStream_t)stream, a);
    else LAUNCH(k_mask_blend, flat_blocks((size_t)n * H * W, dev), 256, (cudaStream_t)stream, a);
    return RCT_OK;
}

int rct_softmax_threshold_mask(const float *logits, int n_classes, size_t npix, int cls, float threshold,
                               uint8_t *mask, void *stream)
{
    if (npix == 0) return RCT_OK;
    if (!logits || !mask || n_classes <= 0 || cls < 0 || cls >= n_classes) return fail(RCT_ERR_ARG, "bad argument");
    int dev, rc = ensure_device(&dev);
    if (rc) return rc;
    LAUNCH(k_softmax_mask, flat_blocks(npix, dev), 256, (cudaStream_t)stream, logits, n_classes, npix, cls, threshold, mask);
    return RCT_OK;
}

int rct_argmax_labels(const float *logits, int n_classes, size_t npix, uint8_t *labels, void *stream)
{
    if (npix == 0) return RCT_OK;
    if (!logits || !labels || n_classes <= 0 || n_classes > 256) return fail(RCT_ERR_ARG, "bad argument");
    int dev, rc = ensure_device(&dev);
    if (rc) return rc;
    LAUNCH(k_argmax_labels, flat_blocks(npix, dev), 256, (cudaStream_t)stream, logits, n_classes, npix, labels);
    return RCT_OK;
}

// ---- f-4: video temporal-consistency step (video/utils.py:89-105, 217-229) ------------------------
static int warp_launch(const uint8_t *stylized, const uint8_t *prev, const float *flow, double alpha, uint8_t *out,
                       int n, int H, int W, int channels, bool blend, void *stream)
{
    if (n <= 0 || H <= 0 || W <= 0 || channels <= 0 || channels > 4) return fail(RCT_ERR_ARG, "bad geometry");
    if (H > 32767 || W > 32767) return fail(RCT_ERR_ARG, "cv2.remap addresses sources with 16-bit coordinates");
    if (!prev || !flow || !out || (blend && !stylized)) return fail(RCT_ERR_ARG, "null pointer");
    int dev, rc = ensure_device(&dev);
    if (rc) return rc;
    WarpArgs g{stylized, prev, flow, out, n, H, W, channels, (float)alpha, (float)(1.0 - alpha)};
    const bool rgb4 = channels == 3 && W % 4 == 0 && (((uintptr_t)flow) & 15) == 0 &&
                      ((((uintptr_t)prev) | ((uintptr_t)out) | ((uintptr_t)stylized)) & 3) == 0;
    if (rgb4) {  // four pixels per thread, vector loads / stores
        const unsigned blocks = flat_blocks((size_t)n * H * W / 4, dev);
        if (blend) LAUNCH(k_warp_blend_rgb4<true>, blocks, 256, (cudaStream_t)stream, g);
        else LAUNCH(k_warp_blend_rgb4<false>, blocks, 256, (cudaStream_t)stream, g);
        return RCT_OK;
    }
    const unsigned blocks = flat_blocks((size_t)n * H * W, dev);
    if (blend) LAUNCH(k_warp_blend<true>, blocks, 256, (cudaStream_t)stream, g);
    else LAUNCH(k_warp_blend<false>, blocks, 256, (cudaStream_t)stream, g);
    return RCT_OK;
}

int rct_warp_image(const uint8_t *image, const float *flow, uint8_t *warped, int n, int H, int W, int channels,
                   void *stream)
{
    return warp_launch(nullptr, image, flow, 0.0, warped, n, H, W, channels, false, stream);
}

int rct_warp_blend(const uint8_t *stylized, const uint8_t *prev_stylized, const float *flow, double alpha, uint8_t *out,
                   int n, int H, int W, int channels, void *stream)
{
    return warp_launch(stylized, prev_stylized, flow, alpha, out, n, H, W, channels, true, stream);
}

int rct_blend_images(const uint8_t *stylized, const uint8_t *warped, double alpha, uint8_t *out, size_t nbytes,
                     void *stream)
{
    if (nbytes == 0) return RCT_OK;
    if (!stylized || !warped || !out) return fail(RCT_ERR_ARG, "null pointer");
    int dev, rc = ensure_device(&dev);
    if (rc) return rc;
    const int vec = ((((uintptr_t)stylized) | ((uintptr_t)warped) | ((uintptr_t)out)) & 15) == 0;
    LAUNCH(k_blend, flat_blocks((nbytes + 15) / 16, dev), 256, (cudaStream_t)stream, stylized, warped, out, nbytes,
           (float)alpha, (float)(1.0 - alpha), vec);
    return RCT_OK;
}

// ---- f-3: baseline JPEG encoder (rct_jpeg.cuh) -------------------------------------------------
size_t rct_jpeg_workspace_bytes(int H, int W, int n)
{
    if (H <= 0 || W <= 0 || n <= 0 || H > 65535 || W > 65535) return 0;
    return jpeg::carve(jpeg::make_geom(H, W, 0), n, nullptr).bytes;
}

size_t rct_jpeg_max_bytes(int H, int W)
{
    if (H <= 0 || W <= 0 || H > 65535 || W > 65535) return 0;
    const jpeg::Geom g = jpeg::make_geom(H, W, 0);
    // header + every entropy-coded byte stuffed + EOI
    return (size_t)jpeg::kHeaderMax + 2u * (((size_t)g.nblk * jpeg::kBlockBitsMax + 7u) / 8u) + 2u;
}

int rct_jpeg_encode_async(const uint8_t *rgb, size_t frame_stride, int n, int H, int W, int quality, void *workspace,
                          uint8_t *out, size_t out_stride, size_t out_cap, unsigned long long *nbytes, void *stream)
{
    if (n == 0) return RCT_OK;
    if (!rgb || !workspace || !out || !nbytes) return fail(RCT_ERR_ARG, "null pointer");
    if (n < 0 || H <= 0 || W <= 0 || H > 65535 || W > 65535) return fail(RCT_ERR_ARG, "JPEG frames are 1..65535 pixels a side");
    if (quality < 1 || quality > 100) return fail(RCT_ERR_ARG, "quality %d outside 1..100", quality);
    if (((uintptr_t)workspace & 255u) != 0) return fail(RCT_ERR_ARG, "workspace must be 256-byte aligned");
    if (out_cap > out_stride && n > 1) return fail(RCT_ERR_ARG, "out_cap exceeds out_stride");
    int dev, rc = ensure_device(&dev);
    if (rc) return rc;
    const jpeg::Geom g = jpeg::make_geom(H, W, frame_stride);
    const jpeg::Workspace ws = jpeg::carve(g, n, workspace);
    int launches = 0;
    cudaError_t e = jpeg::encode_launch(rgb, g, n, quality, ws, out, out_stride, out_cap, nbytes, g_sm_count[dev],
                                        (cudaStream_t)stream, &launches);
    t_launches += launches;
    if (e != cudaSuccess) return fail(RCT_ERR_CUDA, "JPEG encoder launch failed: %s", cudaGetErrorString(e));
    return RCT_OK;
}

// ---- f-3, decode half (rct_jpeg_dec.cuh) ----------------------------------------------------------
int rct_jpeg_decode_info(const uint8_t *file_host, size_t nbytes, int *H, int *W, int *channels, size_t *workspace_bytes)
{
    if (!file_host) return fail(RCT_ERR_ARG, "null pointer");
    jpeg::DecPlan pl;
    if (!jpeg::parse_jpeg(file_host, nbytes, &pl)) return fail(RCT_ERR_ARG, "JPEG decoder: %s", pl.error);
    if (H) *H = pl.f.H;
    if (W) *W = pl.f.W;
    if (channels) *channels = pl.f.ncomp == 3 ? 3 : 1;
    if (workspace_bytes) *workspace_bytes = jpeg::carve_dec(pl.f, nullptr).bytes;
    jpeg::free_plan(&pl);
    return RCT_OK;
}

int rct_jpeg_decode_async(const uint8_t *file_host, size_t nbytes, void *workspace, size_t workspace_bytes, uint8_t *pixels,
                          int flags, void *stream)
{
    if (!file_host || !workspace || !pixels) return fail(RCT_ERR_ARG, "null pointer");
    if (((uintptr_t)workspace & 255u) != 0) return fail(RCT_ERR_ARG, "workspace must be 256-byte aligned");
    jpeg::DecPlan pl;
    if (!jpeg::parse_jpeg(file_host, nbytes, &pl)) return fail(RCT_ERR_ARG, "JPEG decoder: %s", pl.error);
    int dev, rc = ensure_device(&dev);
    if (rc == RCT_OK && jpeg::carve_dec(pl.f, nullptr).bytes > workspace_bytes)
        rc = fail(RCT_ERR_ARG, "workspace of %zu bytes is too small", workspace_bytes);
    if (rc == RCT_OK) {
        int launches = 0;
        cudaError_t e = jpeg::decode_launch(file_host, pl, jpeg::carve_dec(pl.f, workspace), pixels, (cudaStream_t)stream, &launches,
                                            (flags & RCT_JPEG_SERIAL) == 0);
        t_launches += launches;
        if (e != cudaSuccess) rc = fail(RCT_ERR_CUDA, "JPEG decoder launch failed: %s", cudaGetErrorString(e));
    }
    jpeg::free_plan(&pl);
    return rc;
}

// ---- f-3: PNG -> RGB8 on the host (rct_png.h explains why the host) -----------------------------------
int rct_png_info(const uint8_t *file_host, size_t nbytes, int *H, int *W)
{
    if (!file_host) return fail(RCT_ERR_ARG, "null pointer");
    rct_png::Info info;
    if (!rct_png::parse_header(file_host, nbytes, &info)) return fail(RCT_ERR_ARG, "%s", info.error.c_str());
    if (H) *H = info.H;
    if (W) *W = info.W;
    return RCT_OK;
}

int rct_png_decode_host(const uint8_t *file_host, size_t nbytes, uint8_t *rgb_host, size_t rgb_capacity)
{
    if (!file_host || !rgb_host) return fail(RCT_ERR_ARG, "null pointer");
    rct_png::Info info;
    if (!rct_png::parse_header(file_host, nbytes, &info)) return fail(RCT_ERR_ARG, "%s", info.error.c_str());
    if ((size_t)info.H * (size_t)info.W * 3u > rgb_capacity) return fail(RCT_ERR_ARG, "PNG: output buffer too small");
    try {  // the inflate / unfilter buffers are sized by the file's header: no exception may cross the C ABI
        if (!rct_png::decode(file_host, nbytes, &info, rgb_host)) return fail(RCT_ERR_ARG, "%s", info.error.c_str());
    } catch (const std::bad_alloc &) {
        return fail(RCT_ERR_NOMEM, "PNG: out of host memory for a %d x %d image", info.W, info.H);
    }
    return RCT_OK;
}


// ---- f-3 on the e2e path: files in, file out ------------------------------------------------------------
namespace {
__global__ void __launch_bounds__(256) k_gray_to_rgb(const uint8_t *g, uint8_t *rgb, size_t n)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const uint8_t v = g[i];
        rgb[3 * i] = v;
        rgb[3 * i + 1] = v;
        rgb[3 * i + 2] = v;
    }
}
bool is_png(const uint8_t *f, size_t n) { return n >= 8 && f[0] == 0x89 && f[1] == 'P' && f[2] == 'N' && f[3] == 'G'; }

// one image file (host bytes) -> RGB8 pixels at `dst` (device, H*W*3); `scratch` holds the decoder workspace
int decode_file_to_device(const uint8_t *file, size_t nbytes, int H, int W, const char *what, uint8_t *dst, uint8_t *scratch,
                          size_t scratch_bytes, uint8_t *gray_tmp, std::vector<uint8_t> &host_tmp, int sm, cudaStream_t st)
{
    if (is_png(file, nbytes)) {
        rct_png::Info info;
        if (!rct_png::parse_header(file, nbytes, &info)) return fail(RCT_ERR_ARG, "%s: %s", what, info.error.c_str());
        if (info.H != H || info.W != W) return fail(RCT_ERR_ARG, "%s is %dx%d, the plan expects %dx%d", what, info.W, info.H, W, H);
        try {  // sized by the plan / the file's header: no exception may cross the C ABI
            host_tmp.resize((size_t)H * W * 3);
            if (!rct_png::decode(file, nbytes, &info, host_tmp.data())) return fail(RCT_ERR_ARG, "%s: %s", what, info.error.c_str());
        } catch (const std::bad_alloc &) {
            return fail(RCT_ERR_NOMEM, "%s: out of host memory for a %d x %d PNG", what, W, H);
        }
        CK(cudaMemcpyAsync(dst, host_tmp.data(), host_tmp.size(), cudaMemcpyHostToDevice, st));
        CK(cudaStreamSynchronize(st));  // host_tmp is reused for the next file
        return RCT_OK;
    }
    jpeg::DecPlan pl;
    if (!jpeg::parse_jpeg(file, nbytes, &pl)) return fail(RCT_ERR_ARG, "%s: JPEG decoder: %s", what, pl.error);
    int rc = RCT_OK;
    if (pl.f.H != H || pl.f.W != W) rc = fail(RCT_ERR_ARG, "%s is %dx%d, the plan expects %dx%d", what, pl.f.W, pl.f.H, W, H);
    else if (jpeg::carve_dec(pl.f, nullptr).bytes > scratch_bytes) rc = fail(RCT_ERR_ARG, "%s: decoder workspace too small", what);
    if (rc == RCT_OK) {
        int launches = 0;
        const bool gray = pl.f.ncomp != 3;
        cudaError_t e = jpeg::decode_launch(file, pl, jpeg::carve_dec(pl.f, scratch), gray ? gray_tmp : dst, st, &launches, true);
        t_launches += launches;
        if (e != cudaSuccess) rc = fail(RCT_ERR_CUDA, "%s: JPEG decoder launch failed: %s", what, cudaGetErrorString(e));
        else if (gray) {
            k_gray_to_rgb<<<sm * 4, 256, 0, st>>>(gray_tmp, dst, (size_t)H * W);
            ++t_launches;
        }
    }
    jpeg::free_plan(&pl);
    return rc;
}
}  // namespace

size_t rct_files_out_capacity(int H, int W) { return rct_jpeg_max_bytes(H, W); }

int rct_transfer_files_host(rct_plan *p, const uint8_t *content_file, size_t content_bytes, const uint8_t *stylised_file,
                            size_t stylised_bytes, const uint8_t *mask_h, int quality, uint8_t *out_file, size_t out_capacity,
                            size_t *out_bytes, int *status_host)
{
    if (!p || !content_file || !stylised_file || !mask_h || !out_file || !out_bytes) return fail(RCT_ERR_ARG, "null argument");
    if (quality < 1 || quality > 100) return fail(RCT_ERR_ARG, "quality %d outside 1..100", quality);
    DeviceGuard guard(p->device);
    const int H = p->H, W = p->W, sh = p->sh, sw = p->sw;
    const size_t npix = (size_t)H * W, spix = (size_t)sh * sw;
    const size_t cb = align_up(npix * 3), sb = align_up(spix * 3), mb = align_up(npix);
    // device buffers: content | stylised | mask | out pixels (one frame), then decoder / encoder workspaces and the file
    const size_t enc_ws = align_up(rct_jpeg_workspace_bytes(H, W, 1));
    const size_t cap = align_up(rct_jpeg_max_bytes(H, W));
    const size_t big = npix > spix ? npix : spix;
    // decoder workspace bound: coefficients (2 bytes per sample incl. padding to MCUs) + planes + scan + states
    const size_t dec_ws = align_up(big * 16 + (content_bytes > stylised_bytes ? content_bytes : stylised_bytes) * 8 + (1u << 20));
    // grey plane of a one-component JPEG before its expansion to RGB: a region of its own, sized for the larger image
    // (it once borrowed d_o, which holds 3*H*W bytes: a grey stylised file of more than 3*H*W pixels ran into d_dec)
    const size_t gb = align_up(big);
    const size_t need = cb + sb + mb + cb + dec_ws + enc_ws + cap + 256 + gb;
    if (p->files_ws_bytes < need) {
        if (p->files_ws) CK(cudaFree(p->files_ws));
        p->files_ws = nullptr;
        p->files_ws_bytes = 0;
        CK(cudaMalloc((void **)&p->files_ws, need));
        p->files_ws_bytes = need;
    }
    uint8_t *d_c = p->files_ws, *d_s = d_c + cb, *d_m = d_s + sb, *d_o = d_m + mb, *d_dec = d_o + cb, *d_enc = d_dec + dec_ws,
            *d_file = d_enc + enc_ws;
    unsigned long long *d_n = reinterpret_cast<unsigned long long *>(d_file + cap);
    uint8_t *d_gray = d_file + cap + 256;
    cudaStream_t st = p->own_stream;
    PlanScope scope(p, st);
    scope.drain[0] = st;  // on failure: the files / mask / output of the caller may be in flight on st
    int rc = scope.enter();
    if (rc) return rc;
    p->kt_mode = RCT_MODE_MASK;
    std::vector<uint8_t> host_tmp;
    if ((rc = decode_file_to_device(content_file, content_bytes, H, W, "content file", d_c, d_dec, dec_ws, d_gray, host_tmp, p->sm, st))) return rc;
    if ((rc = decode_file_to_device(stylised_file, stylised_bytes, sh, sw, "stylised file", d_s, d_dec, dec_ws, d_gray, host_tmp, p->sm, st))) return rc;
    CK(cudaMemcpyAsync(d_m, mask_h, npix, cudaMemcpyHostToDevice, st));
    if ((rc = ensure_status_frames(p, 1))) return rc;
    FrameSet f{d_c, d_s, d_m, d_o, cb, sb, mb, cb};
    if ((rc = p->front ? phase_front(p, f, 1, RCT_MODE_MASK, 0, 0, npix, st) : phase_hist(p, f, 1, RCT_MODE_MASK, 0, 0, npix, st))) return rc;
    if ((rc = phase_solve(p, 1, st, !p->front))) return rc;
    if ((rc = phase_apply(p, f, 1, RCT_MODE_MASK, 0, 0, npix, p->status_frames, st))) return rc;
    if ((rc = rct_jpeg_encode_async(d_o, cb, 1, H, W, quality, d_enc, d_file, cap, cap, d_n, st))) return rc;
    unsigned long long nb = 0;
    int status = 0;
    CK(cudaMemcpyAsync(&nb, d_n, sizeof nb, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(&status, p->status_frames, sizeof status, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (status_host) *status_host = status;
    *out_bytes = (size_t)nb;
    if (nb > out_capacity) return fail(RCT_ERR_ARG, "output file needs %llu bytes, capacity is %zu", nb, out_capacity);
    CK(cudaMemcpyAsync(out_file, d_file, (size_t)nb, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if ((rc = scope.leave())) return rc;
    return status;
}

// ---- phase timing (RCT_OPT_TIMING) -------------------------------------------------------------
int rct_plan_timing(rct_plan *p, double *ms4, long long *groups, int reset)
{
    if (!p) return fail(RCT_ERR_ARG, "null plan");
    DeviceGuard guard(p->device);
    if (p->ev_busy.size() % 5 != 0) return fail(RCT_ERR_ARG, "timing events out of step (mixed entry points?)");
    for (size_t g = 0; g + 5 <= p->ev_busy.size(); g += 5) {
        CK(cudaEventSynchronize(p->ev_busy[g + 4]));
        for (int k = 0; k < 4; ++k) {
            float ms = 0.f;
            CK(cudaEventElapsedTime(&ms, p->ev_busy[g + k], p->ev_busy[g + k + 1]));
            p->phase_ms[k] += (double)ms;
        }
        ++p->phase_groups;
    }
    for (cudaEvent_t e : p->ev_busy) p->ev_free.push_back(e);
    p->ev_busy.clear();
    if (ms4) memcpy(ms4, p->phase_ms, sizeof p->phase_ms);
    if (groups) *groups = p->phase_groups;
    if (reset) {
        memset(p->phase_ms, 0, sizeof p->phase_ms);
        p->phase_groups = 0;
    }
    return RCT_OK;
}

static const char *const kTkNames[RCT_TK_COUNT] = {
    "begin", "memset", "k_front_count", "k_front_scan", "k_front_part", "k_front_hist", "k_hist", "k_compact",
    "k_lab_moments", "k_pca", "k_keys", "k_plan_fine", "k_fine_hist", "k_plan_parts", "k_scatter", "k_part_sort",
    "k_part_scan", "k_match", "k_apply", "k_keytab"};
const char *rct_kernel_timing_name(int id) { return (id >= 0 && id < RCT_TK_COUNT) ? kTkNames[id] : nullptr; }

int rct_plan_kernel_timing(rct_plan *p, double *ms, long long *groups, int reset)
{
    if (!p) return fail(RCT_ERR_ARG, "null plan");
    DeviceGuard guard(p->device);
    for (size_t i = 0; i < p->kev.size(); ++i) {
        if (p->kev_id[i] == RCT_TK_BEGIN) {
            ++p->kernel_groups;
            continue;
        }
        if (i == 0) continue;
        CK(cudaEventSynchronize(p->kev[i]));
        float t = 0.f;
        CK(cudaEventElapsedTime(&t, p->kev[i - 1], p->kev[i]));
        p->kernel_ms[p->kev_id[i]] += (double)t;
    }
    for (cudaEvent_t e : p->kev) p->ev_free.push_back(e);
    p->kev.clear();
    p->kev_id.clear();
    if (ms) memcpy(ms, p->kernel_ms, sizeof p->kernel_ms);
    if (groups) *groups = p->kernel_groups;
    if (reset) {
        memset(p->kernel_ms, 0, sizeof p->kernel_ms);
        p->kernel_groups = 0;
    }
    return RCT_OK;
}

// ---- introspection -------------------------------------------------------------------------
int rct_plan_info(rct_plan *p, int slot, rct_frame_info *info)
{
    if (!p || !info || slot < 0 || slot >= p->slots) return fail(RCT_ERR_ARG, "bad info request");
    DeviceGuard guard(p->device);
    CK(cudaDeviceSynchronize());
    FrameCtrl *h = new (std::nothrow) FrameCtrl;
    if (!h) return fail(RCT_ERR_NOMEM, "out of host memory");
    cudaError_t e = cudaMemcpy(h, p->ctrl + slot, sizeof(FrameCtrl), cudaMemcpyDeviceToHost);
    if (e == cudaSuccess) {
        info->n_target = (int64_t)h->r[0].n_pix; info->n_source = (int64_t)h->r[1].n_pix;
        info->d_target = h->r[0].d; info->d_source = h->r[1].d;
        memcpy(info->t_mean, h->r[0].pca.mean, 24); memcpy(info->t_comp, h->r[0].pca.comp, 24);
        memcpy(info->s_mean, h->r[1].pca.mean, 24); memcpy(info->s_comp, h->r[1].pca.comp, 24);
        info->status = h->status;
    }
    delete h;
    if (e != cudaSuccess) return fail(RCT_ERR_CUDA, "info copy failed: %s", cudaGetErrorString(e));
    return RCT_OK;
}

int rct_plan_debug_hist(rct_plan *p, int slot, uint32_t *ht, uint32_t *hs)
{
    if (!p || slot < 0 || slot >= p->slots) return fail(RCT_ERR_ARG, "bad slot");
    DeviceGuard guard(p->device);
    CK(cudaDeviceSynchronize());
    unsigned *h = p->hist + (size_t)slot * p->L.hist_stride_u32;
    if (!p->cube_dirty) {  // partition front end: the cube is all zero; rebuild it from the colour lists
        LAUNCH(k_list_to_cube, 1024, 256, 0, p->ctrl, p->jobs, slot, 0, h);
        LAUNCH(k_list_to_cube, 1024, 256, 0, p->ctrl, p->jobs, slot, 1, h + kCube);
        CK(cudaDeviceSynchronize());
    }
    if (ht) CK(cudaMemcpy(ht, h, 4 * (size_t)kCube, cudaMemcpyDeviceToHost));
    if (hs) CK(cudaMemcpy(hs, h + kCube, 4 * (size_t)kCube, cudaMemcpyDeviceToHost));
    if (!p->cube_dirty) CK(cudaMemset(h, 0, 4 * p->L.hist_stride_u32));
    return RCT_OK;
}

int rct_plan_debug_target(rct_plan *p, int slot, uint32_t *colour, double *key, int64_t *j, double *matched,
                          uint32_t *cum)
{
    if (!p || slot < 0 || slot >= p->slots) return fail(RCT_ERR_ARG, "bad slot");
    if (!(p->flags & RCT_FLAG_DEBUG)) return fail(RCT_ERR_ARG, "plan was not created with RCT_FLAG_DEBUG");
    rct_frame_info info;
    int rc = rct_plan_info(p, slot, &info);
    if (rc) return rc;
    size_t d = (size_t)info.d_target;
    if (d == 0 || info.status != 0) return RCT_OK;
    const Layout &L = p->L;
    uint8_t *b = p->ws + L.off_region + L.region_bytes * (size_t)(2 * slot);
    {
        // sorted records: key | payload | part-local cumulative count; globalise with the part bases
        std::vector<ulonglong2> rec(d);
        std::vector<unsigned> pstart(L.maxparts + 1), pbase(L.maxparts + 1);
        FrameCtrl *h = new (std::nothrow) FrameCtrl;
        if (!h) return fail(RCT_ERR_NOMEM, "out of host memory");
        cudaError_t e = cudaMemcpy(h, p->ctrl + slot, sizeof(FrameCtrl), cudaMemcpyDeviceToHost);
        unsigned np = h->r[0].n_parts;
        delete h;
        if (e != cudaSuccess) return fail(RCT_ERR_CUDA, "debug copy failed: %s", cudaGetErrorString(e));
        CK(cudaMemcpy(rec.data(), b + L.r_recP, 16 * d, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(pstart.data(), b + L.r_pstart, 4 * (L.maxparts + 1), cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(pbase.data(), b + L.r_ppix, 4 * (L.maxparts + 1), cudaMemcpyDeviceToHost));
        for (unsigned q = 0; q < np; ++q)
            for (unsigned i = pstart[q]; i < pstart[q + 1]; ++i) {
                if (key) memcpy(&key[i], &rec[i].x, 8);
                if (colour) colour[i] = (uint32_t)rec[i].y;
                if (cum) cum[i] = pbase[q] + (uint32_t)(rec[i].y >> 32);
            }
    }
    if (j) CK(cudaMemcpy(j, p->dbg_j + (size_t)slot * L.dbg_stride, 8 * d, cudaMemcpyDeviceToHost));
    if (matched) CK(cudaMemcpy(matched, p->dbg_m + (size_t)slot * L.dbg_stride, 8 * d, cudaMemcpyDeviceToHost));
    return RCT_OK;
}

int rct_plan_debug_keytab(rct_plan *p, int slot, unsigned long long *stats10, double *max_key_err)
{
    if (!p || !stats10 || slot < 0 || slot >= p->slots) return fail(RCT_ERR_ARG, "bad slot");
    DeviceGuard guard(p->device);
    CK(cudaDeviceSynchronize());
    memset(stats10, 0, 10 * sizeof(unsigned long long));
    if (max_key_err) *max_key_err = 0.0;
    if (!p->kt_live) return RCT_OK;
    const uint8_t *kt0 = p->ws + p->L.off_kt;
    KtParams hp;
    CK(cudaMemcpy(&hp, kt0 + (size_t)slot * kKtBytes, sizeof hp, cudaMemcpyDeviceToHost));
    stats10[0] = hp.valid;
    stats10[1] = hp.n_runs;
    stats10[2] = (unsigned long long)kKtCells;
    stats10[3] = hp.n_flagged;
    if (!hp.valid) return RCT_OK;
    unsigned long long *d_counts = nullptr;
    CK(cudaMalloc((void **)&d_counts, 8 * sizeof(unsigned long long)));
    cudaError_t e = cudaMemset(d_counts, 0, 8 * sizeof(unsigned long long));
    if (e == cudaSuccess) {
        k_kt_verify<<<p->sm * 4, 256>>>(p->ctrl, p->jobs, kt0, p->lut, p->L.lut_stride_u32, slot, d_counts);
        if (max_key_err) k_kt_keyerr<<<p->sm * 8, 256>>>(p->ctrl, kt0, slot, reinterpret_cast<double *>(d_counts + 6));
        e = cudaDeviceSynchronize();
    }
    unsigned long long h[8] = {0};
    if (e == cudaSuccess) e = cudaMemcpy(h, d_counts, sizeof h, cudaMemcpyDeviceToHost);
    cudaFree(d_counts);
    if (e != cudaSuccess) return fail(RCT_ERR_CUDA, "key-table check failed: %s", cudaGetErrorString(e));
    for (int k = 0; k < 6; ++k) stats10[4 + k] = h[k];
    if (max_key_err) memcpy(max_key_err, &h[6], 8);
    return RCT_OK;
}

int rct_plan_debug_lut(rct_plan *p, int slot, uint32_t *lut_host)
{
    if (!p || !lut_host || slot < 0 || slot >= p->slots) return fail(RCT_ERR_ARG, "bad slot");
    DeviceGuard guard(p->device);
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(lut_host, p->lut + (size_t)slot * p->L.lut_stride_u32, 4 * (size_t)kCube, cudaMemcpyDeviceToHost));
    if (p->lut_mode) {  // expand run index + palette to the uint32 RGB per colour the caller expects
        const uint16_t *idx = reinterpret_cast<const uint16_t *>(lut_host);
        uint32_t *pal = new (std::nothrow) uint32_t[kPaletteCap];
        if (!pal) return fail(RCT_ERR_NOMEM, "out of host memory");
        memcpy(pal, lut_host + kPaletteOffsetU32, sizeof(uint32_t) * kPaletteCap);
        for (size_t c = (size_t)kCube; c-- > 0;) lut_host[c] = pal[idx[c]];  // descending: idx[c] lies below lut_host[c]
        delete[] pal;
    }
    return RCT_OK;
}

}  // extern "C"
