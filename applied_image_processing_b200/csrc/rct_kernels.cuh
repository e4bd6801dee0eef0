// rct_kernels.cuh -- sm_100a kernels of the regional colour-transfer path.
//
// Pipeline (one launch group = up to `slots` frames, blockIdx.z = slot, blockIdx.y = region
// where a kernel works per region; region 0 = target/foreground, 1 = source/background):
//
//   k_front_*       pixels -> list of distinct colours + counts per region (a-1..a-3), by colour
//                   partition + shared-memory histograms (rct_front.cuh; the default front end)
//   k_hist          pixels -> two 2^24-bin colour-cube histograms        (row-sharded path: the
//   k_compact       cube   -> list of distinct colours + counts           cubes are what is all-reduced)
//   k_lab_moments   list   -> exact integer moments, lab bounding box    (a-4, a-5)
//   k_pca           moments-> mean_, components_[0], key range           (a-5)
//   k_keys          list   -> float64 projection keys + coarse key histogram
//   k_plan_fine / k_fine_hist / k_plan_parts / k_scatter / k_part_sort / k_part_scan   (rct_sort.cuh)
//                   exact ascending sort of the distinct keys (two-level equalised partition
//                   into shared-memory-sized parts) + cumulative pixel counts
//   k_match         rank -> quantile lookup with np.interp semantics -> colour LUT (a-6, a-7)
//   k_apply         pixels -> output through the LUT, fused with resample + combine (a-8, a-9)
//
// Everything after k_hist works per DISTINCT colour; the per-pixel passes (k_hist,
// k_apply) are pure integer/byte streaming kernels.
#pragma once
#include "rct_math.cuh"

namespace rct {

constexpr int kCoarse = 4096;           // coarse key-histogram bins per region
constexpr unsigned kCube = 1u << 24;    // colour-cube bins
constexpr int kFineTarget = 16384;      // fine bins aimed at per sort job (equalised)
constexpr int kFineMax = kFineTarget + kCoarse;  // upper bound on fine bins per job
#ifndef RCT_PARTT
#define RCT_PARTT 3072
#endif
constexpr int kPartT = RCT_PARTT;            // entries per part aimed at (part < kPartT + largest fine bin)
constexpr int kPartCap = 4096;          // entries a part may have to be sorted in shared memory
constexpr int kFillStride = 32;         // uint32 stride of the per-part fill counters (128 B apart)
constexpr unsigned kPaletteCap = 65536;               // palette LUT: run indices are uint16
constexpr size_t kPaletteOffsetU32 = (size_t)1 << 23;  // the palette sits behind the 2^24 uint16 indices of a slot's LUT

struct RegionCtrl {
    unsigned long long n_pix;           // region pixel count N (non-black pixels)
    unsigned int d;                     // distinct colours (sort entries)
    unsigned int n_fine;                // fine bins in use
    unsigned int n_parts;               // parts in use
    unsigned int skip;                  // 1: the frame has an empty region, nothing to sort
    long long limbs[9][3];              // exact moments: s1[0..2], s2[00,01,02,11,12,22]
    unsigned long long lab_max[3];      // [0]: ord_enc(max key) of the generic match_cdf path
    unsigned long long lab_min[3];      // [0]: ~ord_enc(min key)
    unsigned int cmax[3];               // per-channel maximum of the region's colours
    unsigned int cmin_inv[3];           // 255 - per-channel minimum
    Pca pca;
    double klo, cscale;                 // coarse bin = floor((key-klo)*cscale)
    unsigned int coarse[kCoarse];
    unsigned int fine_base[kCoarse + 1];
};

// colour-partitioned front end (rct_front.cuh): bucket = colour >> kBucketShift
constexpr int kBucketShift = 14;
constexpr int kBucketBins = 1 << kBucketShift;  // colours per bucket = bins of the shared-memory histogram
constexpr int kNB = 1024;                       // buckets per region
#ifndef RCT_CHUNK
#define RCT_CHUNK 65536
#endif
constexpr unsigned kChunk = RCT_CHUNK;              // items one CTA of k_front_hist accumulates

struct FrontCtrl {
    unsigned bcnt[2 * kNB];      // items per (region, bucket)
    unsigned bticket[2 * kNB];   // chunk CTAs done per split bucket
    unsigned bstart[2][kNB + 1]; // exclusive offsets inside the region's item array
    unsigned bcur[2 * kNB];      // write cursors of k_front_part
    unsigned nchunks;            // CTAs of k_front_hist with work
    unsigned pad_[3];
};

struct FrameCtrl {
    RegionCtrl r[2];
    int status;
    unsigned n_palette;  // runs of equal output colour along the sorted target (palette LUT)
    unsigned n_runs;     // run starts k_match recorded for the key table (rct_keytab.cuh)
    int pad_[1];
    FrontCtrl front;
};

// One sort job = one region of one slot (or one array of the generic match_cdf call).
struct Job {
    RegionCtrl *rc;
    const double *keyA;          // unsorted keys
    const unsigned *payA;        // payload (colour index); NULL = identity
    const unsigned *cntA;        // multiplicity; NULL = 1
    ulonglong2 *recP;            // partitioned, then sorted records: x = key bits, y = payload | count << 32
                                 // (count becomes the part-local inclusive cumulative count)
    unsigned *fcnt;              // [kFineMax+1] fine-bin counts, then exclusive offsets (fcnt[n_fine] = d)
    unsigned short *pof;         // [kFineMax] part of each fine bin
    unsigned *pstart;            // [maxparts+1] first entry of each part
    unsigned *ppix;              // [maxparts+1] pixels per part, then exclusive bases (ppix[n_parts] = N)
    unsigned *pfill;             // [maxparts*kFillStride] scatter fill counters
    unsigned cap, maxparts;
};

// ------------------------------------------------------------------------------------------
// streaming loads/stores of 16 RGB pixels (48 bytes) per thread
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 ldg_stream(const uint4 *p)
{
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}
__device__ __forceinline__ void stg_stream(uint4 *p, uint4 v)
{
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y),
                 "r"(v.z), "r"(v.w)
                 : "memory");
}

__device__ __forceinline__ void load16_rgb(const uint8_t *base, size_t p0, int nvalid, bool vec, uint32_t c[16])
{
    if (vec && nvalid == 16) {
        const uint4 *q = reinterpret_cast<const uint4 *>(base + 3 * p0);
        uint4 a = ldg_stream(q), b = ldg_stream(q + 1), d = ldg_stream(q + 2);
        uint32_t w[12] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w, d.x, d.y, d.z, d.w};
#pragma unroll
        for (int m = 0; m < 4; ++m) {
            c[4 * m + 0] = w[3 * m] & 0xffffffu;
            c[4 * m + 1] = __funnelshift_r(w[3 * m], w[3 * m + 1], 24) & 0xffffffu;
            c[4 * m + 2] = __funnelshift_r(w[3 * m + 1], w[3 * m + 2], 16) & 0xffffffu;
            c[4 * m + 3] = w[3 * m + 2] >> 8;
        }
    } else {
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            if (k < nvalid) {
                const uint8_t *p = base + 3 * (p0 + k);
                c[k] = (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16);
            } else {
                c[k] = 0;
            }
        }
    }
}

__device__ __forceinline__ void store16_rgb(uint8_t *base, size_t p0, int nvalid, bool vec, const uint32_t c[16],
                                            unsigned keep_mask /* bit k: write pixel k */)
{
    if (vec && nvalid == 16 && keep_mask == 0xffffu) {
        uint32_t w[12];
#pragma unroll
        for (int m = 0; m < 4; ++m) {
            w[3 * m + 0] = c[4 * m] | (c[4 * m + 1] << 24);
            w[3 * m + 1] = (c[4 * m + 1] >> 8) | (c[4 * m + 2] << 16);
            w[3 * m + 2] = (c[4 * m + 2] >> 16) | (c[4 * m + 3] << 8);
        }
        uint4 *q = reinterpret_cast<uint4 *>(base + 3 * p0);
        stg_stream(q, make_uint4(w[0], w[1], w[2], w[3]));
        stg_stream(q + 1, make_uint4(w[4], w[5], w[6], w[7]));
        stg_stream(q + 2, make_uint4(w[8], w[9], w[10], w[11]));
    } else {
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            if (k < nvalid && ((keep_mask >> k) & 1u)) {
                uint8_t *p = base + 3 * (p0 + k);
                p[0] = (uint8_t)c[k];
                p[1] = (uint8_t)(c[k] >> 8);
                p[2] = (uint8_t)(c[k] >> 16);
            }
        }
    }
}

__device__ __forceinline__ void load16_u8(const uint8_t *base, size_t p0, int nvalid, bool vec, uint32_t m[16])
{
    if (vec && nvalid == 16) {
        uint4 a = ldg_stream(reinterpret_cast<const uint4 *>(base + p0));
        uint32_t w[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
        for (int k = 0; k < 16; ++k) m[k] = (w[k >> 2] >> (8 * (k & 3))) & 255u;
    } else {
#pragma unroll
        for (int k = 0; k < 16; ++k) m[k] = (k < nvalid) ? base[p0 + k] : 0u;
    }
}

// stylised pixels through the Pillow-NEAREST tables (a-1), never materialised
__device__ __forceinline__ void gather16_rgb(const uint8_t *sty, const int *__restrict__ xtab,
                                             const int *__restrict__ ytab, int W, int sw, size_t p0, int nvalid,
                                             uint32_t c[16])
{
    int y = (int)(p0 / (size_t)W);
    int x = (int)(p0 - (size_t)y * (size_t)W);
    size_t rowoff = (size_t)__ldg(ytab + y) * (size_t)sw;
    int prev = -1;       // source x of the previous pixel: up-sampling repeats source pixels, which
    uint32_t pv = 0;     // are then taken from the register instead of three more byte loads
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        if (k < nvalid) {
            const int xs = __ldg(xtab + x);
            if (xs != prev) {
                const uint8_t *p = sty + 3 * (rowoff + (size_t)xs);
                pv = (uint32_t)__ldg(p) | ((uint32_t)__ldg(p + 1) << 8) | ((uint32_t)__ldg(p + 2) << 16);
                prev = xs;
            }
            c[k] = pv;
            if (++x == W) {
                x = 0;
                ++y;
                prev = -1;
                if (k + 1 < nvalid) rowoff = (size_t)__ldg(ytab + y) * (size_t)sw;
            }
        } else {
            c[k] = 0;
        }
    }
}

struct PixArgs {
    const uint8_t *content, *sty, *mask;
    uint8_t *out;
    size_t content_stride, sty_stride, mask_stride, out_stride;  // bytes per frame
    const int *xtab, *ytab;                                      // NULL when no resample
    unsigned *hist;        // slot 0; slot s at + s*hist_stride (uint32 units): [2][2^24]
    size_t hist_stride;
    const unsigned *lut;   // slot 0; per slot [2^24] uint32 RGB, or (lut_mode 1) [2^24] uint16 run indices + palette
    size_t lut_stride;
    int lut_mode;
    const uint8_t *kt;     // key tables of the slots (rct_keytab.cuh), or NULL
    FrameCtrl *ctrl;       // [slots]
    int *status_out;       // [slots] or NULL
    int W, sw;
    int sh_dbg;            // stylised height (RCT_RS_DEBUG builds only)
    size_t p_begin, p_end; // pixel range within a frame (row sharding)
    int mode, label;
    int lpf;               // labels per frame of a fused label pass (slot = frame * lpf + label offset); else 1
    int aggregate;         // warp-aggregate equal colours before the atomic
};

__device__ __forceinline__ void region_flags(int mode, int label, uint32_t m, bool &in_t, bool &in_s)
{
    if (mode == 0) {
        in_t = (m == 0);
        in_s = !in_t;
    } else if (mode == 1) {
        in_t = in_s = (m == (uint32_t)label);
    } else {
        in_t = in_s = true;
    }
}

__device__ __forceinline__ void hist_add(unsigned *hist, uint32_t col, bool valid, int aggregate)
{
    if (aggregate) {
        unsigned lane = threadIdx.x & 31u;
        uint32_t key = valid ? col : (0x80000000u | lane);
        unsigned peers = __match_any_sync(0xffffffffu, key);
        if (valid && (unsigned)(__ffs(peers) - 1) == lane) atomicAdd(hist + col, (unsigned)__popc(peers));
    } else {
        if (valid) atomicAdd(hist + col, 1u);
    }
}

// k_hist: a-1 (resample), a-2 (mask algebra), a-3 (non-black selection) fused into one read
// of the pixels; output = the two colour-cube histograms.
template <bool RESAMPLE>
__global__ void __launch_bounds__(256) k_hist(PixArgs a)
{
    const int slot = blockIdx.y;
    const int frame = slot / a.lpf, label = a.label + slot % a.lpf;  // fused label pass: slot = (frame, label)
    const uint8_t *content = a.content + (size_t)frame * a.content_stride;
    const uint8_t *sty = a.sty + (size_t)frame * a.sty_stride;
    const uint8_t *mask = a.mask ? a.mask + (size_t)frame * a.mask_stride : nullptr;
    unsigned *hist_t = a.hist + (size_t)slot * a.hist_stride;
    unsigned *hist_s = hist_t + kCube;
    const size_t npix = a.p_end - a.p_begin;
    const size_t ngroups = (npix + 15) / 16;
    const bool vec_c = (((uintptr_t)(content + 3 * a.p_begin)) & 15) == 0;
    const bool vec_s = !RESAMPLE && (((uintptr_t)(sty + 3 * a.p_begin)) & 15) == 0;
    const bool vec_m = mask && (((uintptr_t)(mask + a.p_begin)) & 15) == 0;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    const unsigned lane = threadIdx.x & 31u;
    // warp-uniform trip count (hist_add may use full-warp match)
    for (size_t gbase = (size_t)blockIdx.x * blockDim.x + (threadIdx.x & ~31u); gbase < ngroups; gbase += stride) {
        size_t g = gbase + lane;
        bool live = g < ngroups;
        size_t p0 = a.p_begin + 16 * (live ? g : 0);
        int nvalid = live ? (int)min((size_t)16, a.p_end - p0) : 0;
        uint32_t c[16], s[16], m[16];
        load16_rgb(content, p0, nvalid, vec_c, c);
        if (RESAMPLE) gather16_rgb(sty, a.xtab, a.ytab, a.W, a.sw, p0, nvalid, s);
        else load16_rgb(sty, p0, nvalid, vec_s, s);
        if (mask) load16_u8(mask, p0, nvalid, vec_m, m);
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            bool in_t, in_s;
            region_flags(a.mode, label, mask ? m[k] : 0u, in_t, in_s);
            bool ok = k < nvalid;
            hist_add(hist_t, c[k], ok && in_t && c[k] != 0u, a.aggregate);
            hist_add(hist_s, s[k], ok && in_s && s[k] != 0u, a.aggregate);
        }
    }
}

// ------------------------------------------------------------------------------------------
// block-level helpers
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned warp_excl_scan(unsigned v, unsigned &total)
{
    unsigned lane = threadIdx.x & 31u;
    unsigned x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        unsigned y = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= (unsigned)o) x += y;
    }
    total = __shfl_sync(0xffffffffu, x, 31);
    return x - v;
}

// exclusive scan of one value per thread across the block (blockDim.x <= 1024)
__device__ __forceinline__ unsigned block_excl_scan(unsigned v, unsigned *s_warp /*[33]*/, unsigned &block_total)
{
    unsigned wtotal;
    unsigned ex = warp_excl_scan(v, wtotal);
    unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
    unsigned nwarps = (blockDim.x + 31) >> 5;
    __syncthreads();  // protect s_warp reuse
    if (lane == 31) s_warp[warp] = wtotal;
    __syncthreads();
    if (warp == 0) {
        unsigned t = lane < nwarps ? s_warp[lane] : 0u, tt;
        unsigned e = warp_excl_scan(t, tt);
        if (lane < nwarps) s_warp[lane] = e;
        if (lane == 0) s_warp[32] = tt;
    }
    __syncthreads();
    block_total = s_warp[32];
    return ex + s_warp[warp];
}

// k_compact: dense scan of the cube -> (colour, count) list per region; order irrelevant
// (the moments are exact integers, the sort is by exact key).
__global__ void __launch_bounds__(256) k_compact(unsigned *hist0, size_t hist_stride, FrameCtrl *ctrl,
                                                 const Job *jobs)
{
    const int region = blockIdx.y, slot = blockIdx.z;
    const unsigned *hist = hist0 + (size_t)slot * hist_stride + (size_t)region * kCube;
    const Job &job = jobs[slot * 2 + region];
    __shared__ unsigned s_warp[33];
    __shared__ unsigned s_base;
    unsigned bin0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4u;
    uint4 h = *reinterpret_cast<const uint4 *>(hist + bin0);
    unsigned n = (h.x != 0) + (h.y != 0) + (h.z != 0) + (h.w != 0);
    unsigned total;
    unsigned off = block_excl_scan(n, s_warp, total);
    if (total == 0) return;
    if (threadIdx.x == 0) s_base = atomicAdd(&ctrl[slot].r[region].d, total);
    __syncthreads();
    unsigned pos = s_base + off;
    unsigned *col = const_cast<unsigned *>(job.payA), *cnt = const_cast<unsigned *>(job.cntA);
    unsigned hv[4] = {h.x, h.y, h.z, h.w};
#pragma unroll
    for (int k = 0; k < 4; ++k)
        if (hv[k]) {
            col[pos] = bin0 + k;
            cnt[pos] = hv[k];
            ++pos;
        }
}

__device__ __forceinline__ long long warp_sum_ll(long long v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// k_lab_moments: a-4 + the reductions of a-5 over the distinct-colour list.  Multiplicities
// below 128 (virtually all of them) take the 64-bit fast path of rct_math.cuh; the rest, and
// the rare overflow-guard flushes, go through 48-bit limbs in shared memory.  The next
// entry is loaded before the current one is evaluated (the loop is latency-bound otherwise).
#ifndef RCT_LMB
#define RCT_LMB 2
#endif
__global__ void __launch_bounds__(256, RCT_LMB) k_lab_moments(FrameCtrl *ctrl, const Job *jobs)
{
    const int region = blockIdx.y, slot = blockIdx.z;
    const Job &job = jobs[slot * 2 + region];
    RegionCtrl *rc = &ctrl[slot].r[region];
    const unsigned d = rc->d;
    if ((size_t)blockIdx.x * blockDim.x >= d) return;
    __shared__ MathTabs T;
    __shared__ long long s_red[8][28];
    __shared__ unsigned long long s_slow[9][3];
    __shared__ unsigned s_box[8][2];
    if (threadIdx.x < 27) s_slow[threadIdx.x / 3][threadIdx.x % 3] = 0ull;
    stage_tabs(&T);  // ends with a barrier
    MomentsFast f;
    fast_zero(f);
    long long npx = 0;
    int terms = 0;
    auto spill = [&](const i128 *s1, const i128 *s2) {  // exact add into the block's limbs
        long long l[3];
#pragma unroll
        for (int k = 0; k < 9; ++k) {
            split_limbs(k < 3 ? s1[k] : s2[k - 3], l);
#pragma unroll
            for (int j = 0; j < 3; ++j) atomicAdd(&s_slow[k][j], (unsigned long long)l[j]);
        }
    };
    unsigned cmx = 0u, cmn = 0x00ffffffu;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned col = 0, cnt = 0;
    if (e < d) {
        col = job.payA[e];
        cnt = job.cntA[e];
    }
    // The hot loop holds nothing but fast-path entries (its 25 accumulators then stay in fixed register pairs); an
    // entry that cannot take the fast path, or the 128th term, leaves it for the cold code below.
    while (e < d) {
        bool slow = false;
        double lab[3];
        unsigned coln = 0, cntn = 0;
        size_t en = e;
        auto step = [&]() -> bool {  // one entry; false: it has to take the slow path
            en = e + stride;
            coln = cntn = 0;
            if (en < d) {
                coln = job.payA[en];
                cntn = job.cntA[en];
            }
            rgb2lab(col, &T, lab);
            cmx = __vmaxu4(cmx, col);
            cmn = __vminu4(cmn, col);
            npx += cnt;
            if (!(cnt < 128u && fast_add(f, lab, cnt))) return false;
            col = coln;
            cnt = cntn;
            e = en;
            ++terms;
            return true;
        };
#pragma unroll 1
        do {  // two entries per trip: halves the register shuffling at the back edge
            if (!step()) {
                slow = true;
                break;
            }
            if (!(terms < kFastTerms && e < d)) break;
            if (!step()) {
                slow = true;
                break;
            }
        } while (terms < kFastTerms && e < d);
        if (slow) {
            Moments m;
            moments_zero(m);
            moments_add(m, lab, (long long)cnt);
            spill(m.s1, m.s2);
            col = coln;
            cnt = cntn;
            e = en;
        }
        if (terms == kFastTerms) {
            i128 s1[3], s2[6];
            fast_finish(f, s1, s2);
            spill(s1, s2);
            fast_zero(f);
            terms = 0;
        }
    }
    long long v[28];
    {
        i128 s1[3], s2[6];
        fast_finish(f, s1, s2);
#pragma unroll
        for (int k = 0; k < 3; ++k) split_limbs(s1[k], v + 3 * k);
#pragma unroll
        for (int k = 0; k < 6; ++k) split_limbs(s2[k], v + 9 + 3 * k);
    }
    v[27] = npx;
    const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
#pragma unroll
    for (int k = 0; k < 28; ++k) {
        long long s = warp_sum_ll(v[k]);
        if (lane == 0) s_red[warp][k] = s;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        cmx = __vmaxu4(cmx, __shfl_xor_sync(0xffffffffu, cmx, o));
        cmn = __vminu4(cmn, __shfl_xor_sync(0xffffffffu, cmn, o));
    }
    if (lane == 0) {
        s_box[warp][0] = cmx;
        s_box[warp][1] = cmn;
    }
    __syncthreads();
    const unsigned nwarps = blockDim.x >> 5;
    if (threadIdx.x < 9) {  // one moment each: sum limbs over warps, renormalise, publish
        long long l[3];
        for (int j = 0; j < 3; ++j) l[j] = (long long)s_slow[threadIdx.x][j];
        for (unsigned w = 0; w < nwarps; ++w)
            for (int j = 0; j < 3; ++j) l[j] += s_red[w][3 * threadIdx.x + j];
        i128 tot = join_limbs(l);
        split_limbs(tot, l);
        for (int j = 0; j < 3; ++j)
            atomicAdd(reinterpret_cast<unsigned long long *>(&rc->limbs[threadIdx.x][j]), (unsigned long long)l[j]);
    } else if (threadIdx.x == 9) {
        long long n = 0;
        for (unsigned w = 0; w < nwarps; ++w) n += s_red[w][27];
        atomicAdd(&rc->n_pix, (unsigned long long)n);
    } else if (threadIdx.x >= 32 && threadIdx.x < 35) {
        int i = threadIdx.x - 32;
        unsigned a = 0u, b = 255u;
        for (unsigned w = 0; w < nwarps; ++w) {
            a = max(a, (s_box[w][0] >> (8 * i)) & 255u);
            b = min(b, (s_box[w][1] >> (8 * i)) & 255u);
        }
        atomicMax(&rc->cmax[i], a);
        atomicMax(&rc->cmin_inv[i], 255u - b);
    }
}

// k_pca: a-5 on the device (no host round trip): mean_, components_[0], sign rule, and the
// key range used by the bucket sort.  One thread per slot.
__global__ void k_pca(FrameCtrl *ctrl, int slots)
{
    int slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= slots) return;
    FrameCtrl *fc = &ctrl[slot];
    for (int region = 0; region < 2; ++region) {
        RegionCtrl *rc = &fc->r[region];
        if (rc->n_pix == 0) continue;
        i128 s1[3], s2[6];
        for (int k = 0; k < 3; ++k) s1[k] = join_limbs(rc->limbs[k]);
        for (int k = 0; k < 6; ++k) s2[k] = join_limbs(rc->limbs[3 + k]);
        Pca p;
        pca_from_moments((long long)rc->n_pix, s1, s2, &p);
        rc->pca = p;
        // key range by interval arithmetic over the region's colour box: LMS is increasing in
        // every channel (positive matrix), log10 is increasing, the two linear maps take the
        // sign-wise extremes.  (A key a rounding error outside the range is clamped into the
        // edge bins by coarse_bin, which keeps the bin map monotone.)
        double ylo[3], yhi[3];
        for (int i = 0; i < 3; ++i) {
            double a = 0.0, b = 0.0;
            for (int j = 0; j < 3; ++j) {
                a += c_mat[3 * i + j] * g_tabs.unit[255u - rc->cmin_inv[j]];
                b += c_mat[3 * i + j] * g_tabs.unit[rc->cmax[j]];
            }
            ylo[i] = clog10(a < 1e-6 ? 1e-6 : a, &g_tabs);
            yhi[i] = clog10(b < 1e-6 ? 1e-6 : b, &g_tabs);
        }
        double lo = 0.0, hi = 0.0;
        for (int k = 0; k < 3; ++k) {
            double l0 = 0.0, l1 = 0.0;
            for (int i = 0; i < 3; ++i) {
                double m = c_mat[9 + 3 * k + i];
                l0 += fmin(m * ylo[i], m * yhi[i]);
                l1 += fmax(m * ylo[i], m * yhi[i]);
            }
            lo += fmin(p.comp[k] * l0, p.comp[k] * l1);
            hi += fmax(p.comp[k] * l0, p.comp[k] * l1);
        }
        lo -= p.pm;
        hi -= p.pm;
        double span = hi - lo;
        rc->klo = lo;
        rc->cscale = (span > 0.0 && span < 1e300) ? (double)kCoarse / span : 0.0;
    }
    fc->status = (fc->r[0].n_pix == 0) ? 1 : (fc->r[1].n_pix == 0 ? 2 : 0);
    fc->r[0].skip = fc->r[1].skip = (fc->status != 0);
}

__device__ __forceinline__ int coarse_bin(double key, double klo, double cscale, double &t)
{
    t = (key - klo) * cscale;
    int cb = (int)floor(t);
    if (!(t >= 0.0)) cb = 0;  // also catches NaN
    if (cb > kCoarse - 1) cb = kCoarse - 1;
    return cb;
}

// k_keys: projection key per distinct colour (C7) + privatised coarse histogram.
__global__ void __launch_bounds__(256) k_keys(FrameCtrl *ctrl, const Job *jobs)
{
    const int region = blockIdx.y, slot = blockIdx.z;
    const Job &job = jobs[slot * 2 + region];
    RegionCtrl *rc = &ctrl[slot].r[region];
    if (ctrl[slot].status != 0) return;
    const unsigned d = rc->d;
    if ((size_t)blockIdx.x * blockDim.x >= d) return;
    __shared__ MathTabs T;
    __shared__ unsigned s_hist[kCoarse];
    for (int i = threadIdx.x; i < kCoarse; i += blockDim.x) s_hist[i] = 0;
    stage_tabs(&T);
    const double c0 = rc->pca.comp[0], c1 = rc->pca.comp[1], c2 = rc->pca.comp[2], pm = rc->pca.pm;
    const double klo = rc->klo, cscale = rc->cscale;
    double *keyA = const_cast<double *>(job.keyA);
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned col = e < d ? job.payA[e] : 0u;
    while (e < d) {  // the next colour is in flight while this one is evaluated
        const size_t en = e + stride;
        const unsigned coln = en < d ? job.payA[en] : 0u;
        double lab[3], comp[3] = {c0, c1, c2};
        rgb2lab(col, &T, lab);
        double key = project(lab, comp, pm);
        keyA[e] = key;
        double t;
        atomicAdd(&s_hist[coarse_bin(key, klo, cscale, t)], 1u);
        col = coln;
        e = en;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < kCoarse; i += blockDim.x) {
        unsigned v = s_hist[i];
        if (v) atomicAdd(&rc->coarse[i], v);
    }
}

// coarse histogram of an arbitrary key array (generic match_cdf path)
__global__ void __launch_bounds__(256) k_coarse_hist(const Job *jobs)
{
    const Job &job = jobs[blockIdx.y];
    RegionCtrl *rc = job.rc;
    const unsigned d = rc->d;
    if ((size_t)blockIdx.x * blockDim.x >= d) return;
    __shared__ unsigned s_hist[kCoarse];
    for (int i = threadIdx.x; i < kCoarse; i += blockDim.x) s_hist[i] = 0;
    __syncthreads();
    const double klo = rc->klo, cscale = rc->cscale;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < d; e += (size_t)gridDim.x * blockDim.x) {
        double t;
        atomicAdd(&s_hist[coarse_bin(job.keyA[e], klo, cscale, t)], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < kCoarse; i += blockDim.x) {
        unsigned v = s_hist[i];
        if (v) atomicAdd(&rc->coarse[i], v);
    }
}

#include "rct_front.cuh"
#include "rct_keytab_layout.cuh"
#include "rct_sort.cuh"
#include "rct_keytab.cuh"
#include "rct_apply.cuh"
#include "rct_binned.cuh"
#include "rct_adjacent.cuh"
#include "rct_video.cuh"

// ------------------------------------------------------------------------------------------
// the reference's stand-alone element-wise functions
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_rgb_to_lab(const uint8_t *rgb, double *lab, size_t n)
{
    __shared__ MathTabs T;
    stage_tabs(&T);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const uint8_t *p = rgb + 3 * i;
        uint32_t col = (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16);
        double v[3];
        rgb2lab(col, &T, v);
        lab[3 * i] = v[0];
        lab[3 * i + 1] = v[1];
        lab[3 * i + 2] = v[2];
    }
}

__global__ void __launch_bounds__(256) k_lab_to_rgb(const double *lab, uint8_t *rgb, double *rgbf, size_t n)
{
    __shared__ MathTabs T;
    stage_tabs(&T);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        double v[3] = {lab[3 * i], lab[3 * i + 1], lab[3 * i + 2]}, f[3];
        uint32_t c = lab2rgb(v, &T, f);
        rgb[3 * i] = (uint8_t)c;
        rgb[3 * i + 1] = (uint8_t)(c >> 8);
        rgb[3 * i + 2] = (uint8_t)(c >> 16);
        if (rgbf) {
            rgbf[3 * i] = f[0];
            rgbf[3 * i + 1] = f[1];
            rgbf[3 * i + 2] = f[2];
        }
    }
}

__global__ void __launch_bounds__(256) k_pca_inverse(const double *matched, size_t n, const double *mean3,
                                                     const double *comp3, double *lab)
{
    const double m0 = mean3[0], m1 = mean3[1], m2 = mean3[2], c0 = comp3[0], c1 = comp3[1], c2 = comp3[2];
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        double m = matched[i];
        lab[3 * i] = m * c0 + m0;
        lab[3 * i + 1] = m * c1 + m1;
        lab[3 * i + 2] = m * c2 + m2;
    }
}

// apply_pca on arbitrary float64 [n,3] data: moments pass ...
__global__ void __launch_bounds__(256) k_pca_moments(const double *lab, size_t n, RegionCtrl *rc, int *bad)
{
    __shared__ long long s_red[8][28];
    Moments m;
    moments_zero(m);
    bool oob = false;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (size_t)gridDim.x * blockDim.x) {
        double v[3] = {lab[3 * e], lab[3 * e + 1], lab[3 * e + 2]};
        if (!(fabs(v[0]) < 64.0 && fabs(v[1]) < 64.0 && fabs(v[2]) < 64.0)) {
            oob = true;
            continue;
        }
        moments_add(m, v, 1);
    }
    if (oob) *bad = 1;
    long long v[28];
#pragma unroll
    for (int k = 0; k < 3; ++k) split_limbs(m.s1[k], v + 3 * k);
#pragma unroll
    for (int k = 0; k < 6; ++k) split_limbs(m.s2[k], v + 9 + 3 * k);
    v[27] = m.n;
    const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
#pragma unroll
    for (int k = 0; k < 28; ++k) {
        long long s = warp_sum_ll(v[k]);
        if (lane == 0) s_red[warp][k] = s;
    }
    __syncthreads();
    const unsigned nwarps = blockDim.x >> 5;
    if (threadIdx.x < 9) {
        long long l[3] = {0, 0, 0};
        for (unsigned w = 0; w < nwarps; ++w)
            for (int j = 0; j < 3; ++j) l[j] += s_red[w][3 * threadIdx.x + j];
        i128 tot = join_limbs(l);
        split_limbs(tot, l);
        for (int j = 0; j < 3; ++j)
            atomicAdd(reinterpret_cast<unsigned long long *>(&rc->limbs[threadIdx.x][j]), (unsigned long long)l[j]);
    } else if (threadIdx.x == 9) {
        long long nn = 0;
        for (unsigned w = 0; w < nwarps; ++w) nn += s_red[w][27];
        atomicAdd(&rc->n_pix, (unsigned long long)nn);
    }
}

// ... eigen-solve ...
__global__ void k_pca_solve(RegionCtrl *rc, double *mean3, double *comp3)
{
    if (threadIdx.x != 0 || blockIdx.x != 0 || rc->n_pix == 0) return;
    i128 s1[3], s2[6];
    for (int k = 0; k < 3; ++k) s1[k] = join_limbs(rc->limbs[k]);
    for (int k = 0; k < 6; ++k) s2[k] = join_limbs(rc->limbs[3 + k]);
    Pca p;
    pca_from_moments((long long)rc->n_pix, s1, s2, &p);
    rc->pca = p;
    for (int i = 0; i < 3; ++i) {
        mean3[i] = p.mean[i];
        comp3[i] = p.comp[i];
    }
}

// ... and projection (transform): X @ components_.T - mean_ @ components_.T
__global__ void __launch_bounds__(256) k_pca_project(const double *lab, size_t n, const RegionCtrl *rc, double *proj)
{
    const double comp[3] = {rc->pca.comp[0], rc->pca.comp[1], rc->pca.comp[2]};
    const double pm = rc->pca.pm;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (size_t)gridDim.x * blockDim.x) {
        double v[3] = {lab[3 * e], lab[3 * e + 1], lab[3 * e + 2]};
        proj[e] = project(v, comp, pm);
    }
}

// min/max of a key array -> klo / cscale of a job (generic match_cdf path)
__global__ void __launch_bounds__(256) k_key_range(const double *key, size_t n, RegionCtrl *rc)
{
    double a = -1e308, b = 1e308;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (size_t)gridDim.x * blockDim.x) {
        double k = key[e];
        a = fmax(a, k);
        b = fmin(b, k);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        a = fmax(a, __shfl_xor_sync(0xffffffffu, a, o));
        b = fmin(b, __shfl_xor_sync(0xffffffffu, b, o));
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMax(&rc->lab_max[0], ord_enc(a));
        atomicMax(&rc->lab_min[0], ~ord_enc(b));
    }
}

__global__ void k_key_range_finish(RegionCtrl *rc0, RegionCtrl *rc1, unsigned long long n0, unsigned long long n1)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    RegionCtrl *rcs[2] = {rc0, rc1};
    unsigned long long ns[2] = {n0, n1};
    for (int r = 0; r < 2; ++r) {
        RegionCtrl *rc = rcs[r];
        rc->n_pix = ns[r];
        rc->d = (unsigned)ns[r];
        double hi = ord_dec(rc->lab_max[0]), lo = ord_dec(~rc->lab_min[0]);
        double span = hi - lo;
        rc->klo = lo;
        rc->cscale = (span > 0.0 && span < 1e300) ? (double)kCoarse / span : 0.0;
    }
}

}  // namespace rct
