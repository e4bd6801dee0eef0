// rct_front.cuh -- colour-partitioned front end: pixels -> list of distinct colours + counts.
// Included by rct_kernels.cuh inside namespace rct (needs PixArgs, FrameCtrl, Job, load16_*).
//
// The colour-cube front end (memset + k_hist + k_compact) pays a random 4-byte atomic into a
// 64 MiB table per pixel: 12x the pixel bytes in DRAM traffic on noisy images plus a fixed
// 256 MiB of memset / scan per frame.  This one only streams:
//
//   k_front_count   pixels -> items (region, colour) -> per-bucket item counts; bucket = the top
//                   10 bits of the 24-bit colour, 1024 buckets per region
//   k_front_scan    exclusive bucket offsets + the chunk -> bucket map of k_front_hist
//   k_front_part    pixels -> items written to their bucket (one reservation per CTA tile and
//                   bucket; order inside a bucket is irrelevant: histograms commute)
//   k_front_hist    one CTA per bucket chunk: 16384-bin shared-memory histogram of the low 14
//                   colour bits -> (colour, count) entries appended to the region's list.
//                   Buckets above kChunk items are split over several CTAs that merge through
//                   the bucket's 64 KiB window of the colour cube (last CTA done compacts and
//                   zeroes the window again: the cube stays all-zero between uses).
//
// Items are exactly the pixels the reference selects (a-2 mask algebra, a-3 non-black rule), so
// the (colour, count) list equals the non-zero bins of the colour-cube histograms bit for bit.
#pragma once

constexpr unsigned kInvalidItem = 0xffffffffu;

struct FrontArgs {
    unsigned *items0;        // item array of (slot 0, region 0); (slot, region) at + (2*slot+region)*items_stride
    size_t items_stride;     // in uint32 units
    unsigned *cmap0;         // chunk map of slot 0: entry = (region*kNB + bucket) | chunk << 11
    size_t cmap_stride;
    unsigned cmap_cap;
    unsigned items_cap;      // entries per item array
};

// items of 16 pixels: colour | region << 24, or kInvalidItem.  MASK mode: every pixel belongs to
// exactly one region, one pass; LABEL / ALL: pass 0 = target items, pass 1 = source items.
__device__ __forceinline__ void front_items(int mode, int label, const uint32_t c[16], const uint32_t s[16],
                                            const uint32_t m[16], bool has_mask, int nvalid, int pass, uint32_t it[16])
{
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        bool in_t, in_s;
        region_flags(mode, label, has_mask ? m[k] : 0u, in_t, in_s);
        uint32_t v = kInvalidItem;
        if (k < nvalid) {
            if (mode == 0) {
                uint32_t col = in_t ? c[k] : s[k];
                if (col != 0u) v = col | (in_t ? 0u : (1u << 24));
            } else if (pass == 0) {
                if (in_t && c[k] != 0u) v = c[k];
            } else {
                if (in_s && s[k] != 0u) v = s[k] | (1u << 24);
            }
        }
        it[k] = v;
    }
}

// Run lengths of equal buckets among a thread's 16 consecutive items (smooth images: one run per
// thread, one shared-memory atomic instead of 16).  Packed 4 bits per item: length - 1.
__device__ __forceinline__ unsigned long long item_runs(const uint32_t it[16])
{
    unsigned long long lens = 0ull;
    unsigned run = 0u;
#pragma unroll
    for (int k = 15; k >= 0; --k) {
        bool same = false;
        if (k < 15) same = (it[k + 1] >> kBucketShift) == (it[k] >> kBucketShift);
        run = same ? run + 1u : 0u;  // invalid items never compare equal to a valid neighbour's bucket...
        if (it[k] == kInvalidItem) run = 0u;  // ...and never start or extend a run
        lens |= (unsigned long long)run << (4 * k);
    }
    return lens;
}
__device__ __forceinline__ unsigned run_len(unsigned long long lens, int k) { return ((unsigned)(lens >> (4 * k)) & 15u) + 1u; }

__device__ __forceinline__ bool run_leader(const uint32_t it[16], int k)
{
    if (it[k] == kInvalidItem) return false;
    if (k == 0) return true;
    return (it[k - 1] >> kBucketShift) != (it[k] >> kBucketShift);
}

__device__ __forceinline__ void count_runs(const uint32_t it[16], unsigned long long lens, unsigned *s_cnt)
{
#pragma unroll
    for (int k = 0; k < 16; ++k)
        if (run_leader(it, k)) atomicAdd(&s_cnt[it[k] >> kBucketShift], run_len(lens, k));
}

template <bool RESAMPLE>
__device__ __forceinline__ void front_load(const PixArgs &a, const uint8_t *content, const uint8_t *sty,
                                           const uint8_t *mask, size_t p0, int nvalid, bool vec_c, bool vec_s,
                                           bool vec_m, uint32_t c[16], uint32_t s[16], uint32_t m[16])
{
    load16_rgb(content, p0, nvalid, vec_c, c);
    if (RESAMPLE) gather16_rgb(sty, a.xtab, a.ytab, a.W, a.sw, p0, nvalid, s);
    else load16_rgb(sty, p0, nvalid, vec_s, s);
    if (mask) load16_u8(mask, p0, nvalid, vec_m, m);
}

template <bool RESAMPLE, bool TWO_PASS>
#ifndef RCT_FCB
#define RCT_FCB 4
#endif
__global__ void __launch_bounds__(256, RCT_FCB) k_front_count(PixArgs a)
{
    const int slot = blockIdx.y;
    const int frame = slot / a.lpf, label = a.label + slot % a.lpf;  // fused label pass: slot = (frame, label)
    const uint8_t *content = a.content + (size_t)frame * a.content_stride;
    const uint8_t *sty = a.sty + (size_t)frame * a.sty_stride;
    const uint8_t *mask = a.mask ? a.mask + (size_t)frame * a.mask_stride : nullptr;
    __shared__ unsigned s_cnt[2 * kNB];
    for (int i = threadIdx.x; i < 2 * kNB; i += blockDim.x) s_cnt[i] = 0;
    __syncthreads();
    const size_t npix = a.p_end - a.p_begin;
    const size_t ngroups = (npix + 15) / 16;
    const bool vec_c = (((uintptr_t)(content + 3 * a.p_begin)) & 15) == 0;
    const bool vec_s = !RESAMPLE && (((uintptr_t)(sty + 3 * a.p_begin)) & 15) == 0;
    const bool vec_m = mask && (((uintptr_t)(mask + a.p_begin)) & 15) == 0;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x; g < ngroups; g += stride) {
        const size_t p0 = a.p_begin + 16 * g;
        const int nvalid = (int)min((size_t)16, a.p_end - p0);
        uint32_t it0[16], it1[16];
        {
            uint32_t c[16], s[16], m[16];
            front_load<RESAMPLE>(a, content, sty, mask, p0, nvalid, vec_c, vec_s, vec_m, c, s, m);
            front_items(a.mode, label, c, s, m, mask != nullptr, nvalid, 0, it0);
            if (TWO_PASS) front_items(a.mode, label, c, s, m, mask != nullptr, nvalid, 1, it1);
        }
        count_runs(it0, item_runs(it0), s_cnt);
        if (TWO_PASS) count_runs(it1, item_runs(it1), s_cnt);
    }
    __syncthreads();
    unsigned *bc = a.ctrl[slot].front.bcnt;
    for (int i = threadIdx.x; i < 2 * kNB; i += blockDim.x) {
        unsigned v = s_cnt[i];
        if (v) atomicAdd(&bc[i], v);
    }
}

// one CTA of kNB threads per slot
__global__ void __launch_bounds__(kNB) k_front_scan(FrameCtrl *ctrl, FrontArgs f)
{
    const int slot = blockIdx.x;
    FrontCtrl &fc = ctrl[slot].front;
    unsigned *cmap = f.cmap0 + (size_t)slot * f.cmap_stride;
    __shared__ unsigned s_warp[33];
    const unsigned t = threadIdx.x;
    unsigned chunk_carry = 0;
    for (int r = 0; r < 2; ++r) {
        const unsigned c = fc.bcnt[r * kNB + t];
        unsigned total, tch;
        const unsigned ex = block_excl_scan(c, s_warp, total);
        fc.bstart[r][t] = ex;
        fc.bcur[r * kNB + t] = ex;
        if (t == kNB - 1) fc.bstart[r][kNB] = total;
        const unsigned nch = (c + kChunk - 1) / kChunk;
        const unsigned cex = block_excl_scan(nch, s_warp, tch) + chunk_carry;
        for (unsigned j = 0; j < nch; ++j)
            if (cex + j < f.cmap_cap) cmap[cex + j] = (unsigned)(r * kNB + t) | (j << 11);
        chunk_carry += tch;
    }
    if (t == 0) fc.nchunks = chunk_carry < f.cmap_cap ? chunk_carry : f.cmap_cap;
}

#ifndef RCT_FPT
#define RCT_FPT 512
#endif
constexpr int kPartThreads = RCT_FPT;
template <bool RESAMPLE, bool TWO_PASS>
#ifndef RCT_FPB
#define RCT_FPB (1024 / kPartThreads)
#endif
__global__ void __launch_bounds__(kPartThreads, RCT_FPB) k_front_part(PixArgs a, FrontArgs f)
{
    const int slot = blockIdx.y;
    const int frame = slot / a.lpf, label = a.label + slot % a.lpf;  // fused label pass: slot = (frame, label)
    const uint8_t *content = a.content + (size_t)frame * a.content_stride;
    const uint8_t *sty = a.sty + (size_t)frame * a.sty_stride;
    const uint8_t *mask = a.mask ? a.mask + (size_t)frame * a.mask_stride : nullptr;
    FrontCtrl &fc = a.ctrl[slot].front;
    unsigned *items_t = f.items0 + (size_t)(2 * slot) * f.items_stride;
    unsigned *items_s = items_t + f.items_stride;
    __shared__ unsigned s_cnt[2 * kNB];
    __shared__ unsigned s_base[2 * kNB];
    const size_t npix = a.p_end - a.p_begin;
    const size_t ngroups = (npix + 15) / 16;
    const size_t ntiles = (ngroups + kPartThreads - 1) / kPartThreads;
    const bool vec_c = (((uintptr_t)(content + 3 * a.p_begin)) & 15) == 0;
    const bool vec_s = !RESAMPLE && (((uintptr_t)(sty + 3 * a.p_begin)) & 15) == 0;
    const bool vec_m = mask && (((uintptr_t)(mask + a.p_begin)) & 15) == 0;
    constexpr int npass = TWO_PASS ? 2 : 1;  // MASK mode: one item per pixel
    for (size_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const size_t g = tile * kPartThreads + threadIdx.x;
        const bool live = g < ngroups;
        const size_t p0 = a.p_begin + 16 * (live ? g : 0);
        const int nvalid = live ? (int)min((size_t)16, a.p_end - p0) : 0;
        uint32_t it0[16], it1[16];
        {
            uint32_t c[16], s[16], m[16];
            front_load<RESAMPLE>(a, content, sty, mask, p0, nvalid, vec_c, vec_s, vec_m, c, s, m);
            front_items(a.mode, label, c, s, m, mask != nullptr, nvalid, 0, it0);
            if (TWO_PASS) front_items(a.mode, label, c, s, m, mask != nullptr, nvalid, 1, it1);
        }
        for (int pass = 0; pass < npass; ++pass) {
            for (int i = threadIdx.x; i < 2 * kNB; i += kPartThreads) s_cnt[i] = 0;
            __syncthreads();
            uint32_t it[16];
#pragma unroll
            for (int k = 0; k < 16; ++k) it[k] = (TWO_PASS && pass) ? it1[k] : it0[k];
            // rank of every item inside its (tile, bucket) run: the count atomic returns it (one
            // shared-memory atomic per run of equal buckets), packed two 16-bit ranks per register
            const unsigned long long lens = item_runs(it);
            unsigned rk[8];
            {
                unsigned base = 0, off = 0;
#pragma unroll
                for (int k = 0; k < 16; ++k) {
                    if (run_leader(it, k)) {
                        base = atomicAdd(&s_cnt[it[k] >> kBucketShift], run_len(lens, k));
                        off = 0;
                    } else {
                        ++off;
                    }
                    const unsigned rank = (base + off) & 0xffffu;  // < tile size (8192)
                    if (k & 1) rk[k >> 1] |= rank << 16;
                    else rk[k >> 1] = rank;
                }
            }
            __syncthreads();
            // one reservation per bucket the tile touches
            for (int i = threadIdx.x; i < 2 * kNB; i += kPartThreads) {
                const unsigned n = s_cnt[i];
                if (n) s_base[i] = atomicAdd(&fc.bcur[i], n);
            }
            __syncthreads();
#pragma unroll
            for (int k = 0; k < 16; ++k) {
                if (it[k] != kInvalidItem) {
                    const unsigned pos = s_base[it[k] >> kBucketShift] + ((rk[k >> 1] >> (16 * (k & 1))) & 0xffffu);
                    unsigned *dst = (it[k] >> 24) ? items_s : items_t;
                    if (pos < f.items_cap) dst[pos] = it[k] & 0x00ffffffu;
                }
            }
            __syncthreads();
        }
    }
}

#ifndef RCT_FHT
#define RCT_FHT 512
#endif
constexpr int kFrontHistThreads = RCT_FHT;
constexpr int kSmallSlotsLog2 = 11, kSmallSlots = 1 << kSmallSlotsLog2;  // open-addressing table of the small-bucket path
#ifndef RCT_SMALLB
#define RCT_SMALLB 1024
#endif
constexpr int kSmallBucketItems = RCT_SMALLB;                             // items a bucket may hold to take it (<= kSmallSlots / 2)
constexpr size_t kSmemFrontHist = (size_t)kBucketBins * 4;
__global__ void __launch_bounds__(kFrontHistThreads) k_front_hist(FrameCtrl *ctrl, const Job *jobs, FrontArgs f,
                                                                  unsigned *hist0, size_t hist_stride)
{
    const int slot = blockIdx.y;
    FrontCtrl &fc = ctrl[slot].front;
    if (blockIdx.x >= fc.nchunks) return;
    const unsigned ent = (f.cmap0 + (size_t)slot * f.cmap_stride)[blockIdx.x];
    const unsigned rb = ent & (2u * kNB - 1u), chunk = ent >> 11;
    const unsigned r = rb / kNB, b = rb % kNB;
    const unsigned start = fc.bstart[r][b], end = fc.bstart[r][b + 1];
    const unsigned nch = (end - start + kChunk - 1) / kChunk;
    const unsigned i0 = start + chunk * kChunk;
    const unsigned i1 = (end - i0 > kChunk) ? i0 + kChunk : end;
    extern __shared__ unsigned s_hist[];  // [kBucketBins]
    __shared__ unsigned s_wbase[kFrontHistThreads / 32];
    __shared__ unsigned s_gbase, s_last;
    const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    if (nch == 1 && i1 - i0 <= (unsigned)kSmallBucketItems) {
        // Small bucket (label maps, sparse colour ranges): zeroing and sweeping 16 384 bins for a few hundred items
        // was 2.0 of the 14 ms of the 8-label step.  An open-addressing table of 2 048 slots in the first 16 KB of
        // the same shared memory instead: slot = (bin + 1) | count << 15 ... kept as two words (key, count).
        unsigned *s_key = s_hist, *s_cnt = s_hist + kSmallSlots;
        for (unsigned i = tid; i < 2u * kSmallSlots; i += kFrontHistThreads) s_hist[i] = 0u;
        __syncthreads();
        const unsigned *items = f.items0 + (size_t)(2 * slot + r) * f.items_stride;
        for (unsigned i = i0 + tid; i < i1; i += kFrontHistThreads) {
            const unsigned key = (__ldcs(items + i) & (unsigned)(kBucketBins - 1)) + 1u;  // 0 = empty slot
            unsigned h = (key * 2654435761u) >> (32 - kSmallSlotsLog2);
            for (;;) {
                const unsigned prev = atomicCAS(&s_key[h], 0u, key);
                if (prev == 0u || prev == key) {
                    atomicAdd(&s_cnt[h], 1u);
                    break;
                }
                h = (h + 1u) & (unsigned)(kSmallSlots - 1);  // at most 1 024 distinct keys in 2 048 slots: terminates
            }
        }
        __syncthreads();
        unsigned nz_warp = 0;
        for (unsigned j = 0; j < (unsigned)kSmallSlots / kFrontHistThreads; ++j)
            nz_warp += __popc(__ballot_sync(0xffffffffu, s_key[j * kFrontHistThreads + tid] != 0u));
        if (lane == 0) s_wbase[warp] = nz_warp;
        __syncthreads();
        RegionCtrl *rcs = &ctrl[slot].r[r];
        if (tid == 0) {
            unsigned run = 0;
            for (int w = 0; w < kFrontHistThreads / 32; ++w) {
                const unsigned t = s_wbase[w];
                s_wbase[w] = run;
                run += t;
            }
            s_gbase = run ? atomicAdd(&rcs->d, run) : 0u;
        }
        __syncthreads();
        const Job &jobs_ = jobs[2 * slot + r];
        unsigned *col = const_cast<unsigned *>(jobs_.payA), *cnt = const_cast<unsigned *>(jobs_.cntA);
        unsigned pos = s_gbase + s_wbase[warp];
        for (unsigned j = 0; j < (unsigned)kSmallSlots / kFrontHistThreads; ++j) {
            const unsigned sl = j * kFrontHistThreads + tid;
            const unsigned k = s_key[sl];
            const unsigned bal = __ballot_sync(0xffffffffu, k != 0u);
            if (k) {
                const unsigned q = pos + __popc(bal & ((1u << lane) - 1u));
                if (q < jobs_.cap) {
                    col[q] = (b << kBucketShift) | (k - 1u);
                    cnt[q] = s_cnt[sl];
                }
            }
            pos += __popc(bal);
        }
        return;
    }
    uint4 *s_hist4 = reinterpret_cast<uint4 *>(s_hist);
    for (unsigned i = tid; i < (unsigned)kBucketBins / 4; i += kFrontHistThreads) s_hist4[i] = make_uint4(0u, 0u, 0u, 0u);
    __syncthreads();
    const unsigned *items = f.items0 + (size_t)(2 * slot + r) * f.items_stride;
    {
        constexpr unsigned kMask = (unsigned)(kBucketBins - 1);
        unsigned a0 = (i0 + 3u) & ~3u, a1 = i1 & ~3u;  // 16-byte aligned body [a0, a1)
        if (a0 > i1) a0 = i1;
        if (a1 < a0) a1 = a0;
        if (i0 + tid < a0) atomicAdd(&s_hist[__ldcs(items + i0 + tid) & kMask], 1u);
        if (a1 + tid < i1) atomicAdd(&s_hist[__ldcs(items + a1 + tid) & kMask], 1u);
        const uint4 *items4 = reinterpret_cast<const uint4 *>(items);
        for (unsigned q = a0 / 4 + tid; q < a1 / 4; q += kFrontHistThreads) {
            const uint4 v = __ldcs(items4 + q);
            atomicAdd(&s_hist[v.x & kMask], 1u);
            atomicAdd(&s_hist[v.y & kMask], 1u);
            atomicAdd(&s_hist[v.z & kMask], 1u);
            atomicAdd(&s_hist[v.w & kMask], 1u);
        }
    }
    __syncthreads();
    if (nch > 1) {
        // merge through the bucket's window of the colour cube; the last CTA to arrive compacts
        unsigned *window = hist0 + (size_t)slot * hist_stride + (size_t)r * kCube + ((size_t)b << kBucketShift);
        for (unsigned i = tid; i < (unsigned)kBucketBins; i += kFrontHistThreads) {
            const unsigned v = s_hist[i];
            if (v) atomicAdd(&window[i], v);
        }
        __threadfence();
        __syncthreads();
        if (tid == 0) s_last = (atomicAdd(&fc.bticket[rb], 1u) == nch - 1u) ? 1u : 0u;
        __syncthreads();
        if (!s_last) return;
        __threadfence();
        for (unsigned i = tid; i < (unsigned)kBucketBins; i += kFrontHistThreads) {
            s_hist[i] = __ldcg(&window[i]);
            window[i] = 0u;
        }
        __syncthreads();
    }
    // compaction: bins are taken in warp-interleaved order (the list order is irrelevant)
    unsigned nz_warp = 0;
    for (unsigned j = 0; j < (unsigned)kBucketBins / kFrontHistThreads; ++j)
        nz_warp += __popc(__ballot_sync(0xffffffffu, s_hist[j * kFrontHistThreads + tid] != 0u));
    if (lane == 0) s_wbase[warp] = nz_warp;
    __syncthreads();
    RegionCtrl *rc = &ctrl[slot].r[r];
    if (tid == 0) {
        unsigned run = 0;
        for (int w = 0; w < kFrontHistThreads / 32; ++w) {
            const unsigned t = s_wbase[w];
            s_wbase[w] = run;
            run += t;
        }
        s_gbase = run ? atomicAdd(&rc->d, run) : 0u;
    }
    __syncthreads();
    const Job &job = jobs[2 * slot + r];
    unsigned *col = const_cast<unsigned *>(job.payA), *cnt = const_cast<unsigned *>(job.cntA);
    unsigned pos = s_gbase + s_wbase[warp];
    for (unsigned j = 0; j < (unsigned)kBucketBins / kFrontHistThreads; ++j) {
        const unsigned bin = j * kFrontHistThreads + tid;
        const unsigned v = s_hist[bin];
        const unsigned bal = __ballot_sync(0xffffffffu, v != 0u);
        if (v) {
            const unsigned p = pos + __popc(bal & ((1u << lane) - 1u));
            if (p < job.cap) {
                col[p] = (b << kBucketShift) | bin;
                cnt[p] = v;
            }
        }
        pos += __popc(bal);
    }
}

// parity tests: rebuild the colour-cube histogram of one region from its (colour, count) list
__global__ void __launch_bounds__(256) k_list_to_cube(const FrameCtrl *ctrl, const Job *jobs, int slot, int region,
                                                      unsigned *cube)
{
    const Job &job = jobs[2 * slot + region];
    const unsigned d = ctrl[slot].r[region].d;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < d; e += (size_t)gridDim.x * blockDim.x)
        cube[job.payA[e]] = job.cntA[e];
}
