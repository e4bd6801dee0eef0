// rct_apply.cuh -- the fused apply pass (a-7..a-9) as a bulk-async (TMA 1-D) staged pipeline.
// Included by rct_kernels.cuh inside namespace rct (needs PixArgs, FrameCtrl).
//
// k_apply (rct_kernels.cuh) keeps a thread's 112 input bytes in registers across two dependent
// round trips (pixels, then LUT gathers) and is latency-bound at 1/3 occupancy.  Here one elected
// thread streams content / stylised / mask tiles into a 3-stage shared-memory ring with
// cp.async.bulk (completion on an mbarrier), all threads resolve their 16 pixels against the
// colour LUT, and the output tile leaves through a 2-stage ring with cp.async.bulk stores:
// loads of tiles i+1, i+2 and the store of tile i-1 are in flight while tile i is looked up.
// Tiles are walked slot-major by the whole grid, so one frame's LUT is hot in L2 at a time.
//
// MASK mode, no resample, 16-byte aligned frames only (the reference's own geometry); anything
// else -- and the last npix % 16 pixels -- goes through k_apply.
#pragma once

// k_apply: a-7..a-9 for pixels: target non-black -> LUT[colour]; target black -> black;
// MASK mode non-target -> stylised pixel (background_np); LABEL mode non-target untouched.
// LUTM: 0 colour table, 1 uint16 run index + palette.  (The key table of rct_keytab.cuh is not used here: with the
// resample gathers in the same thread the extra arithmetic made this kernel slower, 0.143 vs 0.115 ms on 8 x 1080p.)
template <bool RESAMPLE, int LUTM = 0>
__global__ void __launch_bounds__(256, 3) k_apply(PixArgs a)
{
    const int slot = blockIdx.y;
    const int frame = slot / a.lpf, label = a.label + slot % a.lpf;  // fused label pass: slot = (frame, label)
    const uint8_t *content = a.content + (size_t)frame * a.content_stride;
    const uint8_t *sty = a.sty + (size_t)frame * a.sty_stride;
    const uint8_t *mask = a.mask ? a.mask + (size_t)frame * a.mask_stride : nullptr;
    uint8_t *out = a.out + (size_t)frame * a.out_stride;
    const unsigned *lut = a.lut + (size_t)slot * a.lut_stride;
    const int status = a.ctrl[slot].status;
    if (a.status_out && blockIdx.x == 0 && threadIdx.x == 0) a.status_out[slot] = status;
    const size_t npix = a.p_end - a.p_begin;
    const size_t ngroups = (npix + 15) / 16;
    const bool vec_c = (((uintptr_t)(content + 3 * a.p_begin)) & 15) == 0;
    const bool vec_s = !RESAMPLE && (((uintptr_t)(sty + 3 * a.p_begin)) & 15) == 0;
    const bool vec_m = mask && (((uintptr_t)(mask + a.p_begin)) & 15) == 0;
    const bool vec_o = (((uintptr_t)(out + 3 * a.p_begin)) & 15) == 0;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x; g < ngroups; g += stride) {
        size_t p0 = a.p_begin + 16 * g;
        int nvalid = (int)min((size_t)16, a.p_end - p0);
        uint32_t c[16], s[16], m[16], o[16];
        load16_rgb(content, p0, nvalid, vec_c, c);
        if (mask) load16_u8(mask, p0, nvalid, vec_m, m);
        if (a.mode == 0) {
            if (RESAMPLE) gather16_rgb(sty, a.xtab, a.ytab, a.W, a.sw, p0, nvalid, s);
            else load16_rgb(sty, p0, nvalid, vec_s, s);
        }
        // every lane gathers (entry 0 where no lookup is wanted) and selects afterwards: gathers
        // inside divergent branches serialise on their latency (see rct_apply.cuh)
        unsigned keep = 0, tgt = 0;
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            bool in_t, in_s;
            region_flags(a.mode, label, mask ? m[k] : 0u, in_t, in_s);
            if (in_t) tgt |= 1u << k;
            if (in_t || a.mode == 0) keep |= 1u << k;
        }
        uint32_t t[16];
#pragma unroll
        for (int k = 0; k < 16; ++k)
        {
            const uint32_t idx = (((tgt >> k) & 1u) && status == 0 && k < nvalid) ? c[k] : 0u;
            t[k] = LUTM == 1 ? (uint32_t)__ldg(reinterpret_cast<const unsigned short *>(lut) + idx) : __ldg(lut + idx);
        }
        if (LUTM == 1) {
#pragma unroll
            for (int k = 0; k < 16; ++k) t[k] = __ldg(lut + kPaletteOffsetU32 + t[k]);
        }
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            uint32_t v = 0;
            if ((tgt >> k) & 1u) {
                if (c[k] != 0u && k < nvalid) v = (status == 0) ? t[k] : c[k];
            } else if (a.mode == 0) {
                v = s[k];
            }
            o[k] = v;
        }
        store16_rgb(out, p0, nvalid, vec_o, o, keep);
    }
}



constexpr int kApThreads = 256;
constexpr int kApTile = kApThreads * 16;  // pixels per tile
constexpr int kApStages = 3;
constexpr int kApOut = 2;

struct ApSmem {
    alignas(128) uint8_t c[kApStages][kApTile * 3];
    alignas(128) uint8_t s[kApStages][kApTile * 3];
    alignas(128) uint8_t m[kApStages][kApTile];
    alignas(128) uint8_t o[kApOut][kApTile * 3];
    alignas(8) unsigned long long full[kApStages];
};

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src_gmem, unsigned bytes, unsigned long long *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void bulk_s2g(void *dst_gmem, const void *src_smem, unsigned bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem), "r"(smem_u32(src_smem)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read()
{
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void bulk_wait()
{
    asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
// LUT gather that does not allocate in L1: with ~200 KB of the SM carved out as shared memory
// the L1 is a few dozen lines, and an allocating gather can only be in flight while it holds one
__device__ __forceinline__ uint32_t ldg_gather(const unsigned *p)
{
    uint32_t v;
    asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}
// L2 eviction policies: the image streams are touched once (evict first), a frame's colour LUT is
// gathered from again and again while the frame is walked (evict last)
__device__ __forceinline__ unsigned long long l2_policy_evict_first()
{
    unsigned long long pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ unsigned long long l2_policy_evict_last()
{
    unsigned long long pol;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ void bulk_g2s_hint(void *dst_smem, const void *src_gmem, unsigned bytes,
                                              unsigned long long *bar, unsigned long long pol)
{
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
            smem_u32(dst_smem)),
        "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)), "l"(pol)
        : "memory");
}
__device__ __forceinline__ uint32_t ldg_hint(const unsigned *p, unsigned long long pol)
{
    uint32_t v;
    asm volatile("ld.global.nc.L2::cache_hint.u32 %0, [%1], %2;" : "=r"(v) : "l"(p), "l"(pol));
    return v;
}
__device__ __forceinline__ uint32_t ldg_u16_hint(const unsigned short *p, unsigned long long pol)
{
    unsigned short v;
    asm volatile("ld.global.nc.L2::cache_hint.u16 %0, [%1], %2;" : "=h"(v) : "l"(p), "l"(pol));
    return (uint32_t)v;
}
__device__ __forceinline__ void stg8_hint(void *p, uint32_t x, uint32_t y, unsigned long long pol)
{
    asm volatile("st.global.L1::no_allocate.L2::cache_hint.v2.u32 [%0], {%1,%2}, %3;" ::"l"(p), "r"(x), "r"(y), "l"(pol)
                 : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }

// 12 little-endian words = 16 RGB pixels <-> 16 packed colours r | g<<8 | b<<16
__device__ __forceinline__ void unpack16(const uint32_t w[12], uint32_t c[16])
{
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        c[4 * q + 0] = w[3 * q] & 0xffffffu;
        c[4 * q + 1] = __funnelshift_r(w[3 * q], w[3 * q + 1], 24) & 0xffffffu;
        c[4 * q + 2] = __funnelshift_r(w[3 * q + 1], w[3 * q + 2], 16) & 0xffffffu;
        c[4 * q + 3] = w[3 * q + 2] >> 8;
    }
}
__device__ __forceinline__ void pack16(const uint32_t c[16], uint32_t w[12])
{
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        w[3 * q + 0] = c[4 * q] | (c[4 * q + 1] << 24);
        w[3 * q + 1] = (c[4 * q + 1] >> 8) | (c[4 * q + 2] << 16);
        w[3 * q + 2] = (c[4 * q + 2] >> 16) | (c[4 * q + 3] << 8);
    }
}

// 16 pixels (12 words each of content / stylised, 4 words of mask) -> 12 output words.
// Target pixels (mask 0) take LUT[colour] (black stays black; a frame with an empty region keeps
// its content), the others the stylised pixel.  The gathers go out in two batches of eight with
// little else live: given too few registers the compiler recycles the address registers of
// loads still in flight and the gathers serialise on their latency (measured: 21 us per tile).
__device__ __forceinline__ void apply16(const uint32_t wc[12], const uint32_t ws[12], const uint32_t wm[4],
                                        const unsigned *lut, int status, uint32_t wo[12])
{
    uint32_t c[16];
    unpack16(wc, c);
    unsigned tgt = 0, want = 0;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        const uint32_t mk = (wm[k >> 2] >> (8 * (k & 3))) & 255u;
        if (mk == 0u) {
            tgt |= 1u << k;
            if (c[k] != 0u && status == 0) want |= 1u << k;
        }
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        uint32_t t[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) t[k] = __ldg(lut + (((want >> (8 * h + k)) & 1u) ? c[8 * h + k] : 0u));
#pragma unroll
        for (int k = 0; k < 8; ++k)
            if ((want >> (8 * h + k)) & 1u) c[8 * h + k] = t[k];
    }
    uint32_t s[16];
    unpack16(ws, s);
#pragma unroll
    for (int k = 0; k < 16; ++k)
        if (!((tgt >> k) & 1u)) c[k] = s[k];
    pack16(c, wo);
}

// a.p_begin and (a.p_end - a.p_begin) are multiples of 16; all frame bases 16-byte aligned.
__global__ void __launch_bounds__(kApThreads, 2) k_apply_tma(PixArgs a, unsigned tiles_per_frame, unsigned nslots)
{
    extern __shared__ __align__(128) uint8_t ap_smem_raw[];
    ApSmem &S = *reinterpret_cast<ApSmem *>(ap_smem_raw);
    const unsigned tid = threadIdx.x;
    const unsigned total = tiles_per_frame * nslots;
    if (a.status_out && blockIdx.x == 0)
        for (unsigned q = tid; q < nslots; q += kApThreads) a.status_out[q] = a.ctrl[q].status;
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < kApStages; ++s) mbar_init(&S.full[s], 1u);
        fence_mbar_init();
        fence_proxy_async();
    }
    __syncthreads();
    const unsigned G = gridDim.x;
    const unsigned cnt = blockIdx.x < total ? (total - blockIdx.x + G - 1) / G : 0u;

    auto tile_geom = [&](unsigned i, unsigned &slot, size_t &p0, unsigned &n) {
        const unsigned T = blockIdx.x + i * G;
        slot = T / tiles_per_frame;
        const unsigned t = T - slot * tiles_per_frame;
        p0 = a.p_begin + (size_t)t * kApTile;
        const size_t left = a.p_end - p0;
        n = left < (size_t)kApTile ? (unsigned)left : (unsigned)kApTile;
    };
    auto issue = [&](unsigned i) {
        unsigned slot, n;
        size_t p0;
        tile_geom(i, slot, p0, n);
        const unsigned st = i % kApStages;
        mbar_expect_tx(&S.full[st], 7u * n);
        bulk_g2s(S.c[st], a.content + (size_t)slot * a.content_stride + 3 * p0, 3u * n, &S.full[st]);
        bulk_g2s(S.s[st], a.sty + (size_t)slot * a.sty_stride + 3 * p0, 3u * n, &S.full[st]);
        bulk_g2s(S.m[st], a.mask + (size_t)slot * a.mask_stride + p0, n, &S.full[st]);
    };

    if (tid == 0) {
        for (unsigned i = 0; i < (unsigned)(kApStages - 1) && i < cnt; ++i) issue(i);
    }
    for (unsigned i = 0; i < cnt; ++i) {
        const unsigned st = i % kApStages, par = (i / kApStages) & 1u, ob = i % kApOut;
        if (tid == 0) {
            if (i + kApStages - 1 < cnt) issue(i + kApStages - 1);  // refills the stage consumed in iteration i-1
            bulk_wait_read<kApOut - 1>();                           // the store that last used o[ob] has read it
        }
        unsigned slot, n;
        size_t p0;
        tile_geom(i, slot, p0, n);
        mbar_wait(&S.full[st], par);
        const bool active = tid * 16u < n;
        uint32_t wc[12], ws[12], wm[4];
        if (active) {
            const uint4 *qc = reinterpret_cast<const uint4 *>(S.c[st] + 48u * tid);
            const uint4 *qs = reinterpret_cast<const uint4 *>(S.s[st] + 48u * tid);
            const uint4 x0 = qc[0], x1 = qc[1], x2 = qc[2], y0 = qs[0], y1 = qs[1], y2 = qs[2];
            const uint4 z = *reinterpret_cast<const uint4 *>(S.m[st] + 16u * tid);
            wc[0] = x0.x; wc[1] = x0.y; wc[2] = x0.z; wc[3] = x0.w; wc[4] = x1.x; wc[5] = x1.y; wc[6] = x1.z; wc[7] = x1.w;
            wc[8] = x2.x; wc[9] = x2.y; wc[10] = x2.z; wc[11] = x2.w;
            ws[0] = y0.x; ws[1] = y0.y; ws[2] = y0.z; ws[3] = y0.w; ws[4] = y1.x; ws[5] = y1.y; ws[6] = y1.z; ws[7] = y1.w;
            ws[8] = y2.x; ws[9] = y2.y; ws[10] = y2.z; ws[11] = y2.w;
            wm[0] = z.x; wm[1] = z.y; wm[2] = z.z; wm[3] = z.w;
        }
        __syncthreads();  // stage st is free for the producer; o[ob] is free for the writers
        if (active) {
            const unsigned *lut = a.lut + (size_t)slot * a.lut_stride;
            const int status = a.ctrl[slot].status;
            uint32_t wo[12];
            apply16(wc, ws, wm, lut, status, wo);
            uint4 *qo = reinterpret_cast<uint4 *>(S.o[ob] + 48u * tid);
            qo[0] = make_uint4(wo[0], wo[1], wo[2], wo[3]);
            qo[1] = make_uint4(wo[4], wo[5], wo[6], wo[7]);
            qo[2] = make_uint4(wo[8], wo[9], wo[10], wo[11]);
        }
        fence_proxy_async();
        __syncthreads();
        if (tid == 0) {
            bulk_s2g(a.out + (size_t)slot * a.out_stride + 3 * p0, S.o[ob], 3u * n);
            bulk_commit();
        }
    }
    if (tid == 0) bulk_wait<0>();
}

// ------------------------------------------------------------------------------------------
// k_apply_warp: the same pass with one independent 2-stage pipeline PER WARP -- no CTA barrier
// anywhere, so the warps of an SM drift apart and their phases (tile load in flight, LUT gathers,
// output stores) overlap instead of marching in lock step.  A warp tile is 32 * PXL pixels
// (PXL = 8: 1.75 KB of input, 16: 3.5 KB).  The prefetch depth is deliberately shallow: every byte
// of streaming traffic queued in L2 / DRAM ahead of a LUT gather adds to the gather's latency, and
// the gathers (not the streams) are what a tile waits for.
//   BULK = true : lane 0 issues three cp.async.bulk copies per tile, completion on the warp's own
//                 mbarrier (TMA 1-D);
//   BULK = false: every lane issues 16-byte cp.async (LDGSTS) copies, cp.async.wait_group.
// Output goes straight from registers to global memory (3 * PXL contiguous bytes per lane).
// ------------------------------------------------------------------------------------------
constexpr int kAwStages = 2;
template <int PXL>
struct AwCfg {
    static constexpr int kWarps = PXL == 8 ? 8 : 7;
    static constexpr int kThreads = kWarps * 32;
    static constexpr int kTile = 32 * PXL;
    static constexpr int kMinBlocks = PXL == 8 ? 4 : 3;
};
template <int PXL, int STG = 2, bool OST = false>
struct AwWarpSmem {
    alignas(128) uint8_t c[STG][32 * PXL * 3];
    alignas(128) uint8_t s[STG][32 * PXL * 3];
    alignas(128) uint8_t m[STG][32 * PXL];
    alignas(128) uint8_t o[OST ? 2 : 1][OST ? 32 * PXL * 3 : 128];  // output tiles of the bulk-store flavour
};
template <int PXL, int STG = 2, bool OST = false>
struct AwSmem {
    AwWarpSmem<PXL, STG, OST> w[AwCfg<PXL>::kWarps];
    alignas(8) unsigned long long full[AwCfg<PXL>::kWarps][STG];
};

__device__ __forceinline__ void cp_async16(void *dst_smem, const void *src_gmem)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst_smem)), "l"(src_gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait()
{
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// 8 pixels: 6 words each of content / stylised, 2 words of mask -> 6 output words
// LUTM: 0 colour -> RGB table, 1 uint16 run index + palette, 2 key table in front of the colour table
// (rct_keytab.cuh: fp32 key -> cell -> colour, flagged cells fall back to the table)
template <bool HINT, int LUTM = 0>
__device__ __forceinline__ void apply8(const uint32_t wc[6], const uint32_t ws[6], const uint32_t wm[2],
                                       const unsigned *lut, int status, uint32_t wo[6], unsigned long long pol_lut,
                                       const KtRegs *kr = nullptr, const KtView *kv = nullptr)
{
    uint32_t c[8], s[8];
    if (LUTM == 2) {
        // a lane whose eight pixels are all background passes the stylised words through: no unpacking at all
        // (regions are contiguous in real masks, so most lanes are all-background or all-target)
        const uint32_t z0 = wm[0], z1 = wm[1];
        if (__vcmpeq4(z0, 0u) == 0u && __vcmpeq4(z1, 0u) == 0u) {
#pragma unroll
            for (int q = 0; q < 6; ++q) wo[q] = ws[q];
            return;
        }
    }
#pragma unroll
    for (int q = 0; q < 2; ++q) {
        c[4 * q + 0] = wc[3 * q] & 0xffffffu;
        c[4 * q + 1] = __funnelshift_r(wc[3 * q], wc[3 * q + 1], 24) & 0xffffffu;
        c[4 * q + 2] = __funnelshift_r(wc[3 * q + 1], wc[3 * q + 2], 16) & 0xffffffu;
        c[4 * q + 3] = wc[3 * q + 2] >> 8;
    }
    unsigned tgt = 0, want = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const uint32_t mk = (wm[k >> 2] >> (8 * (k & 3))) & 255u;
        if (mk == 0u) {
            tgt |= 1u << k;
            if (c[k] != 0u && status == 0) want |= 1u << k;
        }
    }
    uint32_t t[8];
    if (LUTM == 2) {
        if (want) {  // one branch around the whole batch: the eight table reads stay independent
#pragma unroll
            for (int k = 0; k < 8; ++k) t[k] = kt_lookup(*kr, *kv, c[k]);
            unsigned amb = 0;
#pragma unroll
            for (int k = 0; k < 8; ++k)
                if (((want >> k) & 1u) && (t[k] & kKtFlag)) amb |= 1u << k;
            if (amb) {  // flagged cells (a few per cent of the pixels): the exact table
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const bool fb = (amb >> k) & 1u;
                    const unsigned *g = lut + (fb ? c[k] : 0u);
#ifdef RCT_KT_PLAINLDG
                    const uint32_t v = __ldg(g);
#else
                    const uint32_t v = HINT ? ldg_hint(g, pol_lut) : __ldg(g);
#endif
                    if (fb) t[k] = v;
                }
            }
        }
        if (tgt == 0xffu) {  // all eight pixels are target pixels: the stylised words are not needed
#pragma unroll
            for (int k = 0; k < 8; ++k) c[k] = ((want >> k) & 1u) ? t[k] : c[k];
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                wo[3 * q + 0] = c[4 * q] | (c[4 * q + 1] << 24);
                wo[3 * q + 1] = (c[4 * q + 1] >> 8) | (c[4 * q + 2] << 16);
                wo[3 * q + 2] = (c[4 * q + 2] >> 16) | (c[4 * q + 3] << 8);
            }
            return;
        }
    } else if (LUTM) {  // uint16 run index (L2-resident table), then the palette (L1-resident)
        const unsigned short *l16 = reinterpret_cast<const unsigned short *>(lut);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const unsigned short *g = l16 + (((want >> k) & 1u) ? c[k] : 0u);
            t[k] = HINT ? ldg_u16_hint(g, pol_lut) : (uint32_t)__ldg(g);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) t[k] = __ldg(lut + kPaletteOffsetU32 + t[k]);
    } else {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const unsigned *g = lut + (((want >> k) & 1u) ? c[k] : 0u);
            t[k] = HINT ? ldg_hint(g, pol_lut) : __ldg(g);
        }
    }
#pragma unroll
    for (int q = 0; q < 2; ++q) {
        s[4 * q + 0] = ws[3 * q] & 0xffffffu;
        s[4 * q + 1] = __funnelshift_r(ws[3 * q], ws[3 * q + 1], 24) & 0xffffffu;
        s[4 * q + 2] = __funnelshift_r(ws[3 * q + 1], ws[3 * q + 2], 16) & 0xffffffu;
        s[4 * q + 3] = ws[3 * q + 2] >> 8;
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        if ((want >> k) & 1u) c[k] = t[k];
        if (!((tgt >> k) & 1u)) c[k] = s[k];
    }
#pragma unroll
    for (int q = 0; q < 2; ++q) {
        wo[3 * q + 0] = c[4 * q] | (c[4 * q + 1] << 24);
        wo[3 * q + 1] = (c[4 * q + 1] >> 8) | (c[4 * q + 2] << 16);
        wo[3 * q + 2] = (c[4 * q + 2] >> 16) | (c[4 * q + 3] << 8);
    }
}

__device__ __forceinline__ void stg_stream8(void *p, uint32_t x, uint32_t y)
{
    asm volatile("st.global.L1::no_allocate.v2.u32 [%0], {%1,%2};" ::"l"(p), "r"(x), "r"(y) : "memory");
}

__device__ __forceinline__ void bulk_s2g_hint(void *dst_gmem, const void *src_smem, unsigned bytes, unsigned long long pol)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;" ::"l"(dst_gmem),
                 "r"(smem_u32(src_smem)), "r"(bytes), "l"(pol)
                 : "memory");
}

// OST = true: the output tile goes through shared memory and leaves with ONE cp.async.bulk store
// per tile (the register flavour's three 8-byte stores per lane, 24 bytes apart, touch every 32-byte
// sector of the tile three times: ncu shows the L1 -> crossbar request port 83 % busy on image N)
template <bool BULK, int PXL, int MINB, int STG = 2, bool OST = false, int LUTM = 0>
__global__ void __launch_bounds__(AwCfg<PXL>::kThreads, MINB)
    k_apply_warp(PixArgs a, unsigned tiles_per_frame, unsigned nslots)
{
    constexpr int kWarps = AwCfg<PXL>::kWarps, kTile = AwCfg<PXL>::kTile;
    extern __shared__ __align__(128) uint8_t aw_smem_raw[];
    AwSmem<PXL, STG, OST> &S = *reinterpret_cast<AwSmem<PXL, STG, OST> *>(aw_smem_raw);
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    AwWarpSmem<PXL, STG, OST> &W = S.w[warp];
    const unsigned total = tiles_per_frame * nslots;
    if (a.status_out && blockIdx.x == 0)
        for (unsigned q = threadIdx.x; q < nslots; q += AwCfg<PXL>::kThreads) a.status_out[q] = a.ctrl[q].status;
    if (BULK) {
        if (lane == 0) {
#pragma unroll
            for (int s = 0; s < STG; ++s) mbar_init(&S.full[warp][s], 1u);
            fence_mbar_init();
            fence_proxy_async();
        }
        __syncwarp();
    }
    const unsigned GW = gridDim.x * kWarps, w0 = blockIdx.x * kWarps + warp;
    const unsigned cnt = w0 < total ? (total - w0 + GW - 1) / GW : 0u;
    const unsigned long long pol_stream = l2_policy_evict_first(), pol_lut = l2_policy_evict_last();

    auto tile_geom = [&](unsigned i, unsigned &slot, size_t &p0, unsigned &n) {
        const unsigned T = w0 + i * GW;
        slot = T / tiles_per_frame;
        const unsigned t = T - slot * tiles_per_frame;
        p0 = a.p_begin + (size_t)t * kTile;
        const size_t left = a.p_end - p0;
        n = left < (size_t)kTile ? (unsigned)left : (unsigned)kTile;
    };
    auto issue = [&](unsigned i) {
        unsigned slot, n;
        size_t p0;
        tile_geom(i, slot, p0, n);
        const unsigned st = i % STG;
        const uint8_t *gc = a.content + (size_t)slot * a.content_stride + 3 * p0;
        const uint8_t *gs = a.sty + (size_t)slot * a.sty_stride + 3 * p0;
        const uint8_t *gm = a.mask + (size_t)slot * a.mask_stride + p0;
        if (BULK) {
            if (lane == 0) {
                mbar_expect_tx(&S.full[warp][st], 7u * n);
                if (PXL == 8) {
                    bulk_g2s_hint(W.c[st], gc, 3u * n, &S.full[warp][st], pol_stream);
                    bulk_g2s_hint(W.s[st], gs, 3u * n, &S.full[warp][st], pol_stream);
                    bulk_g2s_hint(W.m[st], gm, n, &S.full[warp][st], pol_stream);
                } else {
                    bulk_g2s(W.c[st], gc, 3u * n, &S.full[warp][st]);
                    bulk_g2s(W.s[st], gs, 3u * n, &S.full[warp][st]);
                    bulk_g2s(W.m[st], gm, n, &S.full[warp][st]);
                }
            }
        } else {
            const unsigned n3 = 3u * n / 16u, n1 = n / 16u;  // 16-byte chunks
            for (unsigned q = lane; q < n3; q += 32u) {
                cp_async16(W.c[st] + 16u * q, gc + 16u * q);
                cp_async16(W.s[st] + 16u * q, gs + 16u * q);
            }
            if (lane < n1) cp_async16(W.m[st] + 16u * lane, gm + 16u * lane);
        }
    };

    static_assert(BULK || STG == 2, "the cp.async flavour is two-stage");
    unsigned kt_slot = 0xffffffffu;
    bool kt_ok = false;
    KtRegs kr{};
    KtView kv{};
#pragma unroll
    for (unsigned i = 0; i + 1 < (unsigned)STG; ++i)
        if (i < cnt) issue(i);
    if (!BULK) cp_async_commit();
    for (unsigned i = 0; i < cnt; ++i) {
        const unsigned st = i % STG;
        if (i + STG - 1 < cnt) issue(i + STG - 1);  // that stage was released by the __syncwarp ending iteration i-1
        if (BULK) {
            mbar_wait(&S.full[warp][st], (i / STG) & 1u);
        } else {
            cp_async_commit();  // possibly empty: keeps one group per iteration
            cp_async_wait<1>();
            __syncwarp();
        }
        unsigned slot, n;
        size_t p0;
        tile_geom(i, slot, p0, n);
        const bool active = lane * (unsigned)PXL < n;
        const unsigned *lut = a.lut + (size_t)slot * a.lut_stride;
        const int status = a.ctrl[slot].status;
        if (LUTM == 2 && slot != kt_slot) {  // a new frame: its key-table parameters (warp-uniform)
            kt_slot = slot;
            kv = kt_view(a.kt, slot);
            kt_ok = a.kt != nullptr && __ldg(&kv.par->valid) != 0u;
            if (kt_ok) kr = kt_load(kv.par);
        }
        if (PXL == 16) {
            uint32_t wc[12], ws[12], wm[4], wo[12];
            if (active) {
                const uint4 *qc = reinterpret_cast<const uint4 *>(W.c[st] + 48u * lane);
                const uint4 *qs = reinterpret_cast<const uint4 *>(W.s[st] + 48u * lane);
                const uint4 x0 = qc[0], x1 = qc[1], x2 = qc[2], y0 = qs[0], y1 = qs[1], y2 = qs[2];
                const uint4 z = *reinterpret_cast<const uint4 *>(W.m[st] + 16u * lane);
                wc[0] = x0.x; wc[1] = x0.y; wc[2] = x0.z; wc[3] = x0.w; wc[4] = x1.x; wc[5] = x1.y; wc[6] = x1.z;
                wc[7] = x1.w; wc[8] = x2.x; wc[9] = x2.y; wc[10] = x2.z; wc[11] = x2.w;
                ws[0] = y0.x; ws[1] = y0.y; ws[2] = y0.z; ws[3] = y0.w; ws[4] = y1.x; ws[5] = y1.y; ws[6] = y1.z;
                ws[7] = y1.w; ws[8] = y2.x; ws[9] = y2.y; ws[10] = y2.z; ws[11] = y2.w;
                wm[0] = z.x; wm[1] = z.y; wm[2] = z.z; wm[3] = z.w;
            }
            __syncwarp();  // the stage is read: the next iteration may refill it
            if (active) {
                apply16(wc, ws, wm, lut, status, wo);
                uint4 *qo = reinterpret_cast<uint4 *>(a.out + (size_t)slot * a.out_stride + 3 * p0 + 48u * lane);
                stg_stream(qo, make_uint4(wo[0], wo[1], wo[2], wo[3]));
                stg_stream(qo + 1, make_uint4(wo[4], wo[5], wo[6], wo[7]));
                stg_stream(qo + 2, make_uint4(wo[8], wo[9], wo[10], wo[11]));
            }
        } else {
            uint32_t wc[6], ws[6], wm[2], wo[6];
            if (active) {
                const uint2 *qc = reinterpret_cast<const uint2 *>(W.c[st] + 24u * lane);
                const uint2 *qs = reinterpret_cast<const uint2 *>(W.s[st] + 24u * lane);
                const uint2 x0 = qc[0], x1 = qc[1], x2 = qc[2], y0 = qs[0], y1 = qs[1], y2 = qs[2];
                const uint2 z = *reinterpret_cast<const uint2 *>(W.m[st] + 8u * lane);
                wc[0] = x0.x; wc[1] = x0.y; wc[2] = x1.x; wc[3] = x1.y; wc[4] = x2.x; wc[5] = x2.y;
                ws[0] = y0.x; ws[1] = y0.y; ws[2] = y1.x; ws[3] = y1.y; ws[4] = y2.x; ws[5] = y2.y;
                wm[0] = z.x; wm[1] = z.y;
            }
            __syncwarp();
            if (OST) {
                static_assert(!OST || (BULK && PXL == 8), "bulk-store flavour: bulk loads, 256-pixel tiles");
                const unsigned ob = i & 1u;
                if (active) apply8<true, (LUTM == 2 ? 0 : LUTM)>(wc, ws, wm, lut, status, wo, pol_lut);
                if (lane == 0) bulk_wait_read<1>();  // the store of tile i-2 has read o[ob]
                __syncwarp();
                if (active) {
                    uint2 *so = reinterpret_cast<uint2 *>(W.o[ob] + 24u * lane);
                    so[0] = make_uint2(wo[0], wo[1]);
                    so[1] = make_uint2(wo[2], wo[3]);
                    so[2] = make_uint2(wo[4], wo[5]);
                }
                fence_proxy_async();
                __syncwarp();
                if (lane == 0) {
                    bulk_s2g_hint(a.out + (size_t)slot * a.out_stride + 3 * p0, W.o[ob], 3u * n, pol_stream);
                    bulk_commit();
                }
            } else if (active) {
                if (LUTM == 2 && kt_ok) apply8<BULK, 2>(wc, ws, wm, lut, status, wo, pol_lut, &kr, &kv);
                else apply8<BULK, (LUTM == 2 ? 0 : LUTM)>(wc, ws, wm, lut, status, wo, pol_lut);  // the cp.async flavour runs without L2 hints (A/B)
                uint8_t *qo = a.out + (size_t)slot * a.out_stride + 3 * p0 + 24u * lane;
                if (BULK) {
                    stg8_hint(qo, wo[0], wo[1], pol_stream);
                    stg8_hint(qo + 8, wo[2], wo[3], pol_stream);
                    stg8_hint(qo + 16, wo[4], wo[5], pol_stream);
                } else {
                    stg_stream8(qo, wo[0], wo[1]);
                    stg_stream8(qo + 8, wo[2], wo[3]);
                    stg_stream8(qo + 16, wo[4], wo[5]);
                }
            }
        }
    }
    if (!BULK) cp_async_wait<0>();
    if (OST && lane == 0) bulk_wait<0>();
}

// ------------------------------------------------------------------------------------------
// k_apply_warp_rs: the per-warp bulk-async pipeline WITH the a-1 NEAREST resample of the
// stylised image (the reference's usual geometry: AdaIN output at 512-short-side scale, content
// at any size).  A 256-pixel tile of the output covers at most two output rows (W >= 256); for
// each, the pixels x in [xa, xb] read the contiguous source span xtab[xa] .. xtab[xb] of source
// row ytab[y] -- staged by one cp.async.bulk per row segment (16-byte aligned outwards) next to
// the content and mask tiles, on the same mbarrier.  The nearest gather then runs against shared
// memory; the resized image is never materialised and every stylised byte crosses L2 -> SM once
// per tile instead of once per output pixel.
// Host guarantees (phase_apply): MASK mode, W >= 256, frame bases / strides / stylised frame size
// 16-byte aligned, and every possible pair of segments fits kRsCap bytes.
// ------------------------------------------------------------------------------------------
constexpr int kRsCap = 2048;
struct RsWarpSmem {
    alignas(128) uint8_t c[kAwStages][256 * 3];
    alignas(128) uint8_t s[kAwStages][kRsCap];
    alignas(128) uint8_t m[kAwStages][256];
};
struct RsSmem {
    RsWarpSmem w[8];
    alignas(8) unsigned long long full[8][kAwStages];
};

struct RsSeg {  // one row segment of a tile
    unsigned npx;       // output pixels in the segment (0 = none)
    unsigned x_first;   // first output x
    unsigned xs_first;  // xtab[x_first]
    unsigned smem_off;  // where the staged span starts in the stage's s[] buffer ...
    unsigned delta;     // ... and the byte offset of source pixel xs_first inside it
    unsigned bytes;     // staged bytes (multiple of 16)
    const uint8_t *g;   // 16-byte aligned global source address
};

__device__ __forceinline__ void rs_segment(const uint8_t *sty, const int *__restrict__ xtab, const int *__restrict__ ytab,
                                           int sw, unsigned y, unsigned xa, unsigned npx, unsigned smem_off, RsSeg &sg)
{
    sg.npx = npx;
    sg.x_first = xa;
    sg.smem_off = smem_off;
    if (npx == 0) {
        sg.bytes = 0;
        sg.delta = 0;
        sg.xs_first = 0;
        sg.g = sty;
        return;
    }
    const unsigned xs0 = (unsigned)__ldg(xtab + xa), xs1 = (unsigned)__ldg(xtab + xa + npx - 1);
    const size_t b0 = 3 * ((size_t)__ldg(ytab + y) * (size_t)sw + xs0), b1 = b0 + 3 * (size_t)(xs1 - xs0 + 1);
    const size_t g0 = b0 & ~(size_t)15, g1 = (b1 + 15) & ~(size_t)15;
    sg.xs_first = xs0;
    sg.delta = (unsigned)(b0 - g0);
    sg.bytes = (unsigned)(g1 - g0);
    sg.g = sty + g0;
}

__global__ void __launch_bounds__(256, 4) k_apply_warp_rs(PixArgs a, unsigned tiles_per_frame, unsigned nslots)
{
    constexpr int kWarps = 8, kTile = 256;
    extern __shared__ __align__(128) uint8_t rs_smem_raw[];
    RsSmem &S = *reinterpret_cast<RsSmem *>(rs_smem_raw);
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    RsWarpSmem &Wm = S.w[warp];
    const unsigned total = tiles_per_frame * nslots;
    if (a.status_out && blockIdx.x == 0)
        for (unsigned q = threadIdx.x; q < nslots; q += 256u) a.status_out[q] = a.ctrl[q].status;
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < kAwStages; ++s) mbar_init(&S.full[warp][s], 1u);
        fence_mbar_init();
        fence_proxy_async();
    }
    __syncwarp();
    const unsigned GW = gridDim.x * kWarps, w0 = blockIdx.x * kWarps + warp;
    const unsigned cnt = w0 < total ? (total - w0 + GW - 1) / GW : 0u;
    const unsigned long long pol_stream = l2_policy_evict_first(), pol_lut = l2_policy_evict_last();
    const unsigned Wd = (unsigned)a.W;

    // tile i of this warp: frame, first pixel, pixel count, and its (at most two) row segments
    auto tile_geom = [&](unsigned i, unsigned &slot, size_t &p0, unsigned &n, RsSeg &sa, RsSeg &sb) {
        const unsigned T = w0 + i * GW;
        slot = T / tiles_per_frame;
        const unsigned t = T - slot * tiles_per_frame;
        p0 = a.p_begin + (size_t)t * kTile;
        const size_t left = a.p_end - p0;
        n = left < (size_t)kTile ? (unsigned)left : (unsigned)kTile;
        const unsigned y = (unsigned)(p0 / Wd), x = (unsigned)(p0 - (size_t)y * Wd);
        // pixels of the tile that lie in its first row.  64-bit on purpose: written as min(n, W - x) in
        // 32-bit arithmetic the optimised build produced na == n for row-crossing tiles (illegal address on
        // the B200; an instrumented build was correct) -- tools/rs_debug.py reproduces it
        const size_t row_end = ((size_t)y + 1u) * (size_t)Wd;
        const unsigned na = (p0 + (size_t)n <= row_end) ? n : (unsigned)(row_end - p0);
#ifdef RCT_RS_DEBUG
        if (na + x > Wd) printf("RS GEOM Wd %u a.W %d x %u y %u n %u na %u\n", Wd, a.W, x, y, n, na);
#endif
        const uint8_t *sty = a.sty + (size_t)slot * a.sty_stride;
        rs_segment(sty, a.xtab, a.ytab, a.sw, y, x, na, 0u, sa);
        rs_segment(sty, a.xtab, a.ytab, a.sw, y + 1u, 0u, n - na, sa.bytes, sb);
    };
    auto issue = [&](unsigned i) {  // lane 0 only
        unsigned slot, n;
        size_t p0;
        RsSeg sa, sb;
        tile_geom(i, slot, p0, n, sa, sb);
        const unsigned st = i % kAwStages;
        unsigned long long *bar = &S.full[warp][st];
#ifdef RCT_RS_DEBUG
        {
            const size_t sty_bytes = 3 * (size_t)a.sw * (size_t)a.sh_dbg;
            const uint8_t *sty0 = a.sty + (size_t)slot * a.sty_stride;
            const bool bad = sa.bytes == 0 || (sa.bytes & 15) || (sb.bytes & 15) || sa.bytes + sb.bytes > (unsigned)kRsCap ||
                             (((uintptr_t)sa.g) & 15) || (((uintptr_t)sb.g) & 15) || sa.g < sty0 ||
                             sa.g + sa.bytes > sty0 + sty_bytes || (sb.bytes && (sb.g < sty0 || sb.g + sb.bytes > sty0 + sty_bytes)) ||
                             (n & 15) || n == 0 || n > 256 || p0 + n > a.p_end;
            if (bad)
                printf("RS BAD tile p0=%llu n=%u slot=%u A(npx %u x %u xs %u off %u d %u b %u g+%lld) B(npx %u b %u off %u g+%lld) sty_bytes %llu\n",
                       (unsigned long long)p0, n, slot, sa.npx, sa.x_first, sa.xs_first, sa.smem_off, sa.delta, sa.bytes,
                       (long long)(sa.g - sty0), sb.npx, sb.bytes, sb.smem_off, (long long)(sb.g - sty0), (unsigned long long)sty_bytes);
        }
#endif
        mbar_expect_tx(bar, 4u * n + sa.bytes + sb.bytes);
        bulk_g2s_hint(Wm.c[st], a.content + (size_t)slot * a.content_stride + 3 * p0, 3u * n, bar, pol_stream);
        bulk_g2s_hint(Wm.m[st], a.mask + (size_t)slot * a.mask_stride + p0, n, bar, pol_stream);
        bulk_g2s(Wm.s[st] + sa.smem_off, sa.g, sa.bytes, bar);  // the (small) source image is re-read: default policy
        if (sb.bytes) bulk_g2s(Wm.s[st] + sb.smem_off, sb.g, sb.bytes, bar);
    };

    unsigned kt_slot = 0xffffffffu;
    bool kt_ok = false;
    KtRegs kr{};
    KtView kv{};
    if (cnt > 0 && lane == 0) issue(0);
    for (unsigned i = 0; i < cnt; ++i) {
        const unsigned st = i % kAwStages;
        if (i + 1 < cnt && lane == 0) issue(i + 1);  // stage (i+1)%2 was released by the __syncwarp ending iteration i-1
        unsigned slot, n;
        size_t p0;
        RsSeg sa, sb;
        tile_geom(i, slot, p0, n, sa, sb);
        const bool active = lane * 8u < n;
        // source x of this lane's 8 pixels (the tables are L1-resident) while the tile is in flight
        unsigned soff[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const unsigned q = lane * 8u + (unsigned)k;
            unsigned off = 0u;
            if (q < n) {
                const bool in_a = q < sa.npx;
                const RsSeg &sg = in_a ? sa : sb;
                const unsigned x = sg.x_first + (in_a ? q : q - sa.npx);
                off = sg.smem_off + sg.delta + 3u * ((unsigned)__ldg(a.xtab + x) - sg.xs_first);
            }
            soff[k] = off;
        }
#ifdef RCT_RS_DEBUG
#pragma unroll
        for (int k = 0; k < 8; ++k)
            if (soff[k] + 2u >= (unsigned)kRsCap) printf("RS BAD soff %u lane %u k %d p0 %llu n %u\n", soff[k], lane, k, (unsigned long long)p0, n);
#endif
        mbar_wait(&S.full[warp][st], (i / kAwStages) & 1u);
        const unsigned *lut = a.lut + (size_t)slot * a.lut_stride;
        const int status = a.ctrl[slot].status;
        if (slot != kt_slot) {  // a new frame: its key table (warp-uniform)
            kt_slot = slot;
            kv = kt_view(a.kt, slot);
            kt_ok = a.kt != nullptr && __ldg(&kv.par->valid) != 0u;
            if (kt_ok) kr = kt_load(kv.par);
        }
        uint32_t wc[6], ws[6], wm[2], wo[6];
        if (active) {
            const uint2 *qc = reinterpret_cast<const uint2 *>(Wm.c[st] + 24u * lane);
            const uint2 x0 = qc[0], x1 = qc[1], x2 = qc[2];
            const uint2 z = *reinterpret_cast<const uint2 *>(Wm.m[st] + 8u * lane);
            wc[0] = x0.x; wc[1] = x0.y; wc[2] = x1.x; wc[3] = x1.y; wc[4] = x2.x; wc[5] = x2.y;
            wm[0] = z.x; wm[1] = z.y;
            uint32_t sp[8];
            const uint8_t *sb8 = Wm.s[st];
#pragma unroll
            for (int k = 0; k < 8; ++k)
                sp[k] = (uint32_t)sb8[soff[k]] | ((uint32_t)sb8[soff[k] + 1] << 8) | ((uint32_t)sb8[soff[k] + 2] << 16);
#pragma unroll
            for (int q = 0; q < 2; ++q) {  // repack as the 6 little-endian words apply8 expects
                ws[3 * q + 0] = sp[4 * q] | (sp[4 * q + 1] << 24);
                ws[3 * q + 1] = (sp[4 * q + 1] >> 8) | (sp[4 * q + 2] << 16);
                ws[3 * q + 2] = (sp[4 * q + 2] >> 16) | (sp[4 * q + 3] << 8);
            }
        }
        __syncwarp();  // the stage is read: the next iteration may refill it
        if (active) {
            if (kt_ok) apply8<true, 2>(wc, ws, wm, lut, status, wo, pol_lut, &kr, &kv);
            else apply8<true>(wc, ws, wm, lut, status, wo, pol_lut);
            uint8_t *qo = a.out + (size_t)slot * a.out_stride + 3 * p0 + 24u * lane;
            stg8_hint(qo, wo[0], wo[1], pol_stream);
            stg8_hint(qo + 8, wo[2], wo[3], pol_stream);
            stg8_hint(qo + 16, wo[4], wo[5], pol_stream);
        }
    }
}

// ------------------------------------------------------------------------------------------
#ifndef RCT_AKB
#define RCT_AKB 4
#endif
// k_apply_kt: the default fused apply pass (MASK mode, same geometry, 16-byte aligned frames): the per-warp
// two-stage bulk-async pipeline of k_apply_warp<true, 8, 4> with the key table (rct_keytab.cuh) in front of the
// colour LUT -- and a lean tile loop.  With the LUT gathers gone the pass is issue-bound (ncu r02: issue slots 70 %
// busy, DRAM 58 %), and 230 of the ~540 warp instructions per tile were loop overhead: two integer divisions for
// the tile geometry, per-tile reloads of the frame's status / LUT base / key-table parameters.  Here the geometry
// of the tile to prefetch advances incrementally (no division), the tile consumed next inherits it, and everything
// that depends on the frame alone is reloaded only when the frame changes.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256, RCT_AKB) k_apply_kt(PixArgs a, unsigned tiles_per_frame, unsigned nslots)
{
    constexpr int kWarps = 8, kTile = 256;
    extern __shared__ __align__(128) uint8_t aw_smem_raw[];
    AwSmem<8> &S = *reinterpret_cast<AwSmem<8> *>(aw_smem_raw);
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    AwWarpSmem<8> &W = S.w[warp];
    const unsigned total = tiles_per_frame * nslots;
    if (a.status_out && blockIdx.x == 0)
        for (unsigned q = threadIdx.x; q < nslots; q += 256u) a.status_out[q] = a.ctrl[q].status;
    if (lane == 0) {
        mbar_init(&S.full[warp][0], 1u);
        mbar_init(&S.full[warp][1], 1u);
        fence_mbar_init();
        fence_proxy_async();
    }
    __syncwarp();
    const unsigned GW = gridDim.x * kWarps;
    unsigned Tn = blockIdx.x * kWarps + warp;  // global index of the next tile to prefetch
    if (Tn >= total) return;
    const unsigned long long pol_stream = l2_policy_evict_first(), pol_lut = l2_policy_evict_last();
    const unsigned npix = (unsigned)(a.p_end - a.p_begin);  // per frame, a multiple of 16 (< 2^32: plan limit)
    unsigned slot_n = Tn / tiles_per_frame, t_n = Tn - slot_n * tiles_per_frame;  // the only division

    auto issue = [&](unsigned st, unsigned slot, unsigned t) {
        if (lane == 0) {
            const unsigned off = t * (unsigned)kTile, left = npix - off, n = left < (unsigned)kTile ? left : (unsigned)kTile;
            const size_t p0 = a.p_begin + off;
            mbar_expect_tx(&S.full[warp][st], 7u * n);
            bulk_g2s_hint(W.c[st], a.content + (size_t)slot * a.content_stride + 3 * p0, 3u * n, &S.full[warp][st], pol_stream);
            bulk_g2s_hint(W.s[st], a.sty + (size_t)slot * a.sty_stride + 3 * p0, 3u * n, &S.full[warp][st], pol_stream);
            bulk_g2s_hint(W.m[st], a.mask + (size_t)slot * a.mask_stride + p0, n, &S.full[warp][st], pol_stream);
        }
    };
    auto advance = [&]() {
        Tn += GW;
        t_n += GW;
        while (t_n >= tiles_per_frame) {
            t_n -= tiles_per_frame;
            ++slot_n;
        }
    };
    unsigned slot_c = slot_n, t_c = t_n;  // the tile consumed next
    issue(0u, slot_n, t_n);
    advance();
    // per-frame state, reloaded when the frame changes
    unsigned cur_slot = 0xffffffffu;
    const unsigned *lut = nullptr;
    uint8_t *out = nullptr;
    int status = 0;
    bool kt_ok = false;
    KtRegs kr{};
    KtView kv{};
    for (unsigned i = 0;; ++i) {
        const unsigned st = i & 1u;
        const bool more = Tn < total;
        const unsigned slot_x = slot_n, t_x = t_n;
        if (more) {  // the other stage was released by the __syncwarp ending the previous iteration
            issue(st ^ 1u, slot_n, t_n);
            advance();
        }
        mbar_wait(&S.full[warp][st], (i >> 1) & 1u);
        if (slot_c != cur_slot) {  // warp-uniform
            cur_slot = slot_c;
            lut = a.lut + (size_t)slot_c * a.lut_stride;
            out = a.out + (size_t)slot_c * a.out_stride + 3 * a.p_begin;
            status = a.ctrl[slot_c].status;
            kv = kt_view(a.kt, slot_c);
            kt_ok = a.kt != nullptr && __ldg(&kv.par->valid) != 0u;
            if (kt_ok) kr = kt_load(kv.par);
        }
        const unsigned off = t_c * (unsigned)kTile, left = npix - off, n = left < (unsigned)kTile ? left : (unsigned)kTile;
        const bool active = lane * 8u < n;
        uint32_t wc[6], ws[6], wm[2], wo[6];
        if (active) {
            const uint2 *qc = reinterpret_cast<const uint2 *>(W.c[st] + 24u * lane);
            const uint2 *qs = reinterpret_cast<const uint2 *>(W.s[st] + 24u * lane);
            const uint2 x0 = qc[0], x1 = qc[1], x2 = qc[2], y0 = qs[0], y1 = qs[1], y2 = qs[2];
            const uint2 z = *reinterpret_cast<const uint2 *>(W.m[st] + 8u * lane);
            wc[0] = x0.x; wc[1] = x0.y; wc[2] = x1.x; wc[3] = x1.y; wc[4] = x2.x; wc[5] = x2.y;
            ws[0] = y0.x; ws[1] = y0.y; ws[2] = y1.x; ws[3] = y1.y; ws[4] = y2.x; ws[5] = y2.y;
            wm[0] = z.x; wm[1] = z.y;
        }
        __syncwarp();  // the stage is read: the next iteration may refill it
        if (active) {
            if (kt_ok) apply8<true, 2>(wc, ws, wm, lut, status, wo, pol_lut, &kr, &kv);
            else apply8<true, 0>(wc, ws, wm, lut, status, wo, pol_lut);
            uint8_t *qo = out + 3u * (size_t)off + 24u * lane;
            stg8_hint(qo, wo[0], wo[1], pol_stream);
            stg8_hint(qo + 8, wo[2], wo[3], pol_stream);
            stg8_hint(qo + 16, wo[4], wo[5], pol_stream);
        }
        if (!more) break;
        slot_c = slot_x;
        t_c = t_x;
    }
}
