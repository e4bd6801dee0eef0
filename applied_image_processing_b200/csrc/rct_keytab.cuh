// rct_keytab.cuh -- the KEY TABLE: an exact accelerator in front of the colour -> RGB LUT of the apply pass.
// Included by rct_kernels.cuh inside namespace rct (after rct_sort.cuh).
//
// Why.  The apply pass resolves every target pixel through LUT[colour], a random 4-byte gather into a 64 MiB table
// per frame: on noisy images every gather drags a 32-byte sector out of DRAM (ncu r01: 776 MB read for 232 MB of
// pixels) and the pass sits at 0.26 of the HBM roofline.  But the output colour is a function of ONE scalar: along
// the sorted target keys the outputs form a few hundred runs of equal RGB (a-7: all recoloured pixels lie on the
// curve mean + m * component).  So
//     out(colour) = run_colour[ number of run thresholds <= key(colour) ].
// The exact key is fp64 (C2/C3/C7, 3 log10 per colour: per-pixel evaluation is what the LUT avoids), but an fp32
// key k~ with a PROVEN error bound eps decides the run whenever no threshold lies within eps of it.  The key axis,
// from the smallest to the largest target key of the frame, is cut into 65537 uniform cells; a cell whose eps-widened interval
// meets no threshold is PURE and stores the run colour, every other cell is flagged and the pixel falls back to
// the exact LUT gather.  ~96 % of the gathers become one read of a 256 KB L1/L2-resident table plus ~30 fp32
// instructions; results are bit-identical by construction, and k_kt_verify (tests) re-checks every distinct target
// colour of a frame against the LUT.
//
//   k_match        (rct_sort.cuh) records a run start (key, sorted index, colour) whenever the output colour changes
//                  along the sorted target (the predecessor of a part's first entry is re-evaluated by the part's
//                  own CTA, so there is no cross-CTA dependency, no ordering and no forced start per part), and the
//                  smallest / largest target key by atomic max
//   k_kt_mark      per run start: flags the cells its eps-widened threshold can reach and seeds the first cell
//                  after them with (sorted index, colour) by a 64-bit atomic max -- no sort needed
//   k_kt_fill      one CTA per slot: prefix maximum of the seeds along the cells = colour of the run each pure cell
//                  lies in; flagged cells and cells before the first run get kKtFlag; fp32 parameters
//   kt_lookup      (apply kernels) fp32 key -> cell -> colour | flag
//
// Error budget (key units unless noted).  k~ = w . lg2(M' rgb) - pm in fp32:
//   LMS: 3 products + 2 adds, inputs exact (integers < 256), M' = RGB_TO_LMS / 255 rounded to fp32, the canonical
//        unit[k] = fp32(k / 255) rounded once more: relative error <= 6 * 2^-24                -> 5.2e-7 in lg2
//   lg2.approx.f32: absolute error <= 2^-22 on [0.5, 2], relative <= 2^-22 elsewhere; |lg2| <= 13.3 (LMS >= 9.4e-5
//        for every non-black colour)                                                           -> 3.2e-6
//   dot with w (|w_i| <= log10(2) * |column i of LMS_TO_LAB|_2 = 0.301, |w|_1 <= 0.91), 3 fma on magnitudes <= 16
//        (the constant pmk = pm + kmin = the smallest w . lg2 of the frame enters as the addend of the first fma: its
//        rounding is one of the four):  4 * 2^-24 * 16                                            -> 3.8e-6
//   total <= 0.91 * 3.7e-6 + 3.8e-6 = 7.2e-6 < kKtEpsKey = 2^-16 = 1.53e-5.  k_kt_keyerr measures the true maximum
//   over all 2^24 colours for a frame's parameters (tests assert it stays below kKtEpsKey / 2).
// Cell coordinate c = (k~ - kmin) * tscale in fp32, 0 <= c <= 65536, the cells spanning [smallest, largest] target
//   key (+0.1 %): rounding of the product and of tscale <= 2 * 2^-24 * 65536 = 0.008 cells, round-to-nearest ties
//   included: kKtEpsC = 1/64 cell.
#pragma once

struct KtView {
    const KtParams *par;
    const unsigned *cells;
};
__device__ __forceinline__ KtView kt_view(const uint8_t *kt0, size_t slot)
{
    const uint8_t *b = kt0 + slot * kKtBytes;
    return KtView{reinterpret_cast<const KtParams *>(b), reinterpret_cast<const unsigned *>(b + kKtOffCells)};
}

struct KtRegs {
    float w[3], pmk, tscale;   // the matrix M' = RGB_TO_LMS / 255 is the same for every frame: __constant__ c_ktm
};
__device__ __forceinline__ KtRegs kt_load(const KtParams *p)
{
    KtRegs r;
#pragma unroll
    for (int i = 0; i < 3; ++i) r.w[i] = __ldg(&p->w[i]);
    r.pmk = __ldg(&p->pmk);
    r.tscale = __ldg(&p->tscale);
    return r;
}
// (w . lg2(M' rgb)) - pmk in fp32 = key - kmin: the arithmetic whose error kKtEpsKey bounds
__device__ __forceinline__ float kt_dot(const KtRegs &r, uint32_t col)
{
    // bytes -> floats without a conversion instruction: 0x4B0000kk is 8388608 + kk
    const float fr = __uint_as_float(__byte_perm(col, 0x4B000000u, 0x7440)) - 8388608.f;
    const float fg = __uint_as_float(__byte_perm(col, 0x4B000000u, 0x7441)) - 8388608.f;
    const float fb = __uint_as_float(__byte_perm(col, 0x4B000000u, 0x7442)) - 8388608.f;
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const float lms = fmaf(fb, c_ktm[3 * i + 2], fmaf(fg, c_ktm[3 * i + 1], fr * c_ktm[3 * i]));
        float l;
        asm("lg2.approx.f32 %0, %1;" : "=f"(l) : "f"(lms));
        acc = fmaf(l, r.w[i], i == 0 ? -r.pmk : acc);
    }
    return acc;
}
// cell of a colour: round-to-nearest by the magic-number add (no conversion instruction)
__device__ __forceinline__ unsigned kt_cell(const KtRegs &r, uint32_t col)
{
    float c = kt_dot(r, col) * r.tscale;
#if RCT_KT_NOCLAMP
    // no clamp: coordinates a little outside [0, 65536] (error bound) land in flagged entries, see kKtCellsAlloc
    return __float_as_uint(c + 8388608.f) & 0x1ffffu;
#else
    c = fminf(fmaxf(c, 0.f), (float)(kKtCells - 1));      // also catches NaN (black, never looked up)
    return __float_as_uint(c + 8388608.f) & 0x1ffffu;     // 0 .. 65536
#endif
}
__device__ __forceinline__ unsigned kt_lookup(const KtRegs &r, const KtView &v, uint32_t col)
{
    return __ldg(v.cells + kt_cell(r, col));
}

// ---- builder ---------------------------------------------------------------------------------------------------
// The cells span [kmin, kmax + 0.1 %] of the frame's target keys (k_match leaves both in KtParams by atomic max):
// exact coordinate of a key = (key - kmin) * scale.  Keys above kmax do not exist; the top cells belong to the last run.
struct KtGeom {
    double kmin, scale;
    bool ok;
};
__device__ __forceinline__ KtGeom kt_geom(const FrameCtrl *fc, const KtParams *par, int interleave)
{
    KtGeom g;
    g.ok = fc->status == 0 && interleave && fc->n_runs > 0 && fc->n_runs <= (unsigned)kKtRunCap && par->kmax_enc != 0ull;
    g.kmin = g.scale = 0.0;
    if (g.ok) {
        const double kmax = ord_dec(par->kmax_enc);
        g.kmin = ord_dec(~par->kmin_enc_inv);
        const double span = (kmax - g.kmin) * 1.001;
        g.ok = span > 0.0 && span < 1e300;
        if (g.ok) g.scale = (double)(kKtCells - 1) / span;
        if (g.ok && kKtEpsKey * g.scale + kKtEpsC > kKtSlackCells) g.ok = false;  // a near-flat region: the LUT alone
    }
    return g;
}

// per run start (threshold key thr, sorted index i, colour): the cells c whose widened interval contains thr, i.e.
//   c in [ (thr - epsK - kmin) * scale - epsC - 0.5 , (thr + epsK - kmin) * scale + epsC + 0.5 ],
// are flagged; the first cell after them is seeded with the run
__global__ void __launch_bounds__(256) k_kt_mark(const FrameCtrl *ctrl, uint8_t *kt0, int interleave)
{
    const int slot = blockIdx.y;
    uint8_t *blk = kt0 + (size_t)slot * kKtBytes;
    const KtGeom g = kt_geom(&ctrl[slot], reinterpret_cast<const KtParams *>(blk), interleave);
    if (!g.ok) return;
    unsigned long long *seeds = reinterpret_cast<unsigned long long *>(blk + kKtOffSeeds);
    uint8_t *flags = blk + kKtOffFlags;
    const double *rkey = reinterpret_cast<const double *>(blk + kKtOffRunKey);
    const uint2 *ric = reinterpret_cast<const uint2 *>(blk + kKtOffRunIc);
    const unsigned n = ctrl[slot].n_runs;
    for (unsigned q = blockIdx.x * blockDim.x + threadIdx.x; q < n; q += gridDim.x * blockDim.x) {
        const double thr = rkey[q];
        const uint2 ic = ric[q];
        // (the 1e-6 of a cell covers the fp64 rounding of these two expressions themselves, ~1e-10 of a cell)
        double lo = ceil((thr - kKtEpsKey - g.kmin) * g.scale - kKtEpsC - 0.5 - 1e-6);
        double hi = floor((thr + kKtEpsKey - g.kmin) * g.scale + kKtEpsC + 0.5 + 1e-6);
        if (!(lo > 0.0)) lo = 0.0;
        if (!(hi > 0.0)) hi = 0.0;
        if (lo > (double)(kKtCells - 1)) lo = (double)(kKtCells - 1);
        if (hi > (double)(kKtCells - 1)) hi = (double)(kKtCells - 1);
        const int c_lo = (int)lo, c_hi = (int)hi;
        for (int c = c_lo; c <= c_hi; ++c) flags[c] = 1;
        if (c_hi + 1 < kKtCells) atomicMax(&seeds[c_hi + 1], ((unsigned long long)(ic.x + 1u) << 32) | ic.y);
    }
}

// one CTA of 1024 threads per slot: prefix maximum of the seeds -> cells.  Warp w owns the contiguous chunk of
// kKtChunk cells starting at w * kKtChunk and walks it 32 cells at a time (coalesced), twice: chunk maximum first,
// then -- with the maxima of the chunks before it -- the running maximum per cell.
constexpr int kKtChunk = (kKtCells + 31) / 32;
__global__ void __launch_bounds__(1024) k_kt_fill(const FrameCtrl *ctrl, uint8_t *kt0, int interleave)
{
    const int slot = blockIdx.x;
    uint8_t *blk = kt0 + (size_t)slot * kKtBytes;
    KtParams *par = reinterpret_cast<KtParams *>(blk);
    const KtGeom g = kt_geom(&ctrl[slot], par, interleave);
    if (!g.ok) {
        if (threadIdx.x == 0) {
            par->valid = 0;
            par->n_runs = ctrl[slot].n_runs;
        }
        return;
    }
    const unsigned long long *seeds = reinterpret_cast<const unsigned long long *>(blk + kKtOffSeeds);
    const uint8_t *flags = blk + kKtOffFlags;
    unsigned *cells = reinterpret_cast<unsigned *>(blk + kKtOffCells);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int c0 = warp * kKtChunk, c1 = min(c0 + kKtChunk, kKtCells);
    __shared__ unsigned long long s_w[32];
    unsigned long long mx = 0;
#pragma unroll 8
    for (int c = c0 + lane; c < c1; c += 32) {  // independent loads: all in flight at once
        const unsigned long long s = seeds[c];
        mx = s > mx ? s : mx;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long y = __shfl_xor_sync(0xffffffffu, mx, o);
        mx = y > mx ? y : mx;
    }
    if (lane == 0) s_w[warp] = mx;
    __syncthreads();
    unsigned long long carry = 0;
    for (int w = 0; w < warp; ++w) carry = s_w[w] > carry ? s_w[w] : carry;
    unsigned flagged = 0;
    constexpr int kRows = 8;  // rows of 32 cells loaded together (the scan itself is a dependent chain)
    for (int base = c0; base < c1; base += 32 * kRows) {  // warp-uniform trip count
        unsigned long long sv[kRows];
        unsigned fv[kRows];
#pragma unroll
        for (int r = 0; r < kRows; ++r) {
            const int c = base + 32 * r + lane;
            sv[r] = c < c1 ? seeds[c] : 0ull;
            fv[r] = c < c1 ? (unsigned)flags[c] : 0u;
        }
#pragma unroll
        for (int r = 0; r < kRows; ++r) {
            const int c = base + 32 * r + lane;
            // the seeds of a row rarely number more than one: reduce to the row maximum first, scan only if needed
            unsigned long long inc = sv[r];
            if (__any_sync(0xffffffffu, inc != 0ull)) {
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const unsigned long long y = __shfl_up_sync(0xffffffffu, inc, o);
                    if (lane >= o) inc = y > inc ? y : inc;
                }
            }
            inc = carry > inc ? carry : inc;
            if (c < c1) {
                const unsigned v = (fv[r] || inc == 0ull) ? kKtFlag : (unsigned)inc;
                cells[c] = v;
                flagged += v == kKtFlag;
            }
            carry = __shfl_sync(0xffffffffu, inc, 31);
        }
    }
    for (int c = kKtCells + (int)threadIdx.x; c < kKtCellsAlloc; c += 1024) cells[c] = kKtFlag;
    if (flagged) atomicAdd(&par->n_flagged, flagged);
    if (threadIdx.x == 0) {
        const RegionCtrl *rc = &ctrl[slot].r[0];
        for (int i = 0; i < 9; ++i) par->m[i] = c_ktm[i];
        const double L2 = 0.30102999566398119521;  // log10(2): lg2 -> log10
        for (int i = 0; i < 3; ++i) {
            double a = 0.0;
            for (int j = 0; j < 3; ++j) a += rc->pca.comp[j] * c_mat[9 + 3 * j + i];
            par->w[i] = (float)(a * L2);
        }
        par->pmk = (float)(rc->pca.pm + g.kmin);
        par->tscale = (float)g.scale;
        par->kmin = g.kmin;
        par->scale = g.scale;
        par->n_runs = ctrl[slot].n_runs;
        par->valid = 1;
    }
}

// ---- checks (tests / introspection) --------------------------------------------------------------------------------
// maximum |fp32 key - canonical fp64 key| over ALL 2^24 colours for the slot's parameters (black excluded)
__global__ void __launch_bounds__(256) k_kt_keyerr(const FrameCtrl *ctrl, const uint8_t *kt0, int slot, double *max_err)
{
    __shared__ MathTabs T;
    stage_tabs(&T);
    const KtView v = kt_view(kt0, (size_t)slot);
    const KtRegs r = kt_load(v.par);
    const RegionCtrl *rc = &ctrl[slot].r[0];
    const double comp[3] = {rc->pca.comp[0], rc->pca.comp[1], rc->pca.comp[2]}, pm = rc->pca.pm;
    double worst = 0.0;
    for (unsigned col = blockIdx.x * blockDim.x + threadIdx.x + 1u; col < kCube; col += gridDim.x * blockDim.x) {
        double lab[3];
        rgb2lab(col, &T, lab);
        const double err = fabs((double)kt_dot(r, col) - (project(lab, comp, pm) - v.par->kmin));
        worst = fmax(worst, err);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) worst = fmax(worst, __shfl_xor_sync(0xffffffffu, worst, o));
    if ((threadIdx.x & 31) == 0) atomicMax(reinterpret_cast<unsigned long long *>(max_err), (unsigned long long)__double_as_longlong(worst));
}
// every distinct target colour of the slot through the key table exactly as the apply pass does:
// counts[0] = colours resolved by a pure cell, counts[1] = of those, colours whose cell disagrees with the LUT (must
// be 0), counts[2] = colours sent to the LUT by a flagged cell; pixel-weighted in counts[3..5]
__global__ void __launch_bounds__(256) k_kt_verify(const FrameCtrl *ctrl, const Job *jobs, const uint8_t *kt0, const unsigned *lut0,
                                                   size_t lut_stride, int slot, unsigned long long *counts)
{
    const KtView v = kt_view(kt0, (size_t)slot);
    if (!v.par->valid) return;
    const KtRegs r = kt_load(v.par);
    const Job &jt = jobs[2 * slot];
    const unsigned d = ctrl[slot].r[0].d;
    const unsigned *lut = lut0 + (size_t)slot * lut_stride;
    unsigned long long c[6] = {0, 0, 0, 0, 0, 0};
    for (unsigned e = blockIdx.x * blockDim.x + threadIdx.x; e < d; e += gridDim.x * blockDim.x) {
        const unsigned col = jt.payA[e], cnt = jt.cntA[e];
        const unsigned got = kt_lookup(r, v, col);
        if (got & kKtFlag) {
            ++c[2];
            c[5] += cnt;
        } else {
            ++c[0];
            c[3] += cnt;
            if (got != lut[col]) {
                ++c[1];
                c[4] += cnt;
            }
        }
    }
    for (int k = 0; k < 6; ++k)
        if (c[k]) atomicAdd(&counts[k], c[k]);
}
