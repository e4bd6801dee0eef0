// rct_keytab.cuh -- the KEY TABLE: an exact accelerator in front of the colour -> RGB LUT of the apply pass.
// Included by rct_kernels.cuh inside namespace rct (after rct_sort.cuh).
//
// Why.  The apply pass resolves every target pixel through LUT[colour], a random 4-byte gather into a 64 MiB table
// per frame: on noisy images every gather drags a 32-byte sector out of DRAM (ncu r01: 776 MB read for 232 MB of
// pixels) and the pass sits at 0.26 of the HBM roofline.  But the output colour is a function of ONE scalar: along
// the sorted target keys the outputs form a few thousand runs of equal RGB (a-7: all recoloured pixels lie on the
// curve mean + m * component).  So
//     out(colour) = run_colour[ number of run thresholds <= key(colour) ].
// The exact key is fp64 (C2/C3/C7, 3 log10 per colour: per-pixel evaluation is what the LUT avoids), but an fp32
// key k~ with a PROVEN error bound eps decides the run whenever no threshold lies within eps of it.  The key axis is
// cut into ~64K cells (histogram-equalised through the coarse key histogram the sort builds anyway); a cell whose
// eps-widened interval meets no threshold is PURE and stores the run colour, every other cell is flagged and the
// pixel falls back to the exact LUT gather.  ~90 % of the gathers become two L1/L2-resident table reads
// (32 KB + 272 KB per frame) plus ~40 fp32 instructions; results are bit-identical by construction, and
// k_kt_verify (tests) re-checks every distinct target colour of a frame against the LUT.
//
//   k_match        (rct_sort.cuh) records a run start (key, colour) whenever the output colour changes along the
//                  sorted target; every part starts a run, so no cross-CTA dependency
//   k_kt_sort      one CTA per slot: runs sorted by key (bitonic, shared memory), fp32 parameters, cell directory
//   k_kt_cells     every cell: binary search of its eps-widened key interval in the thresholds -> colour or flag
//   kt_lookup      (apply kernels) fp32 key -> cell -> colour | flag
//
// Error budget (key units unless noted).  k~ = w . lg2(M' rgb) - pm in fp32:
//   LMS: 3 products + 2 adds, inputs exact (integers < 256), M' = RGB_TO_LMS / 255 rounded to fp32, the canonical
//        unit[k] = fp32(k / 255) rounded once more: relative error <= 6 * 2^-24                -> 5.2e-7 in lg2
//   lg2.approx.f32: absolute error <= 2^-22 on [0.5, 2], relative <= 2^-22 elsewhere; |lg2| <= 13.3 (LMS >= 9.4e-5
//        for every non-black colour)                                                           -> 3.2e-6
//   dot with w (|w|_1 <= 1.3), 3 fma + 1 subtraction on magnitudes <= 16: 4 * 2^-24 * 16       -> 3.8e-6
//   total <= 1.3 * 3.7e-6 + 3.8e-6 = 8.6e-6 < kKtEpsKey = 2^-16 = 1.53e-5.  k_kt_keyerr measures the true maximum
//   over all 2^24 colours for a frame's parameters (tests assert it stays below kKtEpsKey / 2).
// Cell coordinate t = k~ * tscale + toff in fp32, |t| <= 4096: rounding of t, tscale and toff <= 3 * 4096 * 2^-24
//   = 7.3e-4 cell units, sub-cell arithmetic and round-to-nearest ties well below that: kKtEpsT = 2^-9 = 1.95e-3.
#pragma once

struct KtView {
    const KtParams *par;
    const uint2 *dir;
    const unsigned *cells;
};
__device__ __forceinline__ KtView kt_view(const uint8_t *kt0, size_t slot)
{
    const uint8_t *b = kt0 + slot * kKtBytes;
    return KtView{reinterpret_cast<const KtParams *>(b), reinterpret_cast<const uint2 *>(b + kKtOffDir),
                  reinterpret_cast<const unsigned *>(b + kKtOffCells)};
}

// the fp32 key of a packed colour (the arithmetic whose error kKtEpsKey bounds)
struct KtRegs {
    float m[9], w[3], pm, tscale, toff;
};
__device__ __forceinline__ KtRegs kt_load(const KtParams *p)
{
    KtRegs r;
#pragma unroll
    for (int i = 0; i < 9; ++i) r.m[i] = __ldg(&p->m[i]);
#pragma unroll
    for (int i = 0; i < 3; ++i) r.w[i] = __ldg(&p->w[i]);
    r.pm = __ldg(&p->pm);
    r.tscale = __ldg(&p->tscale);
    r.toff = __ldg(&p->toff);
    return r;
}
__device__ __forceinline__ float kt_key(const KtRegs &r, uint32_t col)
{
    // bytes -> floats without a conversion instruction: 0x4B0000kk is 8388608 + kk
    const float fr = __uint_as_float(0x4B000000u | (col & 255u)) - 8388608.f;
    const float fg = __uint_as_float(0x4B000000u | ((col >> 8) & 255u)) - 8388608.f;
    const float fb = __uint_as_float(0x4B000000u | (col >> 16)) - 8388608.f;
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const float lms = fmaf(fb, r.m[3 * i + 2], fmaf(fg, r.m[3 * i + 1], fr * r.m[3 * i]));
        float l;
        asm("lg2.approx.f32 %0, %1;" : "=f"(l) : "f"(lms));
        acc = i == 0 ? l * r.w[0] : fmaf(l, r.w[i], acc);
    }
    return acc - r.pm;
}
// cell of an fp32 key: coarse cell by round-to-nearest (magic-number add), sub-cell by the directory
__device__ __forceinline__ unsigned kt_cell(const KtRegs &r, const uint2 *dir, float key)
{
    float t = fmaf(key, r.tscale, r.toff);
    t = fminf(fmaxf(t, 0.f), (float)kCoarse);
    const float tb = t + 8388608.f;
    const unsigned cb = __float_as_uint(tb) & 0x1fffu;   // 0 .. 4096
    const float frac = t - (tb - 8388608.f);             // [-0.5, 0.5]
    const uint2 e = __ldg(dir + cb);
    const float ns = __uint_as_float(e.y);
    const float u = fmaf(frac, ns, 0.5f * ns);           // [0, ns]
    int sub = (int)(__float_as_uint((u - 0.5f) + 8388608.f) & 0x7fffffu);
    if (u < 0.5f) sub = 0;                               // (u - 0.5) + 2^23 may round below 2^23
    const int nsi = (int)ns;
    sub = sub < nsi - 1 ? sub : nsi - 1;
    return e.x + (unsigned)sub;
}
__device__ __forceinline__ unsigned kt_lookup(const KtRegs &r, const KtView &v, uint32_t col)
{
    return __ldg(v.cells + kt_cell(r, v.dir, kt_key(r, col)));
}

// ---- builder ---------------------------------------------------------------------------------------------------
// one CTA of 1024 threads per slot
__global__ void __launch_bounds__(1024) k_kt_sort(FrameCtrl *ctrl, uint8_t *kt0, int interleave)
{
    const int slot = blockIdx.x;
    uint8_t *blk = kt0 + (size_t)slot * kKtBytes;
    KtParams *par = reinterpret_cast<KtParams *>(blk);
    uint2 *dir = reinterpret_cast<uint2 *>(blk + kKtOffDir);
    double *rkey = reinterpret_cast<double *>(blk + kKtOffRunKey);
    unsigned *rcol = reinterpret_cast<unsigned *>(blk + kKtOffRunCol);
    const RegionCtrl *rc = &ctrl[slot].r[0];
    const unsigned n = ctrl[slot].n_runs;
    const bool ok = ctrl[slot].status == 0 && interleave && n > 0 && n <= (unsigned)kKtRunCap && rc->cscale > 0.0 &&
                    rc->d > 0;
    if (!ok) {
        if (threadIdx.x == 0) {
            par->valid = 0;
            par->n_runs = n;
        }
        return;
    }
    extern __shared__ __align__(16) uint8_t kt_smem[];
    double *s_key = reinterpret_cast<double *>(kt_smem);                  // [kKtRunCap]
    unsigned *s_col = reinterpret_cast<unsigned *>(s_key + kKtRunCap);    // [kKtRunCap]
    __shared__ unsigned s_warp[33];
    unsigned np2 = 1;
    while (np2 < n) np2 <<= 1;
    for (unsigned i = threadIdx.x; i < np2; i += blockDim.x) {
        s_key[i] = i < n ? rkey[i] : INFINITY;
        s_col[i] = i < n ? rcol[i] : 0u;
    }
    __syncthreads();
    for (unsigned k = 2; k <= np2; k <<= 1)
        for (unsigned j = k >> 1; j > 0; j >>= 1) {
            for (unsigned i = threadIdx.x; i < np2; i += blockDim.x) {
                const unsigned x = i ^ j;
                if (x > i) {
                    const bool up = (i & k) == 0;
                    const double a = s_key[i], b = s_key[x];
                    if ((a > b) == up) {
                        s_key[i] = b;
                        s_key[x] = a;
                        const unsigned c = s_col[i];
                        s_col[i] = s_col[x];
                        s_col[x] = c;
                    }
                }
            }
            __syncthreads();
        }
    for (unsigned i = threadIdx.x; i < n; i += blockDim.x) {
        rkey[i] = s_key[i];
        rcol[i] = s_col[i];
    }
    // cell directory: coarse cell cb (t in [cb-0.5, cb+0.5)) holds about half of the floor-bins cb-1 and cb of the
    // coarse key histogram; its share of the kKtCellsTarget cells is proportional to that count
    const double per = (double)kKtCellsTarget / (double)rc->d;
    unsigned carry = 0;
    for (int base = 0; base < kKtCoarse; base += 1024) {
        const int cb = base + (int)threadIdx.x;
        unsigned ns = 0;
        if (cb < kKtCoarse) {
            const unsigned lo = cb > 0 ? rc->coarse[cb - 1] : 0u, hi = cb < kCoarse ? rc->coarse[cb] : 0u;
            const double want = ceil(0.5 * ((double)lo + (double)hi) * per);
            ns = want < 1.0 ? 1u : (want > (double)kKtSubMax ? (unsigned)kKtSubMax : (unsigned)want);
        }
        unsigned total;
        const unsigned ex = block_excl_scan(ns, s_warp, total);
        if (cb < kKtCoarse) dir[cb] = make_uint2(carry + ex, __float_as_uint((float)ns));
        carry += total;
    }
    if (threadIdx.x == 0) {
        for (int i = 0; i < 9; ++i) par->m[i] = (float)(c_mat[i] / 255.0);
        const double L2 = 0.30102999566398119521;  // log10(2): lg2 -> log10
        for (int i = 0; i < 3; ++i) {
            double a = 0.0;
            for (int j = 0; j < 3; ++j) a += rc->pca.comp[j] * c_mat[9 + 3 * j + i];
            par->w[i] = (float)(a * L2);
        }
        par->pm = (float)rc->pca.pm;
        par->tscale = (float)rc->cscale;
        par->toff = (float)(-rc->klo * rc->cscale);
        par->n_cells = carry;
        par->n_flagged = 0;
        par->n_runs = n;
        par->n_sorted = n;
        par->valid = 1;
    }
}

// every cell: the exact-key interval that can reach it, widened by the error bounds, against the sorted thresholds
__global__ void __launch_bounds__(256) k_kt_cells(FrameCtrl *ctrl, uint8_t *kt0)
{
    const int slot = blockIdx.y;
    uint8_t *blk = kt0 + (size_t)slot * kKtBytes;
    KtParams *par = reinterpret_cast<KtParams *>(blk);
    if (!par->valid) return;
    const uint2 *dir = reinterpret_cast<const uint2 *>(blk + kKtOffDir);
    unsigned *cells = reinterpret_cast<unsigned *>(blk + kKtOffCells);
    const double *rkey = reinterpret_cast<const double *>(blk + kKtOffRunKey);
    const unsigned *rcol = reinterpret_cast<const unsigned *>(blk + kKtOffRunCol);
    const RegionCtrl *rc = &ctrl[slot].r[0];
    const unsigned n = par->n_sorted;
    const double klo = rc->klo, inv = 1.0 / rc->cscale;
    unsigned flagged = 0;
    for (int cb = blockIdx.x; cb < kKtCoarse; cb += gridDim.x) {
        const uint2 e = dir[cb];
        const unsigned ns = (unsigned)__uint_as_float(e.y);
        for (unsigned sub = threadIdx.x; sub < ns; sub += blockDim.x) {
            const double t_lo = (double)cb - 0.5 + (double)sub / (double)ns - kKtEpsT;
            const double t_hi = (double)cb - 0.5 + (double)(sub + 1) / (double)ns + kKtEpsT;
            // keys clamped at the ends of the coarse range land in the cells that contain t = 0 / t = kCoarse
            const double k_lo = (cb == 0 && t_lo <= 0.0) ? -INFINITY : klo + t_lo * inv - kKtEpsKey;
            const double k_hi = (cb == kCoarse && t_hi >= (double)kCoarse) ? INFINITY : klo + t_hi * inv + kKtEpsKey;
            // a = thresholds <= k_lo, b = thresholds <= k_hi: keys in [k_lo, k_hi] belong to runs a-1 .. b-1
            unsigned lo = 0, hi = n;
            while (lo < hi) {
                const unsigned mid = (lo + hi) >> 1;
                if (rkey[mid] <= k_lo) lo = mid + 1; else hi = mid;
            }
            const unsigned a = lo;
            hi = n;
            while (lo < hi) {
                const unsigned mid = (lo + hi) >> 1;
                if (rkey[mid] <= k_hi) lo = mid + 1; else hi = mid;
            }
            const unsigned b = lo;
            unsigned v = kKtFlag;
            if (a > 0) {  // below the first threshold there is no run: only keys off by the error bound get there
                v = rcol[a - 1];
                for (unsigned q = a; q < b; ++q)
                    if (rcol[q] != v) {
                        v = kKtFlag;
                        break;
                    }
            }
            cells[e.x + sub] = v;
            flagged += v == kKtFlag;
        }
    }
    if (flagged) atomicAdd(&par->n_flagged, flagged);
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
        const double err = fabs((double)kt_key(r, col) - project(lab, comp, pm));
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
