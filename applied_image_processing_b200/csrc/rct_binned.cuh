// rct_binned.cuh -- BINNED CDF matching (BASELINE.json config 5 "1024-bin histograms", north_star subsystems 2-3:
// masked-statistics reduction into shared-memory privatised histograms + CDF build / inverse-lookup kernel).
// Included by rct_kernels.cuh inside namespace rct.
//
// NO COUNTERPART IN THE REFERENCE (match_cdf, LST:99-125, is exact and sort-based): an approximation with its own
// written-down definition, shared with the C oracle (orc_match_binned) and compared with it
// bit for bit (histograms, CDF indices, uint8 output).  Never the parity path.  Its point: the only cross-GPU
// state of a row-sharded image is 2 x 28 integers (exact moments), 4 doubles (key ranges) and 2 x B counters.
//
//   k_bin_keys   per distinct colour: lab, projection key (C2/C3/C7, the exact path's arithmetic) + min / max key
//   k_bin_hist   B-bin histogram of the keys weighted by the colour's pixel count: shared-memory privatised atomics,
//                one flush of the non-zero bins per CTA
//   k_bin_match  exclusive CDFs of both regions in shared memory (CDF build), then per distinct target colour the
//                in-bin rank lerp, the inverse lookup in the source CDF (binary search over B bins), inverse PCA,
//                lab -> RGB -> the colour -> output LUT the apply pass reads
// The histogram of a region lives in RegionCtrl::coarse[0..B) (the exact path's coarse key histogram is not
// needed here), its key range in lab_max[0] / lab_min[0] as order-preserving integers.
#pragma once

constexpr int kBinMax = kCoarse;  // 4096

__global__ void __launch_bounds__(256) k_bin_keys(FrameCtrl *ctrl, const Job *jobs)
{
    const int region = blockIdx.y, slot = blockIdx.z;
    const Job &job = jobs[slot * 2 + region];
    RegionCtrl *rc = &ctrl[slot].r[region];
    if (ctrl[slot].status != 0) return;
    const unsigned d = rc->d;
    if ((size_t)blockIdx.x * blockDim.x >= d) return;
    __shared__ MathTabs T;
    stage_tabs(&T);
    const double comp[3] = {rc->pca.comp[0], rc->pca.comp[1], rc->pca.comp[2]}, pm = rc->pca.pm;
    double *keyA = const_cast<double *>(job.keyA);
    double kmax = -1e308, kmin = 1e308;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < d; e += (size_t)gridDim.x * blockDim.x) {
        double lab[3];
        rgb2lab(job.payA[e], &T, lab);
        const double key = project(lab, comp, pm);
        keyA[e] = key;
        kmax = fmax(kmax, key);
        kmin = fmin(kmin, key);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        kmax = fmax(kmax, __shfl_xor_sync(0xffffffffu, kmax, o));
        kmin = fmin(kmin, __shfl_xor_sync(0xffffffffu, kmin, o));
    }
    if ((threadIdx.x & 31) == 0 && kmax >= kmin) {
        atomicMax(&rc->lab_max[0], ord_enc(kmax));
        atomicMax(&rc->lab_min[0], ~ord_enc(kmin));
    }
}

struct BinRange {
    double lo, hi, scale;
};
__device__ __forceinline__ BinRange bin_range(const RegionCtrl *rc, int bins)
{
    BinRange r;
    r.hi = ord_dec(rc->lab_max[0]);
    r.lo = ord_dec(~rc->lab_min[0]);
    r.scale = r.hi > r.lo ? (double)bins / (r.hi - r.lo) : 0.0;
    return r;
}

__global__ void __launch_bounds__(256) k_bin_hist(FrameCtrl *ctrl, const Job *jobs, int bins)
{
    const int region = blockIdx.y, slot = blockIdx.z;
    const Job &job = jobs[slot * 2 + region];
    RegionCtrl *rc = &ctrl[slot].r[region];
    if (ctrl[slot].status != 0) return;
    const unsigned d = rc->d;
    if ((size_t)blockIdx.x * blockDim.x >= d) return;
    __shared__ unsigned s_hist[kBinMax];
    for (int i = threadIdx.x; i < bins; i += blockDim.x) s_hist[i] = 0;
    __syncthreads();
    const BinRange r = bin_range(rc, bins);
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < d; e += (size_t)gridDim.x * blockDim.x) {
        int b = (int)((job.keyA[e] - r.lo) * r.scale);
        if (b > bins - 1) b = bins - 1;
        atomicAdd(&s_hist[b], job.cntA[e]);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < bins; i += blockDim.x) {
        const unsigned v = s_hist[i];
        if (v) atomicAdd(&rc->coarse[i], v);
    }
}

// CDF build + inverse lookup.  dbg_code[e] = b * 65536 + b' and dbg_matched[e] per entry of the (unsorted)
// target list when the plan was created with RCT_FLAG_DEBUG.
__global__ void __launch_bounds__(256) k_bin_match(FrameCtrl *ctrl, const Job *jobs, unsigned *lut0, size_t lut_stride,
                                                   int bins, double *dbg_matched0, long long *dbg_code0, size_t dbg_stride)
{
    const int slot = blockIdx.y;
    if (ctrl[slot].status != 0) return;
    const Job &jt = jobs[2 * slot];
    const RegionCtrl *rt = &ctrl[slot].r[0], *rs = &ctrl[slot].r[1];
    const unsigned d = rt->d;
    if ((size_t)blockIdx.x * blockDim.x >= d) return;
    __shared__ MathTabs T;
    __shared__ unsigned s_ct[kBinMax + 1], s_cs[kBinMax + 1];  // exclusive cumulative counts, [bins] = N
    __shared__ unsigned s_warp[33];
    stage_tabs(&T);
    {   // exclusive scans of the two histograms (counts fit 32 bits: a plan holds < 2^32 pixels)
        unsigned carry_t = 0, carry_s = 0;
        for (int base = 0; base < bins; base += 256) {
            const int i = base + (int)threadIdx.x;
            const unsigned vt = i < bins ? rt->coarse[i] : 0u, vs = i < bins ? rs->coarse[i] : 0u;
            unsigned tot_t, tot_s;
            const unsigned et = block_excl_scan(vt, s_warp, tot_t);
            const unsigned es = block_excl_scan(vs, s_warp, tot_s);
            if (i < bins) {
                s_ct[i] = carry_t + et;
                s_cs[i] = carry_s + es;
            }
            carry_t += tot_t;
            carry_s += tot_s;
        }
        if (threadIdx.x == 0) {
            s_ct[bins] = carry_t;
            s_cs[bins] = carry_s;
        }
        __syncthreads();
    }
    const BinRange bt = bin_range(rt, bins), bs = bin_range(rs, bins);
    const double w_s = (bs.hi - bs.lo) / (double)bins;
    const double nt = (double)rt->n_pix, ns = (double)rs->n_pix;
    const Pca pca = rt->pca;
    unsigned *lut = lut0 + (size_t)slot * lut_stride;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < d; e += (size_t)gridDim.x * blockDim.x) {
        const double pos = (jt.keyA[e] - bt.lo) * bt.scale;
        int b = (int)pos;
        if (b > bins - 1) b = bins - 1;
        const double f = pos - (double)b;
        const double q = (double)s_ct[b] + f * (double)(s_ct[b + 1] - s_ct[b]);
        const double qs = (q / nt) * ns;
        int lo = 0, hi = bins - 1;  // smallest b' with c_s[b'+1] >= qs
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if ((double)s_cs[mid + 1] >= qs) hi = mid; else lo = mid + 1;
        }
        const unsigned hs = s_cs[lo + 1] - s_cs[lo];
        double f2 = hs ? (qs - (double)s_cs[lo]) / (double)hs : 0.0;
        if (f2 < 0.0) f2 = 0.0;
        if (f2 > 1.0) f2 = 1.0;
        const double m = bs.lo + ((double)lo + f2) * w_s;
        double lab[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) lab[c] = m * pca.comp[c] + pca.mean[c];  // inverse_transform (_base.py:205)
        lut[jt.payA[e]] = lab2rgb(lab, &T, nullptr);
        if (dbg_matched0) {
            dbg_matched0[(size_t)slot * dbg_stride + e] = m;
            dbg_code0[(size_t)slot * dbg_stride + e] = (long long)b * 65536 + lo;
        }
    }
}

// ---- exchange buffer of the row-sharded binned path: contiguous images of the per-region statistics that the ranks
// all-reduce (torch.distributed over NCCL), packed from / unpacked into slot 0's FrameCtrl -----------------------------
//   which 0: int64 [2][28]  = 27 moment limbs + n_pix per region        (SUM)
//   which 1: double [4]     = max_t, -min_t, max_s, -min_s of the keys  (MAX)
//   which 2: uint32 [2][B]  = the two key histograms                    (SUM)
__global__ void k_xchg_pack(FrameCtrl *ctrl, void *xchg, int which, int bins)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (which == 0) {
        if (t < 56) {
            const RegionCtrl *rc = &ctrl[0].r[t / 28];
            const int k = t % 28;
            reinterpret_cast<long long *>(xchg)[t] = k < 27 ? rc->limbs[k / 3][k % 3] : (long long)rc->n_pix;
        }
    } else if (which == 1) {
        if (t < 4) {
            const RegionCtrl *rc = &ctrl[0].r[t / 2];
            // an empty band leaves the fields zero: -inf on both sides, neutral for MAX
            double v;
            if (t % 2 == 0) v = rc->lab_max[0] ? ord_dec(rc->lab_max[0]) : -INFINITY;
            else v = rc->lab_min[0] ? -ord_dec(~rc->lab_min[0]) : -INFINITY;
            reinterpret_cast<double *>(xchg)[t] = v;
        }
    } else {
        if (t < 2 * bins) reinterpret_cast<unsigned *>(xchg)[t] = ctrl[0].r[t / bins].coarse[t % bins];
    }
}
__global__ void k_xchg_unpack(FrameCtrl *ctrl, const void *xchg, int which, int bins)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (which == 0) {
        if (t < 56) {
            RegionCtrl *rc = &ctrl[0].r[t / 28];
            const int k = t % 28;
            const long long v = reinterpret_cast<const long long *>(xchg)[t];
            if (k < 27) rc->limbs[k / 3][k % 3] = v;
            else rc->n_pix = (unsigned long long)v;
        }
    } else if (which == 1) {
        if (t < 4) {
            RegionCtrl *rc = &ctrl[0].r[t / 2];
            const double v = reinterpret_cast<const double *>(xchg)[t];
            if (t % 2 == 0) rc->lab_max[0] = ord_enc(v);
            else rc->lab_min[0] = ~ord_enc(-v);
        }
    } else {
        if (t < 2 * bins) ctrl[0].r[t / bins].coarse[t % bins] = reinterpret_cast<const unsigned *>(xchg)[t];
    }
}
