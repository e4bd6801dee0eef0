// rct_sort.cuh -- exact ascending sort of the per-region key lists and the CDF match.
// Included by rct_kernels.cuh (needs RegionCtrl, Job, coarse_bin, block_excl_scan).
//
// Sort = distribution sort on a monotone image of the float64 key, then an exact local sort:
//
//   k_keys / k_coarse_hist   4096 uniform coarse bins over the key range (exact counts)
//   k_plan_fine              coarse bin with c keys -> ~c*F/d fine sub-bins (histogram
//                            equalisation; F ~ d/256, at most 16384)
//   k_fine_hist              exact fine-bin counts (shared-memory privatised)
//   k_plan_parts             consecutive fine bins grouped into parts of ~2048 entries
//                            (part = floor(exclusive offset / 2048), ids compacted)
//   k_scatter                entries -> their part (one fill counter per part; the write
//                            frontier is one sector per part, so L2 merges the stores)
//   k_part_sort              one CTA per part: counting sort on the estimated rank in shared
//                            memory, insertion sort inside the (tiny) buckets, part-local
//                            inclusive cumulative pixel counts.  Oversized parts / buckets take
//                            a bitonic fallback (shared or global memory) -- slow but exact.
//   k_part_scan              exclusive pixel bases of the parts
//
// Equal keys fall into the same fine bin, hence the same part and the same bucket: tie groups
// never straddle a part.  Final order is (key, payload) ascending, independent of the
// arrival order of the scatter, so results are deterministic.
#pragma once
// (included from inside namespace rct)

// key -> (fine bin, fraction inside the fine bin); lexicographically monotone in key
__device__ __forceinline__ unsigned fine_bin(double key, double klo, double cscale, const unsigned *fine_base,
                                             double &ff)
{
    double t;
    int cb = coarse_bin(key, klo, cscale, t);
    unsigned base = fine_base[cb], nb = fine_base[cb + 1] - base;
    double frac = t - (double)cb;
    double f = frac * (double)nb;
    unsigned sub = 0;
    if (f > 0.0) sub = (f >= (double)nb) ? nb - 1 : (unsigned)f;
    ff = f - (double)sub;
    if (!(ff > 0.0)) ff = 0.0;
    if (ff > 1.0) ff = 1.0;
    return base + sub;
}

// k_plan_fine: 1024 threads, 4 coarse bins each.
__global__ void __launch_bounds__(1024) k_plan_fine(const Job *jobs)
{
    const Job &job = jobs[blockIdx.y];
    RegionCtrl *rc = job.rc;
    if (rc->skip) return;
    __shared__ unsigned s_warp[33];
    const unsigned d = rc->d;
    unsigned F = d / 256u;
    if (F < 1u) F = 1u;
    if (F > (unsigned)kFineTarget) F = (unsigned)kFineTarget;
    unsigned nb[4], sum = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        unsigned c = rc->coarse[threadIdx.x * 4 + k];
        // round(c*F/d) >= 1; sum over bins <= F + kCoarse
        unsigned long long q = d ? ((unsigned long long)c * F + d / 2) / d : 0ull;
        nb[k] = q < 1ull ? 1u : (unsigned)q;
        sum += nb[k];
    }
    unsigned total;
    unsigned off = block_excl_scan(sum, s_warp, total);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        rc->fine_base[threadIdx.x * 4 + k] = off;
        off += nb[k];
    }
    if (threadIdx.x == 1023) {
        rc->fine_base[kCoarse] = total;
        rc->n_fine = total;
    }
}

// k_fine_hist: dynamic shared memory = (kFineMax + kCoarse + 1) * 4 bytes.
constexpr int kFineHistChunk = 65536;  // entries per block
__global__ void __launch_bounds__(512) k_fine_hist(const Job *jobs)
{
    const Job &job = jobs[blockIdx.y];
    RegionCtrl *rc = job.rc;
    if (rc->skip) return;
    const unsigned d = rc->d;
    const size_t e0 = (size_t)blockIdx.x * kFineHistChunk;
    if (e0 >= d) return;
    const size_t e1 = e0 + kFineHistChunk < d ? e0 + kFineHistChunk : d;
    extern __shared__ unsigned s_dyn[];
    unsigned *s_base = s_dyn;                 // [kCoarse+1]
    unsigned *s_hist = s_dyn + kCoarse + 1;   // [n_fine]
    const unsigned nf = rc->n_fine;
    for (unsigned i = threadIdx.x; i <= (unsigned)kCoarse; i += blockDim.x) s_base[i] = rc->fine_base[i];
    for (unsigned i = threadIdx.x; i < nf; i += blockDim.x) s_hist[i] = 0;
    __syncthreads();
    const double klo = rc->klo, cscale = rc->cscale;
    for (size_t e = e0 + threadIdx.x; e < e1; e += blockDim.x) {
        double ff;
        atomicAdd(&s_hist[fine_bin(job.keyA[e], klo, cscale, s_base, ff)], 1u);
    }
    __syncthreads();
    for (unsigned i = threadIdx.x; i < nf; i += blockDim.x) {
        unsigned v = s_hist[i];
        if (v) atomicAdd(&job.fcnt[i], v);
    }
}

// k_plan_parts: exclusive scan of the fine counts (in place), part ids, part starts.
// 1024 threads, kPlanPer consecutive fine bins each.
constexpr int kPlanPer = (kFineMax + 1023) / 1024;
__global__ void __launch_bounds__(1024) k_plan_parts(const Job *jobs)
{
    const Job &job = jobs[blockIdx.y];
    RegionCtrl *rc = job.rc;
    if (rc->skip) return;
    __shared__ unsigned s_warp[33];
    const unsigned nf = rc->n_fine, d = rc->d;
    const unsigned f0 = threadIdx.x * kPlanPer;
    unsigned c[kPlanPer], sum = 0;
#pragma unroll
    for (int k = 0; k < kPlanPer; ++k) {
        c[k] = (f0 + k < nf) ? job.fcnt[f0 + k] : 0u;
        sum += c[k];
    }
    unsigned total;
    unsigned off = block_excl_scan(sum, s_warp, total);
    // part "grid id" of a fine bin = floor(exclusive offset / kPartT); a new part starts where
    // the grid id changes.  prev_g for the first bin of this thread: offset of the bin before.
    unsigned prev_c = (f0 > 0 && f0 - 1 < nf) ? job.fcnt[f0 - 1] : 0u;
    __syncthreads();  // all reads of fcnt done before the in-place overwrite
    unsigned g_prev = (f0 > 0) ? (off - prev_c) / (unsigned)kPartT : 0u;
    unsigned flags = 0, nbound = 0, o = off;
#pragma unroll
    for (int k = 0; k < kPlanPer; ++k) {
        unsigned g = o / (unsigned)kPartT;
        if (f0 + k < nf) {
            job.fcnt[f0 + k] = o;
            if (f0 + k > 0 && g != g_prev) {
                flags |= 1u << k;
                ++nbound;
            }
        }
        g_prev = g;
        o += c[k];
    }
    unsigned tot_b;
    unsigned pid = block_excl_scan(nbound, s_warp, tot_b);  // boundaries before this thread's bins
    o = off;
#pragma unroll
    for (int k = 0; k < kPlanPer; ++k) {
        if (f0 + k < nf) {
            if (flags & (1u << k)) {
                ++pid;
                job.pstart[pid] = o;
            }
            job.pof[f0 + k] = (unsigned short)pid;
        }
        o += c[k];
    }
    if (threadIdx.x == 0) {
        job.pstart[0] = 0;
        job.pstart[tot_b + 1] = d;
        job.fcnt[nf] = d;
        rc->n_parts = tot_b + 1;
    }
}

// One sort record: the float64 key and (payload | count << 32), moved as one 16-byte word.
typedef ulonglong2 Rec;  // x: key bits, y: lo = payload, hi = multiplicity / part-local inclusive cumulative count
__device__ __forceinline__ double rec_key(const Rec &r) { return __longlong_as_double((long long)r.x); }
__device__ __forceinline__ unsigned rec_pay(const Rec &r) { return (unsigned)r.y; }
__device__ __forceinline__ unsigned rec_cnt(const Rec &r) { return (unsigned)(r.y >> 32); }
__device__ __forceinline__ Rec make_rec(double key, unsigned pay, unsigned cnt)
{
    Rec r;
    r.x = (unsigned long long)__double_as_longlong(key);
    r.y = (unsigned long long)pay | ((unsigned long long)cnt << 32);
    return r;
}
__device__ __forceinline__ bool rec_less(const Rec &a, const Rec &b)
{
    double ka = rec_key(a), kb = rec_key(b);
    return ka < kb || (ka == kb && rec_pay(a) < rec_pay(b));
}

// k_scatter: dynamic shared memory = (kCoarse + 1) * 4 + kFineMax * 2 bytes.
// Each thread moves kScatPer entries per trip so that the fill-counter atomics overlap.
constexpr int kScatPer = 4;
__global__ void __launch_bounds__(256) k_scatter(const Job *jobs)
{
    const Job &job = jobs[blockIdx.y];
    RegionCtrl *rc = job.rc;
    if (rc->skip) return;
    const unsigned d = rc->d;
    if ((size_t)blockIdx.x * blockDim.x * kScatPer >= d) return;
    extern __shared__ unsigned s_dyn[];
    unsigned *s_base = s_dyn;  // [kCoarse+1]
    unsigned short *s_pof = reinterpret_cast<unsigned short *>(s_dyn + kCoarse + 1);
    const unsigned nf = rc->n_fine;
    for (unsigned i = threadIdx.x; i <= (unsigned)kCoarse; i += blockDim.x) s_base[i] = rc->fine_base[i];
    for (unsigned i = threadIdx.x; i < nf; i += blockDim.x) s_pof[i] = job.pof[i];
    __syncthreads();
    const double klo = rc->klo, cscale = rc->cscale;
    Rec *recP = job.recP;
    const size_t stride = (size_t)gridDim.x * blockDim.x * kScatPer;
    for (size_t e0 = (size_t)blockIdx.x * blockDim.x * kScatPer + threadIdx.x; e0 < d; e0 += stride) {
        double key[kScatPer];
        unsigned p[kScatPer], pos[kScatPer];
#pragma unroll
        for (int k = 0; k < kScatPer; ++k) {
            size_t e = e0 + (size_t)k * blockDim.x;
            key[k] = e < d ? job.keyA[e] : 0.0;
        }
#pragma unroll
        for (int k = 0; k < kScatPer; ++k) {
            double ff;
            p[k] = s_pof[fine_bin(key[k], klo, cscale, s_base, ff)];
        }
#pragma unroll
        for (int k = 0; k < kScatPer; ++k) {
            size_t e = e0 + (size_t)k * blockDim.x;
            pos[k] = e < d ? atomicAdd(&job.pfill[(size_t)p[k] * kFillStride], 1u) : 0u;
        }
#pragma unroll
        for (int k = 0; k < kScatPer; ++k) {
            size_t e = e0 + (size_t)k * blockDim.x;
            if (e < d) {
                unsigned pay = job.payA ? job.payA[e] : (unsigned)e;
                unsigned cnt = job.cntA ? job.cntA[e] : 1u;
                recP[__ldg(job.pstart + p[k]) + pos[k]] = make_rec(key[k], pay, cnt);
            }
        }
    }
}

// ascending bitonic network over n (any n) records by (key, payload); works on shared or
// global memory (generic pointers); implicit +inf padding above n.
__device__ void block_bitonic(Rec *rec, unsigned n)
{
    for (unsigned long long k = 2; k < 2ull * n; k <<= 1) {
        for (unsigned long long j = k >> 1; j > 0; j >>= 1) {
            for (unsigned i = threadIdx.x; i < n; i += blockDim.x) {
                unsigned long long partner = (j == (k >> 1)) ? ((unsigned long long)i ^ (k - 1)) : ((unsigned long long)i ^ j);
                if (partner > i && partner < n) {
                    Rec a = rec[i], b = rec[partner];
                    if (rec_less(b, a)) {
                        rec[i] = b;
                        rec[partner] = a;
                    }
                }
            }
            __syncthreads();
        }
    }
}

__device__ __forceinline__ unsigned warp_incl_scan(unsigned v)
{
    const unsigned lane = threadIdx.x & 31u;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        unsigned y = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= (unsigned)o) v += y;
    }
    return v;
}

// In-place scan of s[0..n) in shared memory by the whole CTA (bank-conflict free: every
// warp sweeps a contiguous segment 32 elements at a time).  INCLUSIVE selects the flavour;
// GET reads element i, PUT stores the scanned value.  Returns the total.  Two barriers inside,
// one after.
template <bool INCLUSIVE, typename Get, typename Put>
__device__ __forceinline__ unsigned cta_scan(unsigned n, unsigned *s_wtot /*[33]*/, Get get, Put put)
{
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const unsigned seg = ((n + nwarps - 1) / nwarps + 31u) & ~31u;
    const unsigned b0 = warp * seg, b1 = (b0 + seg < n) ? b0 + seg : n;
    unsigned carry = 0;
    for (unsigned base = b0; base < b1; base += 32) {
        unsigned i = base + lane;
        unsigned v = i < b1 ? get(i) : 0u;
        unsigned inc = warp_incl_scan(v);
        if (i < b1) put(i, carry + (INCLUSIVE ? inc : inc - v));
        carry += __shfl_sync(0xffffffffu, inc, 31);
    }
    __syncthreads();  // also orders put() of phase 1 before the get() of phase 2
    if (lane == 0) s_wtot[warp] = carry;
    __syncthreads();
    if (warp == 0) {
        unsigned t = lane < nwarps ? s_wtot[lane] : 0u;
        unsigned inc = warp_incl_scan(t);
        if (lane < nwarps) s_wtot[lane] = inc - t;
        if (lane == 31) s_wtot[32] = inc;
    }
    __syncthreads();
    const unsigned wbase = s_wtot[warp];
    if (wbase)
        for (unsigned base = b0; base < b1; base += 32) {
            unsigned i = base + lane;
            if (i < b1) put(i, get(i) + wbase);
        }
    const unsigned total = s_wtot[32];
    __syncthreads();
    return total;
}

// k_part_sort: dynamic shared memory = kPartCap*16 + 2*kPartCap*4 bytes.  512 threads, up to
// kSortItems records per thread held in registers between the counting and the placing pass.
// `cap` (<= kPartCap) can be lowered by the caller to exercise the fallback.
constexpr int kSortThreads = 512;
constexpr int kSortItems = kPartCap / kSortThreads;
__global__ void __launch_bounds__(kSortThreads) k_part_sort(const Job *jobs, unsigned cap)
{
    const Job &job = jobs[blockIdx.y];
    RegionCtrl *rc = job.rc;
    if (rc->skip) return;
    const unsigned p = blockIdx.x;
    if (p >= rc->n_parts) return;
    const unsigned start = job.pstart[p], n = job.pstart[p + 1] - start;
    if (n == 0) {
        if (threadIdx.x == 0) job.ppix[p] = 0;
        return;
    }
    __shared__ unsigned s_warp[33];
    __shared__ unsigned s_flag;
    Rec *grec = job.recP + start;
    const unsigned tid = threadIdx.x;

    if (n > cap) {
        // fallback: the part does not fit in shared memory -> bitonic network in global memory
        block_bitonic(grec, n);
        unsigned carry = 0;
        for (unsigned base = 0; base < n; base += kSortThreads) {
            unsigned i = base + tid;
            Rec r = i < n ? grec[i] : make_rec(0.0, 0u, 0u);
            unsigned c = rec_cnt(r), total;
            unsigned ex = block_excl_scan(c, s_warp, total);
            if (i < n) grec[i] = make_rec(rec_key(r), rec_pay(r), carry + ex + c);
            carry += total;
        }
        if (tid == 0) job.ppix[p] = carry;
        return;
    }

    extern __shared__ unsigned s_dyn[];
    Rec *srec = reinterpret_cast<Rec *>(s_dyn);                    // [kPartCap]
    unsigned *sc = reinterpret_cast<unsigned *>(srec + kPartCap);  // [2*kPartCap]
    const unsigned nbk = 2 * n;
    for (unsigned i = tid; i < nbk; i += kSortThreads) sc[i] = 0;
    if (tid == 0) s_flag = 0;
    const double klo = rc->klo, cscale = rc->cscale;
    const unsigned *fine_base = rc->fine_base;
    const unsigned *foff = job.fcnt;
    Rec r[kSortItems];
    unsigned b[kSortItems];
#pragma unroll
    for (int k = 0; k < kSortItems; ++k) {
        unsigned i = tid + k * kSortThreads;
        if (i < n) r[k] = grec[i];
    }
#pragma unroll
    for (int k = 0; k < kSortItems; ++k) {
        unsigned i = tid + k * kSortThreads;
        if (i < n) {
            double ff;
            unsigned fb = fine_bin(rec_key(r[k]), klo, cscale, fine_base, ff);
            unsigned o0 = __ldg(foff + fb), o1 = __ldg(foff + fb + 1);
            double est = ((double)(o0 - start) + ff * (double)(o1 - o0)) * 2.0;  // estimated rank x 2
            unsigned bb = (unsigned)est;
            b[k] = bb < nbk ? bb : nbk - 1;
        }
    }
    __syncthreads();
    // pass 1: bucket sizes
#pragma unroll
    for (int k = 0; k < kSortItems; ++k)
        if (tid + k * kSortThreads < n) atomicAdd(&sc[b[k]], 1u);
    __syncthreads();
    cta_scan<false>(nbk, s_warp, [&](unsigned i) { return sc[i]; }, [&](unsigned i, unsigned v) { sc[i] = v; });
    // pass 2: place (sc[b] becomes the end of bucket b)
#pragma unroll
    for (int k = 0; k < kSortItems; ++k)
        if (tid + k * kSortThreads < n) srec[atomicAdd(&sc[b[k]], 1u)] = r[k];
    __syncthreads();
    // exact order inside each bucket (buckets hold ~0.5 records on average)
    for (unsigned bk = tid; bk < nbk; bk += kSortThreads) {
        unsigned lo = bk ? sc[bk - 1] : 0u, hi = sc[bk];
        if (hi - lo > 32u) {
            s_flag = 1;  // clustered keys: leave it to the block-wide network below
            continue;
        }
        for (unsigned i = lo + 1; i < hi; ++i) {
            Rec x = srec[i];
            unsigned j = i;
            while (j > lo && rec_less(x, srec[j - 1])) {
                srec[j] = srec[j - 1];
                --j;
            }
            srec[j] = x;
        }
    }
    __syncthreads();
    if (s_flag) block_bitonic(srec, n);  // ends with a barrier
    // part-local inclusive cumulative counts
    unsigned total = cta_scan<true>(
        n, s_warp, [&](unsigned i) { return (unsigned)(srec[i].y >> 32); },
        [&](unsigned i, unsigned v) { srec[i].y = (srec[i].y & 0xffffffffull) | ((unsigned long long)v << 32); });
    if (tid == 0) job.ppix[p] = total;
    for (unsigned i = tid; i < n; i += kSortThreads) grec[i] = srec[i];
}

// k_part_scan: exclusive scan of the part pixel totals (in place); ppix[n_parts] = N.
__global__ void __launch_bounds__(1024) k_part_scan(const Job *jobs)
{
    const Job &job = jobs[blockIdx.y];
    RegionCtrl *rc = job.rc;
    if (rc->skip) return;
    __shared__ unsigned s_warp[33];
    const unsigned np = rc->n_parts;
    unsigned carry = 0;
    for (unsigned base = 0; base < np; base += 1024) {
        unsigned i = base + threadIdx.x;
        unsigned v = i < np ? job.ppix[i] : 0u, total;
        unsigned ex = block_excl_scan(v, s_warp, total);
        if (i < np) job.ppix[i] = carry + ex;
        carry += total;
    }
    if (threadIdx.x == 0) job.ppix[np] = carry;
}

// ------------------------------------------------------------------------------------------
// a-6: match_cdf with np.linspace / np.interp semantics on the sorted (key, cumulative count)
// representation.  A multiset of n values is d distinct-entry records; the q-th smallest
// value (0-based) is the key of the first entry whose cumulative count exceeds q.
// ------------------------------------------------------------------------------------------
struct Sorted {
    const Rec *rec;                  // sorted; hi word of .y = part-local inclusive cumulative count
    const unsigned *pstart, *pbase;  // [n_parts+1]
    unsigned n_parts, d;
    long long n;
    double inv;                      // 1.0 / (n - 1): the step of np.linspace(0, 1, n)
};

__device__ __forceinline__ Sorted sorted_of(const Job &j)
{
    long long n = (long long)j.rc->n_pix;
    return Sorted{j.recP, j.pstart, j.ppix, j.rc->n_parts, j.rc->d, n, n > 1 ? 1.0 / (double)(n - 1) : 0.0};
}

// np.linspace(0, 1, v.n)[k]  (same arithmetic as lin(), with the reciprocal hoisted)
__device__ __forceinline__ double lin_of(const Sorted &v, long long k)
{
    if (v.n == 1) return 0.0;
    if (k == v.n - 1) return 1.0;
    return (double)k * v.inv;
}

struct Hint {
    unsigned p, i;
    bool valid;
};

__device__ __forceinline__ unsigned cum_local(const Sorted &v, unsigned i)
{
    return (unsigned)(__ldg(reinterpret_cast<const unsigned long long *>(v.rec + i) + 1) >> 32);
}
__device__ __forceinline__ double key_at(const Sorted &v, unsigned i)
{
    return __longlong_as_double((long long)__ldg(reinterpret_cast<const unsigned long long *>(v.rec + i)));
}

// index of the entry holding the q-th smallest value; h carries the last position found
__device__ unsigned os_index(const Sorted &v, long long q, Hint &h)
{
    unsigned p;
    bool near = false;
    if (h.valid && q >= (long long)__ldg(v.pbase + h.p)) {
        // ranks move forward slowly: same part, or one of the next two
        p = h.p;
        int hop = 0;
        while (hop < 3 && q >= (long long)__ldg(v.pbase + p + 1)) {
            ++p;
            ++hop;
        }
        near = q < (long long)__ldg(v.pbase + p + 1);
    }
    if (!near) {
        unsigned lo = 0, hi = v.n_parts - 1;  // last part with pbase <= q
        while (lo < hi) {
            unsigned mid = (lo + hi + 1) >> 1;
            if ((long long)__ldg(v.pbase + mid) <= q) lo = mid; else hi = mid - 1;
        }
        p = lo;
    }
    const unsigned b0 = __ldg(v.pbase + p);
    const unsigned ql = (unsigned)(q - (long long)b0);
    const unsigned ps = __ldg(v.pstart + p), pe = __ldg(v.pstart + p + 1);
    unsigned idx;
    if (h.valid && h.p == p && h.i >= ps && h.i < pe) {
        idx = h.i;
    } else {  // interpolate inside the part (multiplicities are roughly even within a part)
        unsigned tot = __ldg(v.pbase + p + 1) - b0;
        idx = ps + (unsigned)(((unsigned long long)ql * (pe - ps)) / tot);
        if (idx >= pe) idx = pe - 1;
    }
    h.p = p;
    h.valid = true;
#pragma unroll 1
    for (int s = 0; s < 8; ++s) {
        if (cum_local(v, idx) <= ql) {
            ++idx;  // stays < pe: ql < part total = cum_local(pe-1)
        } else if (idx > ps && cum_local(v, idx - 1) > ql) {
            --idx;
        } else {
            h.i = idx;
            return idx;
        }
    }
    unsigned lo = ps, hi = pe - 1;  // smallest idx with cum_local > ql
    while (lo < hi) {
        unsigned mid = (lo + hi) >> 1;
        if (cum_local(v, mid) > ql) hi = mid; else lo = mid + 1;
    }
    h.i = lo;
    return lo;
}

__device__ __forceinline__ double os_value(const Sorted &v, long long q, Hint &h) { return key_at(v, os_index(v, q, h)); }

// np.interp(x, np.linspace(0,1,v.n), sorted(v)) for 0 <= x <= 1
__device__ double interp_lin_os(double x, const Sorted &v, Hint &h)
{
    const long long n = v.n;
    if (n == 1) return os_value(v, 0, h);
    long long m = (long long)(x * (double)(n - 1));
    if (m < 0) m = 0;
    if (m > n - 1) m = n - 1;
    while (m + 1 <= n - 1 && lin_of(v, m + 1) <= x) ++m;
    while (m > 0 && lin_of(v, m) > x) --m;
    if (m >= n - 1) return os_value(v, n - 1, h);
    double xm = lin_of(v, m);
    double fm = os_value(v, m, h);
    if (xm == x) return fm;
    double fm1 = os_value(v, m + 1, h);
    double slope = (fm1 - fm) / (lin_of(v, m + 1) - xm);
    return slope * (x - xm) + fm;
}

// entry i of the sorted target (inside part p); ht / hs: search hints into target / source
__device__ double match_one(const Sorted &t, const Sorted &s, unsigned p, unsigned i, Hint &ht, Hint &hs,
                            long long &j_out)
{
    const double key = key_at(t, i);
    const unsigned pe = __ldg(t.pstart + p + 1);
    unsigned i2 = i;
    while (i2 + 1 < pe && key_at(t, i2 + 1) == key) ++i2;  // ties share the upper rank
    const long long r = (long long)__ldg(t.pbase + p) + (long long)cum_local(t, i2) - 1;
    const long long T = t.n, S = s.n;
    double m;
    long long j;
    if (T == S) {
        j = r;
        m = os_value(s, r, hs);
    } else if (T > S) {  // source stretched to T samples (LST:110-115)
        j = r;
        m = interp_lin_os(lin_of(t, r), s, hs);
    } else {  // target stretched to S samples (LST:116-121): t'[k] = interp(lin(S,k), lin(T,.), t)
        ht.p = p;
        ht.i = i2;
        ht.valid = true;
        // j = last k with t'[k] <= key; t'[j] and t'[j+1] are kept for the interpolation
        double tj = 0.0, tj1 = 0.0;
        bool have_j1 = false;
        if (T == 1) {
            j = S - 1;
        } else {
            double xr = lin_of(t, r);
            long long k = (long long)(xr * (double)(S - 1));
            if (k < 0) k = 0;
            if (k > S - 1) k = S - 1;
            double tk = interp_lin_os(lin_of(s, k), t, ht);
            if (tk <= key) {
                while (k + 1 <= S - 1) {
                    double tn = interp_lin_os(lin_of(s, k + 1), t, ht);
                    if (tn <= key) {
                        ++k;
                        tk = tn;
                    } else {
                        tj1 = tn;
                        have_j1 = true;
                        break;
                    }
                }
            } else {
                while (k > 0 && tk > key) {
                    tj1 = tk;
                    have_j1 = true;
                    --k;
                    tk = interp_lin_os(lin_of(s, k), t, ht);
                }
            }
            j = k;
            tj = tk;
        }
        if (j >= S - 1) {
            m = os_value(s, S - 1, hs);
        } else {
            double sj = os_value(s, j, hs);
            if (tj == key) {
                m = sj;
            } else {
                if (!have_j1) tj1 = interp_lin_os(lin_of(s, j + 1), t, ht);
                double slope = (os_value(s, j + 1, hs) - sj) / (tj1 - tj);
                m = slope * (key - tj) + sj;
            }
        }
    }
    j_out = j;
    return m;
}

// pipeline flavour: jobs[2*slot] = target, jobs[2*slot+1] = source; one CTA per target part.
// Writes the colour -> output-RGB LUT (a-7: inverse_transform, lab -> rgb, clip, truncate).
__global__ void __launch_bounds__(256) k_match(FrameCtrl *ctrl, const Job *jobs, unsigned *lut0, size_t lut_stride,
                                               double *dbg_matched0, long long *dbg_j0, size_t dbg_stride)
{
    const int slot = blockIdx.y;
    if (ctrl[slot].status != 0) return;
    const Job &jt = jobs[2 * slot], &js = jobs[2 * slot + 1];
    const unsigned p = blockIdx.x;
    if (p >= jt.rc->n_parts) return;
    const Sorted t = sorted_of(jt), s = sorted_of(js);
    const unsigned ps = jt.pstart[p], pe = jt.pstart[p + 1];
    if (ps == pe) return;
    __shared__ MathTabs T;
    stage_tabs(&T);
    const Pca pca = jt.rc->pca;
    unsigned *lut = lut0 + (size_t)slot * lut_stride;
    Hint ht{0, 0, false}, hs{0, 0, false};
    for (unsigned i = ps + threadIdx.x; i < pe; i += blockDim.x) {
        long long j;
        double m = match_one(t, s, p, i, ht, hs, j);
        double lab[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) lab[c] = m * pca.comp[c] + pca.mean[c];  // inverse_transform
        lut[rec_pay(__ldg(t.rec + i))] = lab2rgb(lab, &T, nullptr);
        if (dbg_matched0) {
            dbg_matched0[(size_t)slot * dbg_stride + i] = m;
            dbg_j0[(size_t)slot * dbg_stride + i] = j;
        }
    }
}

// generic flavour (rct_match_cdf): jobs[0] = target, jobs[1] = source; payload = input index
__global__ void __launch_bounds__(256) k_match_generic(const Job *jobs, double *matched, long long *jidx)
{
    const Job &jt = jobs[0], &js = jobs[1];
    const unsigned p = blockIdx.x;
    if (p >= jt.rc->n_parts) return;
    const Sorted t = sorted_of(jt), s = sorted_of(js);
    const unsigned ps = jt.pstart[p], pe = jt.pstart[p + 1];
    Hint ht{0, 0, false}, hs{0, 0, false};
    for (unsigned i = ps + threadIdx.x; i < pe; i += blockDim.x) {
        long long j;
        double m = match_one(t, s, p, i, ht, hs, j);
        unsigned pay = rec_pay(__ldg(t.rec + i));
        matched[pay] = m;
        if (jidx) jidx[pay] = j;
    }
}
