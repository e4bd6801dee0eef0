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

// k_scatter: dynamic shared memory = (kCoarse + 1) * 4 + kFineMax * 2 bytes.
__global__ void __launch_bounds__(256) k_scatter(const Job *jobs)
{
    const Job &job = jobs[blockIdx.y];
    RegionCtrl *rc = job.rc;
    if (rc->skip) return;
    const unsigned d = rc->d;
    if ((size_t)blockIdx.x * blockDim.x >= d) return;
    extern __shared__ unsigned s_dyn[];
    unsigned *s_base = s_dyn;  // [kCoarse+1]
    unsigned short *s_pof = reinterpret_cast<unsigned short *>(s_dyn + kCoarse + 1);
    const unsigned nf = rc->n_fine;
    for (unsigned i = threadIdx.x; i <= (unsigned)kCoarse; i += blockDim.x) s_base[i] = rc->fine_base[i];
    for (unsigned i = threadIdx.x; i < nf; i += blockDim.x) s_pof[i] = job.pof[i];
    __syncthreads();
    const double klo = rc->klo, cscale = rc->cscale;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < d; e += (size_t)gridDim.x * blockDim.x) {
        double key = job.keyA[e], ff;
        unsigned p = s_pof[fine_bin(key, klo, cscale, s_base, ff)];
        unsigned pos = __ldg(job.pstart + p) + atomicAdd(&job.pfill[(size_t)p * kFillStride], 1u);
        unsigned pay = job.payA ? job.payA[e] : (unsigned)e;
        unsigned cnt = job.cntA ? job.cntA[e] : 1u;
        job.keyP[pos] = key;
        job.auxP[pos] = (unsigned long long)pay | ((unsigned long long)cnt << 32);
    }
}

__device__ __forceinline__ bool rec_less(double ka, unsigned pa, double kb, unsigned pb)
{
    return ka < kb || (ka == kb && pa < pb);
}

// ascending bitonic network over n (any n) records by (key, payload); works on shared or
// global memory (generic pointers); implicit +inf padding above n.
__device__ void block_bitonic(double *key, unsigned long long *aux, unsigned n)
{
    for (unsigned long long k = 2; k < 2ull * n; k <<= 1) {
        for (unsigned long long j = k >> 1; j > 0; j >>= 1) {
            for (unsigned i = threadIdx.x; i < n; i += blockDim.x) {
                unsigned long long partner = (j == (k >> 1)) ? ((unsigned long long)i ^ (k - 1)) : ((unsigned long long)i ^ j);
                if (partner > i && partner < n) {
                    double ka = key[i], kb = key[partner];
                    unsigned long long aa = aux[i], ab = aux[partner];
                    if (rec_less(kb, (unsigned)ab, ka, (unsigned)aa)) {
                        key[i] = kb;
                        key[partner] = ka;
                        aux[i] = ab;
                        aux[partner] = aa;
                    }
                }
            }
            __syncthreads();
        }
    }
}

// k_part_sort: dynamic shared memory = cap*16 + 2*cap*4 bytes (cap = kPartCap unless lowered
// by the caller for testing the fallback).  512 threads.
constexpr int kSortThreads = 512;
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
    double *gkey = job.keyP + start;
    unsigned long long *gaux = job.auxP + start;
    const unsigned tid = threadIdx.x;

    if (n > cap) {
        // fallback: the part does not fit in shared memory -> bitonic network in global memory
        block_bitonic(gkey, gaux, n);
        unsigned carry = 0;
        for (unsigned base = 0; base < n; base += kSortThreads) {
            unsigned i = base + tid;
            unsigned long long a = i < n ? gaux[i] : 0ull;
            unsigned c = (unsigned)(a >> 32), total;
            unsigned ex = block_excl_scan(c, s_warp, total);
            if (i < n) gaux[i] = (a & 0xffffffffull) | ((unsigned long long)(carry + ex + c) << 32);
            carry += total;
        }
        if (tid == 0) job.ppix[p] = carry;
        return;
    }

    extern __shared__ unsigned s_dyn[];
    double *skey = reinterpret_cast<double *>(s_dyn);                               // [cap]
    unsigned long long *saux = reinterpret_cast<unsigned long long *>(skey + cap);  // [cap]
    unsigned *sc = reinterpret_cast<unsigned *>(saux + cap);                        // [2*cap]
    const unsigned nbk = 2 * n;
    for (unsigned i = tid; i < nbk; i += kSortThreads) sc[i] = 0;
    if (tid == 0) s_flag = 0;
    __syncthreads();
    const double klo = rc->klo, cscale = rc->cscale;
    const unsigned *fine_base = rc->fine_base;
    const unsigned *foff = job.fcnt;
    auto bucket_of = [&](double key) -> unsigned {
        double ff;
        unsigned fb = fine_bin(key, klo, cscale, fine_base, ff);
        unsigned o0 = __ldg(foff + fb), o1 = __ldg(foff + fb + 1);
        double est = ((double)(o0 - start) + ff * (double)(o1 - o0)) * 2.0;
        unsigned b = (unsigned)est;
        return b < nbk ? b : nbk - 1;
    };
    // pass 1: bucket sizes
    for (unsigned i = tid; i < n; i += kSortThreads) atomicAdd(&sc[bucket_of(gkey[i])], 1u);
    __syncthreads();
    // exclusive scan of sc[0..nbk): thread owns kPer consecutive buckets
    {
        const unsigned per = (nbk + kSortThreads - 1) / kSortThreads;
        const unsigned b0 = tid * per;
        unsigned sum = 0;
        for (unsigned k = 0; k < per; ++k)
            if (b0 + k < nbk) sum += sc[b0 + k];
        unsigned total;
        unsigned off = block_excl_scan(sum, s_warp, total);
        for (unsigned k = 0; k < per; ++k)
            if (b0 + k < nbk) {
                unsigned v = sc[b0 + k];
                sc[b0 + k] = off;
                off += v;
            }
    }
    __syncthreads();
    // pass 2: place (sc[b] becomes the end of bucket b)
    for (unsigned i = tid; i < n; i += kSortThreads) {
        double key = gkey[i];
        unsigned long long a = gaux[i];
        unsigned pos = atomicAdd(&sc[bucket_of(key)], 1u);
        skey[pos] = key;
        saux[pos] = a;
    }
    __syncthreads();
    // exact order inside each bucket
    for (unsigned b = tid; b < nbk; b += kSortThreads) {
        unsigned lo = b ? sc[b - 1] : 0u, hi = sc[b];
        if (hi - lo > 32u) {
            s_flag = 1;  // clustered keys: leave it to the block-wide network below
            continue;
        }
        for (unsigned i = lo + 1; i < hi; ++i) {
            double k = skey[i];
            unsigned long long a = saux[i];
            unsigned j = i;
            while (j > lo && rec_less(k, (unsigned)a, skey[j - 1], (unsigned)saux[j - 1])) {
                skey[j] = skey[j - 1];
                saux[j] = saux[j - 1];
                --j;
            }
            skey[j] = k;
            saux[j] = a;
        }
    }
    __syncthreads();
    if (s_flag) block_bitonic(skey, saux, n);  // ends with a barrier
    // part-local inclusive cumulative counts, coalesced write-back
    {
        const unsigned per = (n + kSortThreads - 1) / kSortThreads;
        const unsigned i0 = tid * per;
        unsigned sum = 0;
        for (unsigned k = 0; k < per; ++k)
            if (i0 + k < n) sum += (unsigned)(saux[i0 + k] >> 32);
        unsigned total;
        unsigned off = block_excl_scan(sum, s_warp, total);
        for (unsigned k = 0; k < per; ++k)
            if (i0 + k < n) {
                unsigned long long a = saux[i0 + k];
                off += (unsigned)(a >> 32);
                saux[i0 + k] = (a & 0xffffffffull) | ((unsigned long long)off << 32);
            }
        if (tid == 0) job.ppix[p] = total;
    }
    __syncthreads();
    for (unsigned i = tid; i < n; i += kSortThreads) {
        gkey[i] = skey[i];
        gaux[i] = saux[i];
    }
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
    const double *key;
    const unsigned long long *aux;  // lo: payload, hi: part-local inclusive cumulative count
    const unsigned *pstart, *pbase; // [n_parts+1]
    unsigned n_parts, d;
    long long n;
};

__device__ __forceinline__ Sorted sorted_of(const Job &j)
{
    return Sorted{j.keyP, j.auxP, j.pstart, j.ppix, j.rc->n_parts, j.rc->d, (long long)j.rc->n_pix};
}

struct Hint {
    unsigned p, i;
    bool valid;
};

__device__ __forceinline__ unsigned cum_local(const Sorted &v, unsigned i) { return (unsigned)(__ldg(v.aux + i) >> 32); }

// index of the entry holding the q-th smallest value; h carries the last position found
__device__ unsigned os_index(const Sorted &v, long long q, Hint &h)
{
    unsigned p;
    if (h.valid && q >= (long long)__ldg(v.pbase + h.p) && q < (long long)__ldg(v.pbase + h.p + 1)) {
        p = h.p;
    } else {
        unsigned lo = 0, hi = v.n_parts - 1;  // last part with pbase <= q
        while (lo < hi) {
            unsigned mid = (lo + hi + 1) >> 1;
            if ((long long)__ldg(v.pbase + mid) <= q) lo = mid; else hi = mid - 1;
        }
        p = lo;
        h.valid = false;
    }
    const unsigned ql = (unsigned)(q - (long long)__ldg(v.pbase + p));
    const unsigned ps = __ldg(v.pstart + p), pe = __ldg(v.pstart + p + 1);
    if (h.valid && h.i >= ps && h.i < pe) {  // walk from the last position (ranks move slowly)
        unsigned idx = h.i;
#pragma unroll 1
        for (int s = 0; s < 6; ++s) {
            if (cum_local(v, idx) <= ql) {
                ++idx;  // stays < pe: ql < part total = cum_local(pe-1)
            } else if (idx > ps && cum_local(v, idx - 1) > ql) {
                --idx;
            } else {
                h.i = idx;
                return idx;
            }
        }
    }
    unsigned lo = ps, hi = pe - 1;  // smallest idx with cum_local > ql
    while (lo < hi) {
        unsigned mid = (lo + hi) >> 1;
        if (cum_local(v, mid) > ql) hi = mid; else lo = mid + 1;
    }
    h.p = p;
    h.i = lo;
    h.valid = true;
    return lo;
}

__device__ __forceinline__ double os_value(const Sorted &v, long long q, Hint &h) { return __ldg(v.key + os_index(v, q, h)); }

// np.interp(x, np.linspace(0,1,v.n), sorted(v)) for 0 <= x <= 1
__device__ double interp_lin_os(double x, const Sorted &v, Hint &h)
{
    const long long n = v.n;
    if (n == 1) return os_value(v, 0, h);
    long long m = (long long)(x * (double)(n - 1));
    if (m < 0) m = 0;
    if (m > n - 1) m = n - 1;
    while (m + 1 <= n - 1 && lin(n, m + 1) <= x) ++m;
    while (m > 0 && lin(n, m) > x) --m;
    if (m >= n - 1) return os_value(v, n - 1, h);
    double xm = lin(n, m);
    double fm = os_value(v, m, h);
    if (xm == x) return fm;
    double fm1 = os_value(v, m + 1, h);
    double slope = (fm1 - fm) / (lin(n, m + 1) - xm);
    return slope * (x - xm) + fm;
}

// entry i of the sorted target (inside part p); ht / hs: search hints into target / source
__device__ double match_one(const Sorted &t, const Sorted &s, unsigned p, unsigned i, Hint &ht, Hint &hs,
                            long long &j_out)
{
    const double key = __ldg(t.key + i);
    const unsigned pe = __ldg(t.pstart + p + 1);
    unsigned i2 = i;
    while (i2 + 1 < pe && __ldg(t.key + i2 + 1) == key) ++i2;  // ties share the upper rank
    const long long r = (long long)__ldg(t.pbase + p) + (long long)cum_local(t, i2) - 1;
    const long long T = t.n, S = s.n;
    double m;
    long long j;
    if (T == S) {
        j = r;
        m = os_value(s, r, hs);
    } else if (T > S) {  // source stretched to T samples (LST:110-115)
        j = r;
        m = interp_lin_os(lin(T, r), s, hs);
    } else {  // target stretched to S samples (LST:116-121): t'[k] = interp(lin(S,k), lin(T,.), t)
        ht.p = p;
        ht.i = i2;
        ht.valid = true;
        if (T == 1) {
            j = S - 1;
        } else {
            double xr = lin(T, r);
            long long k = (long long)(xr * (double)(S - 1));
            if (k < 0) k = 0;
            if (k > S - 1) k = S - 1;
            while (k + 1 <= S - 1 && interp_lin_os(lin(S, k + 1), t, ht) <= key) ++k;
            while (k > 0 && interp_lin_os(lin(S, k), t, ht) > key) --k;
            j = k;
        }
        if (j >= S - 1) {
            m = os_value(s, S - 1, hs);
        } else {
            double tj = interp_lin_os(lin(S, j), t, ht);
            double sj = os_value(s, j, hs);
            if (tj == key) {
                m = sj;
            } else {
                double slope = (os_value(s, j + 1, hs) - sj) / (interp_lin_os(lin(S, j + 1), t, ht) - tj);
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
        lut[(unsigned)__ldg(t.aux + i)] = lab2rgb(lab, &T, nullptr);
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
        unsigned pay = (unsigned)__ldg(t.aux + i);
        matched[pay] = m;
        if (jidx) jidx[pay] = j;
    }
}

