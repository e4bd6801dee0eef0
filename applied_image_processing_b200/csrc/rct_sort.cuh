// rct_sort.cuh -- exact ascending sort of the per-region key lists and the CDF match.
// Included by rct_kernels.cuh (needs RegionCtrl, Job, coarse_bin, block_excl_scan).
//
// Sort = distribution sort on a monotone image of the float64 key, then an exact local sort:
//
//   k_keys / k_coarse_hist   4096 uniform coarse bins over the key range (exact counts)
//   k_plan_fine              coarse bin with c keys -> ~c*F/d fine sub-bins (histogram
//                            equalisation; F ~ d/256, at most 16384)
//   k_fine_hist              exact fine-bin counts (shared-memory privatised)
//   k_plan_parts             consecutive fine bins grouped into parts of ~kPartT (3072) entries
//                            (part = floor(exclusive offset / kPartT), ids compacted)
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
#ifndef RCT_FDIV
#define RCT_FDIV 256
#endif
    unsigned F = d / (unsigned)RCT_FDIV;
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
#ifndef RCT_FHC
#define RCT_FHC 65536
#endif
constexpr int kFineHistChunk = RCT_FHC;  // entries per block
#ifndef RCT_FHTH
#define RCT_FHTH 1024
#endif
__global__ void __launch_bounds__(RCT_FHTH) k_fine_hist(const Job *jobs)
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
                job.pfill[(size_t)pid * kFillStride] = o;  // fill counters start at the part's first entry
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

// k_scatter: entries -> their part.  The fine-bin tables (fine_base 16 KB, pof <= 40 KB) are staged
// in shared memory once per CTA (two of the five dependent memory round trips of an entry used
// to be L2 hits on them), and the fill counters are seeded with the part starts by k_plan_parts, so
// the atomic returns the record's final position.  Each thread moves kScatPer entries per trip so
// that the loads and the fill-counter atomics overlap.
#ifndef RCT_SCATPER
#define RCT_SCATPER 4
#endif
constexpr int kScatPer = RCT_SCATPER;
constexpr size_t kSmemScatter = (size_t)(kCoarse + 1) * 4 + (size_t)kFineMax * 2 + 16;
#ifndef RCT_SCT
#define RCT_SCT 384
#endif
#ifndef RCT_SCTB
#define RCT_SCTB 3
#endif
#define RCT_SCT_LB __launch_bounds__(RCT_SCT, RCT_SCTB)
__global__ void RCT_SCT_LB k_scatter(const Job *jobs)
{
    const Job &job = jobs[blockIdx.y];
    RegionCtrl *rc = job.rc;
    if (rc->skip) return;
    const unsigned d = rc->d;
    if ((size_t)blockIdx.x * blockDim.x * kScatPer >= d) return;
    extern __shared__ __align__(16) unsigned s_scat[];
    unsigned *s_base = s_scat;                                                       // [kCoarse + 1]
    unsigned short *s_pof = reinterpret_cast<unsigned short *>(s_scat + kCoarse + 2);  // [n_fine]
    {
        const unsigned nf2 = (rc->n_fine + 1u) / 2u;
        const unsigned *pof32 = reinterpret_cast<const unsigned *>(job.pof);
        unsigned *s_pof32 = s_scat + kCoarse + 2;
        for (unsigned i = threadIdx.x; i <= (unsigned)kCoarse; i += blockDim.x) s_base[i] = rc->fine_base[i];
        for (unsigned i = threadIdx.x; i < nf2; i += blockDim.x) s_pof32[i] = pof32[i];
        __syncthreads();
    }
    const double klo = rc->klo, cscale = rc->cscale;
    Rec *recP = job.recP;
    const size_t stride = (size_t)gridDim.x * blockDim.x * kScatPer;
    for (size_t e0 = (size_t)blockIdx.x * blockDim.x * kScatPer + threadIdx.x; e0 < d; e0 += stride) {
        double key[kScatPer];
        unsigned p[kScatPer], pos[kScatPer], pay[kScatPer], cnt[kScatPer];
#pragma unroll
        for (int k = 0; k < kScatPer; ++k) {
            size_t e = e0 + (size_t)k * blockDim.x;
            key[k] = e < d ? job.keyA[e] : 0.0;
            pay[k] = (e < d && job.payA) ? job.payA[e] : (unsigned)e;
            cnt[k] = (e < d && job.cntA) ? job.cntA[e] : 1u;
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
            if (e < d) recP[pos[k]] = make_rec(key[k], pay[k], cnt[k]);
        }
    }
}

// k_scatter_tiled: the same scatter with one GLOBAL reservation per (CTA tile, part) instead of one per entry.
// A CTA of 1024 threads takes tiles of kScat2Tile = 16 K entries: every entry's part is ranked inside the tile by a
// shared-memory atomic, each part the tile touches reserves its run of the part's slice once, and the records go
// out in runs (~10 consecutive 16-byte records per part on a 4K noise frame: the atomics to L2 drop ~10x and the
// stores become 160-byte runs).  The second sweep re-reads the tile's keys / payloads (L2-resident: 256 KB per CTA)
// rather than holding 16 records per thread in registers.  Shared memory: fine-bin tables as in k_scatter + one
// counter per part (parts <= kScat2MaxParts); jobs with more parts take k_scatter.
constexpr int kScat2Threads = 1024;
constexpr int kScat2Per = 16;
constexpr int kScat2Tile = kScat2Threads * kScat2Per;
constexpr int kScat2MaxParts = 8192;
constexpr size_t kSmemScatter2 = kSmemScatter + (size_t)kScat2MaxParts * 4;
__global__ void __launch_bounds__(kScat2Threads, 1) k_scatter_tiled(const Job *jobs)
{
    const Job &job = jobs[blockIdx.y];
    RegionCtrl *rc = job.rc;
    if (rc->skip) return;
    const unsigned d = rc->d;
    if ((size_t)blockIdx.x * kScat2Tile >= d) return;
    extern __shared__ __align__(16) unsigned s_scat[];
    unsigned *s_base = s_scat;                                                         // [kCoarse + 1]
    unsigned short *s_pof = reinterpret_cast<unsigned short *>(s_scat + kCoarse + 2);  // [n_fine]
    unsigned *s_cnt = s_scat + (kSmemScatter / 4);                                     // [n_parts]: count, then base
    const unsigned np = rc->n_parts;
    {
        const unsigned nf2 = (rc->n_fine + 1u) / 2u;
        const unsigned *pof32 = reinterpret_cast<const unsigned *>(job.pof);
        unsigned *s_pof32 = s_scat + kCoarse + 2;
        for (unsigned i = threadIdx.x; i <= (unsigned)kCoarse; i += blockDim.x) s_base[i] = rc->fine_base[i];
        for (unsigned i = threadIdx.x; i < nf2; i += blockDim.x) s_pof32[i] = pof32[i];
    }
    const double klo = rc->klo, cscale = rc->cscale;
    Rec *recP = job.recP;
    for (size_t t0 = (size_t)blockIdx.x * kScat2Tile; t0 < d; t0 += (size_t)gridDim.x * kScat2Tile) {
        for (unsigned i = threadIdx.x; i < np; i += blockDim.x) s_cnt[i] = 0u;
        __syncthreads();
        unsigned pr[kScat2Per];  // part | rank inside (tile, part) << 16  (rank < 16384)
#pragma unroll
        for (int k = 0; k < kScat2Per; ++k) {
            const size_t e = t0 + (size_t)k * kScat2Threads + threadIdx.x;
            pr[k] = 0xffffffffu;
            if (e < d) {
                double ff;
                const unsigned p = s_pof[fine_bin(job.keyA[e], klo, cscale, s_base, ff)];
                pr[k] = p | (atomicAdd(&s_cnt[p], 1u) << 16);
            }
        }
        __syncthreads();
        for (unsigned i = threadIdx.x; i < np; i += blockDim.x) {
            const unsigned c = s_cnt[i];
            if (c) s_cnt[i] = atomicAdd(&job.pfill[(size_t)i * kFillStride], c);
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < kScat2Per; ++k) {
            const size_t e = t0 + (size_t)k * kScat2Threads + threadIdx.x;
            if (pr[k] != 0xffffffffu) {
                const unsigned pos = s_cnt[pr[k] & 0xffffu] + (pr[k] >> 16);
                recP[pos] = make_rec(job.keyA[e], job.payA ? job.payA[e] : (unsigned)e, job.cntA ? job.cntA[e] : 1u);
            }
        }
        __syncthreads();
    }
}

// k_scatter_runs: the scatter with the tile SORTED BY PART in shared memory before it leaves the SM.
// ncu on k_scatter (4 x 4K U): every entry costs the L1 three uncoalesced sectors (its fill-counter atomic, its 16-byte
// record store, 0.9 of coalesced loads) and the kernel sits at 63 % of the L1's and 45 % of the L2's request
// throughput with the warps waiting on the returning atomics -- request-bound, not byte-bound.  Here a CTA of 1024
// threads takes tiles of kScat3Tile = 6144 entries: (1) every entry is ranked inside (tile, part) by a shared-memory
// atomic, (2) the per-part counts are scanned into tile offsets and each part present reserves its run of the part's
// slice with ONE global atomic, (3) the records are laid out part by part in shared memory, (4) the tile is streamed out
// in order: neighbouring lanes write neighbouring records of the same run (~4 records = 2-3 sectors per run on a 4K
// noise frame instead of one sector per record, and a quarter of the atomics).  The records stay in registers between
// (1) and (3): nothing is read twice.  Jobs with more than kScat3MaxParts parts (a 16K frame) take k_scatter.
constexpr int kScat3Threads = 1024;
#ifndef RCT_SCAT3PER
#define RCT_SCAT3PER 6
#endif
constexpr int kScat3Per = RCT_SCAT3PER;
constexpr int kScat3Tile = kScat3Threads * kScat3Per;
constexpr int kScat3MaxParts = 4096;
constexpr int kScat3CP = kScat3MaxParts / kScat3Threads;  // counters per thread in the scan
constexpr size_t kScat3OffCnt = (kSmemScatter + 15) / 16 * 16;
constexpr size_t kScat3OffRec = kScat3OffCnt + (size_t)kScat3MaxParts * 8;
constexpr size_t kSmemScatter3 = kScat3OffRec + (size_t)kScat3Tile * 18;
__global__ void __launch_bounds__(kScat3Threads, 1) k_scatter_runs(const Job *jobs)
{
    const Job &job = jobs[blockIdx.y];
    RegionCtrl *rc = job.rc;
    if (rc->skip) return;
    const unsigned d = rc->d;
    if ((size_t)blockIdx.x * kScat3Tile >= d) return;
    extern __shared__ __align__(16) unsigned s_scat[];
    __shared__ unsigned s_warp[33];
    unsigned *s_base = s_scat;                                                         // [kCoarse + 1]
    unsigned short *s_pof = reinterpret_cast<unsigned short *>(s_scat + kCoarse + 2);  // [n_fine]
    unsigned *s_cnt = s_scat + kScat3OffCnt / 4;   // [n_parts]: entries of the tile in the part, then its tile offset
    unsigned *s_gb = s_cnt + kScat3MaxParts;       // [n_parts]: first global position of the run minus the tile offset
    Rec *srec = reinterpret_cast<Rec *>(reinterpret_cast<unsigned char *>(s_scat) + kScat3OffRec);  // [kScat3Tile]
    unsigned short *spart = reinterpret_cast<unsigned short *>(srec + kScat3Tile);                    // [kScat3Tile]
    const unsigned np = rc->n_parts, tid = threadIdx.x;
    {
        const unsigned nf2 = (rc->n_fine + 1u) / 2u;
        const unsigned *pof32 = reinterpret_cast<const unsigned *>(job.pof);
        unsigned *s_pof32 = s_scat + kCoarse + 2;
        for (unsigned i = tid; i <= (unsigned)kCoarse; i += kScat3Threads) s_base[i] = rc->fine_base[i];
        for (unsigned i = tid; i < nf2; i += kScat3Threads) s_pof32[i] = pof32[i];
    }
    const double klo = rc->klo, cscale = rc->cscale;
    Rec *recP = job.recP;
    for (size_t t0 = (size_t)blockIdx.x * kScat3Tile; t0 < d; t0 += (size_t)gridDim.x * kScat3Tile) {
        for (unsigned i = tid; i < np; i += kScat3Threads) s_cnt[i] = 0u;
        double key[kScat3Per];
        unsigned pay[kScat3Per], cnt[kScat3Per], pr[kScat3Per];  // pr: part | rank inside (tile, part) << 16
#pragma unroll
        for (int k = 0; k < kScat3Per; ++k) {
            const size_t e = t0 + (size_t)k * kScat3Threads + tid;
            key[k] = e < d ? job.keyA[e] : 0.0;
            pay[k] = (e < d && job.payA) ? job.payA[e] : (unsigned)e;
            cnt[k] = (e < d && job.cntA) ? job.cntA[e] : 1u;
        }
        __syncthreads();  // counters cleared (and, on the first trip, the tables staged)
#pragma unroll
        for (int k = 0; k < kScat3Per; ++k) {
            const size_t e = t0 + (size_t)k * kScat3Threads + tid;
            pr[k] = 0xffffffffu;
            if (e < d) {
                double ff;
                const unsigned p = s_pof[fine_bin(key[k], klo, cscale, s_base, ff)];
                pr[k] = p | (atomicAdd(&s_cnt[p], 1u) << 16);
            }
        }
        __syncthreads();
        unsigned c[kScat3CP], sum = 0;
#pragma unroll
        for (int k = 0; k < kScat3CP; ++k) {
            const unsigned i = tid * kScat3CP + k;
            c[k] = i < np ? s_cnt[i] : 0u;
            sum += c[k];
        }
        unsigned n_tile;
        unsigned off = block_excl_scan(sum, s_warp, n_tile);
#pragma unroll
        for (int k = 0; k < kScat3CP; ++k) {
            const unsigned i = tid * kScat3CP + k;
            if (i < np) {
                s_cnt[i] = off;
                if (c[k]) s_gb[i] = atomicAdd(&job.pfill[(size_t)i * kFillStride], c[k]) - off;
                off += c[k];
            }
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < kScat3Per; ++k)
            if (pr[k] != 0xffffffffu) {
                const unsigned p = pr[k] & 0xffffu, j = s_cnt[p] + (pr[k] >> 16);
                srec[j] = make_rec(key[k], pay[k], cnt[k]);
                spart[j] = (unsigned short)p;
            }
        __syncthreads();
        for (unsigned j = tid; j < n_tile; j += kScat3Threads) recP[s_gb[spart[j]] + j] = srec[j];
        __syncthreads();  // the tile is out before its counters and slots are reused
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

// Shared-memory counters are stored padded (one spare word per 32) so that a thread sweeping
// kSortItems consecutive counters never collides with its warp neighbours on a bank.
__device__ __forceinline__ unsigned pad_idx(unsigned i) { return i + (i >> 5); }

// In-place scan of the padded counters sc[0..n), n <= kSortItems * blockDim.x: thread t owns the
// kSortItems consecutive counters starting at kSortItems*t.  Returns the total; ends with a barrier.
constexpr int kSortThreads = 512;
constexpr int kSortItems = kPartCap / kSortThreads;
template <bool INCLUSIVE>
__device__ __forceinline__ unsigned cta_scan_items(unsigned *sc, unsigned n, unsigned *s_warp /*[33]*/)
{
    const unsigned i0 = threadIdx.x * kSortItems;
    unsigned v[kSortItems], sum = 0;
#pragma unroll
    for (int k = 0; k < kSortItems; ++k) {
        v[k] = (i0 + k < n) ? sc[pad_idx(i0 + k)] : 0u;
        sum += v[k];
    }
    unsigned total;
    unsigned off = block_excl_scan(sum, s_warp, total);
#pragma unroll
    for (int k = 0; k < kSortItems; ++k) {
        if (i0 + k < n) sc[pad_idx(i0 + k)] = INCLUSIVE ? off + v[k] : off;
        off += v[k];
    }
    __syncthreads();
    return total;
}

// k_part_sort: dynamic shared memory = kSmemPartSort bytes.  512 threads, up to kSortItems
// records per thread held in registers between the counting and the placing pass; one bucket
// per record on average.  `cap` (<= kPartCap) can be lowered by the caller to exercise the
// fallback.
__global__ void __launch_bounds__(kSortThreads, 2) k_part_sort(const Job *jobs, unsigned cap)
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
    unsigned *sc = reinterpret_cast<unsigned *>(srec + kPartCap);  // [pad_idx(kPartCap) + 1]
    for (unsigned i = tid; i <= pad_idx(n); i += kSortThreads) sc[i] = 0;
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
            double est = (double)(o0 - start) + ff * (double)(o1 - o0);  // estimated rank in the part
            unsigned bb = (unsigned)est;
            b[k] = bb < n ? bb : n - 1;
        }
    }
    __syncthreads();
    // pass 1: bucket sizes
#pragma unroll
    for (int k = 0; k < kSortItems; ++k)
        if (tid + k * kSortThreads < n) atomicAdd(&sc[pad_idx(b[k])], 1u);
    __syncthreads();
    cta_scan_items<false>(sc, n, s_warp);
    // pass 2: place (sc[b] becomes the end of bucket b)
#pragma unroll
    for (int k = 0; k < kSortItems; ++k)
        if (tid + k * kSortThreads < n) srec[atomicAdd(&sc[pad_idx(b[k])], 1u)] = r[k];
    __syncthreads();
    // exact order inside each bucket (one record per bucket on average): every record counts
    // the members of its bucket that precede it -- no data movement, little divergence
    unsigned fp[kSortItems];
#pragma unroll
    for (int k = 0; k < kSortItems; ++k) {
        fp[k] = 0;
        if (tid + k * kSortThreads < n) {
            unsigned lo = b[k] ? sc[pad_idx(b[k] - 1)] : 0u, hi = sc[pad_idx(b[k])];
            if (hi - lo > 32u) {
                s_flag = 1;  // clustered keys: leave it to the block-wide network below
            } else {
                unsigned rank = lo;
                for (unsigned m = lo; m < hi; ++m) rank += rec_less(srec[m], r[k]) ? 1u : 0u;
                fp[k] = rank;
            }
        }
    }
    __syncthreads();
    if (s_flag) {
        block_bitonic(srec, n);  // ends with a barrier
    } else {
#pragma unroll
        for (int k = 0; k < kSortItems; ++k)
            if (tid + k * kSortThreads < n) srec[fp[k]] = r[k];
        __syncthreads();
    }
    // part-local inclusive cumulative counts (the counters are free again)
    for (unsigned i = tid; i < n; i += kSortThreads) sc[pad_idx(i)] = rec_cnt(srec[i]);
    __syncthreads();
    unsigned total = cta_scan_items<true>(sc, n, s_warp);
    if (tid == 0) job.ppix[p] = total;
    for (unsigned i = tid; i < n; i += kSortThreads) {
        Rec x = srec[i];
        x.y = (x.y & 0xffffffffull) | ((unsigned long long)sc[pad_idx(i)] << 32);
        grec[i] = x;
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
// Ranks are carried in the integer type R: int when both regions hold fewer than 2^31 pixels (every real frame; the
// 64-bit compares, conversions and register pairs of long long ranks were a third of k_match's instructions), long long
// otherwise.  Same values, same floating-point operations in the same order, for either type.
template <class R>
struct SortedT {
    const Rec *rec;                  // sorted; hi word of .y = part-local inclusive cumulative count
    const unsigned *pstart, *pbase;  // [n_parts+1]
    unsigned n_parts, d;
    R n;
    double inv;                      // 1.0 / (n - 1): the step of np.linspace(0, 1, n)
};
typedef SortedT<long long> Sorted;

template <class R>
__device__ __forceinline__ SortedT<R> sorted_as(const Job &j)
{
    const R n = (R)j.rc->n_pix;
    return SortedT<R>{j.recP, j.pstart, j.ppix, j.rc->n_parts, j.rc->d, n, n > 1 ? 1.0 / (double)(n - 1) : 0.0};
}
__device__ __forceinline__ Sorted sorted_of(const Job &j) { return sorted_as<long long>(j); }

// np.linspace(0, 1, v.n)[k]  (same arithmetic as lin(), with the reciprocal hoisted)
template <class R>
__device__ __forceinline__ double lin_of(const SortedT<R> &v, R k)
{
    if (v.n == 1) return 0.0;
    if (k == v.n - 1) return 1.0;
    return (double)k * v.inv;
}

struct Hint {
    unsigned p, i;
    bool valid;
    bool fresh;  // ignore .i: start every search from the interpolated position (interleaved traversal)
};

template <class R>
__device__ __forceinline__ unsigned cum_local(const SortedT<R> &v, unsigned i)
{
    return (unsigned)(__ldg(reinterpret_cast<const unsigned long long *>(v.rec + i) + 1) >> 32);
}
template <class R>
__device__ __forceinline__ double key_at(const SortedT<R> &v, unsigned i)
{
    return __longlong_as_double((long long)__ldg(reinterpret_cast<const unsigned long long *>(v.rec + i)));
}

// index of the entry holding the q-th smallest value; h carries the last position found
template <class R>
__device__ unsigned os_index(const SortedT<R> &v, R q, Hint &h)
{
    unsigned p;
    bool near = false;
    if (h.valid && q >= (R)__ldg(v.pbase + h.p)) {
        // ranks move forward slowly: same part, or one of the next two
        p = h.p;
        int hop = 0;
        while (hop < 3 && q >= (R)__ldg(v.pbase + p + 1)) {
            ++p;
            ++hop;
        }
        near = q < (R)__ldg(v.pbase + p + 1);
    }
    if (!near) {
        unsigned lo = 0, hi = v.n_parts - 1;  // last part with pbase <= q
        while (lo < hi) {
            unsigned mid = (lo + hi + 1) >> 1;
            if ((R)__ldg(v.pbase + mid) <= q) lo = mid; else hi = mid - 1;
        }
        p = lo;
    }
    const unsigned b0 = __ldg(v.pbase + p);
    const unsigned ql = (unsigned)(q - (R)b0);
    const unsigned ps = __ldg(v.pstart + p), pe = __ldg(v.pstart + p + 1);
    // entries per rank inside this part (multiplicities are roughly even within a part)
    const float scale = __fdividef((float)(pe - ps), (float)(__ldg(v.pbase + p + 1) - b0));
    unsigned idx;
    if (!h.fresh && h.valid && h.p == p && h.i >= ps && h.i < pe) {
        idx = h.i;
    } else {
        idx = ps + (unsigned)((float)ql * scale);
        if (idx >= pe) idx = pe - 1;
    }
    h.p = p;
    h.valid = true;
#pragma unroll 1
    for (int s = 0; s < 8; ++s) {  // secant steps on the cumulative counts
        const unsigned c = cum_local(v, idx);
        if (c <= ql) {
            unsigned step = (unsigned)((float)(ql + 1u - c) * scale);
            idx += step ? step : 1u;
            if (idx >= pe) idx = pe - 1;  // ql < part total = cum_local(pe-1)
        } else if (idx > ps) {
            const unsigned cp = cum_local(v, idx - 1);
            if (cp <= ql) {
                h.i = idx;
                return idx;
            }
            unsigned step = (unsigned)((float)(cp - ql) * scale);
            if (!step) step = 1u;
            idx = (idx - ps > step) ? idx - step : ps;
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

template <class R>
__device__ __forceinline__ double os_value(const SortedT<R> &v, R q, Hint &h) { return key_at(v, os_index(v, q, h)); }

// values at ranks q and q+1 with one search (q+1 < v.n)
template <class R>
__device__ __forceinline__ void os_pair(const SortedT<R> &v, R q, Hint &h, double &v0, double &v1)
{
    unsigned idx = os_index(v, q, h);
    v0 = key_at(v, idx);
    R g_end = (R)__ldg(v.pbase + h.p) + (R)cum_local(v, idx);  // first rank after idx
    v1 = (q + 1 < g_end) ? v0 : key_at(v, idx + 1);
}

// np.interp's bracket for 0 <= x <= 1 over np.linspace(0, 1, v.n), v.n >= 2: the largest m with lin(m) <= x, starting
// from the estimate floor(x * (n - 1)).  lin() is non-decreasing in its index, so either the upward or the downward
// walk runs, never both; xm = lin(m), and x1 = lin(m + 1) when m < n - 1.
template <class R>
__device__ __forceinline__ R lin_bracket(const SortedT<R> &v, double x, double &xm, double &x1)
{
    const R n = v.n;
    R m = (R)(x * (double)(n - 1));
    if (m < 0) m = 0;
    if (m > n - 1) m = n - 1;
    xm = lin_of(v, m);
    x1 = 1.0;
    if (xm <= x) {
        while (m + 1 <= n - 1) {
            x1 = lin_of(v, (R)(m + 1));
            if (!(x1 <= x)) break;
            ++m;
            xm = x1;
        }
    } else {
        while (m > 0 && xm > x) {
            --m;
            x1 = xm;
            xm = lin_of(v, m);
        }
    }
    return m;
}

// np.interp(x, np.linspace(0,1,v.n), sorted(v)) for 0 <= x <= 1
template <class R>
__device__ double interp_lin_os(double x, const SortedT<R> &v, Hint &h)
{
    const R n = v.n;
    if (n == 1) return os_value(v, (R)0, h);
    double xm, x1;
    const R m = lin_bracket(v, x, xm, x1);
    if (m >= n - 1) return os_value(v, (R)(n - 1), h);
    double fm, fm1;
    os_pair(v, m, h, fm, fm1);
    if (xm == x) return fm;
    double slope = (fm1 - fm) / (x1 - xm);
    return slope * (x - xm) + fm;
}

// The neighbourhood of one target entry in rank space: ranks [g_lo, g_hi) hold `key`, rank g_hi
// holds k_next, rank g_lo-1 holds k_prev.  The stretched-target evaluations of a-6 only ever
// look at ranks next to the entry's own, so they resolve here without any search.
template <class R>
struct TWinT {
    R g_lo, g_hi;
    double key, k_prev, k_next;
};

// value at rank q by a full search, as a real function call: for call sites that (almost) never execute, so that the
// search is not inlined a dozen times into the kernel (k_match was 210 KB of SASS, i-cache misses showed in the stalls)
#ifndef RCT_OSCOLD
#define RCT_OSCOLD 0
#endif
#if RCT_OSCOLD
#define RCT_COLD_FN __noinline__
#else
#define RCT_COLD_FN __forceinline__
#endif
template <class R>
__device__ RCT_COLD_FN double os_value_cold(const Rec *rec, const unsigned *pstart, const unsigned *pbase, unsigned n_parts, R q)
{
    SortedT<R> v{rec, pstart, pbase, n_parts, 0u, (R)0, 0.0};
    Hint h{0, 0, false, true};
    return os_value(v, q, h);
}

template <class R>
__device__ __forceinline__ double t_value(const SortedT<R> &t, const TWinT<R> &w, R m, Hint &ht)
{
    if (m >= w.g_lo && m < w.g_hi) return w.key;
    if (m == w.g_hi) return w.k_next;
    if (m == w.g_lo - 1) return w.k_prev;
    return os_value_cold<R>(t.rec, t.pstart, t.pbase, t.n_parts, m);
}

// t'[.] at x: np.interp(x, np.linspace(0,1,T), sorted target), T >= 2
template <class R>
__device__ double interp_target(double x, const SortedT<R> &t, const TWinT<R> &w, Hint &ht)
{
    const R n = t.n;
    double xm, x1;
    const R m = lin_bracket(t, x, xm, x1);
    if (m >= n - 1) return t_value(t, w, (R)(n - 1), ht);
    double fm = t_value(t, w, m, ht);
    if (xm == x) return fm;
    double fm1 = t_value(t, w, (R)(m + 1), ht);
    double slope = (fm1 - fm) / (x1 - xm);
    return slope * (x - xm) + fm;
}

// entry i of the sorted target (inside part p); ht / hs: search hints into target / source
template <class R>
__device__ double match_one(const SortedT<R> &t, const SortedT<R> &s, unsigned p, unsigned i, Hint &ht, Hint &hs,
                            long long &j_out)
{
    const double key = key_at(t, i);
    const unsigned ps = __ldg(t.pstart + p), pe = __ldg(t.pstart + p + 1);
    unsigned i2 = i;
    while (i2 + 1 < pe && key_at(t, i2 + 1) == key) ++i2;  // ties share the upper rank
    const R base = (R)__ldg(t.pbase + p);
    const R g_hi = base + (R)cum_local(t, i2);
    const R r = g_hi - 1;
    const R T = t.n, S = s.n;
    double m;
    R j;
    if (T == S) {
        j = r;
        m = os_value(s, r, hs);
    } else if (T > S) {  // source stretched to T samples (LST:110-115)
        j = r;
        m = interp_lin_os(lin_of(t, r), s, hs);
    } else {  // target stretched to S samples (LST:116-121): t'[k] = interp(lin(S,k), lin(T,.), t)
        // j = last k with t'[k] <= key; t'[j] and t'[j+1] are kept for the interpolation
        double tj = 0.0, tj1 = 0.0;
        bool have_j1 = false;
        if (T == 1) {
            j = S - 1;
        } else {
            unsigned i1 = i;
            while (i1 > ps && key_at(t, i1 - 1) == key) --i1;
            TWinT<R> w;
            w.key = key;
            w.g_hi = g_hi;
            w.g_lo = base + (i1 > ps ? (R)cum_local(t, i1 - 1) : (R)0);
            w.k_prev = i1 > 0 ? key_at(t, i1 - 1) : key;
            w.k_next = i2 + 1 < t.d ? key_at(t, i2 + 1) : key;
            ht.p = p;
            ht.i = i2;
            ht.valid = true;
            double xr = lin_of(t, r);
            R k = (R)(xr * (double)(S - 1));
            if (k < 0) k = 0;
            if (k > S - 1) k = S - 1;
            double tk = interp_target(lin_of(s, k), t, w, ht);
            if (tk <= key) {
                while (k + 1 <= S - 1) {
                    double tn = interp_target(lin_of(s, (R)(k + 1)), t, w, ht);
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
                    tk = interp_target(lin_of(s, k), t, w, ht);
                }
            }
            j = k;
            tj = tk;
            if (j < S - 1 && tj != key && !have_j1) tj1 = interp_target(lin_of(s, (R)(j + 1)), t, w, ht);
        }
        if (j >= S - 1) {
            m = os_value(s, (R)(S - 1), hs);
        } else {
            double sj, sj1;
            os_pair(s, j, hs, sj, sj1);
            if (tj == key) {
                m = sj;
            } else {
                double slope = (sj1 - sj) / (tj1 - tj);
                m = slope * (key - tj) + sj;
            }
        }
    }
    j_out = (long long)j;
    return m;
}

// ------------------------------------------------------------------------------------------
// Palette LUT (RCT_OPT_LUT = 1).  Every recoloured target pixel lies on the one-dimensional curve
// mean + m * component (a-7), and m is non-decreasing along the sorted target entries, so the
// output colours along the sorted order form a few thousand runs of equal RGB (each channel is a
// sum of three exponentials in m: at most three monotone pieces of at most 256 uint8 steps).
// The colour -> RGB table therefore factors into a uint16 run index per colour (32 MiB instead of
// 64 MiB per frame: it stays L2-resident while the apply pass walks the frame) and a palette of
// the run colours (a few KB, L1-resident).
//   k_match      writes the output colour of sorted entry i to rgb_sorted[i]
//   k_pal_count  run starts per part          -> pfill[part] (the scatter counters are dead by now)
//   k_pal_scan   exclusive scan over the parts -> run index of each part's first entry; n_palette
//   k_pal_write  lut16[colour] = run index, palette[run] = colour
// rgb_sorted lives in the target job's unsorted-key array (dead after k_scatter).
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned *pal_rgb_sorted(const Job &jt)
{
    return reinterpret_cast<unsigned *>(const_cast<double *>(jt.keyA));
}
__device__ __forceinline__ bool pal_run_start(const unsigned *rgb, unsigned i)
{
    return i == 0u || __ldg(rgb + i) != __ldg(rgb + i - 1);
}

constexpr int kPalThreads = 256;
__global__ void __launch_bounds__(kPalThreads) k_pal_count(FrameCtrl *ctrl, const Job *jobs)
{
    const int slot = blockIdx.y;
    if (ctrl[slot].status != 0) return;
    const Job &jt = jobs[2 * slot];
    const unsigned p = blockIdx.x;
    if (p >= jt.rc->n_parts) return;
    const unsigned ps = jt.pstart[p], pe = jt.pstart[p + 1];
    const unsigned *rgb = pal_rgb_sorted(jt);
    unsigned c = 0;
    for (unsigned i = ps + threadIdx.x; i < pe; i += kPalThreads) c += pal_run_start(rgb, i) ? 1u : 0u;
    __shared__ unsigned s_w[kPalThreads / 32];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31u) == 0) s_w[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned t = 0;
        for (int w = 0; w < kPalThreads / 32; ++w) t += s_w[w];
        jt.pfill[(size_t)p * kFillStride] = t;
    }
}

// one CTA of 1024 threads per slot
__global__ void __launch_bounds__(1024) k_pal_scan(FrameCtrl *ctrl, const Job *jobs)
{
    const int slot = blockIdx.x;
    if (ctrl[slot].status != 0) return;
    const Job &jt = jobs[2 * slot];
    __shared__ unsigned s_warp[33];
    const unsigned np = jt.rc->n_parts;
    unsigned carry = 0;
    for (unsigned base = 0; base < np; base += 1024) {
        const unsigned i = base + threadIdx.x;
        unsigned v = i < np ? jt.pfill[(size_t)i * kFillStride] : 0u, total;
        const unsigned ex = block_excl_scan(v, s_warp, total);
        if (i < np) jt.pfill[(size_t)i * kFillStride] = carry + ex;
        carry += total;
    }
    if (threadIdx.x == 0) {
        ctrl[slot].n_palette = carry;
        if (carry > kPaletteCap) ctrl[slot].status = 3;  // cannot happen for a-7's curve; apply keeps the content, hosts raise
    }
}

__global__ void __launch_bounds__(kPalThreads) k_pal_write(FrameCtrl *ctrl, const Job *jobs, unsigned *lut0, size_t lut_stride)
{
    const int slot = blockIdx.y;
    if (ctrl[slot].status != 0) return;
    const Job &jt = jobs[2 * slot];
    const unsigned p = blockIdx.x;
    if (p >= jt.rc->n_parts) return;
    const unsigned ps = jt.pstart[p], pe = jt.pstart[p + 1];
    const unsigned *rgb = pal_rgb_sorted(jt);
    unsigned short *lut16 = reinterpret_cast<unsigned short *>(lut0 + (size_t)slot * lut_stride);
    unsigned *palette = lut0 + (size_t)slot * lut_stride + kPaletteOffsetU32;
    __shared__ unsigned s_warp[33];
    unsigned run = jt.pfill[(size_t)p * kFillStride];  // runs started before this part
    for (unsigned i0 = ps; i0 < pe; i0 += kPalThreads) {
        const unsigned i = i0 + threadIdx.x;
        const bool live = i < pe;
        const unsigned f = (live && pal_run_start(rgb, i)) ? 1u : 0u;
        unsigned total;
        const unsigned ex = block_excl_scan(f, s_warp, total);
        if (live) {
            const unsigned idx = run + ex + f - 1u;  // runs started up to and including i, minus one
            lut16[rec_pay(__ldg(jt.recP + i))] = (unsigned short)idx;
            if (f) palette[idx] = __ldg(rgb + i);
        }
        run += total;
    }
}

// pipeline flavour: jobs[2*slot] = target, jobs[2*slot+1] = source; one CTA per target part.
// Writes the colour -> output-RGB LUT (a-7: inverse_transform, lab -> rgb, clip, truncate).
#ifndef RCT_MATCHT
#define RCT_MATCHT 128
#endif
constexpr int kMatchThreads = RCT_MATCHT;
template <int LUTM, class R>
__device__ __forceinline__ void match_body_r(FrameCtrl *ctrl, const Job *jobs, unsigned *lut0, size_t lut_stride,
                                             double *dbg_matched0, long long *dbg_j0, size_t dbg_stride, int interleave,
                                             uint8_t *kt0, const MathTabs *Tp)
{
    const int slot = blockIdx.y;
    const Job &jt = jobs[2 * slot], &js = jobs[2 * slot + 1];
    const unsigned p = blockIdx.x;
    const SortedT<R> t = sorted_as<R>(jt), s = sorted_as<R>(js);
    const unsigned ps = jt.pstart[p], pe = jt.pstart[p + 1];
    const MathTabs &T = *Tp;
    const Pca pca = jt.rc->pca;
    unsigned *lut = lut0 + (size_t)slot * lut_stride;
    unsigned *rgb_sorted = pal_rgb_sorted(jt);
    Hint ht{0, 0, false, false}, hs{0, 0, false, false};
    // interleave: the CTA sweeps the part 128 consecutive entries at a time, so neighbouring lanes
    // read neighbouring target and source records (coalesced, L1-resident) and every source
    // search starts from the position interpolated inside the source part.  Otherwise every
    // thread takes one contiguous run of entries and carries its search position along.
    hs.fresh = interleave != 0;
    auto one_in = [&](unsigned part, unsigned i) -> unsigned {
        long long j;
        double m = match_one(t, s, part, i, ht, hs, j);
        double lab[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) lab[c] = m * pca.comp[c] + pca.mean[c];  // inverse_transform
        // palette LUT: the colour goes to its sorted position (k_pal_* turn the runs of equal colours
        // into indices); otherwise straight into the colour-indexed table
        unsigned *dst = LUTM ? rgb_sorted + i : lut + rec_pay(__ldg(t.rec + i));
        const unsigned rgb = lab2rgb(lab, &T, nullptr);
        *dst = rgb;
        if (dbg_matched0) {
            dbg_matched0[(size_t)slot * dbg_stride + i] = m;
            dbg_j0[(size_t)slot * dbg_stride + i] = j;
        }
        return rgb;
    };
    auto one = [&](unsigned i) -> unsigned { return one_in(p, i); };
    if (!interleave) {
        const unsigned per = (pe - ps + kMatchThreads - 1) / kMatchThreads;
        const unsigned first = ps + threadIdx.x * per;
        const unsigned last = (first + per < pe) ? first + per : pe;
        for (unsigned i = first; i < last; ++i) one(i);
        return;
    }
    if (LUTM || !kt0) {
        for (unsigned i = ps + threadIdx.x; i < pe; i += kMatchThreads) one(i);
        return;
    }
    // Run starts for the key table (rct_keytab.cuh): wherever the output colour differs from the previous sorted
    // entry's, (key, sorted index, colour) is appended to the slot's run list.  Each warp takes groups of 31
    // consecutive entries and lane 0 re-evaluates the entry before the group -- for the part's first group that is the
    // last entry of the previous non-empty part -- so the comparison with the predecessor is a warp shuffle: no shared
    // memory, no CTA barrier (the kernel is latency-bound; a barrier per sweep cost 5-10 % of it), no dependency
    // between CTAs, 3 % redundant evaluations.  Only the region's very first entry opens a run unconditionally.
    uint8_t *blk = kt0 + (size_t)slot * kKtBytes;
    KtParams *par = reinterpret_cast<KtParams *>(blk);
    double *run_key = reinterpret_cast<double *>(blk + kKtOffRunKey);
    uint2 *run_ic = reinterpret_cast<uint2 *>(blk + kKtOffRunIc);
    if (threadIdx.x == 0) {  // extremes of the target keys: the key table's cell range
        atomicMax(&par->kmin_enc_inv, ~ord_enc(key_at(t, ps)));
        atomicMax(&par->kmax_enc, ord_enc(key_at(t, pe - 1)));
    }
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    constexpr unsigned kWarps = kMatchThreads / 32;
    for (unsigned g0 = ps + 31u * warp; g0 < pe; g0 += 31u * kWarps) {  // warp-uniform
        const unsigned i = g0 + lane - 1u;                               // lane 0: the entry before the group
        const bool live = (lane > 0u || g0 > 0u) && i < pe;
        unsigned rgb = 0xffffffffu;
        if (live) {  // lane 0 stores the same values its owner stored: benign
            unsigned part = p;
            if (i < ps) {  // the predecessor of the part's first entry lives in the previous non-empty part
                part = p - 1u;
                while (part > 0u && __ldg(t.pstart + part) > i) --part;
            }
            rgb = one_in(part, i);
        }
        const unsigned prev = __shfl_up_sync(0xffffffffu, rgb, 1);
        if (lane > 0u && live && rgb != prev) {  // entry 0 of the region sees the sentinel: the first run
            const unsigned idx = atomicAdd(&ctrl[slot].n_runs, 1u);
            if (idx < (unsigned)kKtRunCap) {
                run_key[idx] = key_at(t, i);
                run_ic[idx] = make_uint2(i, rgb);  // the sorted index orders the runs without a sort
            }
        }
    }
}

template <int LUTM>
__device__ __forceinline__ void match_body(FrameCtrl *ctrl, const Job *jobs, unsigned *lut0, size_t lut_stride,
                                           double *dbg_matched0, long long *dbg_j0, size_t dbg_stride, int interleave,
                                           uint8_t *kt0)
{
    const int slot = blockIdx.y;
    if (ctrl[slot].status != 0) return;
    const Job &jt = jobs[2 * slot], &js = jobs[2 * slot + 1];
    const unsigned p = blockIdx.x;
    if (p >= jt.rc->n_parts) return;
    if (jt.pstart[p] == jt.pstart[p + 1]) return;
    __shared__ MathTabs T;
    stage_tabs(&T);
    // block-uniform: ranks fit 32-bit integers unless a region holds 2^31 pixels or more
    if ((jt.rc->n_pix | js.rc->n_pix) < (1ull << 31))
        match_body_r<LUTM, int>(ctrl, jobs, lut0, lut_stride, dbg_matched0, dbg_j0, dbg_stride, interleave, kt0, &T);
    else
        match_body_r<LUTM, long long>(ctrl, jobs, lut0, lut_stride, dbg_matched0, dbg_j0, dbg_stride, interleave, kt0, &T);
}

#ifndef RCT_MATCHB
#define RCT_MATCHB 12
#endif
#if RCT_MATCHB > 0
#define RCT_MATCH_LB __launch_bounds__(kMatchThreads, RCT_MATCHB)
#else
#define RCT_MATCH_LB __launch_bounds__(kMatchThreads)
#endif
__global__ void RCT_MATCH_LB k_match(FrameCtrl *ctrl, const Job *jobs, unsigned *lut0, size_t lut_stride,
                                                         double *dbg_matched0, long long *dbg_j0, size_t dbg_stride,
                                                         int interleave, uint8_t *kt0)
{
    match_body<0>(ctrl, jobs, lut0, lut_stride, dbg_matched0, dbg_j0, dbg_stride, interleave, kt0);
}
// palette flavour (RCT_OPT_LUT = 1): output colours go to their sorted position, see k_pal_*
__global__ void __launch_bounds__(kMatchThreads, 7) k_match_pal(FrameCtrl *ctrl, const Job *jobs, unsigned *lut0,
                                                                size_t lut_stride, double *dbg_matched0, long long *dbg_j0,
                                                                size_t dbg_stride, int interleave)
{
    match_body<1>(ctrl, jobs, lut0, lut_stride, dbg_matched0, dbg_j0, dbg_stride, interleave, nullptr);
}

// generic flavour (rct_match_cdf): jobs[0] = target, jobs[1] = source; payload = input index
__global__ void __launch_bounds__(kMatchThreads) k_match_generic(const Job *jobs, double *matched, long long *jidx)
{
    const Job &jt = jobs[0], &js = jobs[1];
    const unsigned p = blockIdx.x;
    if (p >= jt.rc->n_parts) return;
    const Sorted t = sorted_of(jt), s = sorted_of(js);
    const unsigned ps = jt.pstart[p], pe = jt.pstart[p + 1];
    Hint ht{0, 0, false, false}, hs{0, 0, false, false};
    const unsigned per = (pe - ps + kMatchThreads - 1) / kMatchThreads;
    const unsigned i0 = ps + threadIdx.x * per, i1 = (i0 + per < pe) ? i0 + per : pe;
    for (unsigned i = i0; i < i1; ++i) {
        long long j;
        double m = match_one(t, s, p, i, ht, hs, j);
        unsigned pay = rec_pay(__ldg(t.rec + i));
        matched[pay] = m;
        if (jidx) jidx[pay] = j;
    }
}
