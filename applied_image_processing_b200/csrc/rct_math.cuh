// rct_math.cuh -- "canonical arithmetic"  on the device.
//
// Restates, in a fixed operation order, the float64 arithmetic of the reference's
// rgb_to_lab_pixels (LST:67-77), lab_to_rgb_pixels (LST:80-89), sklearn's PCA projection
// (LST:92-96) and its covariance_eigh solver.  This translation unit is compiled with
// -fmad=false so that only the fma() calls written here fuse; the C oracle is an
// independent implementation of the same written-down sequence.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "rct_tables.h"

namespace rct {

// Reinhard matrices (LST:12-19): RGB_TO_LMS, LMS_TO_LAB, LAB_TO_LMS, LMS_TO_RGB, row-major.
__constant__ double c_mat[36];
// RGB_TO_LMS / 255 in fp32, for the key table's fp32 key (rct_keytab.cuh); uploaded with c_mat
__constant__ float c_ktm[9];

// Tables live in global memory (filled once per device) and are staged into shared memory
// by kernels that evaluate log10 / 10^x per lane with divergent indices.
struct MathTabs {
    double log_inv[128];
    double log_t10[128];
    double exp2_t[128];
    double unit[256];
};
__device__ MathTabs g_tabs;

__device__ __forceinline__ void stage_tabs(MathTabs *s)
{
    const double *src = reinterpret_cast<const double *>(&g_tabs);
    double *dst = reinterpret_cast<double *>(s);
    for (int i = threadIdx.x; i < (int)(sizeof(MathTabs) / sizeof(double)); i += blockDim.x) dst[i] = src[i];
    __syncthreads();
}

// C3: log10(x), x >= 1e-6.
__device__ __forceinline__ double clog10(double x, const MathTabs *T)
{
    unsigned long long u = (unsigned long long)__double_as_longlong(x);
    int e = (int)(u >> 52) - 1023;
    unsigned long long mant = u & 0x000fffffffffffffULL;
    int i = (int)(mant >> 45);
    double m = __longlong_as_double((long long)(mant | 0x3ff0000000000000ULL));
    double r = fma(m, T->log_inv[i], -1.0);
    double q = RCT_LOG1P_C7;
    q = fma(q, r, RCT_LOG1P_C6);
    q = fma(q, r, RCT_LOG1P_C5);
    q = fma(q, r, RCT_LOG1P_C4);
    q = fma(q, r, RCT_LOG1P_C3);
    q = fma(q, r, RCT_LOG1P_C2);
    double p = fma(r * r, q, r);
    double hi = fma((double)e, RCT_LOG10_2, T->log_t10[i]);
    return fma(p, RCT_LOG10_E, hi);
}

// C4: 10^y.
__device__ __forceinline__ double cpow10(double y, const MathTabs *T)
{
    double t = y * RCT_LOG2_10_HI;
    if (!(t < 1020.0)) return t != t ? t : __longlong_as_double(0x7ff0000000000000LL);
    if (t < -1020.0) return 0.0;
    double tl = fma(y, RCT_LOG2_10_HI, -t) + y * RCT_LOG2_10_LO;
    double kd = rint(t * 128.0);
    long long k = (long long)kd;
    double rr = fma(kd, -0.0078125, t) + tl;
    int j = (int)(k & 127);
    long long qe = k >> 7;
    double w = rr * RCT_LN2;
    double s = RCT_EXPM1_C6;
    s = fma(s, w, RCT_EXPM1_C5);
    s = fma(s, w, RCT_EXPM1_C4);
    s = fma(s, w, RCT_EXPM1_C3);
    s = fma(s, w, RCT_EXPM1_C2);
    double pw = fma(w * w, s, w);
    double base = T->exp2_t[j];
    double val = fma(base, pw, base);
    double sc = __longlong_as_double((long long)((unsigned long long)(qe + 1023) << 52));
    return val * sc;
}

__device__ __forceinline__ double dot3(const double *row, double a, double b, double c)
{
    return fma(c, row[2], fma(b, row[1], a * row[0]));
}

// C2: rgb_to_lab_pixels for one colour (r | g<<8 | b<<16).
__device__ __forceinline__ void rgb2lab(uint32_t col, const MathTabs *T, double lab[3])
{
    double r = T->unit[col & 255u], g = T->unit[(col >> 8) & 255u], b = T->unit[(col >> 16) & 255u];
    double y[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        double x = dot3(c_mat + 3 * i, r, g, b);
        if (x < 1e-6) x = 1e-6;
        y[i] = clog10(x, T);
    }
#pragma unroll
    for (int i = 0; i < 3; ++i) lab[i] = dot3(c_mat + 9 + 3 * i, y[0], y[1], y[2]);
}

// C9: lab_to_rgb_pixels for one lab triple; returns packed r | g<<8 | b<<16.
__device__ __forceinline__ uint32_t lab2rgb(const double lab[3], const MathTabs *T, double *rgbf)
{
    double w[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) w[i] = cpow10(dot3(c_mat + 18 + 3 * i, lab[0], lab[1], lab[2]), T);
    uint32_t out = 0;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        double c = dot3(c_mat + 27 + 3 * i, w[0], w[1], w[2]);
        if (c < 0.0) c = 0.0;
        if (c > 1.0) c = 1.0;
        if (rgbf) rgbf[i] = c;
        out |= ((uint32_t)(c * 255.0) & 255u) << (8 * i);
    }
    return out;
}

// C5: exact integer moments.
typedef __int128 i128;

struct Moments {  // per-thread accumulators
    i128 s1[3];
    i128 s2[6];
    long long n;
};

__device__ __forceinline__ void moments_zero(Moments &m)
{
#pragma unroll
    for (int i = 0; i < 3; ++i) m.s1[i] = 0;
#pragma unroll
    for (int i = 0; i < 6; ++i) m.s2[i] = 0;
    m.n = 0;
}

__device__ __forceinline__ void moments_add(Moments &m, const double lab[3], long long cnt)
{
    long long q[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) q[i] = __double2ll_rn(lab[i] * 0x1p40);
    m.n += cnt;
#pragma unroll
    for (int i = 0; i < 3; ++i) m.s1[i] += (i128)q[i] * cnt;
    int k = 0;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = i; j < 3; ++j) m.s2[k++] += (i128)q[i] * q[j] * cnt;
}

// Fast exact accumulation for small multiplicities (cnt < 128), all in UNSIGNED 32x32->64 multiply-adds.
// The fixed-point value is taken with a bias, q' = q + 2^44 >= 0 (|lab| < 11 < 2^4, so |q| < 2^44): the fused
// fma(lab, 2^40, 2^52 + 2^44) rounds lab * 2^40 to the nearest integer, ties to even -- the same integer as
// __double2ll_rn(lab * 2^40), the scaling being exact and the magic constant even -- and leaves q' in the low 52 bits
// of the result.  q' = qh * 2^24 + ql with qh < 2^22, ql < 2^24, so with u = cnt*qh < 2^29, v = cnt*ql < 2^31
//   cnt*q'_i*q'_j = A*2^48 + B*2^24 + C,  A = u_i*qh_j < 2^51,  B = u_i*ql_j + v_i*qh_j < 2^54,  C = v_i*ql_j < 2^55
// and none of the 64-bit partial sums can overflow within kFastTerms = 128 additions.  The bias comes out again,
// exactly, when the partial sums are turned into 128-bit integers (fast_finish): with w = sum of cnt,
//   s1_i = s1'_i - 2^44 w,   s2_ij = s2'_ij - 2^44 (s1_i + s1_j) - 2^88 w.
// Same integers as moments_add.  A value outside the biased range (cannot happen for a-4's clamped logarithms) makes
// fast_add return false and the entry takes moments_add.  (Signed operands cost three times the instructions: the
// compiler has no signed 32x32+64 multiply-add.)
constexpr int kFastTerms = 128;
struct MomentsFast {
    unsigned long long U[3], V[3];
    unsigned long long A[6], B[6], C[6];
    unsigned w;
};

__device__ __forceinline__ void fast_zero(MomentsFast &f)
{
#pragma unroll
    for (int i = 0; i < 3; ++i) f.U[i] = f.V[i] = 0ull;
#pragma unroll
    for (int i = 0; i < 6; ++i) f.A[i] = f.B[i] = f.C[i] = 0ull;
    f.w = 0u;
}

__device__ __forceinline__ bool fast_add(MomentsFast &f, const double lab[3], unsigned cnt /* < 128 */)
{
    unsigned qh[3], ql[3];
    bool ok = true;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const double x = fma(lab[i], 0x1p40, 0x1p52 + 0x1p44);
        const unsigned hi = (unsigned)__double2hiint(x), lo = (unsigned)__double2loint(x);
        ok = ok && (hi >> 14) == (0x43300000u >> 14);  // exponent of 2^52 and q' < 2^46
        qh[i] = __funnelshift_r(lo, hi, 24) & 0x3fffffu;
        ql[i] = lo & 0xffffffu;
    }
    if (!ok) return false;
    unsigned u[3], v[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        u[i] = cnt * qh[i];
        v[i] = cnt * ql[i];
        f.U[i] += u[i];
        f.V[i] += v[i];
    }
    int k = 0;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = i; j < 3; ++j) {
            f.A[k] += (unsigned long long)u[i] * qh[j];
            f.B[k] += (unsigned long long)u[i] * ql[j] + (unsigned long long)v[i] * qh[j];
            f.C[k] += (unsigned long long)v[i] * ql[j];
            ++k;
        }
    f.w += cnt;
    return true;
}

// exact (unbiased) values of the nine moments held in f: s1[0..2], s2[00,01,02,11,12,22]
__device__ __forceinline__ void fast_finish(const MomentsFast &f, i128 s1[3], i128 s2[6])
{
    const i128 w = (i128)f.w;
#pragma unroll
    for (int i = 0; i < 3; ++i) s1[i] = ((i128)f.U[i] << 24) + (i128)f.V[i] - (w << 44);
    int k = 0;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = i; j < 3; ++j) {
            s2[k] = ((i128)f.A[k] << 48) + ((i128)f.B[k] << 24) + (i128)f.C[k] - ((s1[i] + s1[j]) << 44) - (w << 88);
            ++k;
        }
}

__device__ __forceinline__ double i128_to_double(i128 v)
{
    long long hi = (long long)(v >> 64);
    unsigned long long lo = (unsigned long long)v;
    return (double)hi * 0x1p64 + (double)lo;
}

// 128-bit value <-> three limbs (48,48,rest) so that many partial values can be summed
// with plain 64-bit integer atomics without carries.
__device__ __forceinline__ void split_limbs(i128 v, long long l[3])
{
    l[0] = (long long)((unsigned long long)v & 0xffffffffffffULL);
    l[1] = (long long)((unsigned long long)(v >> 48) & 0xffffffffffffULL);
    l[2] = (long long)(v >> 96);
}
__device__ __forceinline__ i128 join_limbs(const long long l[3])
{
    return (i128)l[0] + ((i128)l[1] << 48) + ((i128)l[2] << 96);
}

// C6: cyclic Jacobi on a symmetric 3x3, fixed rotation order; plain unfused arithmetic.
__device__ inline void jacobi3(double a[3][3], double v[3][3])
{
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) v[i][j] = (i == j) ? 1.0 : 0.0;
    const int PQ[3][2] = {{0, 1}, {0, 2}, {1, 2}};
    for (int sweep = 0; sweep < 16; ++sweep) {
        if (a[0][1] == 0.0 && a[0][2] == 0.0 && a[1][2] == 0.0) break;
        for (int r = 0; r < 3; ++r) {
            int p = PQ[r][0], q = PQ[r][1];
            double apq = a[p][q];
            if (apq == 0.0) continue;
            double theta = (a[q][q] - a[p][p]) / (2.0 * apq);
            double t = 1.0 / (fabs(theta) + sqrt(theta * theta + 1.0));
            if (theta < 0.0) t = -t;
            double c = 1.0 / sqrt(t * t + 1.0);
            double s = t * c;
            double app = a[p][p], aqq = a[q][q];
            a[p][p] = app - t * apq;
            a[q][q] = aqq + t * apq;
            a[p][q] = a[q][p] = 0.0;
            int k = 3 - p - q;
            double akp = a[k][p], akq = a[k][q];
            a[k][p] = a[p][k] = c * akp - s * akq;
            a[k][q] = a[q][k] = s * akp + c * akq;
            for (int i = 0; i < 3; ++i) {
                double vip = v[i][p], viq = v[i][q];
                v[i][p] = c * vip - s * viq;
                v[i][q] = s * vip + c * viq;
            }
        }
    }
}

struct Pca {
    double mean[3];
    double comp[3];
    double pm;
};

// sklearn covariance_eigh + svd_flip(u_based_decision=False), from exact moments.
__device__ inline void pca_from_moments(long long n_, const i128 s1[3], const i128 s2[6], Pca *out)
{
    double n = (double)n_;
    double g[6];
    for (int i = 0; i < 3; ++i) out->mean[i] = (i128_to_double(s1[i]) / n) * 0x1p-40;
    for (int k = 0; k < 6; ++k) g[k] = i128_to_double(s2[k]) * 0x1p-80;
    double a[3][3], v[3][3];
    int k = 0;
    for (int i = 0; i < 3; ++i)
        for (int j = i; j < 3; ++j) {
            double c = g[k++] - (n * out->mean[i]) * out->mean[j];
            c = (n_ > 1) ? c / (n - 1.0) : 0.0;
            a[i][j] = a[j][i] = c;
        }
    jacobi3(a, v);
    int best = 0;
    for (int i = 1; i < 3; ++i)
        if (a[i][i] > a[best][best]) best = i;
    double c3[3] = {v[0][best], v[1][best], v[2][best]};
    int am = 0;
    for (int i = 1; i < 3; ++i)
        if (fabs(c3[i]) > fabs(c3[am])) am = i;
    if (c3[am] < 0.0)
        for (int i = 0; i < 3; ++i) c3[i] = -c3[i];
    for (int i = 0; i < 3; ++i) out->comp[i] = c3[i];
    out->pm = fma(out->mean[2], c3[2], fma(out->mean[1], c3[1], out->mean[0] * c3[0]));
}

// C7: projection key.
__device__ __forceinline__ double project(const double lab[3], const double comp[3], double pm)
{
    return fma(lab[2], comp[2], fma(lab[1], comp[1], lab[0] * comp[0])) - pm;
}

// C8 helper: np.linspace(0, 1, n)[k]
__device__ __forceinline__ double lin(long long n, long long k)
{
    if (n == 1) return 0.0;
    if (k == n - 1) return 1.0;
    return (double)k * (1.0 / (double)(n - 1));
}

// order-preserving map double -> uint64 (for atomicMax-based min/max)
__device__ __forceinline__ unsigned long long ord_enc(double d)
{
    unsigned long long u = (unsigned long long)__double_as_longlong(d);
    return (u & 0x8000000000000000ULL) ? ~u : (u | 0x8000000000000000ULL);
}
__device__ __forceinline__ double ord_dec(unsigned long long e)
{
    unsigned long long u = (e & 0x8000000000000000ULL) ? (e & 0x7fffffffffffffffULL) : ~e;
    return __longlong_as_double((long long)u);
}

}  // namespace rct
