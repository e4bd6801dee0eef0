// rct_video.cuh -- SURVEY.md 8f row f-4: the temporal-consistency step of the video driver
// (video/utils.py:89-105 warp_image, :217-229 normalize_image / blend_images; called per frame at
// video/utils.py:355-359), the caller shape of the 1024-frame batch.
//
//   k_warp_blend<BLEND>   out = blend_images(stylized, warp_image(prev, flow), alpha) in ONE pass:
//                         flow (2 x float32 planes) + stylised in, uint8 out, the previous stylised
//                         frame gathered through the read-only path (4 taps, 2-D local).
//                         BLEND = false writes the warped frame itself (warp_image alone).
//   k_blend               blend_images alone: flat over bytes, 16 per thread.
//
// warp_image is cv2.remap(INTER_LINEAR, BORDER_REFLECT) on uint8, which OpenCV evaluates in FIXED
// POINT (imgwarp.cpp remap / remapBilinear): coordinates rounded to 1/32 px (cvRound = round half
// to even of the float32 product), weights (32-fx)(32-fy)*32 ... with sum 2^15, result
// (sum + 2^14) >> 15 -- integer arithmetic, reproduced bit for bit.  blend_images is float32:
// alpha * (s/255) + (1-alpha) * (w/255), times 255, clip, truncate; the two products take 256
// values each and come from shared-memory tables (no division in the loop).
#pragma once
// (included from inside namespace rct)

__device__ __forceinline__ int cv_reflect(int p, int n)
{
    if (n == 1) return 0;
    while ((unsigned)p >= (unsigned)n) p = p < 0 ? -p - 1 : n - 1 - (p - n);  // fedcba|abcdefgh|hgfedcb
    return p;
}

// cvRound(v * 32): SSE cvtss2si semantics (round half to even; out of range / NaN -> INT_MIN)
__device__ __forceinline__ int cv_fix5(float v)
{
    const float t = __fmul_rn(v, 32.f);
    if (!(fabsf(t) < 2147483648.f)) return (int)0x80000000;
    return __float2int_rn(t);
}

__device__ __forceinline__ int clamp_short(int v) { return v < -32768 ? -32768 : (v > 32767 ? 32767 : v); }

struct WarpArgs {
    const uint8_t *stylized;  // [n,H,W,C] (BLEND only)
    const uint8_t *prev;      // [n,H,W,C] image to warp
    const float *flow;        // [n,2,H,W]
    uint8_t *out;             // [n,H,W,C]
    int n, H, W, C;
    float a, b;               // float32(alpha), float32(1 - alpha)
};

__device__ __forceinline__ void blend_tables(float a, float b, float *s_a, float *s_b)
{
    for (int i = threadIdx.x; i < 256; i += blockDim.x) {
        const float u = __fdiv_rn((float)i, 255.0f);  // normalize_image: uint8 -> float32 / 255.0
        s_a[i] = __fmul_rn(a, u);
        s_b[i] = __fmul_rn(b, u);
    }
    __syncthreads();
}
__device__ __forceinline__ uint32_t blend_u8(uint32_t s, uint32_t w, const float *s_a, const float *s_b)
{
    float r = __fmul_rn(__fadd_rn(s_a[s], s_b[w]), 255.f);
    r = r < 0.f ? 0.f : (r > 255.f ? 255.f : r);
    return (uint32_t)(int)r;  // astype(uint8) of a value in [0,255]: truncation
}

// one bilinear sample of cv2.remap at map (mx, my): 3 interleaved channels packed r | g<<8 | b<<16
__device__ __forceinline__ uint32_t remap3(const uint8_t *img, int H, int W, float mx, float my)
{
    const int sx = cv_fix5(mx), sy = cv_fix5(my);
    const int ix = clamp_short(sx >> 5), iy = clamp_short(sy >> 5), fx = sx & 31, fy = sy & 31;
    int x0 = ix, x1 = ix + 1, y0 = iy, y1 = iy + 1;
    if ((unsigned)ix >= (unsigned)(W - 1) || (unsigned)iy >= (unsigned)(H - 1)) {  // a tap leaves the image
        x0 = cv_reflect(ix, W); x1 = cv_reflect(ix + 1, W);
        y0 = cv_reflect(iy, H); y1 = cv_reflect(iy + 1, H);
    }
    const int w00 = (32 - fx) * (32 - fy) * 32, w01 = fx * (32 - fy) * 32, w10 = (32 - fx) * fy * 32, w11 = fx * fy * 32;
    const uint8_t *p00 = img + ((size_t)y0 * W + x0) * 3, *p01 = img + ((size_t)y0 * W + x1) * 3;
    const uint8_t *p10 = img + ((size_t)y1 * W + x0) * 3, *p11 = img + ((size_t)y1 * W + x1) * 3;
    uint32_t out = 0;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const int v = (w00 * (int)__ldg(p00 + c) + w01 * (int)__ldg(p01 + c) + w10 * (int)__ldg(p10 + c) +
                       w11 * (int)__ldg(p11 + c) + (1 << 14)) >> 15;
        out |= (uint32_t)v << (8 * c);
    }
    return out;
}

// RGB fast path: W % 4 == 0, flow 16-byte aligned, images 4-byte aligned.  Four pixels per thread:
// two float4 flow loads, three 32-bit stylised loads, three 32-bit stores; the taps go through
// the read-only path (neighbouring pixels share cache lines).
template <bool BLEND>
__global__ void __launch_bounds__(256) k_warp_blend_rgb4(WarpArgs g)
{
    __shared__ float s_a[256], s_b[256];
    if (BLEND) blend_tables(g.a, g.b, s_a, s_b);
    const size_t plane = (size_t)g.H * g.W, quads = plane / 4, total = quads * (size_t)g.n;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int f = (int)(i / quads);
        const size_t q = i - (size_t)f * quads, p = 4 * q;
        const int y = (int)(p / g.W), x = (int)(p - (size_t)y * g.W);
        const float *fl = g.flow + (size_t)f * 2 * plane;
        const float4 fx4 = __ldcs(reinterpret_cast<const float4 *>(fl + p));
        const float4 fy4 = __ldcs(reinterpret_cast<const float4 *>(fl + plane + p));
        const float fxs[4] = {fx4.x, fx4.y, fx4.z, fx4.w}, fys[4] = {fy4.x, fy4.y, fy4.z, fy4.w};
        const uint8_t *img = g.prev + (size_t)f * plane * 3;
        uint32_t c[4];
#pragma unroll
        for (int k = 0; k < 4; ++k)
            c[k] = remap3(img, g.H, g.W, (float)((double)(x + k) + (double)fxs[k]), (float)((double)y + (double)fys[k]));
        if (BLEND) {
            const uint32_t *sw = reinterpret_cast<const uint32_t *>(g.stylized + ((size_t)f * plane + p) * 3);
            const uint32_t w0 = __ldcs(sw), w1 = __ldcs(sw + 1), w2 = __ldcs(sw + 2);
            const uint32_t s[4] = {w0 & 0xffffffu, __funnelshift_r(w0, w1, 24) & 0xffffffu,
                                   __funnelshift_r(w1, w2, 16) & 0xffffffu, w2 >> 8};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                uint32_t o = 0;
#pragma unroll
                for (int ch = 0; ch < 3; ++ch)
                    o |= blend_u8((s[k] >> (8 * ch)) & 255u, (c[k] >> (8 * ch)) & 255u, s_a, s_b) << (8 * ch);
                c[k] = o;
            }
        }
        uint32_t *ow = reinterpret_cast<uint32_t *>(g.out + ((size_t)f * plane + p) * 3);
        __stcs(ow, c[0] | (c[1] << 24));
        __stcs(ow + 1, (c[1] >> 8) | (c[2] << 16));
        __stcs(ow + 2, (c[2] >> 16) | (c[3] << 8));
    }
}

template <bool BLEND>
__global__ void __launch_bounds__(256) k_warp_blend(WarpArgs g)
{
    __shared__ float s_a[256], s_b[256];
    if (BLEND) blend_tables(g.a, g.b, s_a, s_b);
    const size_t plane = (size_t)g.H * g.W, total = plane * (size_t)g.n;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int f = (int)(i / plane);
        const size_t p = i - (size_t)f * plane;
        const int y = (int)(p / g.W), x = (int)(p - (size_t)y * g.W);
        const float *fl = g.flow + (size_t)f * 2 * plane;
        // (x + flow).astype(float32): int64 + float32 promotes to float64 first (video/utils.py:98-99)
        const float mx = (float)((double)x + (double)__ldcs(fl + p));
        const float my = (float)((double)y + (double)__ldcs(fl + plane + p));
        const int sx = cv_fix5(mx), sy = cv_fix5(my);
        const int ix = clamp_short(sx >> 5), iy = clamp_short(sy >> 5), fx = sx & 31, fy = sy & 31;
        const int x0 = cv_reflect(ix, g.W), x1 = cv_reflect(ix + 1, g.W);
        const int y0 = cv_reflect(iy, g.H), y1 = cv_reflect(iy + 1, g.H);
        const int w00 = (32 - fx) * (32 - fy) * 32, w01 = fx * (32 - fy) * 32, w10 = (32 - fx) * fy * 32,
                  w11 = fx * fy * 32;
        const uint8_t *img = g.prev + (size_t)f * plane * g.C;
        const uint8_t *r0 = img + (size_t)y0 * g.W * g.C, *r1 = img + (size_t)y1 * g.W * g.C;
        const uint8_t *sty = BLEND ? g.stylized + i * g.C : nullptr;
        uint8_t *o = g.out + i * g.C;
        for (int c = 0; c < g.C; ++c) {
            const int v = (w00 * (int)__ldg(r0 + x0 * g.C + c) + w01 * (int)__ldg(r0 + x1 * g.C + c) +
                           w10 * (int)__ldg(r1 + x0 * g.C + c) + w11 * (int)__ldg(r1 + x1 * g.C + c) + (1 << 14)) >> 15;
            o[c] = BLEND ? (uint8_t)blend_u8(__ldcs(sty + c), (uint32_t)v, s_a, s_b) : (uint8_t)v;
        }
    }
}

// nbytes % 16 == 0 part through uint4; tail by bytes
__global__ void __launch_bounds__(256) k_blend(const uint8_t *stylized, const uint8_t *warped, uint8_t *out, size_t nbytes,
                                               float a, float b, int vec)
{
    __shared__ float s_a[256], s_b[256];
    blend_tables(a, b, s_a, s_b);
    const size_t nvec = vec ? nbytes / 16 : 0;
    const uint4 *s4 = reinterpret_cast<const uint4 *>(stylized), *w4 = reinterpret_cast<const uint4 *>(warped);
    uint4 *o4 = reinterpret_cast<uint4 *>(out);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (size_t)gridDim.x * blockDim.x) {
        const uint4 s = __ldcs(s4 + i), w = __ldcs(w4 + i);
        const uint32_t sw[4] = {s.x, s.y, s.z, s.w}, ww[4] = {w.x, w.y, w.z, w.w};
        uint32_t r[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            uint32_t acc = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k)
                acc |= blend_u8((sw[q] >> (8 * k)) & 255u, (ww[q] >> (8 * k)) & 255u, s_a, s_b) << (8 * k);
            r[q] = acc;
        }
        __stcs(o4 + i, make_uint4(r[0], r[1], r[2], r[3]));
    }
    for (size_t i = nvec * 16 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nbytes;
         i += (size_t)gridDim.x * blockDim.x)
        out[i] = (uint8_t)blend_u8(stylized[i], warped[i], s_a, s_b);
}
