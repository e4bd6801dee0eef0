// rct_adjacent.cuh -- the two per-pixel steps either side of the hot path:
//
//   k_mask_blend      f-1: the mask-blend tail of AdaIN inference (Style_3DGS/AdaIN/test.py:222-236):
//                     mask   -> F.interpolate(mode="nearest") to the content size,
//                     output -> F.interpolate(mode="bilinear", align_corners=False),
//                     stylized = content * (1 - mask) + output * mask,
//                     fused into one pass over planar float32 NCHW tensors; optionally also emits
//                     what torchvision.utils.save_image would quantise (uint8 HWC) so the colour
//                     transfer can consume it without the lossy JPEG round trip.
//   k_softmax_mask    f-2: DeepLab post-processing (LST:178-188): softmax over the C class planes,
//                     class-0 probability > threshold -> uint8 {0,1}
//   k_argmax_labels   f-2 generalised: per-pixel arg-max label map (first index on ties)
#pragma once
// (included from inside namespace rct)

// torch's nearest source index (UpSampleKernel: nearest_idx): min(floorf(dst * scale), in - 1),
// scale = (float)in / out, all in float32
__device__ __forceinline__ int torch_nearest_idx(int dst, float scale, int in)
{
    int s = (int)floorf(__fmul_rn((float)dst, scale));
    return s < in - 1 ? s : in - 1;
}

// torch's bilinear source coordinate, align_corners=False (area_pixel_compute_source_index)
__device__ __forceinline__ void torch_linear_taps(int dst, float scale, int in, int &i0, int &i1, float &l0, float &l1)
{
    float src = __fsub_rn(__fmul_rn(scale, __fadd_rn((float)dst, 0.5f)), 0.5f);
    if (src < 0.f) src = 0.f;
    i0 = (int)src;
    if (i0 > in - 1) i0 = in - 1;
    i1 = i0 + (i0 < in - 1 ? 1 : 0);
    l1 = __fsub_rn(src, (float)i0);
    l0 = __fsub_rn(1.f, l1);
}

struct BlendArgs {
    const float *content;  // [N,3,H,W]
    const float *output;   // [N,3,oh,ow]
    const float *mask;     // [mh,mw] (shared by the batch, as in the reference) or [N,mh,mw]
    float *stylized;       // [N,3,H,W]
    uint8_t *rgb8;         // optional [N,H,W,3]
    int N, H, W, oh, ow, mh, mw;
    size_t mask_stride;    // elements between the masks of consecutive images (0 = shared)
};

__global__ void __launch_bounds__(256) k_mask_blend(BlendArgs a)
{
    const float sy_o = (float)a.oh / (float)a.H, sx_o = (float)a.ow / (float)a.W;
    const float sy_m = (float)a.mh / (float)a.H, sx_m = (float)a.mw / (float)a.W;
    const size_t plane = (size_t)a.H * a.W, oplane = (size_t)a.oh * a.ow;
    const size_t total = plane * (size_t)a.N;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int n = (int)(i / plane);
        const size_t p = i - (size_t)n * plane;
        const int y = (int)(p / a.W), x = (int)(p - (size_t)y * a.W);
        const float m = a.mask[(size_t)n * a.mask_stride + (size_t)torch_nearest_idx(y, sy_m, a.mh) * a.mw +
                               torch_nearest_idx(x, sx_m, a.mw)];
        int y0, y1, x0, x1;
        float ly0, ly1, lx0, lx1;
        torch_linear_taps(y, sy_o, a.oh, y0, y1, ly0, ly1);
        torch_linear_taps(x, sx_o, a.ow, x0, x1, lx0, lx1);
        const float om = __fsub_rn(1.f, m);
        uint32_t packed = 0;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const float *o = a.output + ((size_t)n * 3 + c) * oplane;
            const float v00 = o[(size_t)y0 * a.ow + x0], v01 = o[(size_t)y0 * a.ow + x1];
            const float v10 = o[(size_t)y1 * a.ow + x0], v11 = o[(size_t)y1 * a.ow + x1];
            const float top = __fadd_rn(__fmul_rn(lx0, v00), __fmul_rn(lx1, v01));
            const float bot = __fadd_rn(__fmul_rn(lx0, v10), __fmul_rn(lx1, v11));
            const float up = __fadd_rn(__fmul_rn(ly0, top), __fmul_rn(ly1, bot));
            const float cv = a.content[((size_t)n * 3 + c) * plane + p];
            const float r = __fadd_rn(__fmul_rn(cv, om), __fmul_rn(up, m));
            a.stylized[((size_t)n * 3 + c) * plane + p] = r;
            // torchvision.utils.save_image: mul(255).add_(0.5).clamp_(0, 255).to(uint8)
            float q = __fadd_rn(__fmul_rn(r, 255.f), 0.5f);
            q = q < 0.f ? 0.f : (q > 255.f ? 255.f : q);
            packed |= (uint32_t)(uint8_t)q << (8 * c);
        }
        if (a.rgb8) {
            uint8_t *d = a.rgb8 + 3 * i;
            d[0] = (uint8_t)packed;
            d[1] = (uint8_t)(packed >> 8);
            d[2] = (uint8_t)(packed >> 16);
        }
    }
}

// W % 4 == 0 and 16-byte aligned planes: four pixels of one row per thread -- the row's taps are
// computed once, content / stylized move as float4, rgb8 as three 32-bit words.  Same arithmetic,
// same order of operations as k_mask_blend.
__global__ void __launch_bounds__(256) k_mask_blend4(BlendArgs a)
{
    const float sy_o = (float)a.oh / (float)a.H, sx_o = (float)a.ow / (float)a.W;
    const float sy_m = (float)a.mh / (float)a.H, sx_m = (float)a.mw / (float)a.W;
    const size_t plane = (size_t)a.H * a.W, oplane = (size_t)a.oh * a.ow, quads = plane / 4;
    const size_t total = quads * (size_t)a.N;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int n = (int)(i / quads);
        const size_t p = 4 * (i - (size_t)n * quads);
        const int y = (int)(p / a.W), x = (int)(p - (size_t)y * a.W);
        int y0, y1;
        float ly0, ly1;
        torch_linear_taps(y, sy_o, a.oh, y0, y1, ly0, ly1);
        const float *mrow = a.mask + (size_t)n * a.mask_stride + (size_t)torch_nearest_idx(y, sy_m, a.mh) * a.mw;
        int x0[4], x1[4];
        float lx0[4], lx1[4], m[4], om[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            torch_linear_taps(x + k, sx_o, a.ow, x0[k], x1[k], lx0[k], lx1[k]);
            m[k] = __ldg(mrow + torch_nearest_idx(x + k, sx_m, a.mw));
            om[k] = __fsub_rn(1.f, m[k]);
        }
        uint32_t packed[4] = {0, 0, 0, 0};
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const float *o = a.output + ((size_t)n * 3 + c) * oplane;
            const float *r0 = o + (size_t)y0 * a.ow, *r1 = o + (size_t)y1 * a.ow;
            const float4 cv4 = __ldcs(reinterpret_cast<const float4 *>(a.content + ((size_t)n * 3 + c) * plane + p));
            const float cv[4] = {cv4.x, cv4.y, cv4.z, cv4.w};
            float r[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float top = __fadd_rn(__fmul_rn(lx0[k], __ldg(r0 + x0[k])), __fmul_rn(lx1[k], __ldg(r0 + x1[k])));
                const float bot = __fadd_rn(__fmul_rn(lx0[k], __ldg(r1 + x0[k])), __fmul_rn(lx1[k], __ldg(r1 + x1[k])));
                const float up = __fadd_rn(__fmul_rn(ly0, top), __fmul_rn(ly1, bot));
                r[k] = __fadd_rn(__fmul_rn(cv[k], om[k]), __fmul_rn(up, m[k]));
                float q = __fadd_rn(__fmul_rn(r[k], 255.f), 0.5f);
                q = q < 0.f ? 0.f : (q > 255.f ? 255.f : q);
                packed[k] |= (uint32_t)(uint8_t)q << (8 * c);
            }
            __stcs(reinterpret_cast<float4 *>(a.stylized + ((size_t)n * 3 + c) * plane + p), make_float4(r[0], r[1], r[2], r[3]));
        }
        if (a.rgb8) {
            uint32_t *d = reinterpret_cast<uint32_t *>(a.rgb8 + 3 * ((size_t)n * plane + p));
            d[0] = packed[0] | (packed[1] << 24);
            d[1] = (packed[1] >> 8) | (packed[2] << 16);
            d[2] = (packed[2] >> 16) | (packed[3] << 8);
        }
    }
}

// logits [C, npix] float32 (one image; the reference handles one image per call, LST:176).
// The decision is DEFINED as the float64 evaluation p_cls = 1 / sum_i exp(x_i - x_cls) > threshold
// (it differs from a float32 softmax only inside a ~1e-7 band around the threshold).  Evaluating
// 21 double-precision exponentials per pixel is 6x slower than streaming the 84 bytes, so the sum
// is first taken in float32 (relative error < 1e-5) and only pixels whose float32 sum lands within
// 1e-3 (relative) of 1/threshold are re-evaluated in float64: same decisions, HBM-bound.
__device__ __forceinline__ uint8_t softmax_decide64(const float *logits, int C, size_t npix, size_t p, float xc,
                                                    float threshold)
{
    double den = 0.0;
    for (int c = 0; c < C; ++c) den += exp((double)logits[(size_t)c * npix + p] - (double)xc);
    return (1.0 / den > (double)threshold) ? 1 : 0;
}

__global__ void __launch_bounds__(256) k_softmax_mask(const float *logits, int C, size_t npix, int cls, float threshold,
                                                      uint8_t *mask)
{
    const bool fast_ok = threshold > 1e-6f && threshold < 1.0f;
    const float inv_thr = fast_ok ? 1.0f / threshold : 0.f;
    const size_t nquad = ((((uintptr_t)logits | (uintptr_t)mask) & 15) == 0 && (npix & 3) == 0) ? npix / 4 : 0;
    const size_t stride = (size_t)gridDim.x * blockDim.x, t0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    // four pixels per thread: one 16-byte load per class plane, one 4-byte store
    for (size_t q = t0; q < nquad; q += stride) {
        const float4 xc = __ldcs(reinterpret_cast<const float4 *>(logits + (size_t)cls * npix) + q);
        float d0 = 0.f, d1 = 0.f, d2 = 0.f, d3 = 0.f;
        for (int c = 0; c < C; ++c) {
            const float4 x = __ldcs(reinterpret_cast<const float4 *>(logits + (size_t)c * npix) + q);
            d0 += expf(x.x - xc.x);
            d1 += expf(x.y - xc.y);
            d2 += expf(x.z - xc.z);
            d3 += expf(x.w - xc.w);
        }
        const float den[4] = {d0, d1, d2, d3}, xcs[4] = {xc.x, xc.y, xc.z, xc.w};
        uint32_t packed = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            uint8_t m;
            if (fast_ok && fabsf(den[k] - inv_thr) > 1e-3f * inv_thr) m = den[k] < inv_thr ? 1 : 0;  // also den = inf
            else m = softmax_decide64(logits, C, npix, 4 * q + k, xcs[k], threshold);
            packed |= (uint32_t)m << (8 * k);
        }
        reinterpret_cast<uint32_t *>(mask)[q] = packed;
    }
    for (size_t p = nquad * 4 + t0; p < npix; p += stride)
        mask[p] = softmax_decide64(logits, C, npix, p, logits[(size_t)cls * npix + p], threshold);
}

__global__ void __launch_bounds__(256) k_argmax_labels(const float *logits, int C, size_t npix, uint8_t *labels)
{
    for (size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x; p < npix; p += (size_t)gridDim.x * blockDim.x) {
        float best = logits[p];
        int arg = 0;
        for (int c = 1; c < C; ++c) {
            const float v = logits[(size_t)c * npix + p];
            if (v > best) {  // first index on ties (torch.argmax)
                best = v;
                arg = c;
            }
        }
        labels[p] = (uint8_t)arg;
    }
}
