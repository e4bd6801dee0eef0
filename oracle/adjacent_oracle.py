"""TEST INFRASTRUCTURE -- numpy restatement of the two steps either side of the hot path
(SURVEY.md 8f).  Only tests/ may import this; the product never does.

f-1: Style_3DGS/AdaIN/test.py:222-236 (torch F.interpolate nearest + bilinear align_corners=False,
     blend), following torch's index rules in float32 (ATen UpSample.h: nearest_idx,
     area_pixel_compute_source_index).  Pinned in tests/test_adjacent.py against torch's own CPU
     F.interpolate, which is the reference's dependency and is present on both machines.
f-2: Style_3DGS/localized_style_transfer.py:178-188 (softmax over classes, class 0 > threshold)."""
import numpy as np

f32 = np.float32


def nearest_idx(out_size, in_size):
    scale = f32(in_size) / f32(out_size)
    idx = np.floor(np.arange(out_size, dtype=f32) * scale).astype(np.int64)
    return np.minimum(idx, in_size - 1)


def linear_taps(out_size, in_size):
    scale = f32(in_size) / f32(out_size)
    src = scale * (np.arange(out_size, dtype=f32) + f32(0.5)) - f32(0.5)
    src = np.maximum(src, f32(0)).astype(f32)
    i0 = np.minimum(src.astype(np.int64), in_size - 1)
    i1 = i0 + (i0 < in_size - 1)
    l1 = (src - i0.astype(f32)).astype(f32)
    l0 = (f32(1) - l1).astype(f32)
    return i0, i1, l0, l1


def adain_mask_blend(content, output, mask):
    """content [N,3,H,W], output [N,3,oh,ow], mask [mh,mw]; float32 -> (stylized f32, uint8 NHWC)."""
    content = np.asarray(content, f32)
    output = np.asarray(output, f32)
    mask = np.asarray(mask, f32)
    N, _, H, W = content.shape
    oh, ow = output.shape[-2:]
    m = mask[nearest_idx(H, mask.shape[0])][:, nearest_idx(W, mask.shape[1])]
    y0, y1, ly0, ly1 = linear_taps(H, oh)
    x0, x1, lx0, lx1 = linear_taps(W, ow)
    top = lx0 * output[:, :, y0][:, :, :, x0] + lx1 * output[:, :, y0][:, :, :, x1]
    bot = lx0 * output[:, :, y1][:, :, :, x0] + lx1 * output[:, :, y1][:, :, :, x1]
    up = (ly0[:, None] * top + ly1[:, None] * bot).astype(f32)
    sty = (content * (f32(1) - m) + up * m).astype(f32)
    q = np.clip(sty * f32(255) + f32(0.5), 0, 255).astype(np.uint8)  # torchvision save_image
    return sty, np.ascontiguousarray(q.transpose(0, 2, 3, 1))


def softmax_mask(logits, threshold=0.5, cls=0):
    """logits [C,H,W] float32 -> (uint8 [1,H,W], float64 class probability)."""
    x = np.asarray(logits, np.float64)
    p = 1.0 / np.exp(x - x[cls]).sum(0)
    return (p > np.float64(f32(threshold))).astype(np.uint8)[None], p


def argmax_labels(logits):
    return np.asarray(logits).argmax(0).astype(np.uint8)


# ---- f-4: video/utils.py:89-105 warp_image (cv2.remap INTER_LINEAR, BORDER_REFLECT) and
# video/utils.py:217-229 normalize_image / blend_images ---------------------------------------------
# cv2.remap on uint8 with float32 maps is FIXED POINT (OpenCV imgwarp.cpp, remap / remapBilinear):
#   sx = cvRound(map_x * 32), sy = cvRound(map_y * 32)      (float32 product, round half to even)
#   ix = saturate_cast<short>(sx >> 5), fx = sx & 31         (same for y)
#   weights = (32-fx)(32-fy), fx(32-fy), (32-fx)fy, fx*fy, each * 32  (BilinearTab_i, sum 2^15)
#   taps outside the image: borderInterpolate(BORDER_REFLECT)  fedcba|abcdefgh|hgfedcb
#   out = (sum w*v + 2^14) >> 15
# Pinned in tests/test_adjacent.py against outputs of cv2 itself (tests/golden/video_f4.npz, and
# the live cv2 when importable).
def reflect_index(p, n):
    p = np.asarray(p, np.int64).copy()
    if n == 1:
        return np.zeros_like(p)
    for _ in range(64):  # borderInterpolate loops until inside
        lo, hi = p < 0, p >= n
        if not (lo.any() or hi.any()):
            break
        p = np.where(lo, -p - 1, p)
        p = np.where(hi, n - 1 - (p - n), p)
    return p


def flow_maps(flow):
    """video/utils.py:92-100: meshgrid (int64) + float32 flow promotes to float64, then astype(float32)."""
    flow = np.asarray(flow, f32)
    h, w = flow.shape[1:]
    x, y = np.meshgrid(np.arange(w), np.arange(h))
    return (x + flow[0]).astype(f32), (y + flow[1]).astype(f32)


def remap_linear_reflect(image, map_x, map_y):
    img = np.asarray(image, np.uint8)
    H, W = img.shape[:2]
    sx = np.rint((np.asarray(map_x, f32) * f32(32)).astype(np.float64))
    sy = np.rint((np.asarray(map_y, f32) * f32(32)).astype(np.float64))
    big = 2.0 ** 31
    sx = np.where(np.isfinite(sx) & (np.abs(sx) < big), sx, -big).astype(np.int64)  # cvRound overflow: INT_MIN
    sy = np.where(np.isfinite(sy) & (np.abs(sy) < big), sy, -big).astype(np.int64)
    ix = np.clip(sx >> 5, -32768, 32767)
    iy = np.clip(sy >> 5, -32768, 32767)
    fx, fy = sx & 31, sy & 31
    x0, x1 = reflect_index(ix, W), reflect_index(ix + 1, W)
    y0, y1 = reflect_index(iy, H), reflect_index(iy + 1, H)
    w00, w01 = (32 - fx) * (32 - fy) * 32, fx * (32 - fy) * 32
    w10, w11 = (32 - fx) * fy * 32, fx * fy * 32
    im = img.astype(np.int64)
    if im.ndim == 2:
        im = im[..., None]
    acc = (w00[..., None] * im[y0, x0] + w01[..., None] * im[y0, x1] + w10[..., None] * im[y1, x0]
           + w11[..., None] * im[y1, x1] + (1 << 14)) >> 15
    return acc.astype(np.uint8).reshape(map_x.shape + img.shape[2:])


def warp_image(image, flow):
    mx, my = flow_maps(flow)
    return remap_linear_reflect(image, mx, my)


def blend_images(stylized, warped, alpha):
    """video/utils.py:223-229 on uint8 inputs: float32 throughout (a Python-float alpha is weak
    under NEP 50), clip(blended*255, 0, 255), astype(uint8) = truncation."""
    s = np.asarray(stylized, np.uint8).astype(f32) / f32(255.0)
    w = np.asarray(warped, np.uint8).astype(f32) / f32(255.0)
    blended = f32(alpha) * s + f32(1 - alpha) * w
    return np.clip(blended * f32(255), 0, 255).astype(np.uint8)
