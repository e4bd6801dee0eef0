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
