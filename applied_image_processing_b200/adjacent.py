"""The per-pixel steps either side of the hot path, on the GPU.

f-1  ``adain_mask_blend``: the mask-blend tail of ``adain_inference``
     (reference Style_3DGS/AdaIN/test.py:222-236).
f-2  ``deeplab_background_mask`` / ``deeplab_label_map``: the post-processing of
     ``extract_foreground_deeplab`` (reference Style_3DGS/localized_style_transfer.py:178-188).

The networks themselves (AdaIN VGG/decoder, DeepLabV3) stay stock PyTorch and are out of scope;
these functions take their output tensors.  CUDA tensors in, CUDA tensors out; no CPU fallback."""
from __future__ import annotations

import numpy as np
import torch

from . import _cabi
from ._cabi import check


def _cuda_f32(x, name):
    if isinstance(x, np.ndarray):
        x = torch.from_numpy(np.ascontiguousarray(x))
    if not isinstance(x, torch.Tensor):
        raise TypeError(f"{name} must be a tensor or ndarray")
    if not torch.cuda.is_available():
        raise RuntimeError("applied_image_processing_b200 needs a CUDA device; there is no CPU fallback")
    if x.device.type != "cuda":
        x = x.cuda()
    return x.float().contiguous()


def adain_mask_blend(content, output_img, content_mask, return_uint8=False):
    """AdaIN/test.py:222-236: ``mask = F.interpolate(mask, content.shape[-2:], mode="nearest")``,
    ``output = F.interpolate(output, content.shape[-2:], mode="bilinear", align_corners=False)``,
    ``content * (1 - mask) + output * mask``.  ``content`` [N,3,H,W] float32, ``output_img``
    [N,3,h,w], ``content_mask`` [mh,mw] or [1,mh,mw] (what extract_foreground_deeplab returns).
    ``return_uint8`` adds the uint8 [N,H,W,3] image torchvision's save_image would write."""
    L = _cabi.lib()
    content = _cuda_f32(content, "content")
    output_img = _cuda_f32(output_img, "output_img").to(content.device)
    mask = _cuda_f32(content_mask, "content_mask").to(content.device)
    if content.dim() == 3:
        content = content.unsqueeze(0)
    if output_img.dim() == 3:
        output_img = output_img.unsqueeze(0)
    if mask.dim() == 3 and mask.shape[0] == 1:
        mask = mask[0]
    if mask.dim() != 2:
        raise ValueError("content_mask must be [mh,mw] or [1,mh,mw]")
    n, ch, H, W = content.shape
    if ch != 3 or output_img.shape[:2] != (n, 3):
        raise ValueError("content / output_img must be [N,3,H,W] with equal N")
    oh, ow = output_img.shape[-2:]
    mh, mw = mask.shape
    out = torch.empty_like(content)
    rgb8 = torch.empty((n, H, W, 3), dtype=torch.uint8, device=content.device) if return_uint8 else None
    with torch.cuda.device(content.device):
        check(L.rct_adain_mask_blend(content.data_ptr(), output_img.data_ptr(), mask.data_ptr(), out.data_ptr(),
                                     rgb8.data_ptr() if return_uint8 else None, n, H, W, oh, ow, mh, mw, 0,
                                     torch.cuda.current_stream().cuda_stream))
    return (out, rgb8) if return_uint8 else out


def deeplab_background_mask(logits, threshold=0.5, as_numpy=True):
    """LST:178-188 after the network: ``probs = softmax(output, dim=0)``; ``probs[0] > threshold``
    -> uint8 [1,H,W] (numpy, like the reference, or a CUDA tensor with ``as_numpy=False``)."""
    L = _cabi.lib()
    logits = _cuda_f32(logits, "logits")
    if logits.dim() != 3:
        raise ValueError("logits must be [C,H,W]")
    C_, H, W = logits.shape
    mask = torch.empty((1, H, W), dtype=torch.uint8, device=logits.device)
    with torch.cuda.device(logits.device):
        check(L.rct_softmax_threshold_mask(logits.data_ptr(), C_, H * W, 0, float(threshold), mask.data_ptr(),
                                           torch.cuda.current_stream().cuda_stream))
    return mask.cpu().numpy() if as_numpy else mask


def deeplab_label_map(logits, as_numpy=False):
    """Per-pixel arg-max over the class planes -> uint8 [H,W] label map (the multi-label masks of
    BASELINE.json config 5; no counterpart in the reference, which only thresholds class 0)."""
    L = _cabi.lib()
    logits = _cuda_f32(logits, "logits")
    if logits.dim() != 3 or logits.shape[0] > 256:
        raise ValueError("logits must be [C,H,W] with C <= 256")
    C_, H, W = logits.shape
    labels = torch.empty((H, W), dtype=torch.uint8, device=logits.device)
    with torch.cuda.device(logits.device):
        check(L.rct_argmax_labels(logits.data_ptr(), C_, H * W, labels.data_ptr(),
                                  torch.cuda.current_stream().cuda_stream))
    return labels.cpu().numpy() if as_numpy else labels
