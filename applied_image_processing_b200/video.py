"""the temporal-consistency step of the reference's video driver,
on the GPU, under the reference's own function names (video/utils.py).

``warp_image(image, flow)``            video/utils.py:89-105  (cv2.remap INTER_LINEAR / BORDER_REFLECT)
``normalize_image(image)``             video/utils.py:217-221
``blend_images(stylized, warped, a)``  video/utils.py:223-229
``warp_blend(stylized, prev, flow, a)`` the two fused, as the frame loop uses them
                                       (video/utils.py:355-359): one pass, no warped temporary.

numpy in -> numpy out (the reference's calling convention); CUDA tensors in -> CUDA tensor out.
Images are uint8 HWC (or NHWC batches with flow [N,2,H,W]); flow is float32 [2,H,W], a torch
tensor as ``estimate_optical_flow`` returns it or an ndarray.  Optical flow estimation itself
(cv2.calcOpticalFlowFarneback) stays OpenCV and is out of scope.  No CPU fallback."""
from __future__ import annotations

import numpy as np
import torch

from . import _cabi
from ._cabi import check


def _dev() -> torch.device:
    if not torch.cuda.is_available():
        raise RuntimeError("applied_image_processing_b200 needs a CUDA device; there is no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def _u8(x, name, dev=None):
    """-> (contiguous uint8 CUDA tensor, came_from_numpy)"""
    if isinstance(x, torch.Tensor):
        if x.dtype != torch.uint8:
            raise TypeError(f"{name} must be uint8, got {x.dtype}")
        was_np = x.device.type != "cuda"
        return x.to(dev or _dev()).contiguous(), was_np
    a = np.asarray(x)
    if a.dtype != np.uint8:
        raise TypeError(f"{name} must be uint8, got {a.dtype}")
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev or _dev()), True


def _flow(flow, dev):
    if isinstance(flow, np.ndarray):
        flow = torch.from_numpy(np.ascontiguousarray(flow))
    return flow.to(dev).float().contiguous()


def _geometry(img, flow):
    if img.dim() == 2:
        img = img.unsqueeze(-1)
    batched = img.dim() == 4
    n = img.shape[0] if batched else 1
    H, W, ch = img.shape[-3:]
    want = (n, 2, H, W) if batched else (2, H, W)
    if tuple(flow.shape) != want:
        raise ValueError(f"flow must be {want}, got {tuple(flow.shape)}")
    return n, H, W, ch


def normalize_image(image):
    """video/utils.py:217-221 (host helper: uint8 -> float32 / 255.0, anything else unchanged)."""
    if image.dtype == np.uint8:
        return image.astype(np.float32) / 255.0
    return image


def warp_image(image, flow):
    """video/utils.py:89-105: warp ``image`` (uint8 HWC) by ``flow`` ([2,H,W] float32)."""
    img, was_np = _u8(image, "image")
    fl = _flow(flow, img.device)
    n, H, W, ch = _geometry(img, fl)
    out = torch.empty_like(img)
    with torch.cuda.device(img.device):
        check(_cabi.lib().rct_warp_image(img.data_ptr(), fl.data_ptr(), out.data_ptr(), n, H, W, ch,
                                         torch.cuda.current_stream().cuda_stream))
    return out.cpu().numpy() if was_np else out


def blend_images(stylized, warped, alpha):
    """video/utils.py:223-229 for uint8 images: clip((alpha*s/255 + (1-alpha)*w/255) * 255) -> uint8."""
    s, was_np = _u8(stylized, "stylized")
    w, _ = _u8(warped, "warped", s.device)
    if s.shape != w.shape:
        raise ValueError("stylized and warped must have the same shape")
    out = torch.empty_like(s)
    with torch.cuda.device(s.device):
        check(_cabi.lib().rct_blend_images(s.data_ptr(), w.data_ptr(), float(alpha), out.data_ptr(), s.numel(),
                                           torch.cuda.current_stream().cuda_stream))
    return out.cpu().numpy() if was_np else out


def warp_blend(stylized, prev_stylized, flow, alpha):
    """``blend_images(stylized, warp_image(prev_stylized, flow), alpha)`` (video/utils.py:355-359)
    as one kernel."""
    s, was_np = _u8(stylized, "stylized")
    p, _ = _u8(prev_stylized, "prev_stylized", s.device)
    if s.shape != p.shape:
        raise ValueError("stylized and prev_stylized must have the same shape")
    fl = _flow(flow, s.device)
    n, H, W, ch = _geometry(p, fl)
    out = torch.empty_like(s)
    with torch.cuda.device(s.device):
        check(_cabi.lib().rct_warp_blend(s.data_ptr(), p.data_ptr(), fl.data_ptr(), float(alpha), out.data_ptr(),
                                         n, H, W, ch, torch.cuda.current_stream().cuda_stream))
    return out.cpu().numpy() if was_np else out


def blend_sequence(stylized_frames, flows, alpha=0.7):
    """The frame loop of ``apply_style_transfer*`` (video/utils.py:304-371) without its file I/O and
    stylisation: ``out[0] = stylized[0]``; ``out[i] = blend_images(stylized[i], warp_image(out[i-1],
    flows[i-1]), alpha)`` -- the recurrence is sequential (each frame warps the previous *blended*
    frame), so it is one fused ``warp_blend`` launch per frame on the caller's stream, frames
    staying on the device.  ``stylized_frames`` uint8 [N,H,W,C], ``flows`` float32 [N-1,2,H,W]
    (``flows[i-1]`` = estimate_optical_flow(frame[i-1], frame[i]))."""
    s, was_np = _u8(stylized_frames, "stylized_frames")
    if s.dim() != 4:
        raise ValueError("stylized_frames must be [N,H,W,C]")
    fl = _flow(flows, s.device)
    n, H, W, ch = s.shape
    if tuple(fl.shape) != (max(n - 1, 0), 2, H, W):
        raise ValueError(f"flows must be {(n - 1, 2, H, W)}, got {tuple(fl.shape)}")
    out = torch.empty_like(s)
    if n == 0:
        return out.cpu().numpy() if was_np else out
    out[0].copy_(s[0])
    L = _cabi.lib()
    with torch.cuda.device(s.device):
        st = torch.cuda.current_stream().cuda_stream
        for i in range(1, n):
            check(L.rct_warp_blend(s[i].data_ptr(), out[i - 1].data_ptr(), fl[i - 1].data_ptr(), float(alpha),
                                   out[i].data_ptr(), 1, H, W, ch, st))
    return out.cpu().numpy() if was_np else out
