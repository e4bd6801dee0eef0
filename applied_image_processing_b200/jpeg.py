"""the write at the end of the reference's pipeline,
``Image.fromarray(combined_image).save(os.path.join(out_dir, "localized_style_transfer_result.jpg"))``
(Style_3DGS/localized_style_transfer.py:241-243), done on the GPU.

``encode_jpeg(image, quality=75)``    uint8 HWC (or NHWC batch) -> bytes (list of bytes for a batch), byte for
                                      byte what Pillow's ``save(..., "JPEG", quality=quality)`` writes
``save_jpeg(image, path, quality=75)`` the reference's call, file on disk

A frame that was transferred on the GPU leaves it as a JFIF stream (about a tenth of H*W*3 bytes at quality 75)
instead of raw pixels, which is what the PCIe-bound end-to-end path wants.  numpy or CUDA tensors in.
No CPU fallback: without the CUDA library this raises."""
from __future__ import annotations

import numpy as np
import torch

from . import _cabi
from ._cabi import check
from .video import _dev, _u8

# first-try output capacity per frame: 1 byte per pixel + header covers every photographic frame at quality <= 95;
# a frame that needs more is re-encoded once with the exact worst-case bound.
_FIRST_TRY_BYTES_PER_PIXEL = 1.0


def _encode(img: torch.Tensor, quality: int, cap: int):
    n, H, W = img.shape[0], img.shape[1], img.shape[2]
    lib = _cabi.lib()
    ws_bytes = lib.rct_jpeg_workspace_bytes(H, W, n)
    if ws_bytes == 0:
        raise ValueError(f"JPEG frames are 1..65535 pixels a side, got {H}x{W}")
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=img.device)
    out = torch.empty((n, cap), dtype=torch.uint8, device=img.device)
    nbytes = torch.empty(n, dtype=torch.int64, device=img.device)
    with torch.cuda.device(img.device):
        check(lib.rct_jpeg_encode_async(img.data_ptr(), 0, n, H, W, int(quality), ws.data_ptr(), out.data_ptr(), cap, cap,
                                        nbytes.data_ptr(), torch.cuda.current_stream().cuda_stream))
    return out, nbytes.cpu().tolist()


def encode_jpeg(image, quality: int = 75):
    """uint8 [H,W,3] -> bytes; [N,H,W,3] -> list of N bytes objects."""
    if not 1 <= int(quality) <= 100:
        raise ValueError(f"quality {quality} outside 1..100")
    img, _ = _u8(image, "image", _dev() if not isinstance(image, torch.Tensor) or image.device.type != "cuda" else image.device)
    single = img.dim() == 3
    if single:
        img = img.unsqueeze(0)
    if img.dim() != 4 or img.shape[-1] != 3:
        raise ValueError(f"image must be [H,W,3] or [N,H,W,3] uint8, got {tuple(img.shape)}")
    n, H, W = img.shape[0], img.shape[1], img.shape[2]
    if n == 0:
        return []
    worst = _cabi.lib().rct_jpeg_max_bytes(H, W)
    cap = min(worst, int(H * W * _FIRST_TRY_BYTES_PER_PIXEL) + 4096)
    out, sizes = _encode(img, quality, cap)
    if max(sizes) > cap:
        out, sizes = _encode(img, quality, worst)
        if max(sizes) > worst:
            raise RuntimeError("JPEG stream exceeds the worst-case bound (internal error)")
    host = out[:, :max(sizes)].cpu().numpy()
    files = [host[k, :sizes[k]].tobytes() for k in range(n)]
    return files[0] if single else files


def save_jpeg(image, path, quality: int = 75):
    """LST:241-243 with the encoder on the GPU: writes the same file as ``Image.fromarray(image).save(path)``."""
    data = encode_jpeg(image, quality)
    if not isinstance(data, bytes):
        raise ValueError("save_jpeg takes one image; use encode_jpeg for a batch")
    with open(path, "wb") as f:
        f.write(data)
