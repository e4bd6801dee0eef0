"""the write at the end of the reference's pipeline,
``Image.fromarray(combined_image).save(os.path.join(out_dir, "localized_style_transfer_result.jpg"))``
(Style_3DGS/localized_style_transfer.py:241-243), done on the GPU.

``encode_jpeg(image, quality=75)``    uint8 HWC (or NHWC batch) -> bytes (list of bytes for a batch), byte for
                                      byte what Pillow's ``save(..., "JPEG", quality=quality)`` writes
``save_jpeg(image, path, quality=75)`` the reference's call, file on disk
``decode_jpeg(data, device=False)``    bytes / path of a baseline JPEG -> uint8 [H,W,3] ([H,W] for a gray file), the
                                      pixels ``np.asarray(Image.open(path))`` gives (LST:200, 219), decoded on the GPU;
                                      ``device=True`` returns the CUDA tensor instead of a numpy array
``jpeg_info(data)``                   (H, W, channels) from the header, host only

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


def _file_bytes(data) -> bytes:
    if isinstance(data, (bytes, bytearray, memoryview)):
        return bytes(data)
    with open(data, "rb") as f:
        return f.read()


def jpeg_info(data):
    """(H, W, channels) of a baseline JPEG; raises ValueError for files the decoder does not take (progressive, CMYK ...)."""
    import ctypes as C

    raw = _file_bytes(data)
    h, w, ch, ws = C.c_int(), C.c_int(), C.c_int(), C.c_size_t()
    buf = (C.c_ubyte * len(raw)).from_buffer_copy(raw)
    lib = _cabi.lib()
    if lib.rct_jpeg_decode_info(buf, len(raw), C.byref(h), C.byref(w), C.byref(ch), C.byref(ws)) != 0:
        raise ValueError(lib.rct_last_error().decode())
    return h.value, w.value, ch.value


# default Huffman decoder for files without restart markers: the self-synchronising parallel one (True; bit-exact
# against Pillow on the B200, profiles/r02_bench_jpeg.json: 4K in 8.3 ms instead of seconds) or one thread per
# restart interval (False)
PARALLEL_DEFAULT = True


def decode_jpeg(data, device: bool = False, parallel=None):
    """LST:200, 219 (`Image.open`): file bytes or path -> pixels, decoded on the GPU, bit-exact with Pillow.
    ``parallel`` selects the self-synchronising Huffman decoder for files without restart markers (same pixels);
    None = the module default."""
    import ctypes as C

    raw = _file_bytes(data)
    lib = _cabi.lib()
    h, w, ch, ws = C.c_int(), C.c_int(), C.c_int(), C.c_size_t()
    buf = (C.c_ubyte * len(raw)).from_buffer_copy(raw)
    if lib.rct_jpeg_decode_info(buf, len(raw), C.byref(h), C.byref(w), C.byref(ch), C.byref(ws)) != 0:
        raise ValueError(lib.rct_last_error().decode())
    dev = _dev()
    work = torch.empty(ws.value, dtype=torch.uint8, device=dev)
    shape = (h.value, w.value, 3) if ch.value == 3 else (h.value, w.value)
    out = torch.empty(shape, dtype=torch.uint8, device=dev)
    use_parallel = PARALLEL_DEFAULT if parallel is None else bool(parallel)
    with torch.cuda.device(dev):
        check(lib.rct_jpeg_decode_async(buf, len(raw), work.data_ptr(), ws.value, out.data_ptr(),
                                        0 if use_parallel else _cabi.JPEG_SERIAL, torch.cuda.current_stream().cuda_stream))
        torch.cuda.current_stream().synchronize()   # `work` and `buf` are released on return
    return out if device else out.cpu().numpy()


_PNG_SIG = b"\x89PNG\r\n\x1a\n"


def decode_png(data):
    """`np.array(Image.open(path).convert("RGB"))` for a PNG file: uint8 [H,W,3].  Runs on the HOST inside the C
    library (zlib inflate and the scanline filters are serial; csrc/rct_png.h), no GPU needed."""
    import ctypes as C

    raw = _file_bytes(data)
    lib = _cabi.lib()
    h, w = C.c_int(), C.c_int()
    buf = (C.c_ubyte * len(raw)).from_buffer_copy(raw)
    if lib.rct_png_info(buf, len(raw), C.byref(h), C.byref(w)) != 0:
        raise ValueError(lib.rct_last_error().decode())
    out = np.empty((h.value, w.value, 3), np.uint8)
    if lib.rct_png_decode_host(buf, len(raw), out.ctypes.data, out.size) != 0:
        raise ValueError(lib.rct_last_error().decode())
    return out


def decode_image(data, device: bool = False):
    """`Image.open(path).convert("RGB")` (LST:200, 219) for the two formats of the reference's inputs: baseline JPEG
    (decoded on the GPU, grey files replicated to three channels) and PNG (decoded on the host, uploaded when
    ``device=True``).  Anything else raises ValueError."""
    raw = _file_bytes(data)
    if raw[:8] == _PNG_SIG:
        px = decode_png(raw)
        return torch.from_numpy(px).to(_dev()) if device else px
    px = decode_jpeg(raw, device=True)
    if px.dim() == 2:
        px = px.unsqueeze(-1).expand(-1, -1, 3).contiguous()
    return px if device else px.cpu().numpy()


def image_info(data):
    """(H, W) of a baseline JPEG or PNG file from its header (host only)."""
    import ctypes as C

    raw = _file_bytes(data)
    if raw[:8] == _PNG_SIG:
        lib = _cabi.lib()
        h, w = C.c_int(), C.c_int()
        buf = (C.c_ubyte * len(raw)).from_buffer_copy(raw)
        if lib.rct_png_info(buf, len(raw), C.byref(h), C.byref(w)) != 0:
            raise ValueError(lib.rct_last_error().decode())
        return h.value, w.value
    return jpeg_info(raw)[:2]


def regional_color_transfer_files(content_file, stylized_file, background_mask, quality: int = 75) -> bytes:
    """The reference's real boundary (LST:191-245: image files in, ``localized_style_transfer_result.jpg`` out) as ONE
    C-ABI call, ``rct_transfer_files_host``: content / stylised file bytes (or paths; baseline JPEG or PNG) and the uint8
    background mask -> decode, NEAREST resample, regional colour transfer, combination, JPEG encode on the device ->
    the bytes of the file Pillow's ``save(path, quality=quality)`` would write for the same pixels."""
    import ctypes as C

    from . import localized_style_transfer as lst

    c_raw, s_raw = _file_bytes(content_file), _file_bytes(stylized_file)
    H, W = image_info(c_raw)
    sh, sw = image_info(s_raw)
    mask = np.ascontiguousarray(background_mask, np.uint8)
    if mask.ndim == 3:
        mask = np.ascontiguousarray(mask[0])
    if mask.shape != (H, W):
        raise ValueError(f"mask is {mask.shape}, the content file is {(H, W)}")
    dev = _dev()
    plan = lst.get_plan(dev, H, W, sh, sw, 1)
    lib = _cabi.lib()
    cap = lib.rct_files_out_capacity(H, W)
    out = np.empty(cap, np.uint8)
    nbytes, status = C.c_size_t(0), C.c_int(0)
    cbuf = (C.c_ubyte * len(c_raw)).from_buffer_copy(c_raw)
    sbuf = (C.c_ubyte * len(s_raw)).from_buffer_copy(s_raw)
    with torch.cuda.device(dev), plan.lock:
        check(lib.rct_transfer_files_host(plan.handle, cbuf, len(c_raw), sbuf, len(s_raw), mask.ctypes.data, int(quality),
                                          out.ctypes.data, cap, C.byref(nbytes), C.byref(status)))
    lst._warn(status.value)
    return out[:nbytes.value].tobytes()
