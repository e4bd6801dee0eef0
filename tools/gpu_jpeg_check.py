"""GPU check of the JPEG encoder kernels through tools/jpeg_selftest (no torch import: starts in a second).
Compares with Pillow byte for byte; prints one line per case and a JSON summary to gpurun_out/jpeg_selftest.json."""
import io, json, os, subprocess, sys, tempfile
import numpy as np
from PIL import Image

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "tools", "jpeg_selftest")


def pillow(a, q):
    b = io.BytesIO()
    Image.fromarray(a).save(b, "JPEG", quality=q)
    return b.getvalue()


def image(kind, h, w, rng):
    if kind == "noise":
        return rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    yy, xx = np.mgrid[0:h, 0:w]
    base = np.stack([(xx * 5 + yy) % 256, (yy * 3) % 256, (xx + yy * 2) % 256], -1)
    return np.clip(base + rng.integers(-12, 12, (h, w, 3)), 0, 255).astype(np.uint8)


def main():
    mode = sys.argv[1] if len(sys.argv) > 1 else "--gpu"
    rng = np.random.default_rng(1)
    cases = [(16, 16, "noise", 75, 1), (1, 1, "noise", 75, 1), (17, 23, "noise", 75, 2), (33, 65, "smooth", 75, 1),
             (100, 75, "noise", 100, 3), (270, 480, "smooth", 75, 2), (1080, 1920, "smooth", 75, 4),
             (2160, 3840, "smooth", 75, 8), (1080, 1920, "noise", 100, 2)]
    rows, ok = [], True
    with tempfile.TemporaryDirectory() as tmp:
        for h, w, kind, q, n in cases:
            a = image(kind, h, w, rng)
            a.tofile(os.path.join(tmp, "in.rgb"))
            p = subprocess.run([EXE, mode, str(h), str(w), os.path.join(tmp, "in.rgb"), os.path.join(tmp, "o.jpg"), str(q),
                                str(n)], capture_output=True, text=True, timeout=120)
            same = p.returncode == 0 and open(os.path.join(tmp, "o.jpg"), "rb").read() == pillow(a, q)
            ok &= same
            rows.append({"H": h, "W": w, "kind": kind, "quality": q, "frames": n, "byte_exact": bool(same),
                         "rc": p.returncode, "out": (p.stdout + p.stderr).strip()})
            print(rows[-1], flush=True)
            os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
            json.dump(rows, open(os.path.join(ROOT, "gpurun_out", "jpeg_selftest.json"), "w"), indent=1)
    print("ALL BYTE-EXACT" if ok else "MISMATCH")
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
