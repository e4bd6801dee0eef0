"""GPU check of the JPEG decoder kernels through tools/jpeg_selftest --gpu-decode (no torch): bit-exact vs Pillow."""
import io, json, os, subprocess, sys, tempfile
import numpy as np
from PIL import Image

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "tools", "jpeg_selftest")


def main():
    mode = sys.argv[1] if len(sys.argv) > 1 else "--gpu-decode"
    rng = np.random.default_rng(2)
    cases = [(16, 16, {}), (1, 1, {}), (2, 3, {}), (17, 23, {}), (33, 65, {"subsampling": 1}), (40, 9, {"subsampling": 0}),
             (100, 75, {"quality": 95}), (70, 90, {"restart_marker_blocks": 3}), (270, 480, {"restart_marker_rows": 1}),
             (1080, 1920, {"restart_marker_rows": 1}), (540, 960, {})]
    rows, ok = [], True
    with tempfile.TemporaryDirectory() as tmp:
        for h, w, kw in cases:
            yy, xx = np.mgrid[0:h, 0:w]
            a = np.clip(np.stack([(xx * 5 + yy) % 256, (yy * 3) % 256, (xx + yy * 2) % 256], -1) + rng.integers(-20, 20, (h, w, 3)), 0, 255).astype(np.uint8)
            b = io.BytesIO()
            Image.fromarray(a).save(b, "JPEG", **kw)
            open(os.path.join(tmp, "i.jpg"), "wb").write(b.getvalue())
            p = subprocess.run([EXE, mode, os.path.join(tmp, "i.jpg"), os.path.join(tmp, "o.raw")], capture_output=True, text=True, timeout=60)
            same = False
            if p.returncode == 0:
                got = np.fromfile(os.path.join(tmp, "o.raw"), np.uint8).reshape(h, w, 3)
                same = bool(np.array_equal(got, np.asarray(Image.open(io.BytesIO(b.getvalue())).convert("RGB"))))
            ok &= same
            rows.append({"H": h, "W": w, "save": kw, "bit_exact": same, "rc": p.returncode, "out": p.stderr.strip()})
            print(rows[-1], flush=True)
            os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
            json.dump(rows, open(os.path.join(ROOT, "gpurun_out", "jpeg_decode_selftest.json"), "w"), indent=1)
    print("ALL BIT-EXACT" if ok else "MISMATCH")
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
