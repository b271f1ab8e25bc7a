"""Golden vectors for the JPEG tests: JPEG files (bytes) written by cv2 / PIL (libjpeg-turbo) in the variants the tile
pipeline can meet, together with what the genuine library -- cv2.imdecode, the call inside upstream LoadImages
[UPSTREAM-RECALLED] -- decodes them to.  python tools/make_golden_jpeg.py -> tests/golden/jpeg_ref.npz
(cv2 4.13.0 / libjpeg-turbo 3.1.2 in the build container)."""
import io
import os

import cv2
import numpy as np
from PIL import Image

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rng = np.random.default_rng(11)


def scene(h, w):
    """Smooth colour field + noise + a few hard-edged shapes (cage-like circles / squares)."""
    a = rng.integers(0, 256, (h // 8 + 2, w // 8 + 2, 3), dtype=np.uint8)
    a = cv2.resize(a, (w, h), interpolation=cv2.INTER_CUBIC).astype(np.int32) + rng.integers(-16, 16, (h, w, 3))
    a = np.ascontiguousarray(np.clip(a, 0, 255).astype(np.uint8))
    for _ in range(4):
        c = (int(rng.integers(0, w)), int(rng.integers(0, h)))
        col = tuple(int(v) for v in rng.integers(0, 256, 3))
        if rng.random() < 0.5:
            cv2.circle(a, c, int(rng.integers(3, max(4, min(h, w) // 4))), col, 2)
        else:
            s = int(rng.integers(3, max(4, min(h, w) // 4)))
            cv2.rectangle(a, c, (c[0] + s, c[1] + s), col, 1)
    return a


def enc_cv2(img, *params):
    ok, buf = cv2.imencode(".jpg", img, list(params))
    assert ok
    return buf.tobytes()


def enc_pil(img, **kw):
    b = io.BytesIO()
    Image.fromarray(img).save(b, "JPEG", **kw)
    return b.getvalue()


cases = {}
for (h, w) in [(96, 128), (64, 64), (37, 53), (17, 16), (16, 17), (100, 3), (31, 6), (1, 1)]:
    img = scene(h, w)
    tag = f"{h}x{w}"
    cases[f"{tag}_q75_420"] = enc_cv2(img, cv2.IMWRITE_JPEG_QUALITY, 75)
    cases[f"{tag}_q95_420_rst"] = enc_cv2(img, cv2.IMWRITE_JPEG_QUALITY, 95, cv2.IMWRITE_JPEG_RST_INTERVAL, 3)
    cases[f"{tag}_q40_420_opt"] = enc_cv2(img, cv2.IMWRITE_JPEG_QUALITY, 40, cv2.IMWRITE_JPEG_OPTIMIZE, 1)
    cases[f"{tag}_q75_444"] = enc_pil(img, quality=75, subsampling=0)
    cases[f"{tag}_gray"] = enc_pil(np.ascontiguousarray(img[:, :, 1]), quality=80)
    cases[f"{tag}_noise_q100"] = enc_pil(rng.integers(0, 256, (h, w, 3), dtype=np.uint8), quality=100)
out = {}
for k, blob in cases.items():
    out["jpg_" + k] = np.frombuffer(blob, np.uint8)
    out["bgr_" + k] = cv2.imdecode(np.frombuffer(blob, np.uint8), cv2.IMREAD_COLOR)
# variants without a device path (must be refused, never decoded some other way)
out["unsupported_progressive"] = np.frombuffer(enc_pil(scene(64, 64), progressive=True), np.uint8)
out["unsupported_422"] = np.frombuffer(enc_pil(scene(64, 64), subsampling=1), np.uint8)
out["unsupported_cmyk"] = np.frombuffer((lambda b: (Image.fromarray(scene(32, 32)).convert("CMYK").save(b, "JPEG"), b.getvalue())[1])(io.BytesIO()), np.uint8)
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "jpeg_ref.npz"), **out)
print(len(cases), "cases,", sum(v.nbytes for v in out.values()), "bytes raw")
