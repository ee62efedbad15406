"""Generates tests/golden/jpeg.npz: JPEG byte strings written by Pillow and the pixels the reference's own decode call
(`Image.open(io.BytesIO(b)).convert("RGB")`, backends/onnxrt_backend.py:478,412) returns for them.  Run here (Pillow 12.2.0 with
libjpeg-turbo); the fixture travels to the GPU box, this script does not need to."""
import io
from pathlib import Path

import numpy as np
from PIL import Image

rng = np.random.default_rng(7)


def synth(h, w, kind):
    yy, xx = np.mgrid[0:h, 0:w]
    if kind == 0:
        a = rng.integers(0, 256, (h, w, 3))
    elif kind == 1:
        a = np.stack([(xx * 3 + yy) % 256, (yy * 5) % 256, (xx * yy) % 256], -1)
    else:
        a = np.stack([128 + 100 * np.sin(xx / 7.0), 128 + 100 * np.cos(yy / 5.0), 128 + 60 * np.sin((xx + yy) / 11.0)], -1)
    return a.astype(np.uint8)


cases = []          # (name, bytes, pixels)
for (h, w) in [(1, 1), (2, 3), (8, 8), (15, 17), (17, 23), (33, 47), (64, 48), (100, 150), (31, 1), (1, 31), (120, 160)]:
    for kind in range(3):
        for sub, sname in ((0, "444"), (1, "422"), (2, "420")):
            q = (30, 75, 95, 100)[(h + w + kind + sub) % 4]
            buf = io.BytesIO()
            Image.fromarray(synth(h, w, kind)).save(buf, "JPEG", quality=q, subsampling=sub)
            cases.append((f"{h}x{w}_k{kind}_{sname}_q{q}", buf.getvalue()))
for (h, w) in [(17, 23), (64, 48), (100, 150)]:
    buf = io.BytesIO(); Image.fromarray(synth(h, w, 2)[..., 0], "L").save(buf, "JPEG", quality=80); cases.append((f"{h}x{w}_gray", buf.getvalue()))
    for tag, kw in (("opt", dict(optimize=True)), ("rstblk3", dict(restart_marker_blocks=3)), ("rstrow1", dict(restart_marker_rows=1))):
        buf = io.BytesIO(); Image.fromarray(synth(h, w, 1)).save(buf, "JPEG", quality=85, **kw); cases.append((f"{h}x{w}_{tag}", buf.getvalue()))
# flavours the decoder must refuse (status 1): progressive, CMYK, PNG
buf = io.BytesIO(); Image.fromarray(synth(40, 56, 1)).save(buf, "JPEG", quality=85, progressive=True); cases.append(("refuse_progressive", buf.getvalue()))
buf = io.BytesIO(); Image.fromarray(synth(40, 56, 1)).convert("CMYK").save(buf, "JPEG", quality=85); cases.append(("refuse_cmyk", buf.getvalue()))
buf = io.BytesIO(); Image.fromarray(synth(40, 56, 1)).save(buf, "PNG"); cases.append(("refuse_png", buf.getvalue()))

out = {"names": np.array([c[0] for c in cases])}
for i, (name, b) in enumerate(cases):
    out[f"bytes_{i}"] = np.frombuffer(b, np.uint8)
    out[f"rgb_{i}"] = np.asarray(Image.open(io.BytesIO(b)).convert("RGB"))
np.savez_compressed(Path(__file__).with_name("jpeg.npz"), **out)
print(len(cases), "cases", sum(len(c[1]) for c in cases), "bytes of JPEG")
