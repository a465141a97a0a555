"""PNG encoder alone on a resident 1080p image: per-kernel device times come from the ncu launch list; this prints
resize + encode per image (CUDA events, back to back) and the file size against zlib."""
import json, os, sys, zlib
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import image_compression_b200 as pkg
from image_compression_b200 import synth
with pkg.Engine(devices=[0], slots_per_gpu=2) as eng:
    for (w, h, c) in [(1920, 1080, 4), (1920, 1080, 3)]:
        img = synth.natural_image(2100, w, h, channels=c)
        hnd = eng.upload_image(img, 0)
        rs = eng.profile_resize_series(hnd, w // 2, h // 2, 10)
        both = eng.profile_resize_series(hnd, w // 2, h // 2, 10, with_png=True)
        data = eng.png_encode(img)
        print(json.dumps({"shape": [w, h, c], "png_bytes": len(data), "raw": img.size,
                          "resize_half_us": round(rs * 1e3, 1), "resize_half_plus_png_quarter_us": round(both * 1e3, 1)}), flush=True)
        eng.free_image(hnd)
