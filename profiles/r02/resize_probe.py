"""Device time of the resize kernel (and resize + PNG encoder) alone: back-to-back launches on one stream
between two CUDA events (icb_profile_resize_series), for the shapes of BASELINE configs[2]."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import image_compression_b200 as pkg
from image_compression_b200 import synth

peak = json.load(open("MEASURED_PEAKS.json"))["hbm_gbs"] if os.path.exists("MEASURED_PEAKS.json") else 6650.0
with pkg.Engine(devices=[0], slots_per_gpu=2) as eng:
    for (w, h, c) in [(3840, 2160, 4), (7680, 4320, 4), (3840, 2160, 3), (7680, 4320, 3), (4000, 3000, 3)]:
        img = synth.natural_image(2000, w, h, channels=c)
        hnd = eng.upload_image(img, 0)
        dw, dh = (1920, 1080) if c == 4 or w != 4000 else (3400, 2550)
        ms = eng.profile_resize_series(hnd, dw, dh, 30)
        touched = (w * h * c if w / dw <= 2.01 else w * h * c // 2) + dw * dh * c
        out = {"src": [w, h, c], "dst": [dw, dh], "resize_us": round(ms * 1e3, 2), "sector_bytes": touched,
               "gbs": round(touched / ms / 1e6, 1), "frac_of_measured_peak": round(touched / ms / 1e6 / peak, 3)}
        if w == 3840 or c == 4:
            out["resize_plus_png_us"] = round(eng.profile_resize_series(hnd, dw, dh, 10, with_png=True) * 1e3, 1)
        print(json.dumps(out), flush=True)
        eng.free_image(hnd)
        if c == 4:       # the same shape with an opaque alpha plane (photographs saved as RGBA): no premultiply / divide
            img2 = img.copy(); img2[:, :, 3] = 255
            hnd = eng.upload_image(img2, 0)
            ms2 = eng.profile_resize_series(hnd, dw, dh, 30)
            print(json.dumps({"src": [w, h, c], "dst": [dw, dh], "alpha": "opaque", "resize_us": round(ms2 * 1e3, 2),
                              "gbs": round(touched / ms2 / 1e6, 1), "frac_of_measured_peak": round(touched / ms2 / 1e6 / peak, 3)}), flush=True)
            eng.free_image(hnd)
