"""Two decode + fit jobs from a 12 MP file, for the ncu launch list of the GPU JPEG reader."""
import io, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import image_compression_b200 as pkg
from image_compression_b200 import synth
from PIL import Image
im = synth.natural_image(3000, 4000, 3000)
bio = io.BytesIO(); Image.fromarray(im).save(bio, "JPEG", quality=92, subsampling=2)
f = np.frombuffer(bio.getvalue(), np.uint8)
with pkg.Engine(devices=[0], slots_per_gpu=1, max_w=4000, max_h=3000) as eng:
    for _ in range(3):
        r = eng.jpg_file_fit(f, 0.25, 1 << 20)
    print(r.status, r.best_q, r.out_len)
