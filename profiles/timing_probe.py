"""In-kernel timestamps of the entropy passes (build with ICB_TIMING=1): where a pass spends its time."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["ICB_TIMING"] = "1"
from image_compression_b200 import build
build.build(force=True)
import image_compression_b200 as pkg
from image_compression_b200 import synth
img = synth.natural_image(1000, 4000, 3000)
with pkg.Engine(devices=[0], slots_per_gpu=1, max_w=4000, max_h=3000) as eng:
    h = eng.upload_image(img, 0)
    for _ in range(3):
        p = eng.profile_jpg_fit_image(h, 0.25, 1048576, 1)
    print(p)
