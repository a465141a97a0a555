"""Concurrent timeline of the fits sharing one GPU (ICB_TIMING build + ICB_DUMP_TIMELINE): six slots kept
busy with resident 12 MP images; at shutdown the library prints, per slot, when the kernels of its LAST fit
started and ended on the GPU's global timer -- how far a pass stretches when other fits' kernels run beside
it.  usage (GPU box): ICB_DUMP_TIMELINE=1 IMGCOMP_B200_LIB=.../libimgcomp_b200_timing.so python profiles/timeline_probe.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import image_compression_b200 as pkg
from image_compression_b200 import synth

SLOTS, N = 6, 96
imgs = [synth.natural_image(1000 + i, 4000, 3000) for i in range(4)]
eng = pkg.Engine(devices=[0], slots_per_gpu=SLOTS, max_w=4000, max_h=3000)
hs = [eng.upload_image(im, 0) for im in imgs]
outs = [eng.alloc_pinned((1 << 20,)) for _ in range(SLOTS)]   # pageable buffers would make every result copy synchronous
tickets = []
for i in range(N):
    if len(tickets) == SLOTS:
        eng.wait(tickets.pop(0))
    tickets.append(eng.submit_jpg_fit_image(hs[i % len(hs)], 0.25, 1 << 20, outs[i % SLOTS]))
for t in tickets:
    eng.wait(t)
eng.close()         # prints the timeline
