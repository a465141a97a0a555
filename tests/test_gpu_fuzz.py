"""-m gpu parity on seeded random cases: odd shapes, every starting quality class (lists built
for q, lists built for the first midpoint and pruned / rebuilt), targets on both sides of every
candidate size.  The oracle decides what is right; nothing here is shape- or quality-specific."""
import numpy as np
import pytest

from image_compression_b200 import synth, LearnedParams

pytestmark = pytest.mark.gpu

QS = [1.0, 0.9, 0.75, 0.6, 0.5, 0.45, 0.44, 0.3, 0.25, 0.1, 0.05, 0.02]


def _case(seed):
    rng = np.random.default_rng(9000 + seed)
    w = int(rng.integers(1, 700))
    h = int(rng.integers(1, 500))
    kind = seed % 3
    if kind == 0:
        img = synth.natural_image(seed, w, h)
    elif kind == 1:
        img = synth.noise_image(seed, w, h)
    else:                                   # flat areas with a few hard edges: long zero runs, EOB-only blocks
        img = np.full((h, w, 3), int(rng.integers(0, 256)), np.uint8)
        for _ in range(4):
            x0, y0 = int(rng.integers(0, w)), int(rng.integers(0, h))
            img[y0: y0 + int(rng.integers(1, 64)), x0: x0 + int(rng.integers(1, 64))] = rng.integers(0, 256, 3)
    q = QS[int(rng.integers(0, len(QS)))]
    return rng, img, q


@pytest.mark.parametrize("seed", range(36))
def test_random_fit_matches_oracle(engine, oracle, seed):
    rng, img, q = _case(seed)
    # a target between the smallest and the largest size the search can see, or just outside
    lo = oracle.encoded_size(img, 0.015625)
    hi = oracle.encoded_size(img, q)
    pick = int(rng.integers(0, 4))
    if pick == 0:
        target = int(rng.integers(lo, max(lo + 1, hi + 1)))
    elif pick == 1:
        target = hi + int(rng.integers(0, 50))            # everything fits: the search climbs to the top
    elif pick == 2:
        target = max(1, lo - int(rng.integers(1, 200)))   # nothing fits at scale 1: the scale loop runs
    else:
        target = oracle.encoded_size(img, q / 2) - int(rng.integers(0, 2))   # on the edge of the first midpoint
    rc, ref, rq, rs, _, _ = oracle.jpg_fit(img, q, target)
    r = engine.jpg_fit(img, q, target)
    assert (r.status, r.best_q, r.scale) == (rc, rq, rs), (img.shape, q, target)
    if rc == 0:
        assert r.data == ref and len(r.data) <= target
        # the learned parameters replay it in one encode (core/ImageCompressionJpg.java:216-238)
        c = engine.jpg_fit(img, q, target, cached=LearnedParams(r.best_q, r.scale))
        assert c.used_cache and c.data == ref


@pytest.mark.parametrize("seed", range(8))
def test_random_sizes_match_oracle(engine, oracle, seed):
    rng, img, _ = _case(100 + seed)
    qs = [float(k) / 128.0 for k in rng.integers(2, 128, 6)]
    assert list(engine.jpg_sizes(img, qs)) == [oracle.encoded_size(img, q) for q in qs]


_FORCED = r"""
import sys
sys.path.insert(0, %r)
import numpy as np
import image_compression_b200 as pkg
from image_compression_b200 import synth
from oracle import oracle
bad = 0
with pkg.Engine(slots_per_gpu=2) as eng:
    for seed, (w, h, q) in enumerate([(640, 480, 1.0), (333, 251, 0.25), (801, 603, 0.75), (97, 1025, 0.5)]):
        img = synth.natural_image(700 + seed, w, h)
        for frac in (0.02, 0.3, 0.7, 1.1):
            target = int(oracle.encoded_size(img, q) * frac)
            rc, ref, rq, rs, _, _ = oracle.jpg_fit(img, q, target)
            r = eng.jpg_fit(img, q, target)
            ok = (r.status, r.best_q, r.scale) == (rc, rq, rs) and (rc != 0 or r.data == ref)
            bad += 0 if ok else 1
print("BAD", bad)
"""


@pytest.mark.parametrize("forced", ["0", "1"])
def test_both_list_strategies_give_the_same_bytes(forced):
    """The host only guesses which way the first midpoint goes (api.cu expect_first_pass_fails); force
    the guess both ways (ICB_PRUNE is read once per process) and compare every result with the oracle."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, ICB_PRUNE=forced)
    out = subprocess.run([sys.executable, "-c", _FORCED % root], env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    assert "BAD 0" in out.stdout, out.stdout[-500:]


@pytest.mark.parametrize("tile", ["32", "64", "128", "256"])
def test_every_entropy_tile_size_gives_the_same_bytes(tile):
    """The entropy tile (blocks per warp and pass) is picked from the image size (common.cuh
    ent_tile_blocks); only time may depend on it.  Force each size (ICB_ENT_TILE is read once per
    process) on images that span several tiles and on ones smaller than a tile, both list strategies."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for prune in ("0", "1"):
        env = dict(os.environ, ICB_ENT_TILE=tile, ICB_PRUNE=prune)
        out = subprocess.run([sys.executable, "-c", _FORCED % root], env=env, capture_output=True, text=True,
                             timeout=600)
        assert out.returncode == 0, out.stderr[-2000:]
        assert "BAD 0" in out.stdout, out.stdout[-500:]
