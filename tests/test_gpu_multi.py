"""One handle driving several GPUs from one process -- the shape of the Java host: one JVM, the file
list split over the GPUs of the box, per-GPU stream work queues (CompressionBatch.java:64-108 becomes
icb_init slots / icb_submit_* / icb_wait).  Needs two or more devices; on a one-GPU box the tests skip
(`gpurun --gpus 2 -- python -m pytest tests/test_gpu_multi.py -m gpu`)."""
import numpy as np
import pytest

from image_compression_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def multi():
    import image_compression_b200 as pkg
    eng = pkg.Engine(slots_per_gpu=3)           # devices=None: every visible GPU
    if eng.device_count < 2:
        eng.close()
        pytest.skip("needs two or more GPUs")
    yield eng
    eng.close()


def test_queue_spreads_fits_over_all_gpus_and_every_result_matches(multi, oracle):
    n_dev = multi.device_count
    imgs = [synth.natural_image(900 + i, 512 + 48 * (i % 5), 384 + 40 * (i % 3)) for i in range(6 * n_dev)]
    target = 24000
    outs = [np.empty(target, np.uint8) for _ in imgs]
    want = [oracle.jpg_fit(im, 0.5, target) for im in imgs]
    seen = set()
    # keep every slot of every GPU busy: submit a window, wait in order
    window = 3 * n_dev
    tickets = []
    results = [None] * len(imgs)
    for i, im in enumerate(imgs):
        if len(tickets) == window:
            j, t = tickets.pop(0)
            results[j] = multi.wait(t)
        tickets.append((i, multi.submit_jpg_fit(im, 0.5, target, outs[i])))
    for j, t in tickets:
        results[j] = multi.wait(t)
    for r, out, (rc, ref, rq, rs, _, _) in zip(results, outs, want):
        assert (r.status, r.best_q, r.scale) == (rc, rq, rs)
        if rc == 0:
            assert bytes(out[:r.out_len]) == ref
        seen.add(r.device)
    assert seen == set(range(n_dev)), "some GPU never took a job: %s" % sorted(seen)


def test_resident_images_run_on_the_gpu_that_holds_them(multi, oracle):
    img = synth.natural_image(950, 800, 600)
    rc, ref, rq, rs, _, _ = oracle.jpg_fit(img, 0.25, 20000)
    for d in range(multi.device_count):
        hnd = multi.upload_image(img, d)
        out = np.empty(20000, np.uint8)
        r = multi.wait(multi.submit_jpg_fit_image(hnd, 0.25, 20000, out))
        multi.free_image(hnd)
        assert r.device == d
        assert (r.status, r.best_q, r.scale) == (rc, rq, rs)
        if rc == 0:
            assert bytes(out[:r.out_len]) == ref


def test_native_executor_with_resident_images_on_every_gpu_never_waits_while_holding_tickets(multi, oracle):
    """Four workers, three tickets each, items bound to the GPU that holds their image: the pattern that left every
    worker waiting for a slot another waiting worker held (bench.py one_process_all_gpus, r02)."""
    n_dev = multi.device_count
    img = synth.natural_image(960, 640, 480)
    rc, ref, rq, rs, _, _ = oracle.jpg_fit(img, 0.25, 16000)
    handles = [multi.upload_image(img, d) for d in range(n_dev)]
    n = 40 * n_dev
    outs = [np.empty(16000, np.uint8) for _ in range(n)]
    items = multi.make_items([dict(kind="jpg_fit_image", image=handles[i % n_dev], init_q=0.25, target=16000, out=outs[i])
                              for i in range(n)])
    for _ in range(5):
        multi.run_list(items, threads=4, inflight=3)
    for r, o in zip(items.results(), outs):
        assert (r.status, r.best_q, r.scale) == (rc, rq, rs)
        if rc == 0:
            assert bytes(o[: r.out_len]) == ref
    for hnd in handles:
        multi.free_image(hnd)


def test_file_jobs_on_every_gpu(multi, oracle):
    """icb_submit_jpg_file_fit through one handle over all GPUs: the reader's kernels (shared-memory opt-in, graphs,
    arenas) are set up per GPU, and every GPU decodes and fits bit-exactly."""
    n_dev = multi.device_count
    img = synth.natural_image(970, 1024, 768)
    data = oracle.encode(img, 0.9)
    px = oracle.jpeg_decode(data)
    rc, ref, rq, rs, _, _ = oracle.jpg_fit(px, 0.25, 30000)
    n = 8 * n_dev
    outs = [np.empty(30000, np.uint8) for _ in range(n)]
    items = multi.make_items([dict(kind="jpg_file_fit", file=data, init_q=0.25, target=30000, out=outs[i]) for i in range(n)])
    multi.run_list(items, threads=2 * n_dev, inflight=2)
    seen = set()
    for r, o in zip(items.results(), outs):
        assert (r.status, r.best_q, r.scale) == (rc, rq, rs)
        assert bytes(o[: r.out_len]) == ref
        seen.add(r.device)
    assert seen == set(range(n_dev)), sorted(seen)
