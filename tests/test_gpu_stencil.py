"""T2: every single-GPU kernel variant, byte-compared with the CPU oracle."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

import b200conv
from b200conv import blur, reference_oracle


def rnd(h, w, ch, seed=0):
    return np.random.default_rng(seed).integers(0, 256, (h, w) if ch == 1 else (h, w, ch), dtype=np.uint8)


def gpu_blur(img, reps, **kw):
    out = blur(torch.from_numpy(img).cuda(), reps, **kw)
    torch.cuda.synchronize()
    return out.cpu().numpy()


SIZES = [(1, 1), (3, 5), (17, 23), (64, 96), (130, 517), (257, 1030)]


@pytest.mark.parametrize("boundary", ["zero", "torus"])
@pytest.mark.parametrize("ch", [1, 3])
@pytest.mark.parametrize("hw", SIZES)
def test_stream_auto_matches_oracle(boundary, ch, hw):
    img = rnd(hw[0], hw[1], ch, seed=hw[0])
    for reps in (0, 1, 2, 3, 7, 13):
        want = reference_oracle(img, reps, boundary)
        got = gpu_blur(img, reps, boundary=boundary)
        assert (got == want).all(), f"reps={reps} differs at {np.argwhere(got != want)[:4]}"


@pytest.mark.parametrize("tb,strip", [(1, 4), (2, 4), (4, 4), (8, 4), (1, 8), (2, 8), (4, 8), (8, 8),
                                      (1, 16), (2, 16), (4, 16)])
@pytest.mark.parametrize("ch", [1, 3])
@pytest.mark.parametrize("boundary", ["zero", "torus"])
def test_stream_variants(tb, strip, ch, boundary):
    img = rnd(150, 700 if strip == 16 else 300, ch, seed=tb * 100 + strip)
    for reps in (tb - 1, tb, tb + 1, 2 * tb + 3):
        if reps < 0:
            continue
        want = reference_oracle(img, reps, boundary)
        got = gpu_blur(img, reps, boundary=boundary, kernel="stream", tb=tb, strip_bytes=strip)
        assert (got == want).all(), f"tb={tb} strip={strip} reps={reps}: {np.argwhere(got != want)[:4]}"


@pytest.mark.parametrize("tb,strip", [(1, 4), (4, 4), (8, 4), (2, 8), (8, 8), (1, 16), (2, 16), (4, 16)])
@pytest.mark.parametrize("ch", [1, 3])
@pytest.mark.parametrize("boundary", ["zero", "torus"])
def test_tma_variants(tb, strip, ch, boundary):
    """Same pipeline fed by cp.async.bulk.tensor.2d tiles + mbarriers instead of LDGs."""
    img = rnd(150, 700 if strip == 16 else 300, ch, seed=tb * 100 + strip + 1)
    for reps in (tb, 2 * tb + 3):
        want = reference_oracle(img, reps, boundary)
        got = gpu_blur(img, reps, boundary=boundary, kernel="tma", tb=tb, strip_bytes=strip)
        assert (got == want).all(), f"tma tb={tb} strip={strip} reps={reps}: {np.argwhere(got != want)[:4]}"


def test_tma_long_ring_wraps_many_times():
    img = rnd(1200, 1100, 3, seed=21)  # several ring revolutions per warp
    want = reference_oracle(img, 12, "zero")
    for chunk_rows in (0, 400):
        assert (gpu_blur(img, 12, kernel="tma", tb=4, chunk_rows=chunk_rows) == want).all()


@pytest.mark.parametrize("chunk_rows", [8, 16, 37])
def test_stream_chunking_and_graph(chunk_rows):
    img = rnd(200, 1100, 3, seed=9)
    want = reference_oracle(img, 40, "zero")
    for g in (0, 1):
        got = gpu_blur(img, 40, kernel="stream", tb=4, chunk_rows=chunk_rows, use_graph=g)
        assert (got == want).all()


@pytest.mark.parametrize("boundary", ["zero", "torus"])
@pytest.mark.parametrize("ch", [1, 3])
def test_naive_kernel(boundary, ch):
    img = rnd(45, 77, ch, seed=4)
    for reps in (1, 2, 5):
        assert (gpu_blur(img, reps, boundary=boundary, kernel="naive") == reference_oracle(img, reps, boundary)).all()


@pytest.mark.parametrize("name", ["box", "edge"])
def test_float_filters_bit_exact(name):
    img = rnd(40, 52, 3, seed=7)
    for boundary in ("zero", "torus"):
        want = reference_oracle(img, 3, boundary, filter=name)
        assert (gpu_blur(img, 3, boundary=boundary, filter=name) == want).all()


@pytest.mark.parametrize("ch", [1, 3])
def test_cuda_compat_mode(ch):
    img = rnd(33, 41, ch, seed=11)
    for reps in (0, 1, 2, 3, 6):
        want = reference_oracle(img, reps, "cuda-compat")
        out, iters = blur(torch.from_numpy(img).cuda(), reps, boundary="cuda-compat", return_iterations=True)
        assert iters == 2 * (reps // 2)
        assert (out.cpu().numpy() == want).all()


def test_cpu_tensor_roundtrip_and_40_reps():
    img = rnd(252, 192, 3, seed=2)
    t = torch.from_numpy(img).pin_memory()
    out = blur(t, 40)
    assert not out.is_cuda
    assert (out.numpy() == reference_oracle(img, 40, "zero")).all()


def test_launch_counter_moves():
    n0 = b200conv.total_launches()
    gpu_blur(rnd(64, 64, 1), 8, kernel="stream", tb=4, use_graph=0)
    assert b200conv.total_launches() - n0 == 2


def test_blur_file_program(tmp_path, monkeypatch):
    from b200conv import blur_file
    from b200conv.utils import read_raw, write_raw
    monkeypatch.chdir(tmp_path)
    img = rnd(120, 96, 3, seed=3)
    write_raw("w.raw", img)
    loop_ms, total_ms = blur_file("w.raw", 96, 120, "rgb", 10)
    assert loop_ms > 0 and total_ms >= loop_ms
    assert (read_raw("blur_w.raw", 96, 120, 3) == reference_oracle(img, 10, "zero")).all()


def test_torch_ops_binding():
    """torch.ops.b200conv.blur — the thin C++ PyTorch extension over the same core."""
    from b200conv.ops import load_torch_ops

    ops = load_torch_ops()
    img = rnd(70, 90, 3, seed=13)
    out = ops.blur(torch.from_numpy(img).cuda(), 6, "torus", "gaussian", 4, "auto")
    torch.cuda.synchronize()
    assert (out.cpu().numpy() == reference_oracle(img, 6, "torus")).all()
    assert ops.rows_division(4, 2520, 1920) == 2


def test_blur_pipeline_overlapped_steps():
    """BlurPipeline: several host images in flight on two engines/streams, each result exact."""
    from b200conv import BlurPipeline

    imgs = [rnd(200, 300, 3, seed=30 + i) for i in range(5)]
    srcs = [torch.from_numpy(x).pin_memory() for x in imgs]
    dsts = [torch.empty_like(s).pin_memory() for s in srcs]
    pipe = BlurPipeline(300, 200, 3, repetitions=12, depth=2)
    for s, d in zip(srcs, dsts):
        pipe.submit(s, d)
    pipe.wait()
    pipe.close()
    for x, d in zip(imgs, dsts):
        assert (d.numpy() == reference_oracle(x, 12, "zero")).all()


def test_run_inside_a_callers_stream_capture():
    # the engine normally wraps long runs in its own CUDA graph; when the caller is already
    # capturing the stream it must fall back to plain launches that land in the caller's graph
    from b200conv import Engine
    img = np.random.default_rng(11).integers(0, 256, (256, 512, 3), dtype=np.uint8)
    eng = Engine(512, 256, 3, tb=4)
    src = torch.from_numpy(img).cuda()
    out = torch.empty_like(src)
    eng.load(src)
    eng.run(4)                                            # first launch (kernel attribute set-up) outside the capture
    eng.load(src)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        eng.run(40)                                       # 10 passes: above the engine's own graph threshold
    g.replay()
    eng.store(out)
    torch.cuda.synchronize()
    assert (out.cpu().numpy() == reference_oracle(img, 40, "zero")).all()
    eng.close()


@pytest.mark.parametrize("boundary", ["zero", "torus"])
@pytest.mark.parametrize("ch", [1, 3])
def test_persistent_kernel_and_per_pass_launches_agree(boundary, ch):
    """One persistent kernel with per-tile dependencies (the default) against one launch per pass."""
    img = rnd(700, 2100, ch, seed=77)
    for reps, opts in ((43, dict(tb=4)), (30, dict(tb=4, chunk_rows=16)), (17, dict(tb=8, strip_bytes=8)),
                       (64, dict(tb=4, halo_depth=16, chunk_rows=24))):
        want = reference_oracle(img, reps, boundary)
        for persistent in (1, 0):
            got = gpu_blur(img, reps, boundary=boundary, kernel="stream", persistent=persistent, **opts)
            assert (got == want).all(), f"persistent={persistent} {opts}: {np.argwhere(got != want)[:4]}"


def test_persistent_kernel_is_one_launch_per_run():
    img = rnd(512, 1024, 3, seed=78)
    eng = b200conv.Engine(1024, 512, 3, tb=4, persistent=1)
    eng.load(torch.from_numpy(img).cuda())
    eng.exchange()
    l0 = eng.launches
    eng.run(400)
    assert eng.launches - l0 == 1
    eng.run(400)
    eng.run(203)   # 50 passes in the table + k=2 and k=1 launched on their own
    out = torch.empty(img.shape, dtype=torch.uint8, device="cuda")
    eng.store(out)
    torch.cuda.synchronize()
    assert (out.cpu().numpy() == reference_oracle(img, 1003, "zero")).all()
    eng.close()
