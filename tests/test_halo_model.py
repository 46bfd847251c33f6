"""T3 without GPUs: the threaded host model of the halo / convergence protocol.

One thread per rank, the engine's epoch-flag and tagged-slot state machine restated with
std::atomic (csrc/core/halo_emul.cpp).  The result must equal the 1-rank oracle on every grid,
with ranks skewed by random sleeps; the same sources built with -fsanitize=thread must report no
data race; and the single-slot-set variant of the convergence exchange (what the GPU kernel used
before the parity fix) must be caught as a deadlock.
"""
import shutil
import subprocess
from pathlib import Path

import numpy as np
import pytest

from b200conv import B2CError, reference_oracle
from b200conv.ops import threaded_halo_model, threaded_persist_model, threaded_resident_model

ROOT = Path(__file__).resolve().parent.parent


def rnd(h, w, ch, seed=0):
    return np.random.default_rng(seed).integers(0, 256, (h, w) if ch == 1 else (h, w, ch), dtype=np.uint8)


@pytest.mark.parametrize("grid", [(1, 1), (1, 2), (2, 1), (2, 2), (2, 4), (4, 2), (3, 3)])
@pytest.mark.parametrize("boundary", ["zero", "torus"])
@pytest.mark.parametrize("ch", [1, 3])
def test_every_grid_gives_the_one_rank_result(grid, boundary, ch):
    img = rnd(24, 24, ch, seed=1)                    # 24 divides by 1, 2, 3 and 4
    for tb, reps in ((1, 5), (4, 11)):
        out, iters = threaded_halo_model(img, reps, grid, boundary, tb=tb, max_delay_us=200, seed=tb)
        assert iters == reps
        assert (out == reference_oracle(img, reps, boundary)).all(), (grid, boundary, ch, tb)


def test_many_epochs_with_skewed_ranks():
    # 250 passes; each rank sleeps a different random time at every point where GPUs may drift apart
    img = rnd(16, 32, 3, seed=2)
    out, iters = threaded_halo_model(img, 500, (2, 4), "torus", tb=2, max_delay_us=300, seed=5)
    assert iters == 500 and (out == reference_oracle(img, 500, "torus")).all()


@pytest.mark.parametrize("grid,boundary,reps", [((2, 2), "zero", 9), ((1, 4), "torus", 9), ((2, 4), "torus", 12),
                                                ((4, 2), "zero", 5), ((1, 2), "torus", 1)])
def test_images_back_to_back_without_a_barrier(grid, boundary, reps):
    # the ShardedPipeline claim: a rank may upload and exchange the next image while its neighbours
    # are still in the last pass of the previous one (odd pass counts flip the buffer parity per image)
    img = rnd(16, 32, 3, seed=7)
    outs, iters = threaded_halo_model(img, reps, grid, boundary, tb=4, max_delay_us=300, seed=11, images=4)
    assert iters == reps
    for j, out in enumerate(outs):
        assert (out == reference_oracle(img ^ np.uint8((37 * j) & 255), reps, boundary)).all(), j


@pytest.mark.parametrize("grid,boundary", [((2, 2), "zero"), ((2, 4), "torus"), ((1, 2), "torus"), ((3, 3), "zero"),
                                           ((1, 1), "torus")])
@pytest.mark.parametrize("tb,halo_depth,reps", [(4, 8, 21), (4, 12, 30), (4, 16, 37), (2, 6, 13), (1, 3, 7), (8, 16, 40)])
def test_one_exchange_per_group_of_passes(grid, boundary, tb, halo_depth, reps):
    # a rank keeps a halo `halo_depth` deep and exchanges it once per `halo_depth` iterations; the
    # passes in between recompute the shrinking halo themselves and write the frames of both
    # buffers, which is what the ack epoch protects (skewed ranks, odd remainders, short last group)
    img = rnd(48, 48, 3, seed=12)
    out, iters = threaded_halo_model(img, reps, grid, boundary, tb=tb, halo_depth=halo_depth, max_delay_us=200, seed=reps)
    assert iters == reps
    assert (out == reference_oracle(img, reps, boundary)).all(), (grid, boundary, tb, halo_depth, reps)


@pytest.mark.parametrize("grid,boundary,halo_depth,reps", [((2, 2), "zero", 0, 23), ((2, 4), "torus", 8, 37),
                                                          ((1, 2), "torus", 16, 50), ((3, 3), "zero", 12, 30)])
def test_epochs_forwarded_by_a_courier_thread(grid, boundary, halo_depth, reps):
    # the rank hands its epochs to a second thread that publishes them to the neighbours (with its own
    # random delay): frames are ordered through two hops, images back to back included
    img = rnd(48, 48, 3, seed=14)
    outs, _ = threaded_halo_model(img, reps, grid, boundary, tb=4, halo_depth=halo_depth, max_delay_us=200, seed=3,
                                  images=3, courier=True)
    for j, out in enumerate(outs):
        assert (out == reference_oracle(img ^ np.uint8((37 * j) & 255), reps, boundary)).all(), j


def test_grouped_exchange_with_images_back_to_back_and_with_checks():
    img = rnd(32, 64, 1, seed=13)
    outs, _ = threaded_halo_model(img, 19, (2, 4), "torus", tb=4, halo_depth=8, max_delay_us=300, seed=2, images=4)
    for j, out in enumerate(outs):
        assert (out == reference_oracle(img ^ np.uint8((37 * j) & 255), 19, "torus")).all(), j
    want = reference_oracle(img, 2000, "zero")
    out, iters = threaded_halo_model(img, 2000, (2, 2), "zero", tb=4, halo_depth=16, check_every=12, max_delay_us=100, seed=4)
    assert iters < 2000 and iters % 12 == 0 and (out == want).all()


def test_halo_depth_is_clamped_to_the_block():
    img = rnd(4, 8, 1, seed=3)                      # 2x2 blocks on a 2x4 grid: tb=8 must clamp to 2
    out, _ = threaded_halo_model(img, 9, (2, 4), "torus", tb=8)
    assert (out == reference_oracle(img, 9, "torus")).all()


@pytest.mark.parametrize("grid", [(1, 1), (1, 2), (2, 2), (1, 4)])
@pytest.mark.parametrize("boundary", ["zero", "torus"])
def test_convergence_is_detected_by_every_rank_at_the_same_check(grid, boundary):
    img = rnd(32, 32, 1, seed=4)
    want = reference_oracle(img, 1000, boundary)
    out, iters = threaded_halo_model(img, 1000, grid, boundary, tb=4, check_every=4, max_delay_us=100, seed=9)
    assert iters < 1000 and iters % 4 == 0
    assert (out == want).all()
    assert (reference_oracle(want, 1, boundary) == want).all()


def test_single_slot_set_loses_a_tag_and_two_sets_do_not():
    # every pass is a check; rank 0 stalls between publishing and reading its slots, so rank 1
    # reaches the next check first.  With one slot set rank 0 then never sees the tag it waits for.
    img = rnd(16, 16, 1, seed=6)
    kw = dict(tb=4, check_every=4, stall_rank=0, stall_ms=100, timeout_ms=1500)
    with pytest.raises(B2CError, match="deadline"):
        threaded_halo_model(img, 16, (1, 2), "torus", single_slot_set=True, **kw)
    out, iters = threaded_halo_model(img, 16, (1, 2), "torus", **kw)
    assert iters == 16 and (out == reference_oracle(img, 16, "torus")).all()


@pytest.mark.parametrize("case", [
    (40, 64, 1, 9, "torus", 4, 148), (130, 488, 1, 13, "zero", 4, 148), (200, 96, 3, 19, "torus", 8, 148),
    (300, 520, 1, 6, "torus", 4, 12), (90, 304, 1, 33, "zero", 16, 148), (61, 8, 1, 7, "torus", 4, 148),
])
def test_resident_cta_protocol_model(case):
    # one thread per CTA of the resident kernel: private window, ring out / halo in through two
    # exchange images chosen by epoch parity, epoch flags, CTAs skewed by random sleeps
    h, w, ch, reps, boundary, k, sms = case
    img = rnd(h, w, ch, seed=8)
    got = threaded_resident_model(img, reps, boundary, k=k, sms=sms, max_delay_us=200, seed=3)
    assert (got == reference_oracle(img, reps, boundary)).all()


@pytest.mark.parametrize("case", [
    # (H, W, ch, reps, k, vw, chunk_rows, workers)
    (96, 200, 1, 24, 4, 40, 8, 3),      # 5 bands x 12 chunks, far fewer workers than tiles
    (61, 90, 3, 16, 4, 60, 7, 16),      # ragged last band and last chunk, more workers than a chunk row
    (40, 64, 1, 32, 8, 64, 40, 4),      # one tile per pass: a pure chain
    (130, 48, 3, 12, 2, 30, 5, 32),     # many thin tiles, workers spread over several passes at once
])
def test_persistent_kernel_queue_model(case):
    # work items pulled from one counter, an item waits for the 3x3 neighbourhood of the pass before,
    # no barrier between passes, random delays
    h, w, ch, reps, k, vw, cr, workers = case
    img = rnd(h, w, ch, seed=21)
    got = threaded_persist_model(img, reps, k=k, vw=vw, chunk_rows=cr, workers=workers, max_delay_us=150, seed=7)
    assert (got == reference_oracle(img, reps, "zero")).all()


TSAN_MAIN = r"""
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "b200conv.h"
int main() {
    const int W = 24, H = 16, reps = 40;
    int bad = 0;
    for (int ch = 1; ch <= 3; ch += 2)
        for (int boundary = 0; boundary <= 1; ++boundary) {
            std::vector<uint8_t> in((size_t)W * H * ch), want(in.size()), got(in.size());
            srand(7);
            for (auto &b : in) b = (uint8_t)rand();
            b2c_oracle_blur(in.data(), want.data(), W, H, ch, reps, boundary, B2C_GAUSSIAN);
            b2c_emul_opts o;
            memset(&o, 0, sizeof o);
            o.tb = 4, o.max_delay_us = 50, o.seed = 3, o.stall_rank = -1;
            int iters = 0;
            int rc = b2c_emulate_halo_threads(in.data(), got.data(), W, H, ch, reps, boundary, 2, 4, &o, &iters);
            bad += rc != B2C_OK || memcmp(want.data(), got.data(), got.size()) != 0;
            {   // four images back to back with no barrier in between (ShardedPipeline)
                b2c_emul_opts p = o;
                p.images = 4;
                std::vector<uint8_t> many(in.size() * 4), x(in.size());
                rc = b2c_emulate_halo_threads(in.data(), many.data(), W, H, ch, 9, boundary, 2, 4, &p, &iters);
                bad += rc != B2C_OK;
                for (int j = 0; j < 4; ++j) {
                    for (size_t i = 0; i < in.size(); ++i) x[i] = (uint8_t)(in[i] ^ (uint8_t)(37 * j));
                    b2c_oracle_blur(x.data(), want.data(), W, H, ch, 9, boundary, B2C_GAUSSIAN);
                    bad += memcmp(want.data(), many.data() + (size_t)j * in.size(), in.size()) != 0;
                }
            }
            o.check_every = 4;   // convergence exchange under the race detector as well
            rc = b2c_emulate_halo_threads(in.data(), got.data(), W, H, ch, 400, boundary, 2, 2, &o, &iters);
            b2c_oracle_blur(in.data(), want.data(), W, H, ch, iters, boundary, B2C_GAUSSIAN);
            bad += rc != B2C_OK || memcmp(want.data(), got.data(), got.size()) != 0;
        }
    for (int boundary = 0; boundary <= 1; ++boundary)      // one exchange per group of passes: data + ack epochs
        for (int hd = 8; hd <= 16; hd += 4) {
            const int W2 = 64, H2 = 32, ch = 3, reps2 = 45;
            std::vector<uint8_t> in((size_t)W2 * H2 * ch), want(in.size()), many(in.size() * 3), x(in.size());
            srand(11);
            for (auto &b : in) b = (uint8_t)rand();
            b2c_emul_opts o;
            memset(&o, 0, sizeof o);
            o.tb = 4, o.max_delay_us = 50, o.seed = 5, o.stall_rank = -1, o.halo_depth = hd, o.images = 3;
            o.courier = hd != 12;   // two of the three depths through courier threads
            int iters = 0;
            int rc = b2c_emulate_halo_threads(in.data(), many.data(), W2, H2, ch, reps2, boundary, 2, 4, &o, &iters);
            bad += rc != B2C_OK;
            for (int j = 0; j < 3; ++j) {
                for (size_t i = 0; i < in.size(); ++i) x[i] = (uint8_t)(in[i] ^ (uint8_t)(37 * j));
                b2c_oracle_blur(x.data(), want.data(), W2, H2, ch, reps2, boundary, B2C_GAUSSIAN);
                bad += memcmp(want.data(), many.data() + (size_t)j * in.size(), in.size()) != 0;
            }
        }
    for (int boundary = 0; boundary <= 1; ++boundary) {   // CTA protocol of the resident kernel, 3 x 3 tiles
        const int RW_ = 488, RH_ = 130, reps2 = 13;
        std::vector<uint8_t> in((size_t)RW_ * RH_), want(in.size()), got(in.size());
        srand(9);
        for (auto &b : in) b = (uint8_t)rand();
        b2c_oracle_blur(in.data(), want.data(), RW_, RH_, 1, reps2, boundary, B2C_GAUSSIAN);
        int rc = b2c_emulate_resident_threads(in.data(), got.data(), RW_, RH_, 1, reps2, boundary, 4, 148, 100, 5);
        bad += rc != B2C_OK || memcmp(want.data(), got.data(), got.size()) != 0;
    }
    {   // work queue + tile dependencies of the persistent kernel: 5 x 12 tiles per pass, 3 workers, then 16
        const int PW = 200, PH = 96, preps = 24;
        std::vector<uint8_t> in((size_t)PW * PH), want(in.size()), got(in.size());
        srand(13);
        for (auto &b : in) b = (uint8_t)rand();
        b2c_oracle_blur(in.data(), want.data(), PW, PH, 1, preps, B2C_ZERO, B2C_GAUSSIAN);
        for (int workers = 3; workers <= 16; workers += 13) {
            int rc = b2c_emulate_persist_threads(in.data(), got.data(), PW, PH, 1, preps, 4, 40, 8, workers, 100, 3);
            bad += rc != B2C_OK || memcmp(want.data(), got.data(), got.size()) != 0;
        }
    }
    printf("bad=%d\n", bad);
    return bad != 0;
}
"""


@pytest.mark.skipif(shutil.which("g++") is None, reason="needs g++")
def test_model_is_race_free_under_thread_sanitizer(tmp_path):
    main = tmp_path / "main.cpp"
    main.write_text(TSAN_MAIN)
    exe = tmp_path / "halo_tsan"
    srcs = [str(ROOT / "csrc/core" / f) for f in ("halo_emul.cpp", "resident_emul.cpp", "persist_emul.cpp", "oracle.cpp",
                                                  "decomp.cpp")]
    cmd = ["g++", "-std=c++17", "-O1", "-g", "-fsanitize=thread", "-pthread", "-I", str(ROOT / "include"),
           "-I", str(ROOT / "csrc/core"), str(main), *srcs, "-o", str(exe)]
    b = subprocess.run(cmd, capture_output=True, text=True)
    if b.returncode != 0 and "tsan" in (b.stderr or "").lower():
        pytest.skip("libtsan is not installed")
    assert b.returncode == 0, b.stderr[-3000:]
    r = subprocess.run([str(exe)], capture_output=True, text=True, timeout=600,
                       env={"TSAN_OPTIONS": "halt_on_error=0 exitcode=66"})
    assert "ThreadSanitizer" not in r.stderr, r.stderr[-4000:]
    assert r.returncode == 0 and "bad=0" in r.stdout, (r.returncode, r.stdout, r.stderr[-2000:])
