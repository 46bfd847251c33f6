"""The ring-first work order of the streaming kernels (map_work, csrc/cuda/stencil_stream.cuh), run on
the host through b2c_debug_map_work.  The per-direction arrival counters of the halo protocol
(HaloParams.expected) assume that every (band, chunk) is visited exactly once; a gap or a duplicate
would leave a counter short and hang the neighbouring GPUs, so the bijection is checked for every
small plan, together with the property that gives the overlap: ring items come before interior ones."""
import ctypes as C

from b200conv._lib import lib


def order(nb, nc, nbw, nbe):
    out = []
    band, chunk = C.c_int(), C.c_int()
    for i in range(nb * nc):
        assert lib.b2c_debug_map_work(i, nb, nc, nbw, nbe, C.byref(band), C.byref(chunk)) == 0
        out.append((band.value, chunk.value))
    return out


def test_every_plan_is_a_bijection_and_ring_first():
    checked = 0
    for nb in range(1, 10):
        for nc in range(1, 10):
            for nbw in range(1, nb + 1):
                for nbe in range(1, nb + 1):
                    items = order(nb, nc, nbw, nbe)
                    assert sorted(items) == [(b, c) for b in range(nb) for c in range(nc)], (nb, nc, nbw, nbe)
                    ring = [c == 0 or c == nc - 1 or b < nbw or b >= nb - nbe for b, c in items]
                    n_ring = sum(ring)
                    assert all(ring[:n_ring]) and not any(ring[n_ring:]), (nb, nc, nbw, nbe)
                    # what the host publishes as expected pusher counts per direction
                    top = sum(c == 0 for _, c in items)
                    west = sum(b < nbw for b, _ in items)
                    nw = sum(c == 0 and b < nbw for b, c in items)
                    assert (top, west, nw) == (nb, nbw * nc, nbw)
                    checked += 1
    assert checked > 2000


def test_bad_arguments_are_refused():
    band, chunk = C.c_int(), C.c_int()
    assert lib.b2c_debug_map_work(6, 2, 3, 1, 1, C.byref(band), C.byref(chunk)) < 0
    assert lib.b2c_debug_map_work(-1, 2, 3, 1, 1, C.byref(band), C.byref(chunk)) < 0


# ---- the band / chunk plan itself (b2c_debug_band_plan runs choose_plan + make_band_plan on the host)
from hypothesis import given, settings  # noqa: E402
from hypothesis import strategies as st  # noqa: E402


@settings(max_examples=400, deadline=None)
@given(rows=st.integers(1, 5000), cols=st.integers(1, 6000), ch=st.sampled_from([1, 3]), nb=st.booleans(),
       tb=st.sampled_from([0, 1, 2, 4, 8]), strip=st.sampled_from([0, 4, 8, 16]), chunk=st.sampled_from([0, 1, 7, 64, 100000]),
       kernel=st.sampled_from([0, 2, 3]))
def test_plan_invariants_the_kernels_rely_on(rows, cols, ch, nb, tb, strip, chunk, kernel):
    out = (C.c_int * 10)()
    assert lib.b2c_debug_band_plan(rows, cols, ch, int(nb), tb, strip, chunk, kernel, 148, out) == 0
    hlo, vw, nbands, nchunks, chunk_rows, nbw, nbe, K, SB, F = list(out)
    bw = cols * ch
    align = 16 if kernel == 3 else SB
    assert K in (1, 2, 4, 8) and SB in (4, 8, 16) and K * (SB // 4) <= 16        # register budget
    if nb:
        assert K <= min(rows, cols) or K == 1                                   # a K-deep halo comes from one neighbour
        assert F % K == 0 and K <= F <= 32 and (F <= min(rows, cols) or F == K)
    # window of a band: [band*vw - hlo, +32*SB); it must hold K*CH bytes of overlap on both sides
    assert vw > 0 and hlo % align == 0 and vw % align == 0
    assert K * ch <= hlo <= 32
    assert 32 * SB - hlo - vw >= K * ch
    assert nbands == -(-bw // vw) and 1 <= chunk_rows <= rows
    last = rows - (nchunks - 1) * chunk_rows          # the last chunk takes the remainder
    assert nchunks >= 1 and 1 <= last < 2 * chunk_rows + 64
    if nb:
        # the whole F-deep ring must sit in the chunks / bands that push it, and the chunks next to
        # the first / last one must not reach into a frame (they do not wait for the flags)
        assert chunk_rows >= min(max(F, 2 * K), rows)
        assert nchunks == 1 or last >= max(F, 2 * K)
        assert nbw == nbands or nbw * vw >= F * ch
        assert nbe == nbands or (nbands - nbe) * vw <= bw - F * ch
        assert 1 <= nbw <= nbands and 1 <= nbe <= nbands
    else:
        assert nchunks == -(-rows // chunk_rows)


@settings(max_examples=300, deadline=None)
@given(rows=st.integers(8, 3000), cols=st.integers(8, 3000), ch=st.sampled_from([1, 3]),
       tb=st.sampled_from([0, 1, 2, 4, 8]), strip=st.sampled_from([0, 4, 8, 16]), chunk=st.sampled_from([0, 1, 9, 64]),
       halo_depth=st.sampled_from([0, 4, 8, 12, 16, 32]), reps=st.integers(1, 90))
def test_pass_groups_of_a_rank_with_neighbours(rows, cols, ch, tb, strip, chunk, halo_depth, reps):
    """One halo exchange per group of passes: what every pass of a group must satisfy so that the
    kernels' ring-membership rules (first/last chunk, first/last bands) cover every frame access."""
    WAIT_DATA, POST_ACK, WAIT_ACK, PUSH = 1, 2, 4, 8
    buf = (C.c_int * (16 * 256))()
    n = lib.b2c_debug_pass_plans(rows, cols, ch, 1, tb, strip, chunk, halo_depth, reps, 148, buf, 256)
    assert 0 < n <= 256
    passes = [list(buf[16 * i:16 * i + 12]) for i in range(n)]
    out = (C.c_int * 10)()
    lib.b2c_debug_band_plan(rows, cols, ch, 1, tb, strip, chunk, 0, 148, out)
    bw = cols * ch
    assert sum(p[0] for p in passes) == reps
    group = []
    for k, ext, mode, gend, prow, pbw, nbands, nchunks, chunk_rows, last, nbw, nbe in passes:
        group.append((k, ext, mode))
        xcol = -(-ext * ch // 16) * 16
        assert prow == rows + 2 * ext and pbw == bw + 2 * xcol       # extended into the frame on all four sides
        assert ext + k <= 32 and xcol + 32 <= 128                    # inside the 40-row / 128-byte frame
        assert last == prow - (nchunks - 1) * chunk_rows
        SB = 16 if strip == 0 else strip
        if nchunks > 1:
            # only the first / last chunk may touch frame rows: readers of the pass (k rows beyond the
            # extended block) and, in the pushing pass, the F-deep ring
            assert chunk_rows >= ext + k and last >= ext + k
        if gend:
            n_in = len(group)
            assert ext == 0 and mode & PUSH
            assert group[0][2] & WAIT_DATA
            assert bool(mode & WAIT_ACK) == (n_in >= 2)
            if n_in >= 2:
                assert group[-2][2] & POST_ACK and group[-2][1] == k
            assert sum(g[0] for g in group) <= 32
            assert [g[1] for g in group] == [sum(x[0] for x in group[i + 1:]) for i in range(n_in)]
            for g in group[:-1]:
                assert not g[2] & PUSH
            group = []
    assert group == []


# ---- the persistent kernel's tile dependencies (stencil_stream.cuh, stream_persist_kernel)

def _tiles(ps, ch, win):
    """Every (band, chunk) of a pass with the regions it needs to read and the region it writes, in
    the coordinates of the un-extended block (rows, bytes)."""
    k, ext, mode, gend, prow, pbw, nb, nc, cr, last, nbw, nbe, hlo, vw, xn, xwb = ps
    out = {}
    for c in range(nc):
        r0, r1 = c * cr, (prow if c == nc - 1 else (c + 1) * cr)
        for b in range(nb):
            c0, c1 = b * vw, min((b + 1) * vw, pbw)
            assert c0 - k * ch >= b * vw - hlo and c1 + k * ch <= b * vw - hlo + win      # the window holds what is needed
            write = (r0 - xn, r1 - xn, c0 - xwb, c1 - xwb)
            read = (r0 - k - xn, r1 + k - xn, c0 - k * ch - xwb, c1 + k * ch - xwb)
            out[(b, c)] = (read, write)
    return out


def _overlap(a, b):
    return a[0] < b[1] and b[0] < a[1] and a[2] < b[3] and b[2] < a[3]


@settings(max_examples=120, deadline=None)
@given(rows=st.integers(40, 700), cols=st.integers(40, 900), ch=st.sampled_from([1, 3]),
       chunk=st.sampled_from([0, 9, 33, 64]), strip=st.sampled_from([0, 4, 8, 16]),
       halo_depth=st.sampled_from([0, 8, 16, 32]), has_nb=st.booleans())
def test_a_tile_waits_for_every_tile_it_depends_on(rows, cols, ch, chunk, strip, halo_depth, has_nb):
    """A work item (pass t, band, chunk) of the persistent kernel waits for the 3x3 neighbourhood
    (clamped) of pass t-1 in THAT pass's tiling.  With neighbours the block is extended into the
    frame by a different amount from pass to pass, so the two tilings are shifted against each other
    and may differ by one band / chunk.  Every tile of pass t-1 that wrote what the item reads (RAW),
    or read what the item overwrites (WAR: pass t writes the buffer pass t-1 read), must be in that
    set — checked on the real geometry of the plan."""
    buf = (C.c_int * (16 * 64))()
    n = lib.b2c_debug_pass_plans(rows, cols, ch, int(has_nb), 4, strip, chunk, halo_depth, 48, 148, buf, 64)
    assert 0 < n <= 64
    passes = [list(buf[16 * i:16 * i + 16]) for i in range(n)]
    passes = [p for p in passes if p[0] == 4]          # the table of the persistent kernel: passes with k == K
    out = (C.c_int * 10)()
    lib.b2c_debug_band_plan(rows, cols, ch, int(has_nb), 4, strip, chunk, 0, 148, out)
    win = 32 * out[8]
    for prev, cur in zip(passes, passes[1:]):
        tp, tc = _tiles(prev, ch, win), _tiles(cur, ch, win)
        pnb, pnc = prev[6], prev[7]
        for (b, c), (read, write) in tc.items():
            waits = {(min(max(b + db, 0), pnb - 1), min(max(c + dc, 0), pnc - 1)) for db in (-1, 0, 1) for dc in (-1, 0, 1)}
            for key, (pread, pwrite) in tp.items():
                if _overlap(pwrite, read) or _overlap(pread, write):
                    assert key in waits, (prev, cur, (b, c), key)
