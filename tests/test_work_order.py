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
