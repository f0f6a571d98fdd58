"""slam_get_cov_block_static: the allocation-free variant of slam_get_cov_block (static device scratch of
SLAM_COV_BLOCK_STATIC_MAX doubles; no counterpart in the reference).

Called through the C ABI (ctypes on libslamcuda.so) and checked against the existing entry point on two calls in a
row -- the second window is smaller and elsewhere, so contents the shared array kept from the first call would show.
Both entry points run the same gather kernel, so the results must be bit-equal (stricter than the 1e-9 the parity
tests use).  This file is named to be collected last: a failure here must not
hide the parity tests.  It builds its own state, so it passes alone in a fresh process as well as after the others."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SENTINEL = -12345.678


def _state(seed, nf):
    rng = np.random.default_rng(seed)
    n = 3 + 2 * nf
    x = np.concatenate([[50.0, 50.0, 0.3], rng.uniform(20, 80, size=2 * nf)])
    L = rng.normal(0, 0.05, size=(n, 8))
    P = L @ L.T + np.diag(rng.uniform(0.01, 0.1, size=n))
    P = np.asfortranarray((P + P.T) * 0.5)
    return x, P


def _create(lib, check, dptr, x, P, spare=17):
    h = C.c_void_p()
    check(lib.slam_create(0, len(x) + spare, C.byref(h)))
    check(lib.slam_set_state(h, dptr(x), dptr(P), len(x), len(x)))
    return h


def _both(lib, check, dptr, h, i0, ni, j0, nj, ldo):
    """The window through the existing entry point and through the static one, each into a sentinel-filled
    ldo x nj column-major buffer."""
    a = np.full((ldo, nj), SENTINEL, order="F")
    b = np.full((ldo, nj), SENTINEL, order="F")
    check(lib.slam_get_cov_block(h, i0, ni, j0, nj, dptr(a), ldo))
    check(lib.slam_get_cov_block_static(h, i0, ni, j0, nj, dptr(b), ldo))
    return a, b


def test_static_cov_block_matches_the_allocating_entry_point():
    from slamcuda._ffi import lib, check, dptr, iptr, COV_BLOCK_STATIC_MAX, ERR_CAPACITY
    x, P = _state(21, 200)
    n = len(x)
    h = _create(lib, check, dptr, x, P)
    # call 1: a large window across the diagonal (54000 of the 65536 scratch elements)
    a, b = _both(lib, check, dptr, h, 5, 180, 40, 300, 180)
    assert np.array_equal(a, b)
    assert np.array_equal(b, P[5:185, 40:340])                   # and it IS the window of the state that was set
    # call 2, straight after: smaller, elsewhere, below the diagonal (mirrored reads), ldo > ni
    a, b = _both(lib, check, dptr, h, 300, 60, 10, 90, 77)
    assert np.array_equal(a, b)
    assert np.array_equal(b[:60], P[300:360, 10:100])
    assert (b[60:] == SENTINEL).all()                            # rows ni..ldo of `out` are the caller's
    # exactly at the limit: 256 x 256
    assert 256 * 256 == COV_BLOCK_STATIC_MAX
    a, b = _both(lib, check, dptr, h, 100, 256, 3, 256, 256)
    assert np.array_equal(a, b) and np.array_equal(b, P[100:356, 3:259])
    # one element past it: rejected with a clear error, nothing written
    out = np.full((257, 256), SENTINEL, order="F")
    rc = lib.slam_get_cov_block_static(h, 0, 257, 0, 256, dptr(out), 257)
    msg = lib.slam_last_error().decode()
    assert rc == ERR_CAPACITY, (rc, msg)
    assert "SLAM_COV_BLOCK_STATIC_MAX" in msg and "65536" in msg and "257 x 256" in msg
    assert (out == SENTINEL).all()
    # the other argument checks are the existing entry point's
    assert lib.slam_get_cov_block_static(h, 0, 4, n - 2, 4, dptr(out), 257) < 0     # window outside the state
    assert lib.slam_get_cov_block_static(h, 0, 8, 0, 8, dptr(out), 4) < 0           # ldo < ni
    assert lib.slam_get_cov_block_static(None, 0, 1, 0, 1, dptr(out), 257) < 0      # NULL handle
    check(lib.slam_get_cov_block_static(h, 7, 0, 7, 0, dptr(out), 257))             # empty window: fine, no write
    assert (out == SENTINEL).all()
    # the state moves (an update), the static path follows it
    ids = np.array([3, 77, 150], dtype=np.int64)
    zf = np.empty((2, 3), order="F")
    for c, i in enumerate(ids):
        dx, dy = x[3 + 2 * (i - 1)] - x[0], x[4 + 2 * (i - 1)] - x[1]
        zf[:, c] = [np.hypot(dx, dy) + 0.05, np.arctan2(dy, dx) - x[2] + 0.004]
    R = np.asfortranarray(np.diag([0.01, (np.pi / 180) ** 2]))
    check(lib.slam_update(h, dptr(zf), 3, dptr(R), iptr(ids)))
    a, b = _both(lib, check, dptr, h, 0, 120, 0, n, 120)
    assert np.array_equal(a, b)
    assert not np.array_equal(b, P[0:120, :])                    # (the update did change these rows)
    check(lib.slam_destroy(h))


def test_static_cov_block_is_shared_between_handles():
    """Two handles on one device use the same array one after the other (sequential calls are the supported use;
    overlapping ones race, as the header says): each gets its own state's window."""
    from slamcuda._ffi import lib, check, dptr
    x1, P1 = _state(5, 60)
    x2, P2 = _state(6, 90)
    h1 = _create(lib, check, dptr, x1, P1)
    h2 = _create(lib, check, dptr, x2, P2)
    for _ in range(2):
        a, b = _both(lib, check, dptr, h1, 2, 100, 0, 123, 100)
        assert np.array_equal(a, b) and np.array_equal(b, P1[2:102, 0:123])
        a, b = _both(lib, check, dptr, h2, 50, 40, 20, 160, 40)
        assert np.array_equal(a, b) and np.array_equal(b, P2[50:90, 20:180])
    check(lib.slam_destroy(h1))
    check(lib.slam_destroy(h2))


def test_static_cov_block_per_device_scratch():
    """The static array exists once per GPU and its device address is resolved per device (cudaGetSymbolAddress with that
    device current): handles on devices 0 and 1 interleaved, each window equal to the allocating entry point's and to
    the state set on that device.  Skipped on a one-GPU box."""
    import slamcuda
    from slamcuda._ffi import lib, check, dptr
    if slamcuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    states = [_state(31, 70), _state(32, 110)]
    hs = []
    for dev, (x, P) in enumerate(states):
        h = C.c_void_p()
        check(lib.slam_create(dev, len(x) + 5, C.byref(h)))
        check(lib.slam_set_state(h, dptr(x), dptr(P), len(x), len(x)))
        hs.append(h)
    for rnd in range(3):
        for dev in (1, 0, 1, 0):
            x, P = states[dev]
            i0, ni, j0, nj = 3 + rnd, 90 + 10 * dev, rnd, 120 + 7 * dev
            a, b = _both(lib, check, dptr, hs[dev], i0, ni, j0, nj, ni)
            assert np.array_equal(a, b) and np.array_equal(b, P[i0:i0 + ni, j0:j0 + nj]), (rnd, dev)
    for h in hs:
        check(lib.slam_destroy(h))
