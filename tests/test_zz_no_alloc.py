"""alegpu_loss_static — the allocation-free variant of alegpu_loss (static device scratch inside the
library; no counterpart in the reference).

Collected last on purpose (`zz`): a failure here hides nothing.
The GPU test calls the new entry point through the C ABI and compares it with the existing entry point
(alegpu_loss) at the tolerance the loss tests use (test_gpu_parity.py::test_loss_sizes), on two calls
in a row of different sizes and masks, so stale scratch contents of the first call would show up in
the second.  It holds whether this file runs alone in a fresh process or after the other tests."""
import ctypes as C

import numpy as np
import pytest

import alegpu
import cases

KEYS = ("clipped_losses", "value_losses", "entropies", "total_losses", "ratio", "dlogits", "dvalues")


def _args(lc, n):
    mk = lc["masks"] | (np.arange(n) == 0).astype(np.uint8)  # at least one counted row
    return (lc["logits"], lc["old_log_probs"], lc["actions"], lc["advantages"], lc["values"],
            lc["returns"], mk, 0.2, 0.5, 0.01)


def _same(got, want, tag):
    for k in KEYS:
        np.testing.assert_allclose(got[k], want[k], rtol=1e-5, atol=1e-6, err_msg=f"{tag}: {k}")
    assert abs(got["loss"] - want["loss"]) <= 1e-5 * abs(want["loss"]) + 1e-6, tag


@pytest.mark.gpu
def test_loss_static_matches_loss_on_two_calls_in_a_row():
    # call 1: 64 blocks of partials; call 2: 2 blocks, other masks -> the 62 stale partials and the old
    # count of call 1 must not leak into call 2; call 3 repeats call 1 (bit-identical to itself)
    shapes = [(16384, 4), (257, 4), (16384, 4), (100, 18)]
    first = None
    for i, (n, A) in enumerate(shapes):
        a = _args(cases.loss_case(n, A, seed=n + A), n)
        got = alegpu.losses.compute_static(*a)
        want = alegpu.losses.compute(*a)
        _same(got, want, f"call {i} (n={n}, A={A})")
        assert np.isfinite(got["loss"])
        if i == 0:
            first = got
        if i == 2:
            for k in KEYS:
                assert np.array_equal(got[k], first[k]), k
            assert got["loss"] == first["loss"]


@pytest.mark.gpu
def test_loss_static_at_its_maximum_size():
    n = alegpu.losses.LOSS_STATIC_MAX_ROWS
    a = _args(cases.loss_case(n, 4, seed=1), n)
    _same(alegpu.losses.compute_static(*a), alegpu.losses.compute(*a), "max rows")


def test_loss_static_rejects_sizes_beyond_the_static_scratch():
    # argument checks come before any CUDA call, so this runs without a GPU
    lib = alegpu.lib()
    none = [None] * 8
    rc = lib.alegpu_loss_static(None, None, None, None, None, None, None,
                                C.c_int64(alegpu.losses.LOSS_STATIC_MAX_ROWS + 1), C.c_int64(4),
                                C.c_float(.1), C.c_float(.5), C.c_float(.01), *none, None)
    assert rc == 1
    msg = lib.alegpu_last_error()
    assert b"exceeds the static scratch" in msg and b"1048576" in msg
    with pytest.raises(ValueError):
        alegpu._check(lib.alegpu_loss_static(None, None, None, None, None, None, None, C.c_int64(4),
                                             C.c_int64(99), C.c_float(.1), C.c_float(.5),
                                             C.c_float(.01), *none, None))
