"""GPU test of rtb_color_to_u8_static, the entry point that allocates nothing (no
counterpart in the reference).  Collected last on purpose (file name).

The conversion is integer-exact on both paths (the same kernel), so the bar is equality -- the bar of every other test here."""
import numpy as np
import pytest

import scenes as S

pytestmark = pytest.mark.gpu


def test_static_scratch_entry_point_equals_the_allocating_one(built):
    import rtb200
    v, c = S.random_soup(8, seed=1)
    sc = rtb200.Scene(v, c)                       # a scene of its own: the test must also pass alone, in a fresh process
    rng = np.random.default_rng(2026)
    # two calls in a row, different sizes and contents: the second must not see anything of the first (shared static scratch)
    first = rng.uniform(0.0, 1.1, 700001).astype(np.float32)
    second = np.concatenate([rng.uniform(0.0, 0.02, 1234), [0.0, 1.0, 0.5, 0.25, np.nan, -1.0, 7.0, np.inf]]).astype(np.float32)
    for lin in (first, second, first[:5], second):
        want = sc.color_to_u8(lin)
        got = sc.color_to_u8_static(lin)
        assert got.dtype == np.uint8 and got.shape == lin.shape
        assert (got == want).all()
    assert list(sc.color_to_u8_static(np.array([0.0, 1.0, 0.5, 0.25, 0.75, 0.01], np.float32))) == [0, 255, 186, 135, 223, 31]   # SURVEY a-12
    # exactly the maximum is accepted, one more is a clear error, zero values is a no-op
    full = rng.uniform(0.0, 1.0, rtb200.binding.COLOR_STATIC_MAX).astype(np.float32)
    assert (sc.color_to_u8_static(full) == sc.color_to_u8(full)).all()
    with pytest.raises(rtb200.RtbError, match="static scratch"):
        sc.color_to_u8_static(np.zeros(rtb200.binding.COLOR_STATIC_MAX + 1, np.float32))
    assert sc.color_to_u8_static(np.zeros(0, np.float32)).size == 0
    # and the allocating path still works after the failed call
    assert (sc.color_to_u8_static(second) == sc.color_to_u8(second)).all()
