"""Known-answer tests for the oracle's Philox4x32-10 (Random123 kat_vectors) and
the definition of the product's noise stream."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "oracle"))
import vae_oracle as O  # noqa: E402


def _kat(ctr, key):
    return [int(v) for v in O.philox4x32_10(np.array(ctr, dtype=np.uint32), np.array(key, dtype=np.uint32))]


def test_random123_known_answers():
    assert _kat([0, 0, 0, 0], [0, 0]) == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    assert _kat([0xffffffff] * 4, [0xffffffff] * 2) == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    assert _kat([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0]) == \
        [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]


def test_stream_is_offset_consistent():
    a = O.philox_uint32(1265, 0, 16)
    b = O.philox_uint32(1265, 4, 12)
    assert np.array_equal(a[4:], b)
    c = O.philox_uint32(1265, 0, 8, first_quad=8)
    assert np.array_equal(a[8:], c)


def test_normal_moments():
    x = O.philox_normal(1265, 0, 1 << 20).astype(np.float64)
    assert abs(x.mean()) < 4e-3 and abs(x.std() - 1) < 3e-3
    assert abs((x ** 3).mean()) < 2e-2 and abs((x ** 4).mean() - 3) < 5e-2
    assert np.isfinite(x).all()
