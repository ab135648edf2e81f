"""The numpy model of the device RNG against Random123's published known-answer vectors for Philox4x32-10, and the
model of the on-device gridgen against the properties of the reference's gridgen (Utils/gridmaker.py:107-139)."""
import numpy as np

from oracle.philox import gen_maps, philox4x32_10

KAT = [
    ((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
    ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
    ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
     (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1)),
]


def test_philox_known_answers():
    for ctr, key, want in KAT:
        got = philox4x32_10(np.array(ctr, np.uint32), np.array(key, np.uint32))
        assert tuple(int(v) for v in got) == want
    ctr = np.array([k[0] for k in KAT], np.uint32)
    key = np.array([k[1] for k in KAT], np.uint32)
    assert np.array_equal(philox4x32_10(ctr, key), np.array([k[2] for k in KAT], np.uint32))     # batched form


def test_gen_maps_model_properties():
    g = gen_maps(6, 37, 70, 0.25, seed=3)
    assert len(g) == 6 and g[0].shape == (37, 70) and all(set(np.unique(x)) <= {-1.0, 1.0} for x in g)
    frac = np.mean([np.mean(x < 0) for x in g])
    assert abs(frac - 0.25) < 5 * 0.0035                                  # sigma of 15 540 Bernoulli cells
    assert all(np.array_equal(a, b) for a, b in zip(g, gen_maps(9, 37, 70, 0.25, seed=3)))      # table size
    assert not np.array_equal(g[0], gen_maps(1, 37, 70, 0.25, seed=4)[0])                       # seed
    assert np.array_equal(gen_maps(1, 37, 64, 0.25, seed=3)[0], g[0][:, :64])                   # a prefix in y
    assert (gen_maps(2, 9, 9, 0.0, seed=1)[1] > 0).all() and (gen_maps(2, 9, 9, 1.0, seed=1)[1] < 0).all()
    hi = gen_maps(1, 5, 8, 0.5, seed=(7 << 32) | 1)[0]
    assert not np.array_equal(hi, gen_maps(1, 5, 8, 0.5, seed=1)[0])      # the high seed word is part of the key
