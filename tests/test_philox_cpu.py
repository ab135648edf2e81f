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


def test_draw_starts_model_properties():
    from oracle.philox import draw_starts
    rng = np.random.RandomState(1)
    walls = rng.random_sample((5, 12, 9)) < 0.3
    seen_maps = set()
    for gid in range(40):
        mi, cells = draw_starts(walls, 4, seed=77, gid=gid, episode=0)
        seen_maps.add(mi)
        assert cells.shape == (4, 2) and len({tuple(c) for c in cells}) == 4          # distinct
        assert not walls[mi][cells[:, 0], cells[:, 1]].any()                          # walkable
        again = draw_starts(walls, 4, seed=77, gid=gid, episode=0)
        assert again[0] == mi and np.array_equal(again[1], cells)                     # a pure function of its key
        assert not np.array_equal(draw_starts(walls, 4, seed=77, gid=gid, episode=1)[1], cells) or \
            not np.array_equal(draw_starts(walls, 4, seed=78, gid=gid, episode=0)[1], cells)
    assert len(seen_maps) == 5                                                        # every map is drawn
    tight = np.ones((1, 4, 4), bool)
    tight[0, 2, 1] = tight[0, 3, 3] = False                                           # two walkable cells, three robots
    mi, cells = draw_starts(tight, 3, seed=1, gid=0, episode=0)
    assert {tuple(c) for c in cells[:2]} == {(2, 1), (3, 3)} and tuple(cells[2]) == (0, 0)
