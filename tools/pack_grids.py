"""Pack the reference's PNG map sets into small committed fixtures (build container only).

The maps are INPUT DATA of every BASELINE config (Grids/bg2_*), not reference source.  Each
output data/grids/<set>.npz holds, for the files of /root/reference/Grids/<set> in SORTED
filename order:  ``names`` (str), ``shape`` (W, L), ``bits`` = np.packbits(grid > 0) per map
(1 = free / +1, 0 = obstacle / -1; verified: every PNG is strictly {0, 255}).
``marl_coverage_b200.maps.load_grid_set`` turns them back into the float64 +-1 grids
``Utils/gridmaker.py:gridload`` produces.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import ref_harness as rh  # noqa: E402
from PIL import Image  # noqa: E402


def main():
    out = os.path.join(ROOT, "data", "grids")
    os.makedirs(out, exist_ok=True)
    for name, limit in (("bg2_100x100", None), ("bg2_256x256", None), ("bg2_1073x1073", None),
                        ("circulartest", None)):
        d = os.path.join(rh.REFERENCE_ROOT, "Grids", name)
        files = sorted(os.listdir(d))[:limit]
        bits, shape = [], None
        for f in files:
            img = np.array(Image.open(os.path.join(d, f)))
            assert set(np.unique(img)) <= {0, 255}, f
            g = np.clip(img.astype(float) - 1, -1, 1)
            shape = g.shape
            bits.append(np.packbits((g > 0).ravel()))
        p = os.path.join(out, name + ".npz")
        np.savez_compressed(p, names=np.array(files), shape=np.array(shape, dtype=np.int32),
                            bits=np.stack(bits))
        print(name, len(files), shape, "%.1f kB" % (os.path.getsize(p) / 1e3))
    # the nine 15x15 maps gridload(None) builds in code (Utils/gridmaker.py:24-80): 6 train + 3 test
    sys.path.insert(0, rh.REFERENCE_ROOT)
    rh._install_stubs()
    from Utils.gridmaker import gridload
    with rh.quiet():
        train, test = gridload(None)
    grids = train + test
    names = ["train%d" % (k + 1) for k in range(len(train))] + ["test%d" % (k + 1) for k in range(len(test))]
    p = os.path.join(out, "builtin15.npz")
    np.savez_compressed(p, names=np.array(names), shape=np.array(grids[0].shape, dtype=np.int32),
                        bits=np.stack([np.packbits((g > 0).ravel()) for g in grids]),
                        n_train=np.int32(len(train)))
    print("builtin15", len(grids), grids[0].shape, "%.1f kB" % (os.path.getsize(p) / 1e3))


if __name__ == "__main__":
    main()
