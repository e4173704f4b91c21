"""Generate tests/golden/masks_reference.npz by running the REFERENCE'S OWN iteration-0 mask
preparation -- lib/utils/image.py::get_pair_mask and lib/utils/mask_dilate.py::mask_dilate,
imported unmodified, np.random seeded -- on the frames tests/helpers.write_pair_files puts on
disk, for the policies the shipped configs use: train / INIT_MASK "box_gt" with and without
MASK_DILATE, test / INIT_MASK "box_rendered" with and without MASK_DILATE; plus mask_dilate alone
on random masks for every `direction`.

    python tests/golden/make_golden_masks.py
"""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, HERE)
REF = "/root/reference"


def main():
    import helpers
    import make_golden_updater as U
    import mx_stub

    mx_stub.install()
    sys.path.insert(0, REF)
    from lib.utils import image
    from lib.utils.mask_dilate import mask_dilate

    rec = {}
    with tempfile.TemporaryDirectory() as tmp:
        pairdb = helpers.write_pair_files(tmp, 41)
        for phase, init in (("train", "box_gt"), ("test", "box_rendered")):
            for dil in (False, True):
                cfg = U.Cfg(SCALES=[(480, 640)], dataset=U.Cfg(DEPTH_FACTOR=1000),
                            TRAIN=U.Cfg(INIT_MASK=init, MASK_DILATE=dil),
                            TEST=U.Cfg(INIT_MASK=init, MASK_DILATE=dil))
                np.random.seed(77)
                mo, mg, mr = image.get_pair_mask(pairdb, cfg, [0] * len(pairdb), phase=phase)
                tag = "%s_%s_dil%d" % (phase, init, dil)
                rec[tag + "/mask_observed"] = np.packbits(np.concatenate(mo).astype(np.uint8))
                rec[tag + "/mask_gt_observed"] = np.packbits(np.concatenate(mg).astype(np.uint8))
                r = np.concatenate(mr)
                assert np.isin(np.concatenate(mo), (0, 1)).all()
                rec[tag + "/mask_rendered"] = r[..., ::4, ::4].astype(np.float32)
                rec[tag + "/mask_rendered_sum"] = r.astype(np.float64).sum(axis=(-2, -1))
    rng = np.random.default_rng(9)
    masks = (rng.random((12, 96, 128)) > 0.97).astype(np.float64)
    masks[3] *= 0.5      # non-binary values stay as they are
    masks[4, 10:20, 10:20] = 3.0  # > 1 is clamped
    outs = []
    np.random.seed(5)
    for m in masks:
        outs.append(mask_dilate(m, max_thickness=10))
    rec["dilate/out"] = np.asarray(outs)
    np.savez_compressed(os.path.join(HERE, "masks_reference.npz"), **rec)
    print("mask golden vectors written: %.0f KB"
          % (os.path.getsize(os.path.join(HERE, "masks_reference.npz")) / 1024))


if __name__ == "__main__":
    main()
