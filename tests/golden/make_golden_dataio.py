"""Generate tests/golden/dataio_reference.npz by running the REFERENCE'S OWN loaders --
lib/utils/image.py get_pair_image / get_gt_observed_depth / get_pair_depth (MASK_INPUTS) and
lib/pair_matching/load_object_points.py, imported unmodified -- on frames written to a temp
directory in the dataset's on-disk formats by tests/helpers.write_pair_files.

    python tests/golden/make_golden_dataio.py
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
PIXEL_MEANS = np.array([102.9801, 115.9465, 122.7717])


def main():
    import helpers
    import make_golden_updater as U
    import mx_stub

    mx_stub.install()
    sys.path.insert(0, REF)
    from lib.pair_matching.load_object_points import load_object_points
    from lib.utils import image

    cfg = U.Cfg(SCALES=[(480, 640)], dataset=U.Cfg(DEPTH_FACTOR=1000, MASK_GT=True),
                network=U.Cfg(PIXEL_MEANS=PIXEL_MEANS, MASK_INPUTS=True),
                TRAIN=U.Cfg(MASK_SYN=False))
    rec = {}
    with tempfile.TemporaryDirectory() as tmp:
        pairdb = helpers.write_pair_files(tmp, 41)
        obs, rend, scale = image.get_pair_image(pairdb, cfg, phase="test")
        gt = image.get_gt_observed_depth(pairdb, cfg, scale, phase="test")
        d_obs, d_rend = image.get_pair_depth(pairdb, cfg, scale, phase="test", random_k=[])
        rec["points"] = load_object_points(os.path.join(tmp, "model", "points.xyz"))
        rec["poses"] = np.stack([np.loadtxt(r["pose_observed"], skiprows=1) for r in pairdb])
    S = 4
    rec["image_observed"] = np.concatenate(obs)[..., ::S, ::S].astype(np.float32)
    rec["image_rendered"] = np.concatenate(rend)[..., ::S, ::S].astype(np.float32)
    rec["depth_gt_observed"] = np.concatenate(gt)[..., ::S, ::S]
    rec["depth_observed"] = np.concatenate(d_obs)[..., ::S, ::S]
    rec["depth_rendered"] = np.concatenate(d_rend)[..., ::S, ::S]
    for k in ("depth_gt_observed", "depth_observed", "depth_rendered"):
        assert rec[k].dtype == np.float32
        rec[k + "_sum"] = np.concatenate({"depth_gt_observed": gt, "depth_observed": d_obs,
                                          "depth_rendered": d_rend}[k]).astype(np.float64).sum(
                                              axis=(-2, -1))
    np.savez_compressed(os.path.join(HERE, "dataio_reference.npz"), **rec)
    print("dataio golden vectors written: %.0f KB"
          % (os.path.getsize(os.path.join(HERE, "dataio_reference.npz")) / 1024))


if __name__ == "__main__":
    main()
