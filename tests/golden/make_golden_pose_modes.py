"""Generate tests/golden/pose_modes_reference.npz: the reference's own
lib/pair_matching/RT_transform.py (RT_transform, calc_RT_delta, calc_se3; imported unmodified,
numpy-2 shim of make_golden.py) for every rot_coord mode the batch updater can be configured
with (MODEL, CAMERA, CAMERA_NEW) and non-trivial translation means / stds, on float32 blobs as
the updater hands them over (batch_updater_py_multi.py:211-312).

    python tests/golden/make_golden_pose_modes.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, HERE)
T_MEANS = np.array([0.01, -0.02, 0.05])
T_STDS = np.array([0.5, 0.8, 1.5])


def main():
    import helpers
    from make_golden import import_reference

    RT, _ = import_reference()
    inp = helpers.make_pose_mode_inputs(51)
    rec = {}
    for mode in ("MODEL", "CAMERA", "CAMERA_NEW"):
        poses, rots, transs, KTs = [], [], [], []
        for b in range(inp["src_pose"].shape[0]):
            refined = RT.RT_transform(inp["src_pose"][b], inp["rot_est"][b], inp["trans_est"][b],
                                      T_MEANS, T_STDS, rot_coord=mode)
            q, t = RT.calc_RT_delta(refined, inp["tgt_pose"][b], T_MEANS, T_STDS, rot_coord=mode,
                                    rot_type="QUAT")
            se3 = np.zeros((3, 4))
            se3[:, :3], se3[:, 3] = RT.calc_se3(refined, inp["tgt_pose"][b])
            poses.append(refined); rots.append(q); transs.append(t)
            KTs.append(np.dot(helpers.K_LM, se3))
        rec[mode + "/pose"], rec[mode + "/rot"] = np.asarray(poses), np.asarray(rots)
        rec[mode + "/trans"], rec[mode + "/KT"] = np.asarray(transs), np.asarray(KTs)
    np.savez_compressed(os.path.join(HERE, "pose_modes_reference.npz"), **rec)
    print("pose-mode golden vectors written")


if __name__ == "__main__":
    main()
