"""Device pose update vs the numpy restatement of the batch-updater pose block."""
import numpy as np
import pytest

import helpers
from oracle import updater_np

pytestmark = pytest.mark.gpu


def test_pose_update(cuda_device):
    import torch

    from mx_deepim_b200 import pipeline

    rng = np.random.default_rng(31)
    B = 32
    src = np.zeros((B, 3, 4), np.float32)
    tgt = np.zeros((B, 3, 4), np.float32)
    for b in range(B):
        src[b], tgt[b] = helpers.random_pose_pair(rng)
    rot = rng.normal(size=(B, 4)).astype(np.float32)
    rot[:, 0] = np.abs(rot[:, 0]) + 2.0  # near-identity, unnormalised (network output)
    trans = rng.normal(0, 0.05, size=(B, 3)).astype(np.float32)
    means, stds = np.zeros(3, np.float32), np.ones(3, np.float32)
    ref_pose, ref_q, ref_t, ref_KT = updater_np.pose_update(src, rot, trans, tgt, helpers.K_LM, means, stds)
    t = lambda a: torch.from_numpy(a).to(cuda_device)
    out = pipeline.pose_update(t(src), t(rot), t(trans), t(tgt), helpers.K_LM)
    torch.cuda.synchronize()
    pose, q, tr, KT = [o.cpu().numpy() for o in out]
    np.testing.assert_allclose(pose, ref_pose, rtol=2e-6, atol=2e-7)
    np.testing.assert_allclose(q, ref_q, rtol=0, atol=2e-6)
    np.testing.assert_allclose(tr, ref_t, rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(KT, ref_KT, rtol=2e-6, atol=2e-4)
    # applying the labels to the refined pose recovers the target pose (consistency property)
    from oracle import pose_np

    for b in range(4):
        rec = pose_np.RT_transform(pose[b].astype(np.float64), q[b], tr[b], means, stds, "CAMERA")
        assert np.abs(rec - tgt[b]).max() < 2e-5


def test_pose_update_non_orthonormal_target_takes_the_jacobi_path(cuda_device):
    """mat2quat of a matrix far from a rotation: the shifted power iteration's guards reject it and
    the Jacobi eigen-solve answers, still matching numpy's eigh (RT_transform.py:446-523)."""
    import torch

    from mx_deepim_b200 import pipeline

    rng = np.random.default_rng(32)
    B = 16
    src = np.zeros((B, 3, 4), np.float32)
    tgt = np.zeros((B, 3, 4), np.float32)
    for b in range(B):
        src[b], tgt[b] = helpers.random_pose_pair(rng)
        if b % 2 == 0:  # sheared, scaled "rotation" blocks
            tgt[b, :, :3] = (tgt[b, :, :3] @ (np.eye(3) + rng.normal(0, 0.4, size=(3, 3)))).astype(np.float32)
    rot = np.tile(np.array([[1, 0, 0, 0]], np.float32), (B, 1))
    trans = np.zeros((B, 3), np.float32)
    means, stds = np.zeros(3, np.float32), np.ones(3, np.float32)
    _, ref_q, ref_t, _ = updater_np.pose_update(src, rot, trans, tgt, helpers.K_LM, means, stds)
    t = lambda a: torch.from_numpy(a).to(cuda_device)
    out = pipeline.pose_update(t(src), t(rot), t(trans), t(tgt), helpers.K_LM)
    torch.cuda.synchronize()
    q = out[1].cpu().numpy()
    np.testing.assert_allclose(q, ref_q, rtol=0, atol=5e-6)
    np.testing.assert_allclose(out[2].cpu().numpy(), ref_t, rtol=1e-5, atol=1e-6)
