"""Host-side logic of the refine-loop feed (no GPU): the synthetic batch's observed-mask policies
and the storage-format packing LoopFeed uploads."""
import numpy as np
import pytest


def _render_fn(meshes):
    from oracle import raster_np

    def fn(cls, poses, pm):
        from mx_deepim_b200 import synth

        B = len(cls)
        depth = np.zeros((B, 1, synth.H, synth.W), np.float32)
        image = np.zeros((B, 3, synth.H, synth.W), np.float32)
        for b in range(B):
            bgr, d = raster_np.render_py(meshes[cls[b]], poses[b, :, :3], poses[b, :, 3],
                                         synth.K_LINEMOD, r_type="mat")
            depth[b, 0] = d
            image[b] = bgr[:, :, [2, 1, 0]].transpose(2, 0, 1) - np.asarray(pm).reshape(3, 1, 1)
        return depth, image

    return fn


@pytest.mark.parametrize("policy", ["hole", "box_gt"])
def test_refine_batch_mask_policies_and_packing(policy):
    from mx_deepim_b200 import pipeline, synth

    rng = np.random.default_rng(3)
    meshes = [synth.make_mesh(rng, level=2, kind=1)]
    B = 2
    host = pipeline.make_refine_batch(meshes, B, 11, render_fn=_render_fn(meshes), mask_policy=policy)
    gt, obs = host["mask_gt_observed"], host["mask_observed"]
    assert gt.sum() > 100
    for b in range(B):
        ys, xs = np.nonzero(gt[b, 0])
        if policy == "box_gt":
            # TRAIN.INIT_MASK box_gt (lib/utils/image.py:330-343): the label mask's bounding
            # rectangle with the reference's exclusive slice end
            want = np.zeros_like(gt[b, 0])
            want[ys.min():ys.max(), xs.min():xs.max()] = 1
            assert np.array_equal(obs[b, 0], want)
        else:
            assert (obs[b, 0] <= gt[b, 0]).all() and obs[b].sum() <= gt[b].sum()
    packed = pipeline.pack_refine_batch(host, pin=False, mask_policy="box_gt" if policy == "box_gt" else "upload")
    assert ("mask_observed" in packed) == (policy != "box_gt")
    assert packed["image_observed"].dtype.itemsize == 1 and tuple(packed["image_observed"].shape) == (B, synth.H, synth.W, 3)
    assert packed["depth_gt_observed"].dtype.itemsize == 2
    # depth survives the 16-bit millimetre round trip exactly (make_refine_batch rounds to 1 mm)
    back = packed["depth_gt_observed"].numpy().view(np.uint16).astype(np.float64) / 1000.0
    assert np.array_equal(back.astype(np.float32), host["depth_gt_observed"][:, 0])
    assert np.array_equal(packed["mask_gt_observed"].numpy(), gt[:, 0].astype(np.uint8))
    with pytest.raises(ValueError):
        pipeline.pack_refine_batch(host, pin=False, mask_policy="nope")
    with pytest.raises(ValueError):
        pipeline.make_refine_batch(meshes, 1, 1, render_fn=_render_fn(meshes), mask_policy="nope")
