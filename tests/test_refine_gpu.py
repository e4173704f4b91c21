"""On-device refine loop vs the numpy restatement of batchUpdaterPyMulti.forward."""
import numpy as np
import pytest

import helpers
from oracle import transform3d_np, updater_np, zoom_np

pytestmark = pytest.mark.gpu


def test_refine_loop_matches_updater_oracle(cuda_device):
    import torch

    from mx_deepim_b200 import pipeline, synth

    rng = np.random.default_rng(11)
    meshes = [synth.make_mesh(rng, level=3, kind=0), synth.make_mesh(rng, level=3, kind=2)]
    B = 3
    loop = pipeline.RefineLoop(meshes, B, cuda_device, seed=5)
    host = pipeline.make_refine_batch(meshes, B, 123, renderer=loop.renderer, occlusion=True)
    loop.set_batch(host)
    pm = synth.PIXEL_MEANS[[2, 1, 0]]
    means, stds = np.zeros(3, np.float32), np.ones(3, np.float32)
    prev_pose = host["src_pose"].copy()
    losses = []
    for it in range(3):
        loop.iterate(it)
        torch.cuda.synchronize()
        rot_est = loop.rot_est.cpu().numpy()
        trans_est = loop.trans_est.cpu().numpy()
        ref = updater_np.updater_forward(meshes, host["class_index"], prev_pose, host["tgt_pose"],
                                         rot_est, trans_est, host["depth_gt_observed"], synth.K_LINEMOD,
                                         pm, means, stds)
        pose = loop.pose[loop.cur].cpu().numpy()
        np.testing.assert_allclose(pose, ref["src_pose"], rtol=2e-6, atol=2e-7)
        np.testing.assert_allclose(loop.rot_gt.cpu().numpy(), ref["rot"], atol=3e-6)
        np.testing.assert_allclose(loop.trans_gt.cpu().numpy(), ref["trans"], rtol=1e-5, atol=2e-6)
        depth = loop.depth_rendered.cpu().numpy()
        fg, rfg = depth > 0, ref["depth_rendered"] > 0
        assert (fg == rfg).mean() >= 0.999
        both = fg & rfg
        rel = np.abs(depth[both] - ref["depth_rendered"][both]) / ref["depth_rendered"][both]
        assert (rel <= 1e-5).mean() >= 0.999
        assert np.array_equal(loop.mask_rendered.cpu().numpy(), (depth > 0.2).astype(np.float32))
        img = loop.image_rendered.cpu().numpy()
        b3 = np.broadcast_to(both, img.shape)
        assert (np.abs(img[b3] - ref["image_rendered"][b3]) < 1e-3).mean() >= 0.995
        valid = loop.step.valid.cpu().numpy()
        assert (valid == ref["flow_weights"][:, :1]).mean() >= 0.999
        v2 = (valid > 0) & (ref["flow_weights"][:, :1] > 0)
        fl = loop.step.flow.cpu().numpy()
        for c in range(2):
            assert np.abs(fl[:, c:c + 1][v2] - ref["flow"][:, c:c + 1][v2]).max() < 2e-3
        # zoom + loss consistent with the oracles fed with the device's own intermediates
        zm = zoom_np.zoom_mask_forward(host["mask_observed"], host["mask_gt_observed"], depth, pose,
                                       synth.K_LINEMOD)
        assert np.array_equal(loop.step.zoom_factor.cpu().numpy(), zm[3])
        # the loss belongs to the pass that produced (rot_est, trans_est): it transforms the
        # model points with the pose BEFORE the update (deepIM_flownet.py:446-499 runs in the
        # network forward on src_pose_i; batch_updater_py_multi.py:211-312 refines afterwards)
        _, _, sums = transform3d_np.pm_loss_forward(
            host["point_cloud_model"], host["point_cloud_observed"], host["point_cloud_weights"],
            rot_est, trans_est, prev_pose, means, stds)
        np.testing.assert_allclose(loop.step.loss_sum.cpu().numpy(), sums, rtol=1e-4)
        losses.append(float(sums.sum()))
        prev_pose = pose
    # the damped ground-truth update converges: the poses approach the targets and the
    # point-matching loss of every pass is smaller than the one before
    err0 = np.abs(host["src_pose"] - host["tgt_pose"]).max()
    err3 = np.abs(prev_pose - host["tgt_pose"]).max()
    assert err3 < 0.5 * err0
    assert losses[0] > losses[1] > losses[2] > 0, losses
    loop.check_status()  # no tile-list overflow in any render of the loop


def test_refine_loop_graph_replay_equals_eager(cuda_device):
    """RefineLoop.capture: the 4-iteration loop as one CUDA graph (programmatic launch edges,
    incremental background included) replays bit-identically to the eager loop, repeatedly."""
    import torch

    from mx_deepim_b200 import pipeline, synth

    rng = np.random.default_rng(12)
    meshes = [synth.make_mesh(rng, level=3, kind=1)]
    B = 4
    loop = pipeline.RefineLoop(meshes, B, cuda_device, seed=6, backward=True)
    host = pipeline.make_refine_batch(meshes, B, 124, renderer=loop.renderer)
    loop.set_batch(host)

    def snap():
        torch.cuda.synchronize()
        return {"depth": loop.depth_rendered.clone(), "image": loop.image_rendered.clone(),
                "mask": loop.mask_rendered.clone(), "pose": loop.pose[loop.cur].clone(),
                "loss": loop.step.loss_sum.clone(), "zf": loop.step.zoom_factor.clone(),
                "zimg": loop.step.zoom_image_rendered.clone(), "flow": loop.step.zoom_flow.clone(),
                "grad": loop.step.grad_rot.clone()}

    loop.set_batch_poses(host)
    loop.run(4)
    eager = snap()
    assert float(eager["mask"].sum()) > 0
    s = torch.cuda.Stream(device=cuda_device)
    with torch.cuda.stream(s):
        loop.capture(host, iterations=4)
        for rep in range(3):
            loop.graph.replay()
            s.synchronize()
            got = snap()
            for k in eager:
                assert torch.equal(got[k], eager[k]), (rep, k)
        # an eager loop in between keeps the renderer's state in step
        loop.set_batch_poses(host)
        loop.run(4)
        loop.graph.replay()
        s.synchronize()
        got = snap()
        for k in eager:
            assert torch.equal(got[k], eager[k]), k


@pytest.mark.parametrize("iters", [3, 4])
def test_refine_loop_replay_is_safe(cuda_device, iters):
    """RefineLoop.replay(): an ODD iteration count (the renderer's ping-pong item lists do not
    return to the same half: two graphs, alternated) and a foreign render through the same
    renderer between replays (the captured pointers would read another frame's item list:
    re-capture) both give the eager loop's results bit for bit, and an eager loop after an odd
    number of replays still does."""
    import torch

    from mx_deepim_b200 import pipeline, synth

    rng = np.random.default_rng(21)
    meshes = [synth.make_mesh(rng, level=3, kind=1), synth.make_mesh(rng, level=3, kind=0)]
    B = 4
    loop = pipeline.RefineLoop(meshes, B, cuda_device, seed=9, backward=True)
    host = pipeline.make_refine_batch(meshes, B, 321, renderer=loop.renderer, occlusion=True)
    loop.set_batch(host)

    def snap():
        torch.cuda.synchronize()
        out = {"depth": loop.depth_rendered.clone(), "image": loop.image_rendered.clone(),
               "mask": loop.mask_rendered.clone(), "pose": loop.pose[loop.cur].clone()}
        for k, v in loop.step.outputs().items():
            if v is not None:
                out[k] = v.clone()
        return out

    def same(a, b, tag):
        for k in a:
            x, y = a[k], b[k]
            if x.dtype == torch.float32:
                x, y = x.view(torch.int32), y.view(torch.int32)
            assert torch.equal(x, y), (tag, k)

    loop.set_batch_poses(host)
    loop.run(iters)
    eager = snap()
    assert float(eager["mask"].sum()) > 0
    s = torch.cuda.Stream(device=cuda_device)
    with torch.cuda.stream(s):
        loop.capture(host, iterations=iters)
        assert len(loop.graphs) == (1 if iters % 2 == 0 else 2)
        for rep in range(5):
            loop.replay()
            same(eager, snap(), ("replay", rep))
        # eager loop right after an odd number of replays: host and device parity must agree
        loop.set_batch_poses(host)
        loop.run(iters)
        same(eager, snap(), "eager-after-replays")
        loop.replay()  # the eager renders moved renderer.serial: this re-captures
        same(eager, snap(), "replay-after-eager")
        # a second user of the renderer between replays (another pose, other output buffers)
        other = loop.renderer.render_batch(loop.cls, loop.obs["tgt_pose"], want=("depth",))
        assert float(other["depth"].sum()) > 0
        loop.replay()
        same(eager, snap(), "replay-after-foreign-render")
        loop.replay()
        same(eager, snap(), "replay-after-foreign-render-2")
    loop.check_status()
    # a NEW batch: set_batch copies into the loop's persistent tensors, so the captured graphs
    # (raw pointers) stay valid and replay the new batch
    host2 = pipeline.make_refine_batch(meshes, B, 654, renderer=loop.renderer)
    loop.set_batch(host2)
    loop.run(iters)
    eager2 = snap()
    assert not torch.equal(eager2["pose"], eager["pose"])
    with torch.cuda.stream(s):
        loop.replay()
        same(eager2, snap(), "replay-new-batch")
        loop.replay()
        same(eager2, snap(), "replay-new-batch-2")
    fresh = pipeline.RefineLoop(meshes, B, cuda_device, seed=9)
    with pytest.raises(RuntimeError):
        fresh.replay()
    with pytest.raises(RuntimeError):
        fresh.set_batch_poses()


def test_sparse_frame_step_equals_dense_step(cuda_device):
    """rc_compare_step_args.frame_renderer (csrc/frame.cu): crop box + GT flow taken from the
    renderer's covered tiles, stale tiles re-zeroed -- every output of every iteration is
    bit-identical to the dense step's, including across loop restarts (the object jumps back to
    its initial pose: the largest set of stale tiles) and after a test-phase iteration."""
    import torch

    from mx_deepim_b200 import pipeline, synth

    rng = np.random.default_rng(13)
    meshes = [synth.make_mesh(rng, level=3, kind=0), synth.make_mesh(rng, level=2, kind=1)]
    B = 5
    loops = [pipeline.RefineLoop(meshes, B, cuda_device, seed=7, backward=True, sparse=sp)
             for sp in (False, True)]
    host = pipeline.make_refine_batch(meshes, B, 125, renderer=loops[0].renderer, occlusion=True)
    for lp in loops:
        lp.set_batch(host)
    keys = ("flow", "flow_weights_valid", "zoom_factor", "bbox", "zoom_mask_rendered",
            "zoom_mask_gt_observed", "zoom_image_rendered", "zoom_flow", "zoom_flow_weights",
            "zoom_trans_gt", "point_matching_loss_sum", "grad_rot")

    def compare(tag):
        torch.cuda.synchronize()
        od, os_ = loops[0].step.outputs(), loops[1].step.outputs()
        for k in keys:
            x, y = od[k], os_[k]
            if x.dtype == torch.float32:  # bit patterns: -0.0 (round(0 - 0.45)) is not +0.0
                x, y = x.view(torch.int32), y.view(torch.int32)
            assert torch.equal(x, y), (tag, k)
        assert torch.equal(loops[0].depth_rendered, loops[1].depth_rendered), tag

    for rep in range(2):
        for lp in loops:
            lp.set_batch_poses(host)
        for it in range(4):
            for lp in loops:
                lp.iterate(it)
            compare((rep, it))
    assert float(loops[1].step.valid.sum()) > 0
    # the resampler really skipped background tiles of the rendered-side planes
    assert int(loops[1].step.zoom_tile_state.sum()) > 0
    # a test-phase iteration renders without a flow pass: the next training step must not trust
    # the flow planes' history
    q = torch.tensor([[1.0, 0.02, -0.01, 0.03]] * B, device=cuda_device)
    t = torch.tensor([[0.01, -0.02, 0.05]] * B, device=cuda_device)
    for lp in loops:
        lp.iterate_test(q, t)
    for it in range(2):
        for lp in loops:
            lp.iterate(it)
        compare(("after_test", it))
    # with event marks (serialised issue order)
    from mx_deepim_b200 import _lib

    for lp in loops:
        marks = [_lib.Event() for _ in range(_lib.RC_STEP_MARKS)]
        lp.iterate(2, marks=marks)
    compare("marks")


def test_refine_loop_side_stream_equals_single_stream(cuda_device):
    """The loss on a second stream beside the rasteriser (and the ZoomTrans backward beside the
    resampler) changes the schedule, not one bit of any output."""
    import torch

    from mx_deepim_b200 import pipeline, synth

    rng = np.random.default_rng(14)
    meshes = [synth.make_mesh(rng, level=3, kind=1), synth.make_mesh(rng, level=2, kind=0)]
    B = 5
    outs = []
    for side in (False, True):
        loop = pipeline.RefineLoop(meshes, B, cuda_device, seed=8, backward=True, side_stream=side)
        host = pipeline.make_refine_batch(meshes, B, 126, renderer=loop.renderer)
        loop.set_batch(host)
        loop.run(3)
        torch.cuda.synchronize()
        loop.check_status()
        outs.append({"depth": loop.depth_rendered.clone(), "pose": loop.pose[loop.cur].clone(),
                     "loss": loop.step.loss_sum.clone(), "zf": loop.step.zoom_factor.clone(),
                     "zimg": loop.step.zoom_image_rendered.clone(),
                     "zflow": loop.step.zoom_flow.clone(), "g_rot": loop.step.grad_rot.clone(),
                     "g_zt": loop.step.grad_zoom_trans.clone(), "rot_est": loop.rot_est.clone()})
    for k in outs[0]:
        assert torch.equal(outs[0][k], outs[1][k]), k


@pytest.mark.parametrize("policy", ["upload", "box_gt"])
def test_loop_feed_from_storage_formats_equals_float_blobs(cuda_device, policy):
    """pipeline.LoopFeed: batches in the dataset's storage formats (u8 BGR, u16 mm depth, u8
    label / mask) from page-locked host memory, decoded on the device, one graph replay per
    loop, refined poses + loss sums back to the host -- equal to the same loop fed with the
    float32 blobs (set_batch).  policy "box_gt": mask_observed is NOT uploaded but derived on the
    device from the label mask (TRAIN.INIT_MASK box_gt, lib/utils/image.py:330-343)."""
    import torch

    from mx_deepim_b200 import pipeline, synth

    rng = np.random.default_rng(15)
    meshes = [synth.make_mesh(rng, level=3, kind=1), synth.make_mesh(rng, level=2, kind=0)]
    B, iters = 4, 4
    mp = "box_gt" if policy == "box_gt" else "hole"
    loop = pipeline.RefineLoop(meshes, B, cuda_device, seed=3, backward=True)
    hosts = [pipeline.make_refine_batch(meshes, B, 500 + i, renderer=loop.renderer, mask_policy=mp,
                                        occlusion=(i == 1)) for i in range(3)]
    want = []
    for h in hosts:
        # the storage formats hold the image rounded to 8 bits: round the float blob the same way
        # so that both paths see the same observed image
        pm = synth.PIXEL_MEANS[[2, 1, 0]].reshape(1, 3, 1, 1)
        h["image_observed"] = (np.clip(np.rint(h["image_observed"].astype(np.float64) + pm), 0, 255)
                               - pm).astype(np.float32)
        loop.set_batch(h)
        loop.run(iters)
        torch.cuda.synchronize()
        want.append({"pose": loop.pose[loop.cur].cpu().numpy().copy(),
                     "loss": loop.step.loss_sum.cpu().numpy().copy(),
                     "mask_observed": loop.obs["mask_observed"].cpu().numpy().copy(),
                     "zimg": loop.step.zoom_image_observed.cpu().numpy().copy(),
                     "zmask": loop.step.zoom_mask_observed.cpu().numpy().copy()})
    packed = [pipeline.pack_refine_batch(h, mask_policy=policy) for h in hosts]
    assert ("mask_observed" in packed[0]) == (policy == "upload")
    feed = pipeline.LoopFeed(loop, packed[0], iterations=iters)
    if policy == "box_gt":
        assert feed.h2d_bytes == pipeline.LoopFeed(loop, pipeline.pack_refine_batch(hosts[0]),
                                                   iterations=iters).h2d_bytes - B * synth.H * synth.W
    s = torch.cuda.Stream(device=cuda_device)
    with torch.cuda.stream(s):
        for rnd in range(2):  # twice through the two staging slots
            for i, pk in enumerate(packed):
                slot = feed.submit(pk)
                feed.done[slot].synchronize()
                np.testing.assert_array_equal(feed.pose_out[slot].numpy(), want[i]["pose"])
                np.testing.assert_array_equal(feed.loss_out[slot].numpy(), want[i]["loss"])
                s.synchronize()
                assert np.array_equal(loop.obs["mask_observed"].cpu().numpy(), want[i]["mask_observed"])
                assert np.array_equal(loop.step.zoom_mask_observed.cpu().numpy(), want[i]["zmask"])
                assert np.array_equal(loop.step.zoom_image_observed.cpu().numpy(), want[i]["zimg"])
    loop.check_status()
