"""The fused compare step (rc_compare_step_dev) against the CPU oracles: forward, the training
backward leg (cfg 5), the overlapped and the serialised issue orders, and CUDA-graph replay."""
import numpy as np
import pytest

from oracle import flow_np, transform3d_np, zoom_np

pytestmark = pytest.mark.gpu

MEANS, STDS = np.zeros(3, np.float32), np.ones(3, np.float32)
LW_PM = 0.1


def _oracle(host, synth):
    ref = {}
    ref["flow"], ref["valid"] = flow_np.flow_kernel_cpu(host["depth_rendered"], host["depth_gt_observed"],
                                                        host["KT"], host["Kinv"])
    zm = zoom_np.zoom_mask_forward(host["mask_observed"], host["mask_gt_observed"],
                                   host["depth_rendered"], host["src_pose"], synth.K_LINEMOD)
    ref["zm"] = zm
    ref["zi"] = zoom_np.zoom_image_with_factor_forward(zm[3], host["image_observed"],
                                                       host["image_rendered"], synth.PIXEL_MEANS[::-1])
    ref["zflow"], ref["zw"] = zoom_np.zoom_flow_forward(zm[3], ref["flow"],
                                                        np.tile(ref["valid"], [1, 2, 1, 1]))
    args = [host[k] for k in ("point_cloud_model", "point_cloud_observed", "point_cloud_weights",
                              "rot_est", "trans_est", "src_pose")]
    ref["loss_map"], _, ref["loss_sum"] = transform3d_np.pm_loss_forward(*args, MEANS, STDS)
    gs = LW_PM / host["point_cloud_model"].shape[2]
    ref["g_rot"], ref["g_tr"] = transform3d_np.pm_loss_backward(*args, MEANS, STDS, gs)
    return ref


def _check(out, ref, backward, zoom_grad=False):
    assert np.array_equal(out["flow_weights_valid"], ref["valid"])
    assert np.array_equal(out["flow"], ref["flow"])
    zm = ref["zm"]
    assert np.array_equal(out["zoom_factor"], zm[3])
    assert np.array_equal(out["bbox"].astype(np.int64), zm[4])
    for k, i in (("zoom_mask_observed", 0), ("zoom_mask_gt_observed", 1), ("zoom_mask_rendered", 2)):
        assert (out[k] != zm[i]).mean() < 1e-5, k
    assert np.abs(out["zoom_image_observed"] - ref["zi"][0]).max() <= 1e-4
    assert np.abs(out["zoom_image_rendered"] - ref["zi"][1]).max() <= 1e-4
    assert np.abs(out["zoom_flow"] - ref["zflow"]).max() <= 4e-4  # 1e-4 px / wx
    assert (out["zoom_flow_weights"] != ref["zw"]).mean() < 1e-5
    np.testing.assert_allclose(out["point_matching_loss"], ref["loss_map"], rtol=1e-4, atol=1e-4)
    np.testing.assert_allclose(out["point_matching_loss_sum"], ref["loss_sum"], rtol=1e-4)
    if backward:
        np.testing.assert_allclose(out["grad_rot"], ref["g_rot"], rtol=5e-4, atol=1e-6)
        np.testing.assert_allclose(out["grad_trans"], ref["g_tr"], rtol=5e-4, atol=1e-6)
        # invZoomTrans.backward (zoom_trans.py:50-76): identity unless b_zoom_grad, then x,y * wx
        g = out["grad_trans"].copy()
        if zoom_grad:
            g[:, 0] *= zm[3][:, 0]
            g[:, 1] *= zm[3][:, 0]
        assert np.array_equal(out["grad_zoom_trans"], g)
    else:
        assert out["grad_rot"] is None


@pytest.mark.parametrize("backward,zoom_grad,serial", [(False, False, False), (True, False, False),
                                                       (True, True, True)])
def test_compare_step_matches_oracles(cuda_device, backward, zoom_grad, serial):
    import torch

    from mx_deepim_b200 import _lib, pipeline, synth

    B = 3
    host = synth.make_compare_batch(B, seed=77, occlusion=True)
    step = pipeline.CompareStep(B, cuda_device, backward=backward, lw_pm=LW_PM, zoom_grad=zoom_grad)
    x = step.to_device(host, non_blocking=False)
    marks = [_lib.Event() for _ in range(_lib.RC_STEP_MARKS)] if serial else None
    step.run(x, marks=marks)
    torch.cuda.synchronize()
    out = {k: (v.cpu().numpy() if v is not None else None) for k, v in step.outputs().items()}
    _check(out, _oracle(host, synth), backward, zoom_grad)
    if serial:
        assert marks[0].elapsed_ms(marks[5]) > 0.0


def test_compare_step_graph_replay_is_deterministic(cuda_device):
    """Captured into a CUDA graph, the step replays bit-identically (no atomics on floats, the
    bbox workspace cleans itself) and still matches the oracle after 5 replays."""
    import torch

    from mx_deepim_b200 import pipeline, synth

    B = 2
    host = synth.make_compare_batch(B, seed=78)
    step = pipeline.CompareStep(B, cuda_device, backward=True, lw_pm=LW_PM)
    x = step.to_device(host, non_blocking=False)
    s = torch.cuda.Stream(device=cuda_device)
    with torch.cuda.stream(s):
        g = step.capture(x)
        first = None
        for _ in range(5):
            g.replay()
            s.synchronize()
            cur = {k: v.clone() for k, v in step.outputs().items() if v is not None}
            if first is None:
                first = cur
            else:
                for k in cur:
                    assert torch.equal(cur[k], first[k]), k
    out = {k: (v.cpu().numpy() if v is not None else None) for k, v in step.outputs().items()}
    _check(out, _oracle(host, synth), True)


@pytest.mark.parametrize("packed", [False, True])
def test_host_feed_overlapped_uploads_keep_batches_apart(cuda_device, packed):
    """pipeline.HostFeed: six different batches submitted back to back through two staging sets
    (upload of batch i+1 on the copy stream while batch i's step runs): every batch's loss sums
    and zoom factors equal the oracle's on that batch's (decoded) inputs."""
    import torch

    from mx_deepim_b200 import pipeline, synth

    B, n = 2, 6
    hosts = [synth.make_compare_batch(B, seed=300 + i, occlusion=bool(i % 2)) for i in range(n)]
    step = pipeline.CompareStep(B, cuda_device)
    means_bgr = np.asarray(synth.PIXEL_MEANS, np.float64).reshape(-1)
    feeds = []
    for h in hosts:
        d = {k: torch.from_numpy(np.ascontiguousarray(h[k])).pin_memory() for k in pipeline.INPUT_KEYS}
        if packed:
            for k in ("image_observed", "image_rendered"):
                rgb = np.clip(np.rint(h[k].astype(np.float64) + means_bgr[::-1].reshape(1, 3, 1, 1)),
                              0, 255).astype(np.uint8)
                d[k] = torch.from_numpy(np.ascontiguousarray(rgb[:, ::-1].transpose(0, 2, 3, 1))).pin_memory()
            for k in ("depth_gt_observed", "depth_rendered"):
                mm = np.rint(h[k][:, 0].astype(np.float64) * 1000.0).astype(np.uint16)
                d[k] = torch.from_numpy(np.ascontiguousarray(mm.view(np.int16))).pin_memory()
                h[k] = (mm.astype(np.float32) / np.float32(1000.0))[:, None]  # what the step sees
            for k in ("mask_observed", "mask_gt_observed"):
                d[k] = torch.from_numpy(np.ascontiguousarray(h[k][:, 0].astype(np.uint8))).pin_memory()
        feeds.append(d)
    feed = pipeline.HostFeed(step, feeds[0], packed=packed, pixel_means_bgr=means_bgr)
    assert feed.h2d_bytes == sum(v.numel() * v.element_size() for v in feeds[0].values())
    snaps = []
    for d in feeds:
        feed.submit(d)
        # stream-ordered snapshots: taken after this batch's step, before the next one's
        snaps.append((step.loss_sum.clone(), step.zoom_factor.clone(), step.valid.clone()))
    torch.cuda.synchronize()
    for i, h in enumerate(hosts):
        zm = zoom_np.zoom_mask_forward(h["mask_observed"], h["mask_gt_observed"], h["depth_rendered"],
                                       h["src_pose"], synth.K_LINEMOD)
        assert np.array_equal(snaps[i][1].cpu().numpy(), zm[3]), i
        args = [h[k] for k in ("point_cloud_model", "point_cloud_observed", "point_cloud_weights",
                               "rot_est", "trans_est", "src_pose")]
        _, _, loss_sum = transform3d_np.pm_loss_forward(*args, MEANS, STDS)
        np.testing.assert_allclose(snaps[i][0].cpu().numpy(), loss_sum, rtol=1e-4)
        _, valid = flow_np.flow_kernel_cpu(h["depth_rendered"], h["depth_gt_observed"], h["KT"], h["Kinv"])
        assert np.array_equal(snaps[i][2].cpu().numpy(), valid), i
    # the pinned read-backs hold the last two batches
    for i in (n - 2, n - 1):
        assert torch.equal(feed.loss[i % 2], snaps[i][0].cpu())
        assert torch.equal(feed.zoom_factor[i % 2], snaps[i][1].cpu())


@pytest.mark.parametrize("B,backward", [(16, True), (64, False)])
def test_compare_step_full_size_batch_permutation_property(cuda_device, B, backward):
    """Size-independent property at BASELINE.json's batch sizes (configs[1] = 16 pairs, the
    refine loop's 64): pairs are independent, so permuting the batch must permute EVERY output
    bit for bit (no cross-pair leakage through shared workspaces, tickets, clusters or tile
    lists), and a pair's outputs must not depend on its position in the batch."""
    import torch

    from mx_deepim_b200 import pipeline, synth

    host = synth.make_compare_batch(B, seed=500 + B, occlusion=True)
    perm = np.random.default_rng(B).permutation(B)
    host_p = {k: (np.ascontiguousarray(v[perm]) if isinstance(v, np.ndarray) and v.ndim >= 1
                  and v.shape[0] == B else v) for k, v in host.items()}
    outs = []
    for h in (host, host_p):
        step = pipeline.CompareStep(B, cuda_device, backward=backward, lw_pm=LW_PM)
        x = step.to_device(h, non_blocking=False)
        step.run(x)
        torch.cuda.synchronize()
        outs.append({k: v.cpu().numpy() for k, v in step.outputs().items() if v is not None})
        del step, x
    assert set(outs[0]) == set(outs[1]) and len(outs[0]) >= 10
    for k, a in outs[0].items():
        b = outs[1][k]
        assert a.shape[0] == B, k
        assert np.array_equal(a[perm].view(np.uint8), b.view(np.uint8)), k
    assert outs[0]["flow_weights_valid"].sum() > 1000 * B  # the batch is not degenerate


def test_compare_step_input_depth_planes(cuda_device):
    """cfg.network.INPUT_DEPTH (deepIM_flownet.py:669-681): ZoomDepth's two planes ride in the
    step's one resampler launch (rc_compare_step_args.depth_observed / zoom_depth_*) and equal
    the ZoomDepth oracle; every other output is unchanged."""
    import torch

    from mx_deepim_b200 import pipeline, synth

    B = 3
    host = synth.make_compare_batch(B, seed=81, occlusion=True)
    step = pipeline.CompareStep(B, cuda_device, input_depth=True)
    x = step.to_device(host, non_blocking=False)
    x["depth_observed"] = x["depth_gt_observed"]
    step.run(x)
    torch.cuda.synchronize()
    out = {k: (v.cpu().numpy() if v is not None else None) for k, v in step.outputs().items()}
    ref = _oracle(host, synth)
    _check(out, ref, False)
    zo, zr = zoom_np.zoom_depth_forward(ref["zm"][3], host["depth_gt_observed"], host["depth_rendered"])
    assert np.abs(out["zoom_depth_observed"] - zo).max() <= 1e-4
    assert np.abs(out["zoom_depth_rendered"] - zr).max() <= 1e-4


def test_compare_step_phases_equal_one_call(cuda_device):
    """rc_compare_step_args.phases: the loss on a second stream, frame -> sample on the first,
    the ZoomTrans backward behind both -- bit-identical to the single call."""
    import torch

    from mx_deepim_b200 import _lib, pipeline, synth

    B = 4
    host = synth.make_compare_batch(B, seed=82)
    one = pipeline.CompareStep(B, cuda_device, backward=True, lw_pm=LW_PM, zoom_grad=True)
    x = one.to_device(host, non_blocking=False)
    one.run(x)
    torch.cuda.synchronize()
    want = {k: v.clone() for k, v in one.outputs().items() if v is not None}
    two = pipeline.CompareStep(B, cuda_device, backward=True, lw_pm=LW_PM, zoom_grad=True)
    a = two.bind(x)
    main, side = torch.cuda.current_stream(), torch.cuda.Stream(device=cuda_device)
    fork, zf, join = torch.cuda.Event(), torch.cuda.Event(), torch.cuda.Event()
    fork.record(main)
    side.wait_event(fork)
    a.phases = _lib.RC_STEP_PHASE_LOSS
    two.run(None, args=a, stream=side)
    a.phases = _lib.RC_STEP_PHASE_FRAME
    two.run(None, args=a)
    zf.record(main)
    a.phases = _lib.RC_STEP_PHASE_SAMPLE
    two.run(None, args=a)
    side.wait_event(zf)
    a.phases = _lib.RC_STEP_PHASE_ZT_BWD
    two.run(None, args=a, stream=side)
    join.record(side)
    main.wait_event(join)
    torch.cuda.synchronize()
    for k, v in two.outputs().items():
        if v is not None:
            assert torch.equal(v, want[k]), k
    a.phases = 16
    with pytest.raises(Exception):
        two.run(None, args=a)
