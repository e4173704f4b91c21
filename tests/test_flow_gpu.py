"""GT flow: sm_100a kernel vs the CPU restatement of flow_kernel (bit-exact), vs the
reference's own .cu compiled for sm_100a (bit-exact), edge cases, and full-size properties."""
import numpy as np
import pytest

import helpers
from oracle import flow_np

pytestmark = pytest.mark.gpu


def _run_host(inp):
    from mx_deepim_b200.flow import gpu_flow_wrapper

    return gpu_flow_wrapper(0)(inp["depth_src"], inp["depth_tgt"], inp["KT"], inp["Kinv"])


def _run_dev(inp, device):
    import torch

    from mx_deepim_b200.flow import gpu_flow_dev

    t = {k: torch.from_numpy(inp[k]).to(device) for k in ("depth_src", "depth_tgt", "KT", "Kinv")}
    flow, valid = gpu_flow_dev(t["depth_src"], t["depth_tgt"], t["KT"], t["Kinv"])
    torch.cuda.synchronize()
    return flow.cpu().numpy(), valid.cpu().numpy()


@pytest.mark.parametrize("batch,seed,occlude", [(1, 0, False), (4, 1, True), (8, 2, False)])
def test_flow_bit_exact_vs_oracle(cuda_device, batch, seed, occlude):
    inp = helpers.make_flow_inputs(batch, seed, occlude)
    ref_flow, ref_valid = flow_np.flow_kernel_cpu(inp["depth_src"], inp["depth_tgt"], inp["KT"],
                                                  inp["Kinv"])
    assert ref_valid.sum() > 500 * batch  # the case is not vacuous
    for flow, valid in (_run_host(inp), _run_dev(inp, cuda_device)):
        assert flow.dtype == np.float32 and flow.shape == (batch, 2, helpers.H, helpers.W)
        assert valid.shape == (batch, 1, helpers.H, helpers.W)
        assert np.array_equal(valid, ref_valid)           # validity mask: bit-exact
        assert np.array_equal(flow.view(np.uint32), ref_flow.view(np.uint32))  # flow: bit-exact


def test_flow_matches_reference_kernel(cuda_device):
    """Device-side pin: the reference gpu_flow_kernel.cu, compiled unmodified for sm_100a."""
    if not flow_np.reference_kernel_available():
        pytest.skip("oracle/_ref/libgpu_flow_ref.so not built (reference sources absent at build)")
    inp = helpers.make_flow_inputs(6, 7, occlude=True)
    ref_flow, ref_valid = flow_np.reference_kernel_gpu(inp["depth_src"], inp["depth_tgt"],
                                                       inp["KT"], inp["Kinv"])
    flow, valid = _run_dev(inp, cuda_device)
    assert np.array_equal(valid, ref_valid)
    assert np.array_equal(flow.view(np.uint32), ref_flow.view(np.uint32))
    # and the CPU restatement agrees with the reference kernel too
    o_flow, o_valid = flow_np.flow_kernel_cpu(inp["depth_src"], inp["depth_tgt"], inp["KT"],
                                              inp["Kinv"])
    assert np.array_equal(o_valid, ref_valid)
    assert np.array_equal(o_flow.view(np.uint32), ref_flow.view(np.uint32))


def test_flow_edge_cases(cuda_device):
    inp = helpers.make_flow_inputs(2, 3)
    # empty source depth -> all zero
    z = dict(inp, depth_src=np.zeros_like(inp["depth_src"]))
    flow, valid = _run_dev(z, cuda_device)
    assert not flow.any() and not valid.any()
    # identity relative pose and tgt == src: zero flow, valid on the whole foreground
    K = helpers.K_LM.astype(np.float64)
    KT = np.tile(np.concatenate([K, np.zeros((3, 1))], axis=1)[None], (2, 1, 1)).astype(np.float32)
    same = dict(inp, depth_tgt=inp["depth_src"].copy(), KT=KT)
    flow, valid = _run_dev(same, cuda_device)
    fg = inp["depth_src"] > 1e-3
    # border pixels can re-project a hair outside [0, W-1] x [0, H-1] (K * Kinv != I in float32)
    inner = (slice(None), slice(None), slice(1, -1), slice(1, -1))
    assert np.array_equal((valid > 0)[inner], fg[inner])
    assert not (valid > 0)[~fg].any()
    assert np.abs(flow).max() < 1e-3
    # ragged width (not a multiple of 4) goes through the host entry point's scalar kernel
    r = {k: v for k, v in inp.items()}
    r["depth_src"] = np.ascontiguousarray(inp["depth_src"][:, :, :, :637])
    r["depth_tgt"] = np.ascontiguousarray(inp["depth_tgt"][:, :, :, :637])
    ref_flow, ref_valid = flow_np.flow_kernel_cpu(r["depth_src"], r["depth_tgt"], r["KT"], r["Kinv"])
    flow, valid = _run_host(r)
    assert np.array_equal(valid, ref_valid) and np.array_equal(flow, ref_flow)


def test_flow_full_size_properties(cuda_device):
    """B = 64 at 640x480: checksum against the oracle and the structural property
    flow != 0 only where valid == 1."""
    inp = helpers.make_flow_inputs(4, 11)
    big = {k: (np.tile(v, (16,) + (1,) * (v.ndim - 1)) if k != "Kinv" else v) for k, v in inp.items()}
    flow, valid = _run_dev(big, cuda_device)
    assert set(np.unique(valid)) <= {0.0, 1.0}
    assert not flow[:, 0][valid[:, 0] == 0].any() and not flow[:, 1][valid[:, 0] == 0].any()
    ref_flow, ref_valid = flow_np.flow_kernel_cpu(inp["depth_src"], inp["depth_tgt"], inp["KT"],
                                                  inp["Kinv"])
    for rep in range(16):
        assert np.array_equal(valid[rep * 4:(rep + 1) * 4], ref_valid)
        assert np.array_equal(flow[rep * 4:(rep + 1) * 4], ref_flow)


def test_host_pointer_flow_is_reentrant(cuda_device):
    """The reference's `_flow` keeps no state between calls (it mallocs and frees per call), so
    engine threads may call it concurrently.  The drop-in shares one staging arena per device:
    four threads hammering gpu_flow with different batch sizes (the arena regrows) must each get
    exactly the oracle's answer."""
    import threading

    from mx_deepim_b200 import flow

    fn = flow.gpu_flow_wrapper(cuda_device.index or 0)
    cases = [helpers.make_flow_inputs(b, 40 + b, occlude=bool(b % 2)) for b in (1, 3, 2, 4)]
    refs = [flow_np.flow_kernel_cpu(c["depth_src"], c["depth_tgt"], c["KT"], c["Kinv"]) for c in cases]
    errors = []

    def worker(i):
        try:
            c = cases[i]
            for _ in range(6):
                f, v = fn(c["depth_src"], c["depth_tgt"], c["KT"], c["Kinv"])
                if not (np.array_equal(f, refs[i][0]) and np.array_equal(v, refs[i][1])):
                    errors.append("thread %d: wrong result" % i)
                    return
        except Exception as exc:  # noqa: BLE001
            errors.append("thread %d: %r" % (i, exc))

    threads = [threading.Thread(target=worker, args=(i,)) for i in range(4)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
