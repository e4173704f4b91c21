"""The C-ABI library loads and exports every symbol include/rendercompare*.h declares
(no compute calls: runs without a GPU)."""
import ctypes
import glob
import os
import re

from conftest import ROOT


def _declared_functions():
    names = []
    for path in sorted(glob.glob(os.path.join(ROOT, "include", "*.h"))):
        text = open(path).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        text = re.sub(r"//[^\n]*", "", text)
        text = re.sub(r"typedef\s+struct\s+\w+\s*\{.*?\}\s*\w+\s*;", "", text, flags=re.S)
        for m in re.finditer(r"\b([A-Za-z_]\w*)\s*\(([^;{]*?)\)\s*;", text, flags=re.S):
            names.append(m.group(1))
    return sorted(set(names))


def test_header_symbols_exported():
    from mx_deepim_b200 import _lib

    lib = ctypes.CDLL(_lib.LIB_PATH)
    declared = _declared_functions()
    assert "rc_flow_dev" in declared and "_flow" in declared and len(declared) >= 15
    missing = [n for n in declared if not hasattr(lib, n)]
    assert not missing, "declared in include/ but not exported: %s" % missing


def test_binding_covers_header():
    from mx_deepim_b200 import _lib

    declared = set(_declared_functions())
    bound = set(_lib.SIGNATURES)
    assert declared == bound, (sorted(declared - bound), sorted(bound - declared))
    lib = _lib.load()
    assert lib.rc_version() == 1
    assert lib.rc_status_string(0) == b"ok"
    assert lib.rc_status_string(3).startswith(b"pointer")


def test_zoom_group_struct_layout():
    from mx_deepim_b200 import _lib

    # two pointers + four int32 + four floats, no padding surprises
    assert ctypes.sizeof(_lib.rc_zoom_group) == 8 + 8 + 16 + 16
    assert _lib.rc_zoom_group.mean.offset == 32


def test_plugins_registered():
    from mx_deepim_b200 import operator, operator_py  # noqa: F401

    assert operator.get_all_registered_operators() == [
        "Transform3D", "ZoomDepth", "ZoomFlow", "ZoomImage", "ZoomImageWithFactor", "ZoomMask",
        "ZoomMaskWithFactor", "ZoomTrans"]
    # argument / output lists as in the reference Props (zoom_mask.py:151-163 ...)
    p = operator.get_prop_class("ZoomMask")(K="[572.4114 0 325.2611 0 573.57043 242.04899 0 0 1]")
    assert p.list_arguments() == ["mask_observed", "mask_gt_observed", "mask_rendered", "src_pose"]
    assert p.list_outputs()[-1] == "zoom_factor"
    assert p.K.dtype.name == "float32" and p.K.shape == (3, 3)
    _, out_shape, _ = p.infer_shape([[2, 1, 480, 640]] * 3 + [[2, 3, 4]])
    assert out_shape == [[2, 1, 480, 640]] * 3 + [[2, 4]]
    f = operator.get_prop_class("ZoomFlow")(b_inv_zoom="True")
    assert f.list_arguments() == ["zoom_factor", "flow"] and f.list_outputs() == ["zoom_flow"]
    t = operator.get_prop_class("Transform3D")(T_means="[0. 0. 0.]", T_stds="[1. 1. 1.]",
                                               rot_coord="CAMERA")
    assert t.list_arguments() == ["point_cloud", "rotation", "translation", "pose_src"]


def test_product_does_not_import_oracle():
    """The product package must never route through the CPU oracle."""
    pkg = os.path.join(ROOT, "mx-deepim_b200")
    offenders = []
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(dirpath, f)).read()
                if re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M):
                    offenders.append(os.path.join(dirpath, f))
    assert not offenders, offenders


def test_gpu_flow_argument_checks():
    """Cython-stub behaviour (gpu_flow.pyx:24-28): wrong dtype / rank raise before any CUDA call."""
    import numpy as np
    import pytest

    from mx_deepim_b200.flow import gpu_flow

    d = np.zeros((1, 1, 8, 8), np.float32)
    with pytest.raises(ValueError):
        gpu_flow(d.astype(np.float64), d, np.zeros((1, 3, 4), np.float32), np.eye(3, dtype=np.float32))
    with pytest.raises(ValueError):
        gpu_flow(d[0], d, np.zeros((1, 3, 4), np.float32), np.eye(3, dtype=np.float32))
    # batch 0 short-circuits without touching the device (gpu_flow.pyx:36)
    e = np.zeros((0, 1, 8, 8), np.float32)
    flow, valid = gpu_flow(e, e, np.zeros((0, 3, 4), np.float32), np.eye(3, dtype=np.float32))
    assert flow.shape == (0, 2, 8, 8) and valid.shape == (0, 1, 8, 8)


def test_struct_layouts_match_header(tmp_path):
    """sizeof / offsetof of every argument struct as gcc sees include/rendercompare.h against
    the ctypes mirrors in _lib.py: a field added on one side only fails here, not on the GPU."""
    import subprocess

    from mx_deepim_b200 import _lib

    structs = {
        "rc_zoom_group": (_lib.rc_zoom_group, ["src", "dst", "channels", "flags", "mean"]),
        "rc_compare_step_args": (_lib.rc_compare_step_args,
                                 ["image_observed", "trans_gt", "flow", "loss_sum", "grad_rot",
                                  "grad_zoom_trans", "pixel_means", "T_stds", "grad_scale",
                                  "rot_coord", "zoom_grad", "batch_size", "num_points", "stream",
                                  "marks"]),
        "rc_mesh_desc": (_lib.rc_mesh_desc, ["vertices", "texture", "n_vertices", "tex_w"]),
        "rc_render_args": (_lib.rc_render_args, ["cls_idx", "status", "K", "pixel_means", "znear",
                                                 "mode", "batch_size", "incremental", "stream", "marks"]),
    }
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "rendercompare.h"',
             'int main(void) {']
    for name, (_, fields) in structs.items():
        lines.append('printf("%s %%zu\\n", sizeof(%s));' % (name, name))
        for f in fields:
            lines.append('printf("%s.%s %%zu\\n", offsetof(%s, %s));' % (name, f, name, f))
    lines += ['return 0;', '}']
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    out = dict(l.split() for l in subprocess.check_output([str(exe)], text=True).splitlines())
    for name, (cls, fields) in structs.items():
        assert ctypes.sizeof(cls) == int(out[name]), name
        for f in fields:
            assert getattr(cls, f).offset == int(out["%s.%s" % (name, f)]), (name, f)
