"""CUDA triangle rasteriser behind the reference's renderer classes.

Drop-in surfaces (same constructor / ``render`` signatures and return values):

  Render_Py(model_dir, classes, K, width, height, zNear, zFar)
      .render(cls_idx, r, t, r_type="quat", K=None) -> (bgr [H,W,3] f32 0..255, depth [H,W] f32 m)
      /root/reference/lib/render_glumpy/render_py_multi.py:49-147
  Render_Py_Light_ModelNet_Multi(model_path_list, texture_path, K, width, height, zNear, zFar,
                                 brightness_ratios=[0.7])
      .render(model_idx, r, t, light_position, light_intensity, brightness_k=0, r_type="quat")
          -> (bgr [H,W,3] u8, depth [H,W] f32)
      /root/reference/lib/render_glumpy/render_py_light_modelnet_multi.py:83-231

plus the batched device fast path ``render_batch`` (torch CUDA tensors in and out, the
post-render block of batch_updater_py_multi.py:279-320 fused in) that the on-device refine
loop uses.  The OpenGL window, glumpy programs and the two glReadPixels per image
(render_py_multi.py:90-106) are replaced by four kernel launches per batch (raster.cu).
"""
from __future__ import annotations

import ctypes
import os

import numpy as np
import torch

from . import _lib
from . import pose as P
from .operator_py import _common as C


# ------------------------------------------------------------------- mesh IO
def load_obj(path, rescale=False):
    """Minimal OBJ reader (v / vt / vn / f, polygons fan-triangulated).  Vertices are
    de-duplicated per (v, vt, vn) index triple, like an indexed GL vertex buffer.
    `rescale=True` centres the mesh and scales its largest extent to 1
    (glumpy.data.objload(rescale=True) as used at render_py_light_modelnet_multi.py:110)."""
    v, vt, vn, faces = [], [], [], []
    with open(path) as fh:
        for line in fh:
            s = line.split()
            if not s:
                continue
            if s[0] == "v":
                v.append([float(x) for x in s[1:4]])
            elif s[0] == "vt":
                vt.append([float(x) for x in s[1:3]])
            elif s[0] == "vn":
                vn.append([float(x) for x in s[1:4]])
            elif s[0] == "f":
                idx = []
                for tok in s[1:]:
                    parts = (tok.split("/") + ["", ""])[:3]
                    idx.append(tuple(int(p) if p else 0 for p in parts))
                for k in range(1, len(idx) - 1):
                    faces.append((idx[0], idx[k], idx[k + 1]))
    table, verts, uvs, nrms, tris = {}, [], [], [], []
    for f in faces:
        tri = []
        for key in f:
            if key not in table:
                table[key] = len(verts)
                vi, ti, ni = key
                verts.append(v[vi - 1 if vi > 0 else vi])
                uvs.append(vt[ti - 1 if ti > 0 else ti] if ti else [0.0, 0.0])
                nrms.append(vn[ni - 1 if ni > 0 else ni] if ni else [0.0, 0.0, 0.0])
            tri.append(table[key])
        tris.append(tri)
    verts = np.asarray(verts, np.float32)
    if rescale:
        lo, hi = verts.min(0), verts.max(0)
        verts = (verts - (lo + hi) / 2) / (hi - lo).max()
    mesh = {"vertices": verts, "uv": np.asarray(uvs, np.float32) if vt else None,
            "normals": np.asarray(nrms, np.float32) if vn else None,
            "triangles": np.asarray(tris, np.int32)}
    if mesh["normals"] is None:
        from .synth import vertex_normals

        mesh["normals"] = vertex_normals(verts, mesh["triangles"]).astype(np.float32)
    return mesh


def load_texture(path):
    """RGB uint8 [H,W,3], row 0 = top of the file (the flip on upload, render_py_multi.py:75-77,
    is applied inside the kernel's texel addressing)."""
    import cv2

    img = cv2.imread(path, cv2.IMREAD_COLOR)
    if img is None:
        raise IOError("cannot read texture %s" % path)
    return np.ascontiguousarray(img[:, :, ::-1])


# ------------------------------------------------------------ device renderer
class DeviceRenderer(object):
    """Owns the rc_renderer handle: meshes + workspaces resident on one GPU."""

    def __init__(self, meshes, K, width=640, height=480, zNear=0.25, zFar=6.0, max_batch=64,
                 device=None):
        if not torch.cuda.is_available():
            raise _lib.RenderCompareError("the rasteriser needs a CUDA device (no CPU fallback)")
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None \
            else torch.device(device)
        self.width, self.height, self.zNear, self.zFar = width, height, zNear, zFar
        self.K = np.asarray(K, np.float32).reshape(3, 3)
        self.max_batch = int(max_batch)
        self.n_meshes = len(meshes)
        self.n_triangles = [int(np.asarray(m["triangles"]).shape[0]) for m in meshes]
        descs = (_lib.rc_mesh_desc * len(meshes))()
        keep = []
        for d, m in zip(descs, meshes):
            v = np.ascontiguousarray(m["vertices"], np.float32)
            t = np.ascontiguousarray(m["triangles"], np.int32)
            tex = m.get("texture")
            tex = np.full((1, 1, 3), 255, np.uint8) if tex is None else np.ascontiguousarray(tex, np.uint8)
            uv = np.ascontiguousarray(m["uv"], np.float32) if m.get("uv") is not None else None
            nr = np.ascontiguousarray(m["normals"], np.float32) if m.get("normals") is not None else None
            keep += [v, t, tex, uv, nr]
            d.vertices, d.triangles, d.texture = v.ctypes.data, t.ctypes.data, tex.ctypes.data
            d.uv = uv.ctypes.data if uv is not None else None
            d.normals = nr.ctypes.data if nr is not None else None
            d.n_vertices, d.n_triangles = v.shape[0], t.shape[0]
            d.tex_h, d.tex_w = tex.shape[0], tex.shape[1]
        self.handle = ctypes.c_void_p()
        lib = _lib.load()
        st = lib.rc_renderer_create(descs, len(meshes), height, width, self.max_batch,
                                    self.device.index or 0, ctypes.byref(self.handle))
        _lib.check(st, "rc_renderer_create")
        self.status = torch.zeros((1,), dtype=torch.int32, device=self.device)
        # host-issued rc_render_dev calls on this handle (CUDA-graph replays do not count): lets
        # an owner of captured graphs notice that somebody else rendered in between
        self.serial = 0

    def __del__(self):
        try:
            if getattr(self, "handle", None):
                _lib.load().rc_renderer_destroy(self.handle)
                self.handle = None
        except Exception:
            pass

    def render_batch(self, cls_idx, poses, K=None, mode=_lib.RC_RENDER_TEXTURED, light_pos=None,
                     light_int=None, brightness_ratio=0.7, pixel_means=(0, 0, 0), mask_thresh=0.2,
                     want=("depth", "image", "mask"), out=None, stream=None, marks=None,
                     incremental=False):
        """cls_idx int32 [B], poses float32 [B,3,4] (device).  Returns a dict of device tensors:
        depth [B,1,H,W]; image [B,3,H,W] = RGB - pixel_means; mask [B,1,H,W]; bgr [B,H,W,3];
        tri_id [B,H,W].  incremental=True: the tensors in `out` still hold this renderer's
        previous frame untouched (a refine loop that owns its buffers), so only the tiles that
        frame covered are re-cleared (rc_render_args.incremental)."""
        B = int(poses.shape[0])
        H, W = self.height, self.width
        out = {} if out is None else out
        f = dict(dtype=torch.float32, device=self.device)
        shapes = {"depth": (B, 1, H, W), "image": (B, 3, H, W), "mask": (B, 1, H, W),
                  "bgr": (B, H, W, 3)}
        for k in set(want) | {"depth"}:
            if k == "tri_id":
                if k not in out:
                    out[k] = torch.empty((B, H, W), dtype=torch.int32, device=self.device)
            elif k not in out:
                out[k] = torch.empty(shapes[k], **f)
        C.check_tensor("cls_idx", cls_idx, dtype=torch.int32, shape=(B,))
        C.check_tensor("poses", poses, shape=(B, 3, 4))
        for k, t in (("light_pos", light_pos), ("light_int", light_int)):
            if t is not None:
                C.check_tensor(k, t, shape=(B, 3))
        for k in set(want) | {"depth"}:
            C.check_tensor(k, out[k], dtype=torch.int32 if k == "tri_id" else torch.float32,
                           shape=(B, H, W) if k == "tri_id" else shapes[k])
        a = _lib.rc_render_args()
        a.cls_idx, a.poses = cls_idx.data_ptr(), poses.data_ptr()
        a.light_pos = light_pos.data_ptr() if light_pos is not None else None
        a.light_int = light_int.data_ptr() if light_int is not None else None
        a.depth = out["depth"].data_ptr()
        a.image_chw = out["image"].data_ptr() if "image" in want else None
        a.bgr_hwc = out["bgr"].data_ptr() if "bgr" in want else None
        a.mask = out["mask"].data_ptr() if "mask" in want else None
        a.tri_id = out["tri_id"].data_ptr() if "tri_id" in want else None
        a.status = self.status.data_ptr()
        Kf = (self.K if K is None else np.asarray(K, np.float32)).reshape(-1)
        for i in range(9):
            a.K[i] = float(Kf[i])
        for i in range(3):
            a.pixel_means[i] = float(pixel_means[i])
        a.znear, a.zfar, a.mask_thresh = self.zNear, self.zFar, mask_thresh
        a.brightness_ratio, a.mode, a.batch_size = brightness_ratio, mode, B
        a.incremental = 1 if incremental else 0
        a.stream = _lib.stream_ptr(stream)
        if marks is not None:
            arr = (ctypes.c_void_p * _lib.RC_RENDER_MARKS)()
            for i, m in enumerate(marks):
                arr[i] = m.handle if m is not None else None
            a.marks = ctypes.cast(arr, ctypes.POINTER(ctypes.c_void_p))
        with torch.cuda.device(self.device):
            st = _lib.load().rc_render_dev(self.handle, ctypes.byref(a))
        self.serial += 1
        _lib.check(st, "rc_render_dev")
        return out

    def note_replay(self, n_renders):
        """A CUDA graph holding `n_renders` captured render_batch calls was just replayed
        (rc_renderer_note_replay keeps the handle's ping-pong parity in step)."""
        _lib.check(_lib.load().rc_renderer_note_replay(self.handle, int(n_renders)),
                   "rc_renderer_note_replay")

    def check_status(self, reset=True):
        """Raise if any render since the last check overflowed its tile lists (raster_scan_kernel
        sets the device flag; the fill / cover passes then clamp and DROP triangles, so depth,
        mask and everything derived from them would be wrong).  Synchronises the device: call it
        where the caller synchronises anyway (Render_Py.render does), at the end of an
        asynchronous loop (RefineLoop.check_status) or once per bench."""
        flag = int(self.status.item())
        if flag:
            if reset:
                self.status.zero_()
            raise _lib.RenderCompareError(
                "rc_render_dev: tile-list capacity exceeded (RC_ERR_CAPACITY): a pair's triangles "
                "cover more tile entries than 8 * max_triangles + 16 * tiles; the frame is incomplete")

    def stats(self, batch):
        out = (ctypes.c_int64 * 3)()
        with torch.cuda.device(self.device):
            _lib.check(_lib.load().rc_render_stats(self.handle, batch, _lib.stream_ptr(), out),
                       "rc_render_stats")
        return {"records": int(out[0]), "tile_entries": int(out[1]), "direct": int(out[2])}

    def _one(self, idx, R, t, K, **kw):
        cls = torch.tensor([int(idx)], dtype=torch.int32, device=self.device)
        pose = np.zeros((1, 3, 4), np.float32)
        pose[0, :, :3], pose[0, :, 3] = R, np.asarray(t).reshape(3)
        o = self.render_batch(cls, torch.from_numpy(pose).to(self.device), K=K,
                              want=("depth", "bgr"), **kw)
        self.check_status()  # synchronises; a dropped triangle must not reach the caller silently
        return o["bgr"][0].cpu().numpy(), o["depth"][0, 0].cpu().numpy()


class Render_Py(object):
    """Multi-class textured renderer (reference render_py_multi.py:16-178)."""

    def __init__(self, model_dir, classes, K, width=640, height=480, zNear=0.25, zFar=6.0,
                 meshes=None, max_batch=64, device=None):
        self.width, self.height, self.zNear, self.zFar = width, height, zNear, zFar
        self.K = K
        self.model_dir = model_dir
        self.classes = classes
        if meshes is None:
            meshes = []
            for cur_class in classes:
                folder = os.path.join(model_dir, cur_class)
                m = load_obj("{}/textured.obj".format(folder), rescale=False)
                m["texture"] = load_texture("{}/texture_map.png".format(folder))
                meshes.append(m)
        self.dev = DeviceRenderer(meshes, K, width, height, zNear, zFar, max_batch, device)

    def render(self, cls_idx, r, t, r_type="quat", K=None):
        R = P.quat2mat(r) if r_type == "quat" else np.asarray(r)
        return self.dev._one(cls_idx, R, t, K)

    def render_batch(self, cls_idx, poses, **kw):
        return self.dev.render_batch(cls_idx, poses, **kw)


class Render_Py_Light_ModelNet_Multi(object):
    """Multi-model Lambert-lit renderer (reference render_py_light_modelnet_multi.py:82-262)."""

    def __init__(self, model_path_list, texture_path, K, width=640, height=480, zNear=0.25,
                 zFar=6.0, brightness_ratios=[0.7], meshes=None, max_batch=64, device=None):
        self.width, self.height, self.zNear, self.zFar = width, height, zNear, zFar
        self.K = K
        self.model_path_list = model_path_list
        self.brightness_ratios = list(brightness_ratios)
        if meshes is None:
            tex = load_texture(texture_path)
            meshes = []
            for path in model_path_list:
                m = load_obj(path, rescale=True)
                m["vertices"] = m["vertices"] / 10.0  # :111
                m["texture"] = tex
                meshes.append(m)
        self.dev = DeviceRenderer(meshes, K, width, height, zNear, zFar, max_batch, device)

    def render(self, model_idx, r, t, light_position, light_intensity, brightness_k=0,
               r_type="quat"):
        R = P.quat2mat(r) if r_type == "quat" else np.asarray(r)
        d = self.dev.device
        lp = torch.tensor(np.asarray(light_position, np.float32).reshape(1, 3), device=d)
        li = torch.tensor(np.asarray(light_intensity, np.float32).reshape(1, 3), device=d)
        bgr, depth = self.dev._one(model_idx, R, t, None, mode=_lib.RC_RENDER_LAMBERT,
                                   light_pos=lp, light_int=li,
                                   brightness_ratio=self.brightness_ratios[brightness_k])
        return bgr.astype(np.uint8), depth

    def render_batch(self, cls_idx, poses, **kw):
        kw.setdefault("mode", _lib.RC_RENDER_LAMBERT)
        return self.dev.render_batch(cls_idx, poses, **kw)


# ------------------------------------------------ the remaining lib/render_glumpy classes
def _load_linemod_model(model_folder):
    m = load_obj("{}/textured.obj".format(model_folder), rescale=False)
    m["texture"] = load_texture("{}/texture_map.png".format(model_folder))
    return m


class Render_Py_Single(object):
    """Single-model textured renderer (reference render_py.py:48-128):
    Render_Py(model_folder, K, ...).render(r, t, r_type="quat", K=None)
        -> (bgr [H,W,3] float32 0..255, depth [H,W] float32 m)."""

    def __init__(self, model_folder, K, width=640, height=480, zNear=0.25, zFar=6.0, mesh=None,
                 max_batch=64, device=None):
        self.width, self.height, self.zNear, self.zFar = width, height, zNear, zFar
        self.K = K
        self.model_folder = model_folder
        mesh = _load_linemod_model(model_folder) if mesh is None else mesh
        self.dev = DeviceRenderer([mesh], K, width, height, zNear, zFar, max_batch, device)

    def render(self, r, t, r_type="quat", K=None):
        R = P.quat2mat(r) if r_type == "quat" else np.asarray(r)
        return self.dev._one(0, R, t, K)


class Render_Py_Depth(Render_Py_Single):
    """render_py_depth.Render_Py (render_py_depth.py:48-128): render(r, t, r_type="quat")
    -> (bgr uint8 = np.round(rgb * 255), depth float32)."""

    def render(self, r, t, r_type="quat"):
        bgr, depth = Render_Py_Single.render(self, r, t, r_type)
        return np.round(bgr).astype(np.uint8), depth


class Render_Py_Light(object):
    """LINEMOD renderer with a point light (reference render_py_light.py:80-213):
    Render_Py_Light(model_folder, K, ..., brightness_ratios=[0.4, 0.3, 0.2])
        .render(r, t, light_position, light_intensity, brightness_k=0, r_type="quat", K=None)
        -> (bgr uint8, depth float32).  Fragment colour t * ((1-b) + b * brightness * I)."""

    def __init__(self, model_folder, K, width=640, height=480, zNear=0.25, zFar=6.0,
                 brightness_ratios=[0.4, 0.3, 0.2], mesh=None, max_batch=64, device=None):
        self.width, self.height, self.zNear, self.zFar = width, height, zNear, zFar
        self.K = K
        self.model_folder = model_folder
        self.brightness_ratios = list(brightness_ratios)
        self.brightness_k = 0
        mesh = _load_linemod_model(model_folder) if mesh is None else mesh
        self.dev = DeviceRenderer([mesh], K, width, height, zNear, zFar, max_batch, device)

    def _lit(self, idx, r, t, light_position, light_intensity, brightness_k, r_type, K, mode):
        R = P.quat2mat(r) if r_type == "quat" else np.asarray(r)
        d = self.dev.device
        lp = torch.tensor(np.asarray(light_position, np.float32).reshape(1, 3), device=d)
        li = torch.tensor(np.asarray(light_intensity, np.float32).reshape(1, 3), device=d)
        self.brightness_k = brightness_k
        bgr, depth = self.dev._one(idx, R, t, K, mode=mode, light_pos=lp, light_int=li,
                                   brightness_ratio=self.brightness_ratios[brightness_k])
        return bgr.astype(np.uint8), depth

    def render(self, r, t, light_position, light_intensity, brightness_k=0, r_type="quat", K=None):
        return self._lit(0, r, t, light_position, light_intensity, brightness_k, r_type, K,
                         _lib.RC_RENDER_LIT_LINEMOD)


class Render_Py_Light_MultiProgram(Render_Py_Light):
    """Multi-class LINEMOD light renderer (reference render_py_light_multi_program.py:84-240):
    Render_Py_Light_MultiProgram(class_name_list, model_folder_dict, K, ...)
        .render(r, t, light_position, light_intensity, class_name, brightness_k=0, r_type="quat")."""

    def __init__(self, class_name_list, model_folder_dict, K, width=640, height=480, zNear=0.25,
                 zFar=6.0, brightness_ratios=[0.4, 0.3, 0.2], meshes=None, max_batch=64,
                 device=None):
        self.width, self.height, self.zNear, self.zFar = width, height, zNear, zFar
        self.K = K
        self.model_folder_dict = model_folder_dict
        self.class_name_list = list(class_name_list)
        self.brightness_ratios = list(brightness_ratios)
        self.brightness_k = 0
        self.class_name = self.class_name_list[-1]
        if meshes is None:
            meshes = [_load_linemod_model(model_folder_dict[c]) for c in self.class_name_list]
        self.dev = DeviceRenderer(meshes, K, width, height, zNear, zFar, max_batch, device)

    def render(self, r, t, light_position, light_intensity, class_name, brightness_k=0,
               r_type="quat"):
        self.class_name = class_name
        return self._lit(self.class_name_list.index(class_name), r, t, light_position,
                         light_intensity, brightness_k, r_type, None, _lib.RC_RENDER_LIT_LINEMOD)


class Render_Py_Light_ModelNet(Render_Py_Light):
    """Single-model ModelNet renderer (reference render_py_light_modelnet.py:82-211):
    Render_Py_Light_ModelNet(model_path, texture_path, K, ..., brightness_ratios=[0.7])
        .render(r, t, light_position, light_intensity, brightness_k=0, r_type="quat")."""

    def __init__(self, model_path, texture_path, K, width=640, height=480, zNear=0.25, zFar=6.0,
                 brightness_ratios=[0.7], mesh=None, max_batch=64, device=None):
        self.width, self.height, self.zNear, self.zFar = width, height, zNear, zFar
        self.K = K
        self.model_path = model_path
        self.brightness_ratios = list(brightness_ratios)
        self.brightness_k = 0
        if mesh is None:
            mesh = load_obj(model_path, rescale=True)
            mesh["vertices"] = mesh["vertices"] / 10.0  # :103
            mesh["texture"] = load_texture(texture_path)
        self.dev = DeviceRenderer([mesh], K, width, height, zNear, zFar, max_batch, device)

    def render(self, r, t, light_position, light_intensity, brightness_k=0, r_type="quat"):
        return self._lit(0, r, t, light_position, light_intensity, brightness_k, r_type, None,
                         _lib.RC_RENDER_LAMBERT)
