"""Generate tests/golden/render_reference.{npz,json} from the REFERENCE'S OWN renderer modules
(/root/reference/lib/render_glumpy/*.py, imported unmodified).

    python tests/golden/make_golden_render.py

glumpy / OpenGL cannot run in this container or on the GPU box, so nothing is drawn here:
`import glumpy` is served by an empty stand-in module (attribute access only) and only the
parts of the renderer classes that are plain Python / numpy are executed:

  * my_compute_calib_proj(K, w, h, zNear, zFar) and _get_view_mtx(R, t) of every class -- the
    projection (with the +0.5 principal-point shift) and view matrices the GL programs receive
    (called unbound with self=None; they do not touch self);
  * the constructor and render() SIGNATURES (parameter names and defaults) of every class: the
    drop-in boundary of SURVEY.md 8(b), compared with mx_deepim_b200/render_glumpy/*.

K is given as float64 (numpy-1.x evaluation, see oracle/raster_np.py) and as float32 (what
numpy 2.x makes of `K[0, 2] + 0.5` on a float32 K: stays float32); both are stored.
"""
import importlib.util
import inspect
import json
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
REF = "/root/reference"
MODULES = {
    "render_py": "Render_Py", "render_py_depth": "Render_Py_depth",
    "render_py_multi": "Render_Py", "render_py_light": "Render_Py_Light",
    "render_py_light_multi_program": "Render_Py_Light_MultiProgram",
    "render_py_light_modelnet": "Render_Py_Light_ModelNet",
    "render_py_light_modelnet_multi": "Render_Py_Light_ModelNet_Multi",
}


class _Anything(types.ModuleType):
    """Module whose every attribute is another inert object (glumpy.app, gl.GL_FLOAT, ...)."""

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return _Anything(self.__name__ + "." + name)

    def __call__(self, *a, **k):
        return None


def install_glumpy_stub():
    if "glumpy" not in sys.modules:
        sys.modules["glumpy"] = _Anything("glumpy")


def load(name):
    if not hasattr(np, "float"):
        np.float = float  # noqa: RT_transform.py:247
    if not hasattr(np, "maximum_sctype"):
        np.maximum_sctype = lambda t: np.longdouble  # noqa: RT_transform.py:246
    if REF not in sys.path:
        sys.path.insert(0, REF)
    install_glumpy_stub()
    spec = importlib.util.spec_from_file_location(
        "_ref_" + name, os.path.join(REF, "lib", "render_glumpy", name + ".py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def signature(fn):
    out = []
    for p in list(inspect.signature(fn).parameters.values())[1:]:  # drop self
        d = None if p.default is inspect.Parameter.empty else repr(p.default)
        out.append([p.name, d])
    return out


def main():
    import helpers

    rng = np.random.default_rng(123)
    poses = [helpers.random_pose_pair(rng)[0] for _ in range(4)]
    rec, meta = {"poses": np.asarray(poses)}, {}
    Ks = {"lm": helpers.K_LM, "other": np.array([[1066.778, 0, 312.9869], [0, 1067.487, 241.3109],
                                                  [0, 0, 1]], np.float32)}
    for name, cls_name in MODULES.items():
        cls = getattr(load(name), cls_name)
        meta[name] = {"class": cls_name, "init": signature(cls.__init__),
                      "render": signature(cls.render)}
        for kn, K in Ks.items():
            for dt in (np.float64, np.float32):
                for (w, h) in ((640, 480), (320, 200)):
                    key = "%s/proj_%s_%s_%dx%d" % (name, kn, np.dtype(dt).name, w, h)
                    rec[key] = np.asarray(cls.my_compute_calib_proj(None, K.astype(dt), w, h, 0.25, 6.0))
        rec[name + "/view"] = np.asarray(
            [cls._get_view_mtx(None, p[:, :3], p[:, 3]) for p in poses])
    np.savez_compressed(os.path.join(HERE, "render_reference.npz"), **rec)
    with open(os.path.join(HERE, "render_reference.json"), "w") as f:
        json.dump(meta, f, indent=1, sort_keys=True)
    print("render golden vectors written: %d arrays" % len(rec))


if __name__ == "__main__":
    main()
