import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch
import helpers
from oracle import raster_np
from mx_deepim_b200 import synth
from mx_deepim_b200.render import Render_Py
K = helpers.K_LM
rng = np.random.default_rng(77)
meshes = [synth.make_mesh(rng, level=3, kind=0), synth.make_mesh(rng, level=4, kind=1), synth.make_mesh(rng, level=4, kind=2, texture_size=128)]
r = Render_Py(None, ["a", "b", "c"], K, meshes=meshes, max_batch=8)
rng = np.random.default_rng(5)
dev = torch.device("cuda:0")
for i in range(6):
    _, pt = synth.random_pose_pair(rng)
    m = i % 3
    out = r.render_batch(torch.tensor([m], dtype=torch.int32, device=dev), torch.from_numpy(pt[None]).to(dev), want=("depth", "bgr", "tri_id"))
    torch.cuda.synchronize()
    depth = out["depth"][0, 0].cpu().numpy(); tri = out["tri_id"][0].cpu().numpy()
    o_bgr, o_depth, o_tri = raster_np.render_py(meshes[m], pt[:, :3], pt[:, 3], K, r_type="mat", want_tri=True)
    both = (depth > 0) & (o_depth > 0)
    rel = np.zeros_like(depth); rel[both] = np.abs(depth[both] - o_depth[both]) / o_depth[both]
    bad = np.argwhere(rel > 1e-6)
    print(i, "mask agree", ((depth > 0) == (o_depth > 0)).mean(), "tri agree", (tri == o_tri).mean(), "n bad depth", len(bad), "max rel", rel.max())
    for (y, x) in bad[:5]:
        print("   px", y, x, "cuda z", depth[y, x], "tri", tri[y, x], "| oracle z", o_depth[y, x], "tri", o_tri[y, x])
