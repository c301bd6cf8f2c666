import sys, numpy as np
sys.path.insert(0, '/root/repo')
from uavmvs_b200 import api, scenes
import oracle
sc = scenes.config1()
obvh = oracle.BVH.build(sc.verts, sc.faces, threads=8)
a = obvh.export()
g = api.BVHTree.create(a.nodes, a.aabbs, a.tris, a.indices)
cloud = api.PointCloud.create(sc.points, sc.normals)
obs = api.VectorArray.create(sc.num_points, 32)
rc, wr = api.Array.create(sc.num_points), api.Array.create(sc.num_points)
api.evaluate_trajectory(sc.views28(), sc.max_distance, sc.target_recon, sc.width, sc.height, g, cloud, obs, rc, wr)
oo = oracle.ObsRays(sc.num_points, 32)
oracle.update_observation_rays(obvh, True, sc.views28(), sc.max_distance, sc.width, sc.height, sc.points, sc.normals, oo)
oracle.process_observation_rays(oo)
ro = oracle.evaluate_observation_rays(oo).astype(np.float64)
rg = rc.download().astype(np.float64)
err = np.abs(rg - ro) / np.maximum(np.maximum(np.abs(rg), np.abs(ro)), 1e-30)
k = np.argsort(err)[-8:]
print("max rel err", err.max(), "n>1e-5", (err > 1e-5).sum(), "n>3e-6", (err > 3e-6).sum())
for i in k: print(i, rg[i], ro[i], err[i], oo.num_rows[i])
