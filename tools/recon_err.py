import numpy as np, sys
sys.path.insert(0,'/root/repo')
import oracle
from uavmvs_b200 import api, scenes
sc = scenes.dense()
ob = oracle.BVH.build(sc.verts, sc.faces); ta = ob.export()
g = api.BVHTree.create(ta.nodes, ta.aabbs, ta.tris, ta.indices)
cloud = api.PointCloud.create(sc.points, sc.normals); obs = api.VectorArray.create(sc.num_points, 32)
r, w = api.Array.create(sc.num_points), api.Array.create(sc.num_points)
api.evaluate_trajectory(sc.views28(), sc.max_distance, sc.target_recon, sc.width, sc.height, g, cloud, obs, r, w)
o = oracle.ObsRays(sc.num_points, 32)
oracle.update_observation_rays(ob, True, sc.views28(), sc.max_distance, sc.width, sc.height, sc.points, sc.normals, o)
oracle.process_observation_rays(o); ro = oracle.evaluate_observation_rays(o)
rg = r.download(); m = ro > 0
rel = np.abs(rg[m]-ro[m])/ro[m]
print('max rel err recon', rel.max(), 'p99', np.percentile(rel, 99), 'n', m.sum(), 'min recon', ro[m].min())
