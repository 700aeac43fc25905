import sys, os, numpy as np
ROOT=os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0,ROOT); sys.path.insert(0,os.path.join(ROOT,'tests'))
import refapi as ref, pathplanning_b200 as pp
from common import load_npz, c2_world_desc
np.set_printoptions(precision=17, linewidth=200)
g = load_npz("edges_c2.npz")
ctx=pp.Context(0); desc,occ,sq,pc=c2_world_desc(g); ctx.upload_map(desc,sq,occ,pc)
w = ref.World(g["lb"], g["ub"], float(g["resolution"]))
for r in g["rects"]: w.add_rectangle(*r)
w.update(); w.set_validator_params(1.0, 0.1)
lb, ub = tuple(g["lb"]), tuple(g["ub"])
start, goal = (-18.0, -18.0, 0.0), (17.0, 17.5, np.pi / 2)
seeds = np.array([5, 6], np.uint64)
out = ctx.rrt_run(np.tile(start, (2, 1)), np.tile(goal, (2, 1)), seeds, lb, ub, dim=3, connection=1, star=1, validator=1, max_iteration=4000, max_nodes=1500, max_connection_distance=1.0, goal_bias=0.05, min_turning_radius=2.0, vp=pp.ValidatorParams(1.0, 0.1))
for i in range(2):
    r = ref.rrtstar_se2_restated(w, lb, ub, start, goal, int(seeds[i]), 4000, 1500, 1.0, 0.05, 1, 2.0)
    n=r["n_nodes"]; same=out["parents"][i,:n]==r["parents"]
    print("tree",i,"n",n,"mismatch",int((~same).sum()))
    for j in np.nonzero(~same)[0][:8]:
        print(" node",j,"gpu parent",out["parents"][i,j],"cost",out["costs"][i,j],"ref parent",r["parents"][j],"cost",r["costs"][j], "pos", r["nodes"][j])
        for pidx in (out["parents"][i,j], r["parents"][j]):
            print("    cand",pidx,"gpucost",out["costs"][i,pidx],"refcost",r["costs"][pidx],"pos",r["nodes"][pidx])
print("---- RS solves between tree-1 nodes")
r = ref.rrtstar_se2_restated(w, lb, ub, start, goal, 6, 4000, 1500, 1.0, 0.05, 1, 2.0)
N=r["nodes"]; G=out["nodes"][1]
pairs=[(1,5),(0,5),(2,6),(0,6),(7,22),(1,22)]
S=np.array([N[a] for a,b in pairs]); T=np.array([N[b] for a,b in pairs])
gs=ctx.rs_solve(S,T,pp.RSCostParams(2.0)); rs=ref.rs_optimal_batch(S,T,2.0)
for k,(a,b) in enumerate(pairs):
    print(a,b,"gpu word",gs["word"][k],"len",gs["length"][k],"ref word",rs["word"][k],"len",rs["length"][k])
S=np.array([G[a] for a,b in pairs]); T=np.array([G[b] for a,b in pairs])
gs=ctx.rs_solve(S,T,pp.RSCostParams(2.0)); rs=ref.rs_optimal_batch(S,T,2.0)
print("with GPU node coordinates")
for k,(a,b) in enumerate(pairs):
    print(a,b,"gpu word",gs["word"][k],"len",gs["length"][k],"ref word",rs["word"][k],"len",rs["length"][k], G[a]-N[a], G[b]-N[b])
