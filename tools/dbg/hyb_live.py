import sys, os, numpy as np
ROOT=os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0,ROOT); sys.path.insert(0,os.path.join(ROOT,'tests'))
import refapi as ref, pathplanning_b200 as pp
from common import load_npz, c2_world_desc
np.set_printoptions(precision=17, linewidth=200)
g = load_npz("edges_c2.npz")
ctx=pp.Context(0); desc,occ,sq,pc=c2_world_desc(g); ctx.upload_map(desc,sq,occ,pc)
planner=pp.HybridAStar(ctx, pp.HybridParams(), pp.ValidatorParams(1.0,0.1))
w = ref.World(g["lb"], g["ub"], float(g["resolution"]))
for r in g["rects"]: w.add_rectangle(*r)
w.update(); w.set_validator_params(1.0, 0.1)
H = ref.Hybrid(w)
rng = np.random.default_rng(99)
starts, goals = [], []
while len(starts) < 200:
    p = np.stack([rng.uniform(-18.5, 18.5, 2), rng.uniform(-18.5, 18.5, 2), rng.uniform(-np.pi, np.pi, 2)], 1)
    if w.is_state_valid(p).all():
        starts.append(p[0]); goals.append(p[1])
starts, goals = np.array(starts), np.array(goals)
seeds = np.arange(200, dtype=np.uint64) + 5000
out = planner.search(starts, goals, seeds, max_path=256, max_log=8192)
for i in [7, 13, 34, 60, 76]:
    r = H.search(starts[i], goals[i], int(seeds[i]))
    res = out["results"][i]; log=out["log"][i]
    n=min(r["n_expanded"], int(res["n_expanded"]))
    d=np.nonzero((log["key"][:n]!=r["exp_keys"][:n,:3]).any(1) | (np.abs(log["pose"][:n]-r["exp_poses"][:n]).max(1)>1e-9))[0]
    print("query",i,"ref",r["status"],r["n_expanded"],r["cost"],"gpu",int(res["status"]),int(res["n_expanded"]),float(res["cost"]),"first diff",d[:1], "goal", goals[i])
    if len(d):
        k=d[0]
        print("  ref",r["exp_keys"][k],r["exp_poses"][k]); print("  gpu",log["key"][k],log["pose"][k])
        # which earlier expansion produced these? compare children of all earlier parents
        H.prepare_goal(goals[i])
        for j in range(max(0,k-400),k):
            par=r["exp_poses"][j]
            # no rng: use seed giving no RS shot unless h<10
            cnt,ch,ph=planner.expand(goals[i], par[None], np.array([1],np.uint64))
            e=H.expand(par,1)
            c=ch[0,:cnt[0]]
            if e["n"]!=cnt[0] or not np.array_equal(c["key"],e["keys"]) or np.abs(c["cost"]-e["costs"]).max()>1e-9 or np.abs(c["heuristic"]-e["heuristics"]).max()>1e-9:
                print("   parent",j,par,"ref n",e["n"],"gpu n",cnt[0])
                print("    ref keys",e["keys"].tolist(),"costs",e["costs"],"h",e["heuristics"],"kinds",e["kinds"])
                print("    gpu keys",c["key"].tolist(),"costs",c["cost"],"h",c["heuristic"],"kinds",c["kind"])
                break
