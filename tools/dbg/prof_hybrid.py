import os, sys, numpy as np
ROOT=os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0,ROOT); sys.path.insert(0,os.path.join(ROOT,'tools'))
import pathplanning_b200 as pp
from pathplanning_b200 import synth
from time_hybrid import c2_queries
n=int(sys.argv[1]) if len(sys.argv)>1 else 2048
ctx=pp.Context(0); cfg=synth.C2
occ=synth.rasterize_rectangles(synth.config_rects(cfg),cfg["lb"],cfg["ub"],cfg["resolution"]); sq=synth.exact_sqdist(occ)
rows,cols,origin=synth.map_geometry(cfg["lb"],cfg["ub"],cfg["resolution"])
ctx.upload_map(pp.MapDesc(rows=rows,cols=cols,resolution=cfg["resolution"],world_origin=origin,lb=cfg["lb"],ub=cfg["ub"]),sq,occ,None)
pl=pp.HybridAStar(ctx,pp.HybridParams(),pp.ValidatorParams(1.0,0.1))
s,g=c2_queries(ctx,n); seeds=(np.arange(n)+1000).astype(np.uint64)
out=pl.search(s,g,seeds,max_path=128)
print(out["results"]["n_expanded"].sum())
