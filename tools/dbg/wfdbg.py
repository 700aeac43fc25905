import sys, numpy as np
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import pathplanning_b200 as pp
from common import load_npz, c2_world_desc
g=load_npz("edges_c2.npz"); w=load_npz("wavefront_c2.npz")
ctx=pp.Context(0)
desc,occ,sq,pc=c2_world_desc(g); ctx.upload_map(desc,sq,occ,pc)
which=[int(a) for a in sys.argv[1:]]
goals=w["goals"][which]
try:
    cost,ex=ctx.wavefront(goals)
    print(which,'reached',[int(e.sum()) for e in ex],'expected',w['reached'][which].tolist())
    if which==[0]: print('mismatch cells', int((cost[0].view(np.uint32)!=w['cost0'].view(np.uint32)).sum()))
except Exception as e:
    print('ERR',e)
