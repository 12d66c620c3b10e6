import sys, os; sys.path.insert(0, os.getcwd()); sys.path.insert(0,"tests")
import numpy as np, oak_b200
from oak_ref import *
orc=Oracle(); n=256
eng=oak_b200.BattleEngine(n)
b,d,r=oracle_sample_batch(orc,0,n)
eng.build_battles(sample_teams(0,n), sample_seeds(0,n))
gb,gd,gr=eng.download()
print(os.environ.get("OAKB200_LIB"), "res", gr[:8], r[:8], flush=True)
bad=np.nonzero((gb!=b).any(axis=1))[0]; print("bad battles", len(bad), flush=True)
if len(bad):
    i=bad[0]
    diff=np.nonzero(gb[i]!=b[i])[0]; print(i, diff.tolist()[:80]); print(gb[i][diff][:40], b[i][diff][:40], flush=True)
