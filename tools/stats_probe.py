import sys, time
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import numpy as np, cpu_libs
from epievo_b200.paths import Tree
from epievo_b200.synthetic import simulate_paths
from epievo_b200.device import Context
cpu=cpu_libs.best_cpu()
rates,T,_,_=cpu_libs.oracle().model_from_params(0.844912, 0.893359, -0.8, -1.8)
tree=Tree.balanced(8,0.05); n=600; R=96
ps=simulate_paths(tree,n,rates,T,seed=77)
ctx=Context(0); ctx.set_tree(tree); ctx.set_model(rates,T)
ref=[];gpu=[]
for c in range(R):
    cpu.load_paths(ps); cpu.seed(1000+c)
    J,D,_,_=cpu.run_mcmc(tree,rates,T,30,20); ref.append(np.concatenate([J[1:].ravel(),D[1:].ravel()]))
    ctx.upload_paths(ps); ctx.reset()
    J,D,_=ctx.run_mcmc(30,20,seed=5000+c); gpu.append(np.concatenate([J[1:].ravel(),D[1:].ravel()]))
a=np.array(ref); b=np.array(gpu)
z=(a.mean(0)-b.mean(0))/np.sqrt(a.var(0,ddof=1)/R+b.var(0,ddof=1)/R+1e-300)
order=np.argsort(-np.abs(z))[:12]
for i in order: print(i, "z=%.2f"%z[i], "ref=%.5g gpu=%.5g"%(a[:,i].mean(), b[:,i].mean()))
print("mean z^2", np.nanmean(z**2))
