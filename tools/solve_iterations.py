import sys
sys.path.insert(0,"monte-carlo-posterior-policy-iteration_b200"); sys.path.insert(0,".")
import numpy as np, torch
from ppi_b200 import ops,_lib
import bench
rng=np.random.default_rng(0)
def t(fn,n=20):
    fn(); torch.cuda.synchronize()
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize(); return e0.elapsed_time(e1)/n*1e3
for K,D in ((1<<20,2048),(4096,840),(1024,840),(65536,1024)):
    # costs like the workloads': Rastrigin of N(1, 0.5 I) draws in D dims ~ Gaussian by the CLT
    c=torch.as_tensor(rng.normal(10.0*D, 3.0*np.sqrt(D), K).astype(np.float32),device="cuda")
    for mode,hyper,name in ((_lib.TEMP_ESS_TARGET,K/64,"ess K/64"),(_lib.TEMP_ESS_TARGET,10,"ess 10"),(_lib.TEMP_LOWER_BOUND,0.9,"lbps .9"),(_lib.TEMP_REPS_KL,0.5,"reps .5")):
        nw,lw,st=ops.solve_temperature(c,mode,hyper)
        us=t(lambda: ops.solve_temperature(c,mode,hyper,sync=False))
        print(f"K={K:8d} {name:9s} iterations={st.iterations:3d} alpha={st.alpha:10.4f} status={st.status} {us:7.1f} us")
