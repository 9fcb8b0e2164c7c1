import sys, time
sys.path.insert(0,'/root/repo')
import numpy as np, torch
from phylo_rs_b200 import capi, synthetic
tree=synthetic.yule_flat(10000, 0x5EED0003)
model=synthetic.gtr_model(alpha=0.5,k=4,p_inv=0.2)
with capi.Context() as ctx:
    synthetic.load_into(ctx, tree, model)
    ctx.simulate_alignment(2048, seed=1)
    host = torch.empty((2048, 10000), dtype=torch.uint8, pin_memory=True)
    host.numpy()[:] = ctx.get_alignment(0, 2048)
    for rep in range(4):
        t0=time.perf_counter(); synthetic.load_into(ctx, tree, model); t1=time.perf_counter()
        v=ctx.log_likelihood_host((host.data_ptr(), 10000, 2048)); t2=time.perf_counter()
        w=ctx.log_likelihood(); t3=time.perf_counter()
        print(f"rep {rep}: load_into {1e3*(t1-t0):.2f} ms, log_likelihood_host {1e3*(t2-t1):.2f} ms, plain re-evaluation {1e3*(t3-t2):.2f} ms")
