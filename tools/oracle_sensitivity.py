"""The oracle against itself with the tables perturbed by 1e-7 relative: how fast fp32 rounding differences grow with the
number of sequential samples (the clipped running-sum dynamics amplify them ~10x per 10 samples on the `small` shape).
Companion of tools/seqcheck.py: a GPU-vs-oracle difference that grows at this rate is rounding, not an ordering bug."""
import sys; import os; R=os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0,R); sys.path.insert(0,os.path.join(R,'tests'))
import numpy as np, oracle
from heat_b200 import synth
from helpers import rel_err
d = synth.generate_named("small"); D,K=64,16
negs = oracle.neg_table(2022,0,0,d.clicks[:1000],K,d.num_items)
a = synth.init_tables(d.num_users,d.num_items,D); b = synth.init_tables(d.num_users,d.num_items,D)
rng=np.random.default_rng(1)
b[0][...] = b[0]*(1+1e-7*rng.standard_normal(b[0].shape)).astype(np.float32)
b[1][...] = b[1]*(1+1e-7*rng.standard_normal(b[1].shape)).astype(np.float32)
oa = oracle.COracle(D,K,d.clicks,d.his_items,d.masks,*a); ob = oracle.COracle(D,K,d.clicks,d.his_items,d.masks,*b)
done=0
for n in [1,3,10,30,33,64,70,100,130,200,300,600,1000]:
    while done<n:
        u,i=int(d.clicks[done,0]),int(d.clicks[done,1]); la=oa.step(u,i,negs[done]); lb=ob.step(u,i,negs[done]); done+=1
    print(n, f"user {rel_err(a[0],b[0]):.2e} item {rel_err(a[1],b[1]):.2e} w0 {rel_err(a[2],b[2]):.2e}", la, lb)
