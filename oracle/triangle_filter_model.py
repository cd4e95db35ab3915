"""CPU model of an edge-keyed triangle filter for node2vec's membership test (analysis only, lives with
the test infrastructure because it uses the oracle's graph generator).

Today walk_n2v_kernel decides "is the proposal c a neighbour of prev" with a Bloom block in prev's row: a second random
DRAM fetch per step.  If the aux word of slot cur -> c carried a 16-bit block over the prevs that close a triangle with it
(N(cur) n N(c), hashed by prev), the load that picks c would answer the test.  This script samples proposals the way
walkers make them at stationarity on the benchmark's RMAT graph and reports how full those blocks are and how often they
would say "maybe" (true common neighbours + false positives -> the exact search of today).

    python oracle/triangle_filter_model.py 20      # RMAT scale, BASELINE generator parameters
"""
import sys
import time

import numpy as np
sys.path.insert(0, str(__import__('pathlib').Path(__file__).resolve().parents[1]))
from oracle import wharf_oracle as wo
wo.build()
scale=int(sys.argv[1]); ehalf=(8<<20)>>(20-scale) if scale<=20 else 8<<20
t=time.time(); off,adj=wo.rmat_graph(scale,ehalf); n=1<<scale; m=adj.shape[0]
rp=np.append(off,np.uint64(m)).astype(np.int64); deg=np.diff(rp); adj=adj.astype(np.int64)
print('graph',n,m,time.time()-t)
rng=np.random.default_rng(1)
# walkers at stationary distribution of a uniform walk: cur ~ degree; prev uniform neighbour of cur (reversible chain); c uniform neighbour of cur
K=300000
eidx=rng.integers(0,m,K)                       # a uniform edge slot (prev -> cur): prev = source row, cur = adj
prev=np.searchsorted(rp,eidx,side='right')-1; cur=adj[eidx]
c=adj[rp[cur]+(rng.random(K)*deg[cur]).astype(np.int64)]
ret=(c==prev)
def row(v): return adj[rp[v]:rp[v+1]]
isnbr=np.zeros(K,bool); t_common=np.zeros(K,np.int64); maybe_tri=np.zeros(K,bool)
h=lambda x:(x*0x9E3779B1>>7)&15
for i in range(K):
    if ret[i]: continue
    rc=row(c[i]); rcur=row(cur[i]); rprev=row(prev[i])
    j=np.searchsorted(rprev,c[i]); isnbr[i]= j<rprev.shape[0] and rprev[j]==c[i]
    common=np.intersect1d(rcur,rc,assume_unique=True)
    t_common[i]=common.shape[0]
    bits=np.zeros(16,bool); bits[h(common)]=True
    maybe_tri[i]=bits[h(prev[i])]
nr=~ret
print('return frac',ret.mean(),'neighbour frac',isnbr[nr].mean())
print('common-neighbour count of slot (cur->c): mean',t_common[nr].mean(),'median',np.median(t_common[nr]),'p90',np.percentile(t_common[nr],90),'p99',np.percentile(t_common[nr],99))
print('triangle filter says maybe:',maybe_tri[nr].mean(),' of which true',(maybe_tri&isnbr)[nr].mean(),' false positives among non-neighbours',(maybe_tri&~isnbr)[nr].sum()/max(1,(~isnbr&nr).sum()))
assert not (isnbr&~maybe_tri)[nr].any()
