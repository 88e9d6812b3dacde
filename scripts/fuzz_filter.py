#!/usr/bin/env python
"""Fuzz of the fp32 triangle filter's per-axis error bounds (csrc/sat17.h) on the HOST build of the predicates
(osr_triangle_pair_predicates; no GPU needed): long thin triangles (aspect 1 .. 1000, sizes 0.01 .. 1) that cross,
touch, share vertices, have parallel edges, lie on a coarse grid, nearly coincide or are parallel at offsets of
1e-9 .. 1e-3 -- the filter may answer "separated" / "intersecting" only where the fp64 SAT-17 agrees.
    python scripts/fuzz_filter.py [seed] [rounds of 100 000 pairs]
Round 2: seeds 1-4 x 80 rounds + seed 123 x 40 rounds = 36 000 000 pairs, 0 contradictions."""
import os, sys, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from puzzletunneldiscovery_b200 import _lib
L=_lib.load()
def run(T):
    n=len(T); Tc=np.ascontiguousarray(T.reshape(n,18)); f=np.zeros(n,np.uint8); e=np.zeros(n,np.uint8)
    assert L.osr_triangle_pair_predicates(Tc.ctypes.data_as(_lib.c_dp), n, f.ctypes.data_as(_lib.c_u8p), e.ctypes.data_as(_lib.c_u8p))==0
    return f,e.astype(bool)
rs=np.random.RandomState(int(sys.argv[1]) if len(sys.argv)>1 else 123)
tot=0; bad=0; stats={}
for rnd in range(int(sys.argv[2]) if len(sys.argv)>2 else 40):
    n=100000
    kind=rnd%8
    # thin triangles: long edge ~ a, short ~ a*ratio
    a=10**rs.uniform(-2,0,size=(n,1,1)); ratio=10**rs.uniform(-3,0,size=(n,1,1))
    base=rs.randn(n,1,3)*0.3
    d1=rs.randn(n,1,3); d1/=np.linalg.norm(d1,axis=2,keepdims=True)
    d2=rs.randn(n,1,3); d2-= (d2*d1).sum(2,keepdims=True)*d1; d2/=np.linalg.norm(d2,axis=2,keepdims=True)
    P=np.concatenate([base, base+a*d1, base+a*rs.rand(n,1,1)*d1+a*ratio*d2],axis=1)
    # Q: another thin triangle crossing / near P's plane
    c=P.mean(1,keepdims=True)
    b2=10**rs.uniform(-2,0,size=(n,1,1)); r2=10**rs.uniform(-3,0,size=(n,1,1))
    g1=rs.randn(n,1,3); g1/=np.linalg.norm(g1,axis=2,keepdims=True)
    g2=rs.randn(n,1,3); g2-=(g2*g1).sum(2,keepdims=True)*g1; g2/=np.linalg.norm(g2,axis=2,keepdims=True)
    off=rs.randn(n,1,3)*a*ratio*rs.choice([0,1e-7,1e-5,1e-3,1e-1,1],size=(n,1,1))
    Q0=c+off - 0.5*b2*g1
    Q=np.concatenate([Q0, Q0+b2*g1, Q0+b2*rs.rand(n,1,1)*g1+b2*r2*g2],axis=1)
    if kind==1: Q[:,0]=P[:,1]                      # shared vertex
    if kind==2:                                    # vertex of Q exactly in P's plane (barycentric)
        w=rs.rand(n,3); w/=w.sum(1,keepdims=True); Q[:,0]=np.einsum('nj,njk->nk',w,P)
    if kind==3: Q[:,1]=Q[:,0]+(P[:,1]-P[:,0])*rs.rand(n,1)   # parallel edges
    if kind==4: P=np.round(P*4096)/4096; Q=np.round(Q*4096)/4096
    if kind==5: Q=P[:,[1,2,0]]+rs.randn(n,1,3)*1e-6          # nearly coincident triangles
    if kind==6: Q=P+ (np.cross(d1,d2))*10**rs.uniform(-9,-3,size=(n,1,1))  # parallel offset
    T=np.concatenate([P,Q],axis=1)
    f,e=run(T)
    b=int(((f==0)&e).sum()+((f==1)&~e).sum()); bad+=b; tot+=n
    stats.setdefault(kind,[]).append(((f==2).mean(), e.mean()))
    if b: print('CONTRADICTION kind',kind,b)
print('pairs',tot,'contradictions',bad)
for k,v in stats.items(): print(k,'unsure %.4f hits %.3f'%tuple(np.mean(v,axis=0)))
