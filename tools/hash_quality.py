"""Bucket statistics of the count stage's slot hash (kc_device.cuh: table_hash) against fmix64 (MurmurHash3's finaliser) on
keys shaped like one partition: k-mers that share a minimizer at varying offsets, many of them one substitution apart.
Prints (largest bucket, buckets over 4 entries, fingerprint collisions inside a bucket) for both. CPU only: python tools/hash_quality.py"""
import numpy as np
M64=np.uint64(0xFFFFFFFFFFFFFFFF)
def mix(x):
    x=x.copy()
    x^=x>>np.uint64(32); x*=np.uint64(0xD6E8FEB86659FD93)
    x^=x>>np.uint64(32); x*=np.uint64(0xD6E8FEB86659FD93)
    x^=x>>np.uint64(32)
    return x
def fmix(k):
    k=k.copy()
    k^=k>>np.uint64(33); k*=np.uint64(0xff51afd7ed558ccd); k^=k>>np.uint64(33); k*=np.uint64(0xc4ceb9fe1a85ec53); k^=k>>np.uint64(33); return k
rng=np.random.default_rng(1)
def keys_shared_minimizer(n,k=21,m=15):
    # k-mers containing a fixed m-mer at random offset, plus near-duplicates (1 substitution)
    M=rng.integers(0,4,m)
    out=[]
    base=[]
    for i in range(n//4):
        a=rng.integers(0,k-m+1)
        s=np.concatenate([rng.integers(0,4,a),M,rng.integers(0,4,k-m-a)])
        base.append(s)
        out.append(s)
        for _ in range(3):
            t=s.copy(); p=rng.integers(0,k); t[p]=(t[p]+rng.integers(1,4))%4; out.append(t)
    out=np.array(out,dtype=np.uint64)
    v=np.zeros(len(out),dtype=np.uint64)
    for i in range(k): v|=out[:,i]<<np.uint64(62-2*i)
    return np.unique(v)
def stats(h,nb):
    b=((h&np.uint64(0xFFFFFFFF))*np.uint64(nb))>>np.uint64(32)
    c=np.bincount(b.astype(np.int64),minlength=nb)
    fp=(h>>np.uint64(45))
    # fingerprint collisions within bucket
    pair=b.astype(np.uint64)*np.uint64(1<<20)+fp
    col=len(pair)-len(np.unique(pair))
    # linear-probe overflow: simulate bucket spill
    spill=0;carry=0
    for x in np.concatenate([c,c]):
        t=x+carry; carry=max(0,t-4); 
    ov=(c>4).sum()
    return c.max(), ov, col
for n in (2000,4400):
    ks=keys_shared_minimizer(n)
    nb=len(ks)//2  # load 0.5 of 4-entry buckets => nbuckets = n/2
    print(n,len(ks),'mix',stats(mix(ks),nb),'fmix',stats(fmix(ks),nb))
