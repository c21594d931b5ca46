import sys, time, torch
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/dgl-1_b200')
from benchmarks import graphs
from dgl_b200.dist import edge_balanced_bounds
torch.set_num_threads(8)
c=graphs.C5_RMAT; n,e=c['n'],c['e']
t0=time.time()
# regenerate with access to the un-permuted ids: replicate graphs.rmat
g=torch.Generator(device='cpu'); g.manual_seed(c['seed'])
a,b,cc,_=(0.57,0.19,0.19,0.05)
src=torch.zeros(e,dtype=torch.int64); dst=torch.zeros(e,dtype=torch.int64)
for _ in range(26):
    u=torch.rand(e,generator=g)
    sbit=(u>=a+b).long(); dbit=((u>=a)&(u<a+b)|(u>=a+b+cc)).long()
    src=src*2+sbit; dst=dst*2+dbit
perm=torch.randperm(1<<26,generator=g)
src_g, dst_g = src % n, dst % n                 # generator order (no permutation)
src_r, dst_r = perm[src] % n, perm[dst] % n     # the benchmark's order
print('gen',time.time()-t0,flush=True)
def stats(src,dst,name):
    indeg=torch.bincount(dst,minlength=n); outdeg=torch.bincount(src,minlength=n)
    ip_in=torch.zeros(n+1,dtype=torch.int64); ip_in[1:]=torch.cumsum(indeg,0)
    ip_out=torch.zeros(n+1,dtype=torch.int64); ip_out[1:]=torch.cumsum(outdeg,0)
    bd=edge_balanced_bounds(ip_in,ip_out,8)
    bt=torch.tensor(bd)
    own_d=torch.bucketize(dst,bt[1:-1],right=True); own_s=torch.bucketize(src,bt[1:-1],right=True)
    local=float((own_d==own_s).float().mean())
    halos=[]
    for p in range(8):
        m=(own_d==p)&(own_s!=p)
        halos.append(int(torch.unique(src[m]).numel()))
    rows=[bd[i+1]-bd[i] for i in range(8)]
    print('%-22s local edges %.3f  halo rows per rank (M): %s  max %.2fM  sum %.1fM  rows/rank(M) %s'%(name,local,[round(h/1e6,2) for h in halos],max(halos)/1e6,sum(halos)/1e6,[round(r/1e6,1) for r in rows]),flush=True)
stats(src_r,dst_r,'random (benchmark)')
stats(src_g,dst_g,'generator order')
# degree-sorted relabel of the benchmark graph (generator-agnostic)
deg=torch.bincount(dst_r,minlength=n)+torch.bincount(src_r,minlength=n)
order=torch.argsort(deg,descending=True,stable=True)
new_id=torch.empty(n,dtype=torch.int64); new_id[order]=torch.arange(n)
stats(new_id[src_r],new_id[dst_r],'degree-sorted')
