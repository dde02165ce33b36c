"""Diagnostics (GPU): survivor counts of the giant-cloud filter rounds vs the number of distinct winners.
    python tools_surv_count.py <n_points> [views]"""
import sys, numpy as np, torch
sys.path.insert(0,'/root/repo')
from scream_b200 import _lib, synth
from scream_b200.render import make_offsets
n=int(sys.argv[1]); V=int(sys.argv[2]) if len(sys.argv)>2 else 8
dev=torch.device('cuda:0'); rng=np.random.default_rng(5)
d=rng.standard_normal((n,3)).astype(np.float32); d/=np.linalg.norm(d,axis=1,keepdims=True)
xyz=torch.from_numpy(d*(rng.random((n,1),dtype=np.float32)**np.float32(1/3))).to(dev)
offs=make_offsets([(n//2,n-n//2)]).to(dev); views=torch.from_numpy(synth.random_views(V,5)).to(dev)
lib=_lib.load(); nws=_lib.workspace_bytes(1,V,64,n); ws=torch.zeros(nws,dtype=torch.uint8,device=dev)
imgs=torch.empty((1,V,2,64,64),device=dev); win=torch.empty((1,V,2,64,64),dtype=torch.int32,device=dev); zr=torch.empty((1,V,2),device=dev)
rc=lib.scream_render_forward(xyz.data_ptr(),offs.data_ptr(),n,n,None,views.data_ptr(),1,V,64,24.0,1,imgs.data_ptr(),win.data_ptr(),zr.data_ptr(),ws.data_ptr(),nws,torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize(); assert rc==0
al=lambda x:(x+255)//256*256
cnt_off=al(V*2*4)+al(V*2*4096*16)
cnt=ws[cnt_off:cnt_off+V*2*4].view(torch.int32).cpu().numpy()
cnt2=ws[cnt_off+V*2*4:cnt_off+V*2*8].view(torch.int32).cpu().numpy()
print("n",n,"round-1 survivors per image: mean",cnt.mean(),"max",cnt.max(),"frac of cloud",cnt.mean()/(n/2))
st=ws[cnt_off+(V*2*8+7)//8*8:cnt_off+(V*2*8+7)//8*8+24].view(torch.int64).cpu().numpy()
print("   voxel-mask passes (point,view pairs):",int(st[2]),"=",st[2]/(n*V),"of all pairs")
print("   cell-test passes (point,view pairs):",int(st[0]),"=",st[0]/(n*V),"of all pairs")
print("   round-2 survivors per image: mean",cnt2.mean(),"max",cnt2.max())
# how many survivors actually win a pixel
w=win[0].cpu().numpy(); print("distinct winners per image (mean)", np.mean([len(np.unique(w[v,c])) for v in range(V) for c in range(2)]))
# where (in the image plane) the round-2 survivors of image (view 0, source cloud) lie, against the winners
S1=(n//16 if n//16>49152 else 49152); S1=min(S1,1<<26); S2=1<<18
thr_off=cnt_off+al(V*2*8+64+48); surv_off=thr_off+al(V*2*4096*45*4); surv2_off=surv_off+al(V*2*S1*4)
ids=ws[surv2_off:surv2_off+int(cnt2[0])*4].view(torch.int32).long()
R=views[0]; src=xyz[:n//2]
def radii(idx):
    p=src[idx]@R.T; return (p[:,0]**2+p[:,1]**2).sqrt()*32
edges=torch.tensor([0,24,28,30,31,31.5,31.8,32.5],device=dev)
rs=radii(ids); rw=radii(torch.unique(win[0,0,0].long()))
print("radius (pixels) histogram edges",edges.tolist())
print("  survivors",torch.histc(rs,bins=1,min=0,max=40).item(),[int(((rs>=edges[i])&(rs<edges[i+1])).sum()) for i in range(len(edges)-1)])
print("  winners  ",[int(((rw>=edges[i])&(rw<edges[i+1])).sum()) for i in range(len(edges)-1)])
