import sys; sys.path.insert(0,'.')
import torch
from fedicra_b200 import ops
g=torch.Generator().manual_seed(0)
B,C,H,W=12,2,256,256
emb=((torch.rand(B,3,H,W,generator=g)*255).round()/255).cuda()
feat=torch.rand(B,C,H*W,generator=g).cuda(); gout=torch.randn(B,C,H*W,generator=g).cuda()
tree,mask=ops.mst_grid(emb,want_mask=True)
order=ops.bfs_grid(mask,H,W)
bands=ops.band_plan(order[3],order[4])
ew=ops.edge_weight_fwd(emb.reshape(B,3,-1),order[0],order[1],50.0)
for _ in range(3):
    out,sv=ops.tree_filter_forward(feat,ew,*order,bands)
    gf,gw=ops.tree_filter_backward(gout,ew,*order,bands,sv,True)
torch.cuda.synchronize()
