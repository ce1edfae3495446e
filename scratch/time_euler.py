import sys; sys.path.insert(0,'.')
import torch
from fedicra_b200 import ops
g=torch.Generator().manual_seed(0)
B,H,W=48,256,256
emb=((torch.rand(B,3,H,W,generator=g)*255).round()/255).cuda()
_,mask=ops.mst_grid(emb,want_mask=True,want_edges=False)
for _ in range(3): ops.bfs_euler(mask,H,W)
torch.cuda.synchronize()
