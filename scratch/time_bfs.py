import sys; sys.path.insert(0,'.')
import torch, time
from fedicra_b200 import ops
def t(fn, n=10):
    fn(); torch.cuda.synchronize()
    a=torch.cuda.Event(enable_timing=True); b=torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); b.synchronize()
    return a.elapsed_time(b)/n
g=torch.Generator().manual_seed(0)
for (B,H,W) in [(12,256,256),(1,256,256),(48,256,256),(12,512,512)]:
    emb=((torch.rand(B,3,H,W,generator=g)*255).round()/255).cuda()
    tree,mask=ops.mst_grid(emb,want_mask=True)
    si,sp,sc,ls,nl=ops.bfs_grid(mask,H,W)
    print(B,H,W,'levels',nl.min().item(),nl.max().item(),'max width',int((ls[:,1:]-ls[:,:-1]).clamp(min=0).max()))
    print('  bfs_grid ms',t(lambda: ops.bfs_grid(mask,H,W)),' mst_grid ms',t(lambda: ops.mst_grid(emb,want_mask=True)), ' bfs_forward ms', t(lambda: ops.bfs_forward(tree)), ' bfs_euler ms', t(lambda: ops.bfs_euler(mask,H,W)))
