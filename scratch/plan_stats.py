import sys; sys.path.insert(0,'.')
import torch
from fedicra_b200 import ops
g=torch.Generator().manual_seed(2022)
for (B,H,W,kind) in [(12,256,256,'img'),(12,256,256,'head'),(4,512,512,'img')]:
    if kind=='img': emb=((torch.rand(B,3,H,W,generator=g)*255).round()/255).cuda()
    else: emb=torch.nn.functional.interpolate(torch.randn(B,2,H//4,W//4,generator=g),size=(H,W),mode='bilinear').cuda()
    _,mask=ops.mst_grid(emb,want_mask=True,want_edges=False)
    si,sp,sc,ls,nl=ops.bfs_euler(mask,H,W)
    bands=ops.band_plan(ls,nl).view(B,-1).cpu()
    ls=ls.cpu()
    for b in range(min(B,3)):
        nb=int(bands[b,0]); cuts=bands[b,1:nb+2]
        pos=ls[b][cuts.long()]
        sizes=(pos[1:]-pos[:-1]); lev=(cuts[1:]-cuts[:-1])
        print(kind,H,'img',b,'nl',int(nl[b]),'nb',nb,'max pos',int(sizes.max()),'mean pos',float(sizes.float().mean()),'max lev',int(lev.max()),'mean lev',float(lev.float().mean()), 'maxwidth', int((ls[b,1:int(nl[b])+1]-ls[b,:int(nl[b])]).max()))
