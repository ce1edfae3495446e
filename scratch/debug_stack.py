import sys; sys.path.insert(0,'.')
import torch, torch.nn.functional as F
from oracle import losses_cpu
from tests.helpers import loss_inputs, rel_err
import fedicra_b200 as fb
def up(a,h,w): return F.interpolate(a,size=(h,w),mode="bilinear",align_corners=False).cuda()
for (B,C,Cin,H,W) in [(12,2,1,64,64),(2,2,1,256,256),(12,2,1,256,256)]:
    image, logits, a1, a2, a3, label = loss_inputs(B,C,Cin,H,W,"quant8")
    for which in ("tel","crf"):
        xr=[t.clone().requires_grad_() for t in (logits,a1,a2,a3)]
        tot,tel,crf,as3=losses_cpu.loss_stack(xr[0],image,xr[1],xr[2],xr[3],label,C,gcrf="shift",dtype=torch.float64)
        (tel if which=="tel" else crf).backward()
        xs=[t.clone().requires_grad_() for t in (logits,a1,a2,a3)]
        img=image.cuda(); three=img.repeat(1,3,1,1); roi=(label==C).cuda(); lg=xs[0].cuda()
        if which=="tel":
            out=fb.MScaleRecurveTreeEnergyLoss()(lg,three,*(up(a,H,W) for a in xs[1:]),roi,0.1)[0]
        else:
            out=fb.ModelLossSemsegGatedCRF()(torch.softmax(lg,1),losses_cpu.KERNELS_DESC,5,img,H,W)["loss"]
        out.backward()
        print(B,H,which,'val',rel_err(out.detach().cpu(),(tel if which=="tel" else crf).detach()),'grads',[rel_err(a.grad,b.grad) if b.grad is not None else None for a,b in zip(xs,xr)], 'max|g|',[float(b.grad.abs().max()) if b.grad is not None else None for b in xr])
