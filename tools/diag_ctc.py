"""Print CTC kernel-vs-oracle error statistics for a sweep of shapes (GPU box diagnostic)."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import ctc_oracle as co
from tests.test_ctc_gpu import make_case
m = importlib.import_module("end-to-end-asr-pytorch_b200.ctc")
dev = torch.device("cuda:0")
cases = [(2,1500,2000,200),(3,12,6,4),(4,50,31,20),(8,400,31,150),(2,300,1500,149),(4,200,5000,40),(2,64,31,60),(2,130,31,64)]
for pair in (1,0):
    for K in (0,4):
        os.environ["B2ASR_CTC_PAIR"]=str(pair); os.environ["B2ASR_CTC_K"]=str(K)
        for (B,T,V,L) in cases:
            x,y,il,tl = make_case(B,T,V,L,seed=B*1000+T)
            _,nll64,g64 = co.reference_call(x,y,il.tolist(),tl,torch.float64)
            _,_,g32 = co.reference_call(x,y,il.tolist(),tl,torch.float32)
            xd = x.to(dev).requires_grad_(True)
            try:
                loss,nll = m.ctc_loss(xd,y.to(dev),il,tl,from_logits=True,batch_first=True,return_nll=True)
                loss.backward(); torch.cuda.synchronize()
            except Exception as e:
                print("pair",pair,"K",K,(B,T,V,L),"EXC",e); continue
            dn = (nll.detach().cpu().double()-nll64).abs().max().item()
            g = xd.grad.cpu().double()
            err=(g-g64).abs(); scale=1.0/(B*tl.clamp_min(1).double())
            tol=1e-4*g64.abs()+2e-6*scale.view(B,1,1)
            print("pair",pair,"K",K,(B,T,V,L),"nll_err %.3e"%dn,"g_err %.3e"%err.max().item(),"gmax %.3e"%g64.abs().max().item(),"worst/tol %.2f"%(err/tol).max().item(), "err/scale %.2e"%(err/scale.view(B,1,1)).max().item(), "aten32 err/scale %.2e"%((g32.double()-g64).abs()/scale.view(B,1,1)).max().item())
