import sys, torch
sys.path.insert(0,'.')
import mddctc
from mddctc import _lib, _rt
L=_lib.lib()
T,B,C=1000,4096,42
x=torch.log_softmax(torch.randn(T,B,C,device='cuda'),-1)
il=torch.full((B,),T,dtype=torch.int64,device='cuda')
am=torch.empty((T,B),dtype=torch.uint8,device='cuda'); tok=torch.empty((B,T),dtype=torch.int32,device='cuda'); ol=torch.empty(B,dtype=torch.int32,device='cuda')
def run(tokens):
    L.mdd_greedy_decode(x.data_ptr(),T,B,C,il.data_ptr(),0,None,am.data_ptr(),tok.data_ptr() if tokens else None,T,ol.data_ptr() if tokens else None,_rt.stream_ptr(x.device))
for tokens in (False, True):
    for _ in range(3): run(tokens)
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(20): run(tokens)
    e1.record(); torch.cuda.synchronize()
    ms=e0.elapsed_time(e1)/20
    print('%s: %.4f ms, %.0f GB/s' % ('argmax+collapse' if tokens else 'argmax only', ms, T*B*(C*4+1)/ms/1e6))
