import torch, ctypes as C
from pvnet_b200 import _lib
L=_lib.lib()
torch.cuda.init(); torch.zeros(1,device='cuda')
s=C.c_void_p(torch.cuda.current_stream().cuda_stream)
for k in (0,1,0,1):
    print("kind",k,L.pvv_probe_tflops(k,4096,s))
