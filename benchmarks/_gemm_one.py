import sys, torch
sys.path.insert(0, '/root/repo')
from multimodaldiscussiontransformer_b200 import ops
M,N,K=238806,3072,768
x=torch.randn(M,K,device='cuda').bfloat16(); w=torch.randn(N,K,device='cuda').bfloat16(); b=torch.randn(N,device='cuda').bfloat16()
for _ in range(3): ops.gemm_bias_act(x,w,b)
for _ in range(2): ops.gemm_bias_act(x,w,b,gelu_act=True)
torch.cuda.synchronize()
