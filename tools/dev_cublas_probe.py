import torch
M, D, F = 256*197, 768, 3072
for (m,n,k) in [(M,D,F),(M,F,D),(M,3*D,D)]:
    a = torch.randn(m,k,device="cuda").bfloat16(); w = torch.randn(n,k,device="cuda").bfloat16()
    for _ in range(2): torch.matmul(a, w.t())
torch.cuda.synchronize()
