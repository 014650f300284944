#!/bin/bash
# which component makes the interpreter die with "terminate called" at exit?
cd "$(dirname "$0")/.."
echo "--- A: gemm only"; python -c "
import torch
from pipe_b200.ops.gemm import gemm_nt
a=torch.randn(256,256,device='cuda').bfloat16(); print(gemm_nt(a,a).shape); torch.cuda.synchronize()"; echo "rc=$?"
echo "--- B: fused layer fwd+bwd, no pipe"; python -c "
import torch
from pipe_b200.models.fused import FusedEncoderLayer
l=FusedEncoderLayer(256,4,512,0.1).cuda(); x=torch.randn(25,3,256,device='cuda').bfloat16().requires_grad_()
l(x).float().sum().backward(); torch.cuda.synchronize(); print('ok')"; echo "rc=$?"
echo "--- C: fused in Pipe"; python -c "
import torch
from pipe_b200 import Pipe
from pipe_b200.models.fused import FusedEncoderLayer
from torch import nn
p=Pipe(nn.Sequential(FusedEncoderLayer(256,4,512,0.1).cuda()),chunks=2); x=torch.randn(4,25,256,device='cuda').bfloat16().requires_grad_()
p(x).local_value().float().sum().backward(); torch.cuda.synchronize(); print('ok')"; echo "rc=$?"
echo "--- D: eager in Pipe"; python -c "
import torch
from pipe_b200 import Pipe
from torch import nn
p=Pipe(nn.Sequential(nn.Linear(8,8).cuda()),chunks=2); x=torch.randn(4,8,device='cuda').requires_grad_()
p(x).local_value().sum().backward(); torch.cuda.synchronize(); print('ok')"; echo "rc=$?"
echo "--- E: FusedClipAdam"; python -c "
import torch
from pipe_b200.models.fused import FusedEncoderLayer
from pipe_b200.ops.optim import FusedClipAdam
l=FusedEncoderLayer(256,4,512,0.1).cuda(); o=FusedClipAdam(l,lr=1e-3,max_norm=0.5); o.zero_grad()
x=torch.randn(25,3,256,device='cuda').bfloat16().requires_grad_()
l(x).float().sum().backward(); o.step(); torch.cuda.synchronize(); print('ok')"; echo "rc=$?"
