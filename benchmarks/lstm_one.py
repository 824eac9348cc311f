import os, sys
sys.path.insert(0, "/root/repo")
import torch
from bigdl_b200.nn.layer import LSTM, Recurrent, Sequential
H = int(sys.argv[1]); B, T = 20, 35
m = Sequential()
for _ in range(2):
    m.add(Recurrent().add(LSTM(H, H)))
m.cuda().training(); m.getParameters()
x = torch.randn(B, T, H, device="cuda").bfloat16(); g = (torch.randn(B, T, H, device="cuda") * 0.01).bfloat16()
for _ in range(3):
    xin = x.clone().requires_grad_(True)
    with torch.enable_grad():
        y = m._call(xin)
    torch.autograd.backward(y, g)
torch.cuda.synchronize()
print("ok")
