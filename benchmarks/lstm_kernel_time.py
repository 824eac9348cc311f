"""Device time of the persistent LSTM kernels alone (CUDA events around back-to-back launches)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bigdl_b200.ops.recurrent import LstmSeqFn

B, T = 20, 35
for H in (648, 650, 1500, 1504):
    xw = (torch.randn(B, T, 4 * H, device="cuda") * 0.5).bfloat16().requires_grad_(True)
    whh = (torch.randn(4 * H, H, device="cuda") / H ** 0.5).requires_grad_(True)
    whh.grad = torch.zeros_like(whh)
    w16 = whh.detach().bfloat16()
    g = (torch.randn(B, T, H, device="cuda") * 0.1).bfloat16()

    def fwd():
        with torch.no_grad():
            return LstmSeqFn.apply(xw.detach(), whh, w16, None, None, False)

    def fb():
        hs, hl, cl = LstmSeqFn.apply(xw, whh, w16, None, None, False)
        hs.backward(g)

    for name, fn in (("fwd", fwd), ("fwd+bwd", fb)):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            fn()
        e1.record()
        torch.cuda.synchronize()
        print(f"H={H} {name}: {e0.elapsed_time(e1) / 20 * 1e3:.0f} us per call ({e0.elapsed_time(e1) / 20 / T * 1e3:.1f} us/step)", flush=True)
