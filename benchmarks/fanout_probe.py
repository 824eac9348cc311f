"""Which backward path do the Inception modules take? Counts kernel launches of one eager step by name."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bigdl_b200.models.inception import Inception_v1_NoAuxClassifier
from bigdl_b200.nn.criterion import ClassNLLCriterion
from bigdl_b200.nn.fusion import fuse
from bigdl_b200.ops import functions as fn, elementwise as ew

m = Inception_v1_NoAuxClassifier(1000).cuda()
fuse(m)
x = torch.randn(16, 3, 224, 224, device="cuda")
y = torch.randint(1, 1001, (16,), device="cuda").float()
crit = ClassNLLCriterion()
cnt = {"colsum": 0, "colsum_sum_relu": 0, "addn": 0, "premask_hit": 0}
for name in ("colsum", "colsum_sum_relu", "addn"):
    orig = getattr(ew, name)
    def wrap(*a, _o=orig, _n=name, **k):
        cnt[_n] += 1
        return _o(*a, **k)
    setattr(ew, name, wrap)
for it in range(3):
    for k in cnt: cnt[k] = 0
    out = m.forward(x)
    loss = crit.forward(out, y)
    m.backward(x, crit.backward(out, y))
    torch.cuda.synchronize()
    print(it, cnt)
