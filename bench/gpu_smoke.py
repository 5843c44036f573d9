import os, sys, time, torch, copy
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch.nn.functional as F
from distributed_model_parallel_b200.models import MobileNetV2
from distributed_model_parallel_b200.models.mobilenetv2 import Block
from distributed_model_parallel_b200.ops import functional as OF
from distributed_model_parallel_b200.optim import FlatFusedSGD
from distributed_model_parallel_b200.train.engine import TrainStep
torch.manual_seed(0)
dev = torch.device("cuda")

def cmp_grads(m, ref, tag):
    worst = 1.0
    for (n, p), q in zip(m.named_parameters(), ref.parameters()):
        if p.grad is None: print("  no grad", n); continue
        cos = F.cosine_similarity(p.grad.flatten().float(), q.grad.flatten().float(), dim=0).item()
        rel = ((p.grad - q.grad).norm() / (q.grad.norm() + 1e-12)).item()
        if cos < 0.98: print("  %s %s cos %.4f rel %.3f |ref| %.3e" % (tag, n, cos, rel, q.grad.norm().item()))
        worst = min(worst, cos)
    print(tag, "worst cosine", worst)

# single blocks
for (cin, cout, e, s) in [(32, 16, 1, 1), (16, 24, 6, 1), (24, 32, 6, 2), (64, 64, 6, 1)]:
    b = Block(cin, cout, e, s).to(dev); r = copy.deepcopy(b)
    x = torch.randn(64, cin, 16, 16, device=dev).bfloat16().float()
    xa = x.clone().requires_grad_(); xb = x.clone().requires_grad_()
    o = b(xa); g = torch.randn_like(o.float()).bfloat16().float(); o.backward(g.to(o.dtype))
    with OF.use_reference_ops():
        orf = r(xb); orf.backward(g)
    print("block", (cin, cout, e, s), "out rel", ((o.float() - orf).norm() / orf.norm()).item(),
          "dx rel", ((xa.grad - xb.grad).norm() / xb.grad.norm()).item())
    cmp_grads(b, r, "block")

m = MobileNetV2().to(dev)
ref = copy.deepcopy(m); ref16 = copy.deepcopy(m)
x = torch.randn(128, 3, 32, 32, device=dev); y = torch.randint(0, 10, (128,), device=dev)
out = m(x); loss = OF.cross_entropy(out, y); loss.backward()
with OF.use_reference_ops():
    out_r = ref(x); loss_r = F.cross_entropy(out_r, y); loss_r.backward()
    with torch.autocast("cuda", dtype=torch.bfloat16):
        out_a = ref16(x); loss_a = F.cross_entropy(out_a.float(), y)
    loss_a.backward()
print("loss ours/fp32/autocast", loss.item(), loss_r.item(), loss_a.item())
print("logit rel err ours", ((out.float() - out_r).norm() / out_r.norm()).item(), "autocast", ((out_a.float() - out_r).norm() / out_r.norm()).item())
cmp_grads(m, ref, "ours-vs-fp32")
cmp_grads(ref16, ref, "autocast-vs-fp32")

mm = MobileNetV2().to(dev)
opt = FlatFusedSGD(mm.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-4)
step = TrainStep(mm, opt, use_graph=False)
for _ in range(3): step(x, y)
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3): step(x, y)
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=25, max_name_column_width=60))
