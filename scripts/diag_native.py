import sys, torch, torch.nn as nn, torch.nn.functional as F
sys.path.insert(0, '.')
from ewdml.models import build_model
from ewdml.ops import layers as L
from ewdml.parallel.flat import FlatModel
DEV = "cuda"
name = sys.argv[1] if len(sys.argv) > 1 else "ResNet50"
bs = int(sys.argv[2]) if len(sys.argv) > 2 else 8
torch.manual_seed(3)
m_ref = build_model(name, 10).to(DEV)
m_nat = build_model(name, 10); m_nat.load_state_dict(m_ref.state_dict())
m_amp = build_model(name, 10).to(DEV); m_amp.load_state_dict(m_ref.state_dict())
flat = FlatModel(m_nat, torch.device(DEV), channels_last=True, with_bf16=True)
x = torch.randn(bs, 3, 32, 32, device=DEV); y = torch.randint(0, 10, (bs,), device=DEV)
for m in (m_ref, m_nat, m_amp):
    m.train()
    for mod in m.modules():
        if isinstance(mod, nn.Dropout): mod.p = 0.0
L.set_compute("torch")
lr = F.cross_entropy(m_ref(x), y); lr.backward()
with torch.autocast("cuda", dtype=torch.bfloat16):
    la = F.cross_entropy(m_amp(x).float(), y)
la.backward()
L.set_compute("native")
flat.zero_grad()
ln = L.cross_entropy(m_nat(x.contiguous(memory_format=torch.channels_last)), y); ln.backward()
torch.cuda.synchronize()
print("loss ref/amp/native", lr.item(), la.item(), ln.item())
for (n, p), pr, pa in zip(m_nat.named_parameters(), m_ref.parameters(), m_amp.parameters()):
    g, gr, ga = p.grad.float().flatten(), pr.grad.float().flatten(), pa.grad.float().flatten()
    c1 = F.cosine_similarity(g, gr, dim=0).item(); c2 = F.cosine_similarity(ga, gr, dim=0).item()
    r1 = (g.norm() / (gr.norm() + 1e-12)).item()
    print(f"{n:40s} native cos {c1:7.4f} (norm ratio {r1:6.3f})   autocast cos {c2:7.4f}")
