import torch, sys
sys.path.insert(0, '.')
from energy_equiv_lattice_gnn_b200 import metrics
fx = torch.load('tests/golden/ref_metrics.pt', weights_only=False)
dev = torch.device('cuda:0')
d = metrics.random_directions()
P, T = (10 * fx['pred']).to(dev), (10 * fx['target']).to(dev)
m, ep, et = metrics.per_lattice_metrics(P, T, d)
rp, rt = torch.linalg.eigvalsh(P.double().cpu()), torch.linalg.eigvalsh(T.double().cpu())
for name, got, ref in (('pred', ep, rp), ('target', et, rt)):
    e = (got.cpu().double() - ref).abs().amax(1) / ref.abs().amax(1)
    bad = (e > 1e-5).nonzero().flatten().tolist()
    print(name, 'bad rows', bad)
    for g in bad[:4]:
        print(g, got[g].cpu().tolist(), ref[g].tolist())
# run again with G=1 for a bad row
