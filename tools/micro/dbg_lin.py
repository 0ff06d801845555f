import os, sys, torch
sys.path.insert(0, "/root/repo")
from energy_equiv_lattice_gnn_b200 import native as nv
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
C, lmax = 32, 4
rows = [2 * l + 1 for l in range(lmax + 1)]
import ast
for npaths in ast.literal_eval(os.environ.get('CFGS', '((1, 1, 1, 1, 1), (2, 2, 2, 2, 2), (5, 8, 10, 10, 9))')):
    lin = nv.make_layout(rows, [p * C for p in npaths]); lout = nv.make_layout(rows, [C] * (lmax + 1))
    woff = [0]
    for p in npaths[:-1]:
        woff.append(woff[-1] + p * C * C)
    W = torch.randn(sum(p * C * C for p in npaths), generator=g).to(dev)
    n = int(os.environ.get('NN', 300))
    x = torch.randn(n, lin.stride, generator=g).to(dev)
    out = nv.block_linear(x, lin, W, woff, [1.0] * 5, lout, n)
    torch.cuda.synchronize()
    xd, Wd = x.double().cpu(), W.double().cpu()
    for b in range(5):
        k, no, r = lin.mul[b], lout.mul[b], lin.rows[b]
        xb = xd[:, lin.off[b]:lin.off[b] + r * k].reshape(n, r, k)
        ref = (xb @ Wd[woff[b]:woff[b] + k * no].reshape(k, no)).reshape(n, r * no)
        got = out[:, lout.off[b]:lout.off[b] + r * no].double().cpu()
        err = (got - ref).abs().max() / ref.abs().max()
        # partial-K hypothesis: which K prefix reproduces got?
        best = None
        for kk in range(32, k + 1, 32):
            refk = (xb[:, :, :kk] @ Wd[woff[b]:woff[b] + k * no].reshape(k, no)[:kk]).reshape(n, r * no)
            e = float((got - refk).abs().max() / ref.abs().max())
            if best is None or e < best[1]:
                best = (kk, e)
        print(npaths[b], "block", b, "K", k, "err %.2e" % err, "best K-prefix", best)
