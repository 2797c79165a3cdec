#!/usr/bin/env python
"""Where does k_stem_wgrad_tc disagree with the definition?  Prints error maxima per (a,t), b, e and channel group."""
import sys, os
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from gaussiank_sgd_b200.ops import native

dev = "cuda:0"
lib = native.cuda_lib()
N, H, W = [int(v) for v in (sys.argv[1:4] if len(sys.argv) > 3 else (1, 32, 32))]
OH, OW = H // 2, W // 2
Hp, Wp = OH + 3, OW + 3
torch.manual_seed(0)
X = torch.zeros(N, Hp, Wp, 12, device=dev)
X[:, 2:-1, 2:-1] = torch.randn(N, OH, OW, 12, device=dev)
gy = torch.randn(N, OH, OW, 64, device=dev)
nsm = lib.gk_stem_wgrad_workspace() // (64 * 192)
os.environ["GK_STEM_WG_DEBUG"] = "1"
work = torch.full((nsm, 128, 192), float("nan"), device=dev)
gw = torch.zeros(64, 3, 7, 7, device=dev)
err = torch.zeros(1, dtype=torch.int32, device=dev)
rc = lib.gk_stem_wgrad(gy.data_ptr(), X.data_ptr(), gw.data_ptr(), work.data_ptr(), N, OH, OW, err.data_ptr(),
                       torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize()
print("rc", rc, "err", int(err))
grid = min(nsm, N * OH)
D = work[:grid].double()
print("nan lanes:", [int(torch.isnan(D[0, l]).any()) for l in range(128)])
print("lane abs max (cta 0):", [round(float(D[0, l].abs().max()), 1) for l in range(128)])
D = torch.nan_to_num(D).sum(0)                                   # 128 lanes x 192
lanes = [(m % 16) + 32 * (m // 16) for m in range(64)]
P = D[lanes].reshape(64, 4, 4, 12)                               # o, a, b, c
R = torch.zeros(64, 4, 4, 12, dtype=torch.float64, device=dev)   # o, a, b, c
for a in range(4):
    for b in range(4):
        R[:, a, b] = torch.einsum("nhwo,nhwc->oc", gy.double(), X[:, a:a + OH, b:b + OW].double())
E = (P - R).abs()
print("scale", float(R.abs().max()), "max err", float(E.max()))
print("per a:", [round(float(E[:, a].max()), 3) for a in range(4)])
print("per b:", [round(float(E[:, :, b].max()), 3) for b in range(4)])
print("per c:", [round(float(E[..., c].max()), 3) for c in range(12)])
print("per o group of 8:", [round(float(E[8 * k:8 * k + 8].max()), 3) for k in range(8)])
# hypotheses
Pf = D[:64].reshape(64, 192) if float(D[16:32].abs().max()) > 0 else P.reshape(64, 192)
Rf = R.reshape(64, 192)
def best_row_match():
    out = []
    for o in range(64):
        d = (Rf - Pf[o:o + 1]).abs().max(1).values
        out.append((int(d.argmin()), round(float(d.min()), 3)))
    return out
print("row o of P matches row .. of R:", best_row_match())
def best_col_match():
    out = []
    for c in range(0, 192):
        d = (Rf - Pf[:, c:c + 1]).abs().max(0).values
        out.append((int(d.argmin()), round(float(d.min()), 2)))
    return out
print("col c of P matches col .. of R:", best_col_match()[:48])
