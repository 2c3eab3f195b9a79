"""Developer check for PB_FAST_BUILD libraries (headline K1 instance only): local operators and condensed system of
HHO(3,2) with fancy stabilisation against the oracle, then the K1 timing."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, scipy.sparse as sp
import proton_b200 as pb
from proton_b200 import PB
from oracle import orc

def frob_rel(a, b): return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300)
k, N = 2, 6
ctx = pb.Context(0)
m = orc.Mesh(N)
hdi = pb.hho_degree_info(k + 1, k)
ctx.mesh(pb.mesh_init_params(N, N)).assemble(hdi, stab_kind=PB.STAB_FANCY, keep_local=True)
lc = ctx.local(PB.LOCAL_LC); oper = ctx.local(PB.LOCAL_OPER); rhs = ctx.local(PB.LOCAL_RHS)
S = ctx.local(PB.LOCAL_SCHUR)
worst = ws = 0.0
for c in range(m.ncells):
    r = m.local(c, k + 1, k, stab=PB.STAB_FANCY)
    worst = max(worst, frob_rel(lc[c], r["data"] + r["stab"]), frob_rel(oper[c][: hdi.rbs - 1], r["oper"]), frob_rel(rhs[c], r["rhs"]))
    K = r["data"] + r["stab"]; C = hdi.cbs
    Sc = K[C:, C:] - K[C:, :C] @ np.linalg.solve(K[:C, :C], K[:C, C:])
    ws = max(ws, frob_rel(S[c], Sc))
print(f"local parity worst {worst:.2e}  condensed block worst {ws:.2e}  ->", "OK" if worst < 1e-11 and ws < 1e-10 else "FAIL", flush=True)
