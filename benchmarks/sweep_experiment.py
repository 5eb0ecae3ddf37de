"""Times single sweep modes (COL_STATS / ROW_LSE) in isolation; used with GFEDNTM_SWEEP_DEBUG to locate
the limiter (bit0: epilogue skips TMEM loads, bit1: MMA issue skipped)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gfedntm_b200.ops import kernels as K
from gfedntm_b200.ops.decoder import make_beta_shadows, refresh_beta_shadows

B, V, Kt = 4096, 200000, int(os.environ.get("K", "200"))
dev = "cuda"
Kp = K.padded_topics(Kt)
thT = (torch.randn(Kp, B, device=dev) * 0.1).to(torch.bfloat16)
beta = torch.randn(Kt, V, device=dev) * 0.3
beta_bf, betaT_bf = make_beta_shadows(Kt, V, dev)
refresh_beta_shadows(beta, beta_bf, betaT_bf)
colc = torch.zeros(K.round_up(V, 128), 4, device=dev)
sbd = torch.zeros(K.round_up(B, 128), 4, device=dev)
lse = torch.zeros(B * 2048, 2, device=dev)
nch = K.default_row_chunks(B, V, Kp=Kp)

def t(fn, n=5):
    fn(); torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / n * 1e3

cs = lambda: K.dec_sweep(K.COL_STATS, B=B, V=V, K=Kt, theta_bf=thT, beta_bf=beta_bf, sb_stream=sbd, colc=colc)
rl = lambda: K.dec_sweep(K.ROW_LSE, B=B, V=V, K=Kt, theta_bf=thT, beta_bf=beta_bf, sb_stream=colc, n_chunks=nch, lse_part=lse)
dc = torch.zeros(B, 2, device=dev); dth = torch.zeros(B, Kp, device=dev)
asp = torch.zeros(2, V, device=dev); dbeta = torch.zeros(Kt, V, device=dev)
nchg = K.default_row_chunks(B, V, second_mma=True, Kp=Kp)
rg = lambda: K.dec_sweep(K.ROW_GRAD, B=B, V=V, K=Kt, theta_bf=thT, beta_bf=beta_bf, sb_stream=colc, n_chunks=nchg, doc_const=dc, dtheta=dth)
cg = lambda: K.dec_sweep(K.COL_GRAD, B=B, V=V, K=Kt, theta_bf=thT, beta_bf=beta_bf, sb_stream=sbd, colc=colc, asp=asp[0], csp=asp[1], dbeta=dbeta)
print(f"   ROW_GRAD {t(rg):.0f} us  COL_GRAD {t(cg):.0f} us")
print(f"debug={os.environ.get('GFEDNTM_SWEEP_DEBUG','0')} K={Kt} COL_STATS {t(cs):.0f} us  ROW_LSE {t(rl):.0f} us")
