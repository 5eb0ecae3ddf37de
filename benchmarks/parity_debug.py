import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
import test_gpu_parity_large as T
from gfedntm_b200.ops.decoder import CscWorkspace, ProdLDADecoder, make_beta_shadows, refresh_beta_shadows

B, V, K = T.B, T.V, T.K
torch.manual_seed(0)
counts = T._corpus(B, seed=11)
batch, sub = T._csr_batch(counts, slice(0, B))
g = torch.Generator(device="cuda").manual_seed(3)
theta = torch.softmax(torch.randn(B, K, device="cuda", generator=g) * 1.5, dim=1)
beta = (torch.rand(K, V, device="cuda", generator=g) - 0.5) * 0.2
beta_bf, betaT_bf = make_beta_shadows(K, V, "cuda")
refresh_beta_shadows(beta, beta_bf, betaT_bf)
dec = ProdLDADecoder(B, V, K, batch.nnz, "cuda")
csc = CscWorkspace(V, batch.nnz, "cuda"); csc.build(batch)
dbeta = torch.zeros(K, V, device="cuda"); loss = torch.zeros(1, dtype=torch.float64, device="cuda")
rm, rv = torch.zeros(V, device="cuda"), torch.ones(V, device="cuda")
rl, dtheta = dec.forward_backward(theta, theta.sum(0), beta_bf, betaT_bf, batch, csc, training=True, running_mean=rm, running_var=rv, dbeta=dbeta, loss=loss)
torch.cuda.synchronize()
for name, bb in (("unrounded beta", beta), ("bf16-rounded beta", beta_bf[:K, :V].float())):
    rl_ref, dth_ref, db_ref = T._oracle_decoder(theta, bb, sub)
    err = (dbeta - db_ref).abs()
    print(name, "dbeta max err/max", float(err.max() / db_ref.abs().max()), "rms rel", float(torch.sqrt((err.double()**2).mean()) / torch.sqrt((db_ref.double()**2).mean())))
    flat = err.flatten().topk(5)
    colcnt = torch.from_numpy(sub.sum(0).A1).cuda()
    for val, idx in zip(flat.values.tolist(), flat.indices.tolist()):
        k, v = idx // V, idx % V
        print(f"   k={k} v={v} got={dbeta[k, v].item():.5f} ref={db_ref[k, v].item():.5f} err={val:.5f} colcount={colcnt[v].item():.0f}")
    e2 = (dtheta[:, :K].double() - dth_ref).abs()
    print("   dtheta max err/max", float(e2.max() / dth_ref.abs().max()), "max|dtheta|", float(dth_ref.abs().max()))
    t5 = e2.flatten().topk(3)
    for val, idx in zip(t5.values.tolist(), t5.indices.tolist()):
        b, k = idx // K, idx % K
        print(f"   b={b} k={k} got={dtheta[b,k].item():.4f} ref={dth_ref[b,k].item():.4f} theta={theta[b,k].item():.4f}")
print("max |dbeta|", float(db_ref.abs().max()), "rms", float(torch.sqrt((db_ref.double()**2).mean())))
