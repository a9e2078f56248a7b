import os, sys
import numpy as np, torch
sys.path.insert(0, '/root/repo')
import gpsig_b200
from torch.profiler import profile, ProfilerActivity
L, d, m, Mz, nb = 100, int(os.environ.get('GPSIG_TIME_D', '8')), 4, 500, 50
TT = m*(m+1)//2
rng = np.random.default_rng(0)
X = np.cumsum(rng.standard_normal((nb, L, d))/np.sqrt(L), axis=1)
Zn = rng.standard_normal((TT, Mz, 2, d))*0.5
dt = torch.float64
k = gpsig_b200.kernels.SignatureRBF(d*L, d, m, lengthscales=np.ones(d))
X2 = torch.as_tensor(X.reshape(nb, L*d), dtype=dt, device='cuda').requires_grad_(True)
Zt = torch.as_tensor(Zn, dtype=dt, device='cuda').requires_grad_(True)
def full_step():
    X2.grad = None; Zt.grad = None
    Kzz, Kzx, Kxx = k.K_tens_n_seq_covs(Zt, X2, full_X_cov=False, increments=True)
    (Kzz.sum() + Kzx.sum() + Kxx.sum()).backward()
for _ in range(3): full_step()
torch.cuda.synchronize()
import time
t0 = time.perf_counter()
for _ in range(20): full_step()
t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
print("20 steps: host issue %.3f ms/step, with sync %.3f ms/step" % ((t1-t0)*50, (t2-t0)*50))
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for _ in range(5): full_step()
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by=os.environ.get("GPSIG_PROF_SORT", "cuda_time_total"), row_limit=40, max_name_column_width=60))
