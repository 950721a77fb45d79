import sys, torch
sys.path.insert(0, '/root/repo')
from conjugategradient_b200 import generators as g, solver as S
n = int(sys.argv[1])
v, rp, c = g.poisson3d7(n)
dev = torch.device('cuda:0')
tv, tr, tc = (torch.as_tensor(a).to(dev) for a in (v, rp, c))
m = S.Mgcg(0)
o = S.default_options(precond=1, nx=n, ny=n, nz=n, norm=2, tolerance=1e-8, max_iteration=500)
import time
for i in range(2):
    t = time.time(); m.setup(tv, tr, tc, o); torch.cuda.synchronize(); print('setup s', time.time() - t)
