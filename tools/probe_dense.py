"""Times one value+gradient evaluation of the logistic-regression program on the GPU."""
import sys, time, json
import numpy as np, torch
sys.path.insert(0, ".")
from fugue_b200 import api, models
N, D, Cn = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
t0 = time.time()
m, dat = models.logistic_regression(N=N, D=D)
P = api.Program(m)
print("build", time.time() - t0, file=sys.stderr)
ext = torch.cuda.ExternalStream(P.stream)
q = torch.randn(D, Cn, dtype=torch.float64, device="cuda") * 0.1
lp = torch.zeros(Cn, dtype=torch.float64, device="cuda")
g = torch.zeros(D, Cn, dtype=torch.float64, device="cuda")
torch.cuda.synchronize()
with torch.cuda.stream(ext):
    for _ in range(2):
        P.logp_grad_dev(q.data_ptr(), 0, Cn, lp.data_ptr(), g.data_ptr())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    R = 5
    for _ in range(R):
        P.logp_grad_dev(q.data_ptr(), 0, Cn, lp.data_ptr(), g.data_ptr())
    e1.record()
    e1.synchronize()
ms = e0.elapsed_time(e1) / R
flops = 4.0 * N * D * Cn
print(json.dumps(dict(N=N, D=D, C=Cn, ms=ms, grad_evals_per_s=Cn / ms * 1e3, dfma_tflops=flops / ms / 1e9,
                      alg_GBps=Cn * (8.0 * N * D + N + 16 * D) / ms / 1e6)))
