"""Where the SMC end-to-end time goes (program build, session + prior pass, ladder, copy-out to pinned host memory, close):
python tools/smc_e2e_probe.py  (one B200, C4 at 2^24 particles)."""
import os, sys, time, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from fugue_b200 import api, models
n = 1 << 24
cfg = api.SMCConfig(rejuvenation_steps=1)
v_pin = torch.zeros(5, n, dtype=torch.float64).pin_memory(); w_pin = torch.zeros(n, dtype=torch.float64).pin_memory(); lw_pin = torch.zeros(n, dtype=torch.float64).pin_memory()
for r in range(5):
    t0 = time.perf_counter()
    P2 = api.Program(models.gaussian_mixture(n_obs=80)[0]); t1 = time.perf_counter()
    S = api.SmcSession(P2, n, cfg, 2000 + r); torch.cuda.synchronize(); t2 = time.perf_counter()
    k = 0
    while S.step() < 1.0: k += 1
    torch.cuda.synchronize(); t3 = time.perf_counter()
    le, ns = C.c_double(), C.c_uint64()
    api.check(api.lib.fc_smc_get(S.h, C.c_void_p(v_pin.data_ptr()), None, C.c_void_p(w_pin.data_ptr()), C.c_void_p(lw_pin.data_ptr()), C.byref(le), C.byref(ns)))
    t4 = time.perf_counter()
    S.close(); P2.close(); t5 = time.perf_counter()
    print(f"program {1e3*(t1-t0):.1f} ms, session(create+prior) {1e3*(t2-t1):.1f}, ladder {1e3*(t3-t2):.1f}, get(D2H 940 MB) {1e3*(t4-t3):.1f}, close {1e3*(t5-t4):.1f}", flush=True)
