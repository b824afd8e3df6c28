"""fc_split_rhat_dev_noalloc (include/fugue_cuda.h): the allocation-free split-R-hat entry point, whose scratch is
a file-scope static __device__ array, against the existing fc_diagnostics_dev on the same device-resident
draws -- bit-equal (same kernels, same order of sums), on two calls in a row with different shapes so that the
second runs over the first call's leftovers. Called through the C ABI with raw device pointers. This file is
collected last on purpose."""
import ctypes as C
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _lib():
    from fugue_b200 import api
    lib = api.lib
    lib.fc_split_rhat_dev_noalloc.argtypes = [C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_void_p, C.c_void_p]
    lib.fc_split_rhat_dev_noalloc.restype = C.c_int32
    return api, lib


def _draws(torch, S, d, Cn, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = torch.randn((S, d, Cn), dtype=torch.float64, device="cuda", generator=g)
    x += torch.linspace(0.0, 0.5, Cn, dtype=torch.float64, device="cuda")          # chains disagree: R-hat > 1
    x *= torch.linspace(0.5, 2.0, d, dtype=torch.float64, device="cuda")[None, :, None]
    return x.contiguous()


def test_noalloc_split_rhat_equals_the_existing_entry_point_twice_in_a_row():
    import torch
    api, lib = _lib()
    torch.cuda.set_device(0)
    for call, (S, d, Cn) in enumerate([(64, 37, 200), (51, 5, 33), (64, 37, 200)]):
        x = _draws(torch, S, d, Cn, seed=call)
        out = torch.full((d,), -7.0, dtype=torch.float64, device="cuda")
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        rc = lib.fc_split_rhat_dev_noalloc(x.data_ptr(), S, d, Cn, out.data_ptr(), None)
        assert rc == 0, lib.fc_last_error().decode()
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before
        want = api.diagnostics_dev(x)["r_hat"]
        got = out.cpu().numpy()
        assert np.all(np.isfinite(got)) and got.min() > 1.0
        assert np.array_equal(got, want), (call, np.abs(got - want).max())
        # and the host entry point on one parameter (fc_split_rhat, diagnostics.rs:230-260): 1e-12 relative
        k = d // 2
        assert abs(got[k] - api.r_hat_f64(x[:, k, :].cpu().numpy())) <= 1e-12 * got[k]


def test_noalloc_split_rhat_rejects_sizes_beyond_the_static_scratch():
    import torch
    api, lib = _lib()
    out = torch.zeros(8, dtype=torch.float64, device="cuda")
    x = torch.zeros(16, dtype=torch.float64, device="cuda")
    # d * 2 * C = 2^21 segments > FC_RHAT_NOALLOC_MAX_SEGMENTS = 2^20: refused before any launch
    rc = lib.fc_split_rhat_dev_noalloc(x.data_ptr(), 4, 1 << 10, 1 << 10, out.data_ptr(), None)
    assert rc != 0
    assert "FC_RHAT_NOALLOC_MAX_SEGMENTS" in lib.fc_last_error().decode()
    torch.cuda.synchronize()
    assert float(out.abs().sum()) == 0.0
    # the largest shape that fits still runs: d * 2 * C = 2^20
    S, d, Cn = 4, 512, 1024
    xs = _draws(torch, S, d, Cn, seed=9)
    o2 = torch.zeros(d, dtype=torch.float64, device="cuda")
    assert lib.fc_split_rhat_dev_noalloc(xs.data_ptr(), S, d, Cn, o2.data_ptr(), None) == 0
    torch.cuda.synchronize()
    assert np.array_equal(o2.cpu().numpy(), api.diagnostics_dev(xs)["r_hat"])
