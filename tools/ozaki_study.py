"""How many exact int8 slices would an error-free ("Ozaki") split of the C2 contractions need to hold the
1e-10 gate? CPU study (numpy, integer arithmetic emulated exactly in int64 / float64), same synthetic data
shape as BASELINE configs[1] on a row subset. For s_x slices of X (fixed data, split once, per-row scaling),
s_b slices of beta (per-chain scaling) and s_r slices of the residual r = y - sigmoid(eta) (fixed-point,
|r| < 1), with b-bit signed slices:
   eta_hat = sum_{i+j <= cut} Xs_i . Bs_j * scale,   g_hat = sum_{i+k <= cut} Xs_i^T . Rs_k * scale
and the relative errors  |eta_hat - eta| / max|eta|,  |g_hat - g| / max(|g_hat|, |g|)  per component (the gate of
tests/test_gpu_fullsize.py). Prints one JSON line per configuration."""
import json, sys
import numpy as np

rng = np.random.default_rng(0)
N, D, C, BITS = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000, 100, 8, 7


def split(a, scale_axis, n_slices, bits=BITS):
    """a = 2^e * sum_i s_i 2^(-bits (i+1)), s_i integers in [-2^(bits-1), 2^(bits-1)] (round-to-nearest per slice:
    the remainder stays exact in float64), e per row / column of `scale_axis`."""
    amax = np.max(np.abs(a), axis=scale_axis, keepdims=True)
    e = np.ceil(np.log2(np.maximum(amax, 1e-300))) + 1
    rem = a / np.exp2(e)            # |rem| < 1/2, exact scaling
    out = []
    for i in range(n_slices):
        rem = rem * float(2 ** bits)
        s = np.rint(rem)
        rem = rem - s               # exact (Sterbenz-like: s is the nearest integer)
        out.append(s)
    return out, e


def study(n_sx, n_sb, n_sr):
    X = rng.standard_normal((N, D))
    beta_true = rng.standard_normal(D) * 0.3
    y = (rng.random(N) < 1 / (1 + np.exp(-(X @ beta_true)))).astype(np.float64)
    B = beta_true[:, None] + 0.05 * rng.standard_normal((D, C))     # chains near the posterior mode: g cancels
    eta = (X.astype(np.longdouble) @ B.astype(np.longdouble))
    r_ld = y[:, None].astype(np.longdouble) - 1 / (1 + np.exp(-eta))
    g = np.asarray(X.T.astype(np.longdouble) @ r_ld, dtype=np.float64)
    Xs, ex = split(X, 1, n_sx)      # per-row scale
    Bs, eb = split(B, 0, n_sb)      # per-chain scale
    cut = max(n_sx, n_sb)
    eta_hat = np.zeros((N, C))
    n_prod = 0
    for i in range(n_sx):
        for j in range(n_sb):
            if i + j >= cut:
                continue
            eta_hat += (Xs[i] @ Bs[j]) * float(2.0 ** (-BITS * (i + j + 2)))   # integer products: exact in float64
            n_prod += 1
    eta_hat *= np.exp2(ex) * np.exp2(eb)
    err_eta = float(np.max(np.abs(eta_hat - np.asarray(eta, dtype=np.float64))) / np.max(np.abs(np.asarray(eta, dtype=np.float64))))
    r = y[:, None] - 1 / (1 + np.exp(-eta_hat))
    Rs, er = split(r, 0, n_sr)      # per-chain scale (|r| < 1)
    XsT, exT = split(X, 0, n_sx)    # the transposed contraction scales per COLUMN of X
    cut2 = max(n_sx, n_sr)
    g_hat = np.zeros((D, C))
    n_prod2 = 0
    for i in range(n_sx):
        for k in range(n_sr):
            if i + k >= cut2:
                continue
            g_hat += (XsT[i].T @ Rs[k]) * float(2.0 ** (-BITS * (i + k + 2)))
            n_prod2 += 1
    g_hat *= np.exp2(exT).T * np.exp2(er)
    rel = np.abs(g_hat - g) / np.maximum(np.abs(g_hat), np.abs(g))
    return dict(rows=N, D=D, chains=C, slice_bits=BITS, slices_X=n_sx, slices_beta=n_sb, slices_r=n_sr, int8_products_eta=n_prod, int8_products_grad=n_prod2,
                eta_rel_err=err_eta, grad_rel_err_max=float(rel.max()), grad_abs_over_sum_abs=float(np.median(np.abs(g)) / np.median(np.abs(X).T @ np.abs(r))),
                passes_1e10=bool(rel.max() <= 1e-10))


for n in (4, 5, 6, 7, 8):
    print(json.dumps(study(n, n, n)), flush=True)
