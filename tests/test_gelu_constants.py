"""The GELU of the estimator's epilogues (`gelu_f`, chatterbox_b200/csrc/s3_gemm.cuh) is relu(x) - |x| 2^P(min(|x|, 6)) with P a
degree-6 fit of log2(erfc(|x| / sqrt 2) / 2).  The reference uses the exact erf form (diffusers `GELU`, approximate="none",
matcha/transformer.py:76-90).  This CPU test reads the seven coefficients out of the header and evaluates the same fp32 arithmetic
in numpy against erf on a dense grid: an edited or mistyped constant fails here without a GPU (the kernel itself is checked by
tests/test_cfm_gpu.py::test_gelu_of_the_epilogues_against_erf)."""
import os
import re

import numpy as np
from scipy.special import erf

HDR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "chatterbox_b200", "csrc", "s3_gemm.cuh")


def _coefficients():
    src = open(HDR).read()
    body = src[src.index("__device__ __forceinline__ float gelu_f(float x)"):]
    body = body[:body.index("return")]
    first = re.search(r"float p = fmaf\(([-0-9.e+]+)f, a, ([-0-9.e+]+)f\);", body)
    rest = re.findall(r"p = fmaf\(p, a, ([-0-9.e+]+)f\);", body)
    assert first and len(rest) == 5, "gelu_f is no longer a degree-6 Horner chain"
    assert "min.NaN.f32" in body and "6.0f" in body
    return [np.float32(first.group(1)), np.float32(first.group(2))] + [np.float32(c) for c in rest]      # highest degree first


def test_gelu_polynomial_matches_erf_gelu_in_fp32():
    c = _coefficients()
    x = np.linspace(-12.0, 12.0, 1_200_001).astype(np.float32)
    a = np.minimum(np.abs(x), np.float32(6.0))
    p = np.full_like(a, c[0])
    for k in c[1:]:
        p = (p * a + k).astype(np.float32)
    h = np.exp2(p).astype(np.float32)
    got = (np.maximum(x, np.float32(0)) - a * h).astype(np.float32)
    want = 0.5 * x.astype(np.float64) * (1.0 + erf(x.astype(np.float64) / np.sqrt(2.0)))
    err = np.abs(got - want)
    print(f"gelu_f constants: max abs error {err.max():.3e} at x = {x[err.argmax()]:.4f}")
    assert err.max() <= 5e-7
    assert abs(float(c[-1]) + 1.0) < 1e-5          # P(0) = log2(1/2)
