"""Exhaustive check (CPU, numpy) that  q = x*r; q = fma(fma(-K, q, x), r, q)  with r = RN(1/K) is the correctly
rounded fp32 quotient x / K for K = 9 and K = 3 over all 2^23 mantissas (binade [1,2); scaling by powers of two is
exact).  csrc/warp_ssim.cuh relies on it to divide the SSIM window sums like ATen's avg_pool2d does, without a
division instruction.

    python tools/check_div_const.py
"""
import numpy as np


def check(k):
    r = np.float32(1.0) / np.float32(k)
    a = (np.arange(1 << 23, dtype=np.uint32) | np.uint32(127 << 23)).view(np.float32)
    q0 = (a * r).astype(np.float32)
    rem = a.astype(np.float64) - np.float64(k) * q0.astype(np.float64)       # exact in fp64
    rem32 = rem.astype(np.float32)
    assert np.all(rem32.astype(np.float64) == rem), 'residual not representable'
    q1 = (rem32.astype(np.float64) * np.float64(r) + q0.astype(np.float64)).astype(np.float32)
    ref = (a / np.float32(k)).astype(np.float32)
    return int((q1 != ref).sum()), int((q0 != ref).sum())


if __name__ == '__main__':
    for k in (9, 3):
        bad, bad0 = check(k)
        print('K=%d: %d mismatches with the correction step (%d without)' % (k, bad, bad0))
        assert bad == 0
