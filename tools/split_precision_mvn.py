"""Numerical experiment for DESIGN.md: could the whitened mvn log-density
    lp = -1/2 |iL (x - mu)|^2 - Z            (runtime.cljc:364-366)
of BASELINE config 5 (d = 64) run on the tcgen05 tensor cores, which have no fp64 kind, and
still meet the 1e-9 relative bar?  The tensor-core arithmetic is EMULATED here in numpy:
operands rounded to the format's significand, products exact, accumulation in fp32 (what
kind::tf32 / kind::f16 accumulate in) - an optimistic model (no intermediate truncation).

  a) 1x / 3x TF32 split (x = x1 + x2 + x3, 11-bit significands), all cross products, fp32 accumulate
  b) the same splits with the partial GEMMs accumulated in fp64 OUTSIDE the tensor core
  c) Ozaki-style slicing: s slices of b bits per operand chosen so that every slice product sum
     over d = 64 terms is exact in the fp32 accumulator, partial GEMMs summed in fp64

python tools/split_precision_mvn.py
"""
import numpy as np


def round_sig(x, bits):
    """round to `bits` significand bits (round to nearest)"""
    m, e = np.frexp(x)
    return np.ldexp(np.round(m * 2.0 ** bits) / 2.0 ** bits, e)


def split(x, bits, parts):
    out, r = [], x.copy()
    for _ in range(parts):
        p = round_sig(r, bits)
        out.append(p)
        r = r - p
    return out


def gemm_fp32acc(A, B):
    # products of low-significand operands are exact in fp64; accumulate in fp32 order-independent model
    return (A.astype(np.float64)[:, :, None] * B.T.astype(np.float64)[None, :, :]).astype(np.float32).sum(1, dtype=np.float32).astype(np.float64)


def main():
    rng = np.random.default_rng(20240603)
    d, n = 64, 2048
    R = 0.5 * np.eye(d) + 0.5 * np.ones((d, d)) / d
    L = np.linalg.cholesky(R)
    iL = np.linalg.inv(L)
    X = rng.normal(size=(n, d)) * 1.5
    ref_w = X @ iL.T
    ref = -0.5 * (ref_w ** 2).sum(1)

    def report(name, W):
        lp = -0.5 * (W ** 2).sum(1)
        print("%-58s max rel err of lp %.2e   (bar 1e-9)" % (name, np.max(np.abs(lp - ref) / np.abs(ref))))

    for parts in (1, 3):
        xs, ls = split(X, 11, parts), split(iL, 11, parts)
        W32 = sum(gemm_fp32acc(a, b) for a in xs for b in ls)
        W64 = sum(a @ b.T for a in xs for b in ls)
        report("%dx TF32 split, fp32 accumulate (%d tensor GEMMs)" % (parts, parts * parts), W32)
        report("%dx TF32 split, partial GEMMs summed in fp64" % parts, W64)
    # Ozaki slicing: FIXED-POINT slices with one exponent per row (x) / per row of iL, b bits each;
    # slice products are integers below 2^(2b), 64 of them sum exactly in the 24-bit accumulator
    # if 2b + 6 <= 24; the partial GEMMs are scaled and summed in fp64 outside the tensor core
    b = 9

    def slices(M, s):
        ex = np.ceil(np.log2(np.abs(M).max(1, keepdims=True)))
        r = M / 2.0 ** ex                         # |r| < 1 row-wise
        out = []
        for k in range(s):
            q = np.trunc(r * 2.0 ** b)            # integer slice, |q| < 2^b
            out.append(q)
            r = r * 2.0 ** b - q
        return out, ex

    for s_ in (3, 4, 5, 6):
        xs, ex = slices(X, s_)
        ls, el = slices(iL, s_)
        W = np.zeros((n, d))
        gemms = 0
        for i, a in enumerate(xs):
            for j, bb in enumerate(ls):
                if i + j < s_:
                    W += gemm_fp32acc(a, bb) * 2.0 ** (-b * (i + j + 2))
                    gemms += 1
        W = W * 2.0 ** ex * 2.0 ** el.T
        report("Ozaki fixed-point, %d slices x %d bits, %d GEMMs (fp32 acc exact)" % (s_, b, gemms), W)

if __name__ == "__main__":
    main()
