"""Fit the full-range polynomials the noise generator uses for the Box-Muller direction
(developer tool; prints the float32 coefficients pasted into csrc/common.cuh and their error over
all 65 536 angles the generator can produce).

    sin x ~ x * P(x^2),  cos x ~ Q(x^2)   on [-pi, pi]

Least squares on Chebyshev nodes, iteratively re-weighted towards the minimax solution (Lawson),
then rounded to float32 and evaluated with float32 Horner steps.
"""
import numpy as np


def lawson(basis, target, iters=200):
    w = np.ones(len(target))
    for _ in range(iters):
        sw = np.sqrt(w)
        c, *_ = np.linalg.lstsq(basis * sw[:, None], target * sw, rcond=None)
        err = np.abs(basis @ c - target)
        w = w * (err + 1e-300)
        w /= w.sum()
    return c


def horner32(x2, coefs):
    acc = np.full_like(x2, np.float32(coefs[-1]), dtype=np.float32)
    for c in coefs[-2::-1]:
        acc = (acc.astype(np.float64) * x2.astype(np.float64) + np.float64(np.float32(c))).astype(np.float32)  # one fused step
    return acc


def main(ns=6, nc=7):
    n = 4001
    x = np.pi * np.cos(np.pi * (np.arange(n) + 0.5) / n)
    x = x[np.abs(x) > 1e-9]
    x2 = x * x
    cs = lawson(np.stack([x2 ** k for k in range(ns)], 1), np.sin(x) / x)
    cc = lawson(np.stack([x2 ** k for k in range(nc)], 1), np.cos(x))
    cs32, cc32 = cs.astype(np.float32), cc.astype(np.float32)
    a = np.arange(65536, dtype=np.float64)
    w = (a.astype(np.uint32) | np.uint32(0x4B000000)).view(np.float32)
    ang = (w.astype(np.float64) * np.float64(np.float32(9.58738019107841e-05)) + np.float64(np.float32(-807.3892822265625))).astype(np.float32)
    exact = (a + 0.5) * 2 * np.pi / 65536 - np.pi
    a2 = (ang.astype(np.float64) * ang.astype(np.float64)).astype(np.float32)
    s = (horner32(a2, cs32).astype(np.float64) * ang.astype(np.float64)).astype(np.float32)
    c = horner32(a2, cc32)
    print("sin coefs:", ", ".join("%.9ef" % v for v in cs32))
    print("cos coefs:", ", ".join("%.9ef" % v for v in cc32))
    print("max |ang - exact|     %.3e" % np.max(np.abs(ang - exact)))
    print("max |sin err|         %.3e" % np.max(np.abs(s - np.sin(exact))))
    print("max |cos err|         %.3e" % np.max(np.abs(c - np.cos(exact))))
    r2 = s.astype(np.float64) ** 2 + c.astype(np.float64) ** 2
    print("|s^2+c^2-1| max %.3e mean %.3e" % (np.max(np.abs(r2 - 1)), np.mean(r2 - 1)))
    print("mean s %.3e  mean c %.3e  mean sc %.3e  mean c^2-s^2 %.3e" % (s.mean(dtype=np.float64), c.mean(dtype=np.float64),
          np.mean(s.astype(np.float64) * c), np.mean(c.astype(np.float64) ** 2 - s.astype(np.float64) ** 2)))


if __name__ == "__main__":
    import sys
    main(*(int(v) for v in sys.argv[1:]))
