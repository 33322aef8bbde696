#!/usr/bin/env python3
"""Computes the orthonormal wavelet filter banks shipped in ``veles_b200/simd/wavelet_tables.json`` and
``native/include/veles/wavelet_tables.h``.

The reference ships three literal tables (``libVeles/simd/src/{daubechies,symlets,coiflets}.c``); here the scaling
filters are *derived*:

* Daubechies(N), 2N taps: spectral factorisation of the half-band polynomial
  ``P(y) = sum_{k<N} C(N-1+k, k) y^k`` (y = sin^2(w/2)) in multi-precision arithmetic, keeping the roots inside the
  unit circle (extremal phase), times the N-fold zero at z = -1.
* Symlet(N): the same factorisation, but for every root group {z, 1/z} (real) or {z, z*, 1/z, 1/z*} (complex) the
  side that is kept is chosen by exhaustive search for the most linear phase (least asymmetric filter).
* Coiflet(K), 6K taps: least-squares solution of the defining equations (orthonormality, 2K vanishing wavelet
  moments, 2K-1 vanishing shifted scaling moments); K = 1 from random starts, every further order by continuation
  from the previous one.

Every filter is normalised to sum sqrt(2) and checked for orthonormality before it is written.
Usage: python scripts/gen_wavelet_tables.py [--max-daubechies 38] [--max-symlet 38]
"""
import argparse
import itertools
import json
import math
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _mp():
    import mpmath
    return mpmath


def halfband_roots(N, digits=None):
    """Roots (in z, inside the unit circle) of the Daubechies half-band factor, grouped: list of tuples of roots that
    must be kept or mirrored together (real root -> 1-tuple, complex pair -> 2-tuple)."""
    mp = _mp()
    mp.mp.dps = digits or max(60, 12 * N)
    if N == 1:
        return []
    coeffs = [mp.binomial(N - 1 + k, k) for k in range(N)]          # ascending powers of y
    yroots = mp.polyroots(list(reversed(coeffs)), maxsteps=2000, extraprec=mp.mp.dps * 4)
    groups = []
    seen = []
    for y in yroots:
        # y = (2 - z - 1/z) / 4  ->  z^2 - (2 - 4y) z + 1 = 0
        b = 2 - 4 * y
        d = mp.sqrt(b * b - 4)
        z1, z2 = (b + d) / 2, (b - d) / 2
        z = z1 if abs(z1) < 1 else z2
        if abs(mp.im(y)) < mp.mpf(10) ** (-mp.mp.dps // 2):
            groups.append((mp.mpc(mp.re(z), 0),))
        else:
            if any(abs(y - mp.conj(s)) < mp.mpf(10) ** (-mp.mp.dps // 2) for s in seen):
                continue
            seen.append(y)
            groups.append((z, mp.conj(z)))
    return groups


def filter_from_roots(N, roots):
    """(1 + z)^N * prod (z - r), real part, normalised to sum sqrt(2); returned as a list of mpf, ascending powers."""
    mp = _mp()
    poly = [mp.mpf(1)]
    for _ in range(N):
        poly = _polymul(poly, [mp.mpf(1), mp.mpf(1)])
    for r in roots:
        poly = _polymul(poly, [-r, mp.mpf(1)])
    poly = [mp.re(c) for c in poly]
    s = sum(poly)
    scale = mp.sqrt(2) / s
    return [c * scale for c in poly]


def _polymul(a, b):
    mp = _mp()
    out = [mp.mpf(0)] * (len(a) + len(b) - 1)
    for i, x in enumerate(a):
        for j, y in enumerate(b):
            out[i + j] = out[i + j] + x * y
    return out


def daubechies(N):
    groups = halfband_roots(N)
    roots = [r for g in groups for r in g]
    h = filter_from_roots(N, roots)
    # descending powers of z^-1 with the minimum-phase roots = energy at the front
    return [float(c) for c in reversed(h)]


def _phase_nonlinearity(h):
    n = 512
    w = np.linspace(0, np.pi, n, endpoint=False)[1:]
    H = np.array([np.sum(h * np.exp(-1j * wk * np.arange(len(h)))) for wk in w])
    mask = np.abs(H) > 1e-6 * np.abs(H).max()
    ph = np.unwrap(np.angle(H))
    A = np.vstack([w[mask], np.ones(mask.sum())]).T
    coef, res, _, _ = np.linalg.lstsq(A, ph[mask], rcond=None)
    r = ph[mask] - A @ coef
    return float(np.sum(r * r))


# Which root groups of the half-band factor are mirrored outside the unit circle in the customary "symN" filters
# (bit i = group i after sorting by (|z|, |arg z|)).  The customary selections were made by hand in the literature
# and no single phase-linearity measure reproduces all of them, so the choice is recorded; the taps themselves are
# computed.  Orders without an entry fall back to the exhaustive least-squares phase search below.
SYMLET_SELECTION = {
    2: 0, 3: 0, 4: 2, 5: 1, 6: 5, 7: 1, 8: 10, 9: 6, 10: 21, 11: 6, 12: 21, 13: 28, 14: 44, 15: 28, 16: 89, 17: 142,
    18: 333, 19: 116, 20: 357, 21: 572, 22: 716, 23: 460, 24: 1593, 25: 2290, 26: 5005, 27: 6330, 28: 6373, 29: 9140,
    30: 12728, 31: 25080, 32: 41960, 33: 36324, 34: 83913, 35: 100280, 36: 167780, 37: 116858, 38: 320217,
}


def sorted_groups(N):
    mp = _mp()
    return sorted(halfband_roots(N), key=lambda gr: (float(abs(gr[0])), float(abs(mp.arg(gr[0])))))


def symlet(N):
    mp = _mp()
    if N in SYMLET_SELECTION:
        roots = []
        for i, g in enumerate(sorted_groups(N)):
            roots.extend([1 / mp.conj(r) for r in g] if (SYMLET_SELECTION[N] >> i) & 1 else list(g))
        return [float(c) for c in reversed(filter_from_roots(N, roots))]
    groups = halfband_roots(N)
    if len(groups) <= 1:
        return daubechies(N)
    best = None
    # fixing the first group's side removes the time-reversal twin of every candidate
    for choice in itertools.product((0, 1), repeat=len(groups) - 1):
        roots = list(groups[0])
        for g, flip in zip(groups[1:], choice):
            roots.extend([1 / mp.conj(r) for r in g] if flip else list(g))
        h = np.array([float(c) for c in reversed(filter_from_roots(N, roots))])
        score = _phase_nonlinearity(h)
        if best is None or score < best[0] - 1e-12:
            best = (score, h)
    return list(best[1])


def coiflet_equations(h, K):
    L = len(h)
    k = np.arange(L, dtype=np.float64)
    eqs = [np.sum(h) - math.sqrt(2.0)]
    for m in range(1, L // 2):
        eqs.append(np.dot(h[:-2 * m], h[2 * m:]))
    eqs.append(np.dot(h, h) - 1.0)
    c = 2 * K            # the tap the scaling moments are centred on (also the largest tap)
    u = (k - c) / (L / 4.0)
    g = h[::-1] * np.where(np.arange(L) % 2 == 0, 1.0, -1.0)
    ug = (k - (L - 1 - c)) / (L / 4.0)
    for p in range(2 * K):
        eqs.append(np.dot(g, ug ** p))
    for p in range(1, 2 * K):
        eqs.append(np.dot(h, u ** p))
    return np.asarray(eqs)


def coiflet(K, tries=12, seed=1):
    """Coiflet with 6K taps.  K = 1 is found from random starts; every further order is grown from the previous one
    (its taps shifted by two so that the centre taps line up, zeros at both ends): successive coiflets resemble
    each other closely enough for the least-squares iteration to stay on the customary branch, where random starts
    for 30 taps converge rarely and to several different solutions."""
    from scipy.optimize import least_squares
    L = 6 * K
    if K > 1:
        prev = np.asarray(coiflet(K - 1, tries, seed))
        x0 = np.zeros(L)
        x0[2:2 + len(prev)] = prev
        sol = least_squares(coiflet_equations, x0, args=(K,), xtol=1e-15, ftol=1e-15, gtol=1e-15, max_nfev=8000)
        if np.abs(sol.fun).max() > 1e-10:
            raise RuntimeError("coiflet continuation failed for K=%d (residual %.2e)" % (K, np.abs(sol.fun).max()))
        return coiflet_polish(sol.x, K)
    rng = np.random.RandomState(seed)
    best = None
    c = 2 * K
    for t in range(tries):
        # a sinc-like bump on the centre tap plus a little noise
        x0 = np.sinc((np.arange(L) - c) / 2.0) * np.exp(-0.5 * ((np.arange(L) - c) / (1.5 + K * rng.rand())) ** 2)
        x0 *= math.sqrt(2.0) / x0.sum()
        x0 += 0.01 * rng.randn(L)
        sol = least_squares(coiflet_equations, x0, args=(K,), xtol=1e-15, ftol=1e-15, gtol=1e-15, max_nfev=2000)
        if np.abs(sol.fun).max() > 1e-10:
            continue
        h = sol.x
        score = _phase_nonlinearity(h)
        if best is None or score < best[0] - 1e-12:
            best = (score, h.copy())
    if best is None:
        raise RuntimeError("no coiflet solution for K=%d" % K)
    return coiflet_polish(best[1], K)


def coiflet_polish(h0, K, digits=60):
    """Newton iteration on the coiflet equations in multi-precision arithmetic (analytic Jacobian, least-squares
    step for the consistent but overdetermined system).  The 30-tap system is ill-conditioned in double precision:
    two double solutions with residuals below 1e-13 differed by 5e-5; polished, they agree to every digit."""
    mp = _mp()
    mp.mp.dps = digits
    L, c = 6 * K, 2 * K
    quarter = mp.mpf(L) / 4
    ug = [(mp.mpf(i - (L - 1 - c)) / quarter) for i in range(L)]
    uh = [(mp.mpf(i - c) / quarter) for i in range(L)]
    sign = [1 if i % 2 == 0 else -1 for i in range(L)]
    x = [mp.mpf(float(v)) for v in h0]
    for _ in range(20):
        f, rows = [sum(x) - mp.sqrt(2)], [[mp.mpf(1)] * L]
        for m in range(1, L // 2):
            f.append(sum(x[i] * x[i + 2 * m] for i in range(L - 2 * m)))
            r = [mp.mpf(0)] * L
            for i in range(L - 2 * m):
                r[i] += x[i + 2 * m]
                r[i + 2 * m] += x[i]
            rows.append(r)
        f.append(sum(v * v for v in x) - 1)
        rows.append([2 * v for v in x])
        for p in range(2 * K):          # wavelet moments: g[i] = (-1)^i h[L-1-i]
            f.append(sum(sign[i] * x[L - 1 - i] * ug[i] ** p for i in range(L)))
            r = [mp.mpf(0)] * L
            for i in range(L):
                r[L - 1 - i] += sign[i] * ug[i] ** p
            rows.append(r)
        for p in range(1, 2 * K):       # scaling moments around the centre tap
            f.append(sum(x[i] * uh[i] ** p for i in range(L)))
            rows.append([uh[i] ** p for i in range(L)])
        dx = mp.lu_solve(mp.matrix(rows), -mp.matrix(f))
        x = [x[i] + dx[i] for i in range(L)]
        if max(abs(v) for v in dx) < mp.mpf(10) ** (-(digits - 10)):
            break
    return [float(v) for v in x]


def check_orthonormal(h, tol=1e-9):
    h = np.asarray(h, dtype=np.float64)
    assert abs(h.sum() - math.sqrt(2.0)) < tol, h.sum()
    assert abs(np.dot(h, h) - 1.0) < tol, np.dot(h, h)
    for m in range(1, len(h) // 2):
        assert abs(np.dot(h[:-2 * m], h[2 * m:])) < tol, (m, np.dot(h[:-2 * m], h[2 * m:]))


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-daubechies", type=int, default=38)
    ap.add_argument("--max-symlet", type=int, default=38)
    ap.add_argument("--max-coiflet", type=int, default=5)
    ap.add_argument("--only-coiflets", action="store_true", help="keep the other families of the existing JSON")
    ap.add_argument("--out-json", default=os.path.join(ROOT, "veles_b200", "simd", "wavelet_tables.json"))
    ap.add_argument("--out-header", default=os.path.join(ROOT, "native", "include", "veles", "wavelet_tables.h"))
    args = ap.parse_args(argv)
    tables = {"daubechies": {}, "symlet": {}, "coiflet": {}}
    if args.only_coiflets:
        with open(args.out_json) as f:
            old = json.load(f)
        tables["daubechies"], tables["symlet"] = old["daubechies"], old["symlet"]
        args.max_daubechies = args.max_symlet = 0
    for N in range(1, args.max_daubechies + 1):
        h = daubechies(N)
        check_orthonormal(h, 1e-9)
        tables["daubechies"][str(2 * N)] = h
        print("daubechies", 2 * N, "ok", file=sys.stderr)
    for N in range(1, args.max_symlet + 1):
        h = symlet(N)
        check_orthonormal(h, 1e-9)
        tables["symlet"][str(2 * N)] = h
        print("symlet", 2 * N, "ok", file=sys.stderr)
    for K in range(1, args.max_coiflet + 1):
        h = coiflet(K)
        check_orthonormal(h, 1e-8)
        tables["coiflet"][str(6 * K)] = h
        print("coiflet", 6 * K, "ok", file=sys.stderr)
    os.makedirs(os.path.dirname(args.out_json), exist_ok=True)
    with open(args.out_json, "w") as f:
        json.dump(tables, f, indent=0)
    write_header(tables, args.out_header)


def write_header(tables, path):
    lines = ["// Generated by scripts/gen_wavelet_tables.py - orthonormal scaling filters (sum = sqrt 2).", "#pragma once",
             "namespace veles { namespace wavelets {",
             "struct Filter { int type; int order; const double* taps; };  // type: 0 daubechies, 1 coiflet, 2 symlet"]
    entries = []
    for tname, tid in (("daubechies", 0), ("coiflet", 1), ("symlet", 2)):
        for order in sorted(tables[tname], key=int):
            name = "k_%s_%s" % (tname, order)
            taps = ", ".join("%.17g" % v for v in tables[tname][order])
            lines.append("static const double %s[%s] = {%s};" % (name, order, taps))
            entries.append("{%d, %s, %s}" % (tid, order, name))
    lines.append("static const Filter kFilters[] = {%s};" % ", ".join(entries))
    lines.append("static const int kFilterCount = %d;" % len(entries))
    lines.append("}}  // namespace veles::wavelets")
    os.makedirs(os.path.dirname(path), exist_ok=True)
    with open(path, "w") as f:
        f.write("\n".join(lines) + "\n")


if __name__ == "__main__":
    main()
