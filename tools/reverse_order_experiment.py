#!/usr/bin/env python
"""Forward-order vs reverse-order backward sweep of one pixel, in float32 (CPU experiment).

The reference walks the splats front to back in BOTH passes and gets the colour from behind by subtraction,
S_i = C_final - prefix_i (main.cpp:623-627); dL/dalpha_i = sum_c dL/dC_c (c_i,c T_i - S_i,c / (1 - alpha_i + 1e-15)).
The north star's variant walks the list back to front. Two ways to do that without per-layer storage:
  (a) 3DGS-style: keep T and the accumulated colour from behind A; going back one layer, T_i = T_{i+1} / (1 - alpha_i),
      A_i = alpha_{i+1} c_{i+1} + (1 - alpha_{i+1}) A_{i+1}; dL/dalpha_i = T_i * dL/dC . (c_i - A_i).
  (b) the same with the transmittances stored by the forward pass (no division), A recurrence unchanged.
alpha is NOT clamped below 1 in the reference (opacity 1 at a pixel centre gives alpha = 1 exactly), and T is only cut
at 1/256. This script measures what that does to (a) and (b) in float32 against a float64 evaluation of the
reference's formula, over random layer stacks with a prescribed maximum alpha.
"""
import numpy as np

f32, f64 = np.float32, np.float64
rng = np.random.default_rng(0)


def stack(n, amax, with_one):
    a = rng.uniform(0.0, amax, n)
    if with_one:
        a[rng.integers(0, n)] = 1.0
    return a, rng.uniform(0, 1, (n, 3)), rng.uniform(0, 1, 3)


def run(dt, a, c, ref):
    """returns (forward-order g, reverse (a) g, reverse (b) g) in dtype dt; layers after the pixel died (T < 1/256) are skipped"""
    a, c, ref = a.astype(dt), c.astype(dt), ref.astype(dt)
    n = len(a)
    T = np.ones(n + 1, dt)
    C = np.zeros(3, dt)
    live = np.zeros(n, bool)
    for i in range(n):
        if T[i] < dt(1 / 256):
            T[i + 1] = T[i]
            continue
        live[i] = True
        C = C + (T[i] * c[i]) * a[i]
        T[i + 1] = T[i] * (dt(1) - a[i])
    dL = C - ref
    # forward order, the reference's formula
    g_f = np.zeros(n, dt)
    P = np.zeros(3, dt)
    for i in range(n):
        if not live[i]:
            continue
        P = P + (T[i] * c[i]) * a[i]
        S = C - P
        g_f[i] = np.dot(dL, c[i] * T[i] - S / ((dt(1) - a[i]) + dt(1e-15)))
    # reverse order
    g_a, g_b = np.zeros(n, dt), np.zeros(n, dt)
    A = np.zeros(3, dt)
    Ta = T[n]
    with np.errstate(divide="ignore", invalid="ignore"):
        for i in range(n - 1, -1, -1):
            if not live[i]:
                continue
            Ta = Ta / (dt(1) - a[i])                      # (a): recover T_i by division
            g_a[i] = Ta * np.dot(dL, c[i] - A)
            g_b[i] = T[i] * np.dot(dL, c[i] - A)          # (b): stored T_i
            A = a[i] * c[i] + (dt(1) - a[i]) * A
    return g_f, g_a, g_b, live


def main():
    print("max alpha | alpha==1 layer | forward-order f32 | reverse (a) T/(1-alpha) f32 | reverse (b) stored T f32   (max rel err of dL/dalpha vs f64 reference formula)")
    for amax, one in [(0.5, False), (0.9, False), (0.99, False), (0.999, False), (0.9999, False), (0.9, True)]:
        e = np.zeros(3)
        bad = np.zeros(3, int)
        for _ in range(2000):
            a, c, ref = stack(24, amax, one)
            g64 = run(f64, a, c, ref)[0]
            gf, ga, gb, live = run(f32, a, c, ref)
            scale = np.abs(g64[live]).mean() + 1e-30
            for k, g in enumerate((gf, ga, gb)):
                d = np.abs(g[live].astype(f64) - g64[live]) / (np.abs(g64[live]) + scale)
                bad[k] += int((~np.isfinite(d)).sum())
                d = d[np.isfinite(d)]
                e[k] = max(e[k], d.max() if len(d) else 0)
        print(f"{amax:9} | {str(one):14} | {e[0]:9.2e} ({bad[0]} nan) | {e[1]:9.2e} ({bad[1]} nan/inf) | {e[2]:9.2e} ({bad[2]} nan/inf)")


if __name__ == "__main__":
    main()
