#!/usr/bin/env python3
"""tools/scan_onepole_error.py -- float32 error of the blocked-scan evaluation of ModTilt's one-pole smoothers
(y[t] = a x[t] + (1 - a) y[t-1]) against the sequential float32 recurrence, both measured against float64.
Tile = 32 frames = 8 lanes x 4 frames: local recurrences from zero, 3-step Kogge-Stone carry scan with B^4, B^8, B^16,
fix-up y[i] += B^(i+1) * carry.  Evidence for DESIGN.md 4.4a (the time-parallel formulation planned for this kernel)."""
import numpy as np
F = np.float32


def sequential(a, x):
    y = np.empty_like(x); s = F(0); b = F(1) - a
    for i, v in enumerate(x):
        s = v * a + s * b; y[i] = s
    return y


def scan(a, x, G=8, c=4):
    b = F(1) - a
    pw = [b]
    for _ in range(c - 1):
        pw.append(pw[-1] * b)                     # B^1 .. B^c
    bc = [pw[-1]]
    while len(bc) < 3:
        bc.append(bc[-1] * bc[-1])                # B^c, B^2c, B^4c
    y = np.empty_like(x); prev = F(0)
    for t0 in range(0, len(x) - G * c + 1, G * c):
        z = np.zeros((G, c), F)
        for j in range(G):
            s = prev if j == 0 else F(0)
            for i in range(c):
                s = x[t0 + j * c + i] * a + s * b; z[j, i] = s
        v = z[:, -1].copy()
        for k, d in enumerate((1, 2, 4)):
            t = np.concatenate([np.zeros(d, F), v[:-d]])
            v = np.where(np.arange(G) >= d, t * bc[k] + v, v).astype(F)
        carry = np.concatenate([[F(0)], v[:-1]]).astype(F)
        for j in range(1, G):
            for i in range(c):
                z[j, i] = z[j, i] + pw[i] * carry[j]
        y[t0:t0 + G * c] = z.reshape(-1); prev = z[-1, -1]
    return y


def main():
    rng = np.random.default_rng(1)
    n = 32 * 1500
    x = (rng.uniform(-1, 1, n) * np.repeat(10 ** (rng.uniform(-60, -6, n // 2400) / 20), 2400)).astype(F) ** 2
    for name, hz in (("a_env 25 Hz", 25.0), ("a_base 1 Hz", 1.0), ("a_trim 0.2 Hz", 0.2), ("a_ratio", None)):
        a = F(0.05) if hz is None else F(1.0) - np.exp(F(-2 * np.pi * hz / 48000.0), dtype=F)
        a64, b64 = float(a), float(F(1) - a)
        ref = np.empty(n); s = 0.0
        for i in range(n):
            s = float(x[i]) * a64 + s * b64; ref[i] = s
        es = np.abs(sequential(a, x) - ref).max() / np.abs(ref).max()
        ep = np.abs(scan(a, x) - ref).max() / np.abs(ref).max()
        print(f"{name:14s} sequential f32 max err / peak = {es:.2e}   blocked scan f32 = {ep:.2e}")


if __name__ == "__main__":
    main()
