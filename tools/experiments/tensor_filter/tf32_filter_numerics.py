#!/usr/bin/env python3
"""How much slack would a TF32 tensor-core version of the closest-hit pre-filter need, and how many more candidates would
it let through?  (Groundwork for DESIGN.md section 9, item 1; numpy only, no GPU, no oracle.)

The pre-filter's discriminant  v = (C.D - O.D)^2 - (|C|^2 - r^2) + 2 C.O - |O|^2  is evaluated three ways on the
149-sphere scene of config 3 against binary64:

  fp32   the shipped FFMA chains (spt_filter.cuh), emulated in float32
  A      two K = 16 GEMMs (tC and the linear part) with hi/lo TF32 splits of both operands, v = tC^2 + na in fp32
  B      ONE GEMM with the square expanded: sum_ij (C_i C_j)(D_i D_j) + C.W + k + q, every factor hi/lo split

TF32 operands are fp32 values with the low 13 mantissa bits cleared; products of two TF32 values are exact in fp32; the
accumulation is modelled pessimistically as fp32 with truncation (round toward zero) after every addition, and
optimistically in binary64.  For each variant: the worst |v - v_exact| over all (ray, sphere) pairs, as an absolute
number and in units of u (A^2 + B^2) (u = 2^-24, A = |C|, B = |O|), and the candidates per ray when the slack is set to
that worst error times two.

    python tools/tf32_filter_numerics.py [--rays 20000]
"""
import argparse
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
U = 2.0 ** -24


def tf32(x):
    """fp32 -> TF32 by truncation (what the tensor core does to an fp32 operand)."""
    b = np.asarray(x, np.float32).view(np.uint32) & np.uint32(0xFFFFE000)
    return b.view(np.float32)


def split(x):
    x = np.asarray(x, np.float32)
    hi = tf32(x)
    lo = tf32((x - hi).astype(np.float32))
    return hi, lo


def rz32(x):
    """binary64 -> fp32 rounded toward zero (pessimistic model of the tensor core's accumulator)."""
    x = np.asarray(x, np.float64)
    f = x.astype(np.float32)
    over = np.abs(f.astype(np.float64)) > np.abs(x)
    f[over] = np.nextafter(f[over], np.float32(0))
    return f


def accumulate(terms, mode):
    """Sum a list of arrays of exact products.  mode 'rz': fp32, truncating after every add; 'f64': binary64."""
    if mode == "f64":
        s = np.zeros_like(terms[0], dtype=np.float64)
        for t in terms:
            s = s + t.astype(np.float64)
        return s
    s = np.zeros_like(terms[0], dtype=np.float32)
    for t in terms:
        s = rz32(s.astype(np.float64) + t.astype(np.float64))
    return s.astype(np.float64)


def make_rays(C, r, n, rng):
    """Secondary-like rays: origins on sphere surfaces and on the ground, directions = normal + shell vector."""
    k = rng.integers(0, len(r), n)
    nrm = rng.normal(size=(n, 3))
    nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    org = C[k] + nrm * r[k][:, None]
    ground = rng.random(n) < 0.5
    g = np.stack([rng.uniform(-25, 25, n), np.zeros(n), rng.uniform(-4, 25, n)], axis=1)
    org[ground] = g[ground]
    nrm[ground] = np.array([0.0, 1.0, 0.0])
    shell = rng.uniform(-0.5, 0.5, size=(n, 3))
    d = nrm + shell
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return org.astype(np.float32), d.astype(np.float32)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rays", type=int, default=20000)
    a = ap.parse_args()
    z = np.load(os.path.join(ROOT, "tests", "golden", "scene_random_seed1.npz"))
    cen = z["centers"].reshape(-1, 4)[:, :3].astype(np.float32)
    rad = z["radii"].astype(np.float32)
    keep = (np.abs(rad) <= 64) & (np.linalg.norm(cen, axis=1) <= 4096)
    C, r = cen[keep], rad[keep]
    rng = np.random.default_rng(5)
    O, D = make_rays(C.astype(np.float64), r.astype(np.float64), a.rays, rng)

    C64, r64, O64, D64 = C.astype(np.float64), r.astype(np.float64), O.astype(np.float64), D.astype(np.float64)
    tc_x = C64 @ D64.T - np.sum(O64 * D64, axis=1)[None, :]                      # (spheres, rays)
    v_x = tc_x ** 2 - (np.sum(C64 ** 2, axis=1) - r64 ** 2)[:, None] + 2 * C64 @ O64.T - np.sum(O64 ** 2, axis=1)[None, :]
    A2B2 = np.sum(C64 ** 2, axis=1)[:, None] + np.sum(O64 ** 2, axis=1)[None, :]
    true_hits = (v_x > 1e-3) & (tc_x > 1e-3)
    print(f"{len(r)} filterable spheres, {a.rays} rays, |C| <= {np.sqrt(np.sum(C64**2,axis=1)).max():.1f}, |O| <= {np.sqrt(np.sum(O64**2,axis=1)).max():.1f}, "
          f"true hits per ray {true_hits.sum() / a.rays:.3f}")

    def report(name, v):
        err = np.abs(v - v_x)
        worst = err.max()
        rel = (err / (U * A2B2)).max()
        slack = 2 * worst
        cand = ((v_x + slack > 1e-3)).sum() / a.rays  # discriminant-only verdict, as shipped
        print(f"{name:34s} worst |err| {worst:9.3g}  = {rel:8.1f} u(A^2+B^2)   candidates/ray at slack 2x worst: {cand:.3f}")

    # ---- fp32 FFMA chains as shipped (round to nearest) ----
    f = np.float32
    nod = -(np.sum(O * D, axis=1, dtype=np.float32))
    tc = np.zeros((len(r), a.rays), np.float32) + nod[None, :]
    na = np.zeros((len(r), a.rays), np.float32) + (-(np.sum(C64 ** 2, axis=1) - r64 ** 2)).astype(np.float32)[:, None]
    for i in range(3):
        tc = (tc.astype(np.float64) + C[:, i:i + 1].astype(np.float64) * D[None, :, i].astype(np.float64)).astype(f)
        na = (na.astype(np.float64) + C[:, i:i + 1].astype(np.float64) * (2 * O[None, :, i]).astype(np.float64)).astype(f)
    v32 = (tc.astype(np.float64) ** 2 + na.astype(np.float64)).astype(f)
    v32 = (v32.astype(np.float64) - np.sum(O.astype(np.float64) ** 2, axis=1).astype(f)[None, :]).astype(f)
    report("fp32 FFMA chains (shipped)", v32.astype(np.float64))
    shipped_slack = 96 * U * (A2B2 + r64[:, None] ** 2)
    print(f"{'':34s} shipped slack 96u(A^2+B^2+r^2): candidates/ray {((v_x + shipped_slack > 1e-3)).sum() / a.rays:.3f}")
    for mult in (160, 256, 512):  # what a worst-case bound for the tensor-core form would ask for
        sl = mult * U * (A2B2 + r64[:, None] ** 2)
        print(f"{'':34s} slack {mult}u(A^2+B^2+r^2): candidates/ray {((v_x + sl > 1e-3)).sum() / a.rays:.3f}")

    Ch, Cl = split(C)
    k_sph = (-(np.sum(C64 ** 2, axis=1) - r64 ** 2)).astype(np.float32)
    kh, kl = split(k_sph)
    for mode in ("rz", "f64"):
        # ---- A: two GEMMs ----
        Dh, Dl = split(D)
        Oh, Ol = split((2 * O).astype(np.float32))
        nodh, nodl = split(nod)
        q = (-np.sum(O64 ** 2, axis=1)).astype(np.float32)
        qh, ql = split(q)
        t_terms, n_terms = [], []
        for i in range(3):
            for cs in (Ch, Cl):
                for ds in (Dh, Dl):
                    t_terms.append(cs[:, i:i + 1].astype(np.float64) * ds[None, :, i].astype(np.float64))
                for os_ in (Oh, Ol):
                    n_terms.append(cs[:, i:i + 1].astype(np.float64) * os_[None, :, i].astype(np.float64))
        ones = np.ones((len(r), 1))
        t_terms += [ones * nodh[None, :].astype(np.float64), ones * nodl[None, :].astype(np.float64)]
        n_terms += [kh[:, None].astype(np.float64) * np.ones((1, a.rays)), kl[:, None].astype(np.float64) * np.ones((1, a.rays)),
                    ones * qh[None, :].astype(np.float64), ones * ql[None, :].astype(np.float64)]
        tcA = accumulate(t_terms, mode).astype(np.float32)
        naA = accumulate(n_terms, mode).astype(np.float32)
        vA = (tcA.astype(np.float64) ** 2 + naA.astype(np.float64)).astype(np.float32)
        report(f"A: two K=16 GEMMs, accumulate {mode}", vA.astype(np.float64))

        # ---- B: one GEMM, square expanded ----
        od = np.sum(O64 * D64, axis=1)
        W = (2 * O64 - 2 * od[:, None] * D64).astype(np.float32)            # linear weights per ray
        qB = (od ** 2 - np.sum(O64 ** 2, axis=1)).astype(np.float32)
        Wh, Wl = split(W)
        qBh, qBl = split(qB)
        terms = []
        for i in range(3):
            for j in range(i, 3):
                cc = (C64[:, i] * C64[:, j] * (1 if i == j else 2)).astype(np.float32)   # per-sphere constant (binary64 on the host)
                dd = (D[:, i].astype(np.float64) * D[:, j].astype(np.float64)).astype(np.float32)
                cch, ccl = split(cc)
                ddh, ddl = split(dd)
                for cs in (cch, ccl):
                    for ds in (ddh, ddl):
                        terms.append(cs[:, None].astype(np.float64) * ds[None, :].astype(np.float64))
        for i in range(3):
            for cs in (Ch, Cl):
                for ws in (Wh, Wl):
                    terms.append(cs[:, i:i + 1].astype(np.float64) * ws[None, :, i].astype(np.float64))
        terms += [kh[:, None].astype(np.float64) * np.ones((1, a.rays)), kl[:, None].astype(np.float64) * np.ones((1, a.rays)),
                  ones * qBh[None, :].astype(np.float64), ones * qBl[None, :].astype(np.float64)]
        vB = accumulate(terms, mode)
        report(f"B: one K={len(terms)} GEMM, accumulate {mode}", vB)


if __name__ == "__main__":
    main()
