// K11 device math — common::camera_models::Camera::liftProjective in float64.
//   PINHOLE         /root/reference/src/common/camera_models/PinholeCamera.cc:358-410, distortion :532-547
//   KANNALA_BRANDT  EquidistantCamera.cc:337-349, backprojectSymmetric :574-652
//   MEI             CataCamera.cc:444-503, distortion :629-644
//   SCARAMUZZA      ScaramuzzaCamera.cc:575-594
// Products and sums are written with __dmul_rn/__dadd_rn where the reference's x86-64 build (no FMA,
// CMakeLists.txt:22-30) evaluates them separately, so results match an IEEE-754 double evaluation.
#pragma once
#include "vs_common.cuh"

namespace vs {

__device__ __forceinline__ double dm(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double da(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double ds(double a, double b) { return __dsub_rn(a, b); }

// d_u of PinholeCamera::distortion / CataCamera::distortion
__device__ __forceinline__ void radtan_distortion(const CameraDev& c, double ux, double uy, double& dx,
                                                  double& dy) {
    const double k1 = c.d[0], k2 = c.d[1], p1 = c.d[2], p2 = c.d[3];
    const double mx2 = dm(ux, ux), my2 = dm(uy, uy), mxy = dm(ux, uy);
    const double rho2 = da(mx2, my2);
    const double rad = da(dm(k1, rho2), dm(dm(k2, rho2), rho2));
    dx = da(da(dm(ux, rad), dm(dm(2.0, p1), mxy)), dm(p2, da(rho2, dm(2.0, mx2))));
    dy = da(da(dm(uy, rad), dm(dm(2.0, p2), mxy)), dm(p1, da(rho2, dm(2.0, my2))));
}

__device__ __forceinline__ void radtan_undistort(const CameraDev& c, double mx_d, double my_d, double& mx_u,
                                                 double& my_u) {
    if (c.no_distortion) {
        mx_u = mx_d;
        my_u = my_d;
        return;
    }
    double dx, dy;
    radtan_distortion(c, mx_d, my_d, dx, dy);
    mx_u = ds(mx_d, dx);
    my_u = ds(my_d, dy);
#pragma unroll 1
    for (int i = 1; i < 8; ++i) {   // "Recursive distortion model", n = 8
        radtan_distortion(c, mx_u, my_u, dx, dy);
        mx_u = ds(mx_d, dx);
        my_u = ds(my_d, dy);
    }
}

// theta + k2 theta^3 + k3 theta^5 + k4 theta^7 + k5 theta^9 - r
__device__ __forceinline__ double kb_poly(const CameraDev& c, double t, double r) {
    const double t2 = t * t;
    return t * (1.0 + t2 * (c.d[0] + t2 * (c.d[1] + t2 * (c.d[2] + t2 * c.d[3])))) - r;
}
__device__ __forceinline__ double kb_dpoly(const CameraDev& c, double t) {
    const double t2 = t * t;
    return 1.0 + t2 * (3.0 * c.d[0] + t2 * (5.0 * c.d[1] + t2 * (7.0 * c.d[2] + t2 * 9.0 * c.d[3])));
}

// Smallest real root >= -1e-10 (clamped to 0) of the odd polynomial; the reference takes it from the
// eigenvalues of the companion matrix.  Newton from theta = r, then a sign scan of [0, theta) guards
// against calibrations where an earlier root exists; bisection refines it if so.
__device__ inline double kb_theta(const CameraDev& c, double r) {
    if (c.kb_npow == 1) return r;
    double t = r;
#pragma unroll 1
    for (int i = 0; i < 40; ++i) {
        const double df = kb_dpoly(c, t);
        if (df == 0.0) break;
        const double step = kb_poly(c, t, r) / df;
        t -= step;
        if (fabs(step) <= 1e-16 * fmax(1.0, fabs(t))) break;
    }
    bool ok = isfinite(t) && t >= -1e-10 && fabs(kb_poly(c, t, r)) <= 1e-9 * fmax(1.0, r);
    const double hi = ok ? t : fmax(4.0, 2.0 * r);
    // with a converged Newton root the scan below only matters if f already changed sign before it: test the 31
    // interior nodes first (independent evaluations, unrolled), and walk them one by one only in that rare case
    bool scan = !ok;
    if (ok && r > 0.0) {
#pragma unroll 8
        for (int j = 1; j < 32; ++j) scan |= kb_poly(c, hi * (double)j / 32.0, r) >= 0.0;
    }
    if (r > 0.0 && scan) {
        double a = 0.0, fa = -r;   // f(0) = -r < 0
#pragma unroll 1
        for (int j = 1; j <= 32; ++j) {
            const double b = hi * (double)j / 32.0;
            const double fb = kb_poly(c, b, r);
            if (fb >= 0.0 && (j < 32 || !ok)) {
                double lo = a, up = b;
#pragma unroll 1
                for (int k = 0; k < 100; ++k) {
                    const double m = 0.5 * (lo + up);
                    if (m == lo || m == up) break;
                    if (kb_poly(c, m, r) >= 0.0) up = m; else lo = m;
                }
                // polish with two Newton steps inside the bracket
                double x = 0.5 * (lo + up);
                for (int k = 0; k < 2; ++k) {
                    const double df = kb_dpoly(c, x);
                    if (df != 0.0) {
                        const double xn = x - kb_poly(c, x, r) / df;
                        if (xn >= a && xn <= b) x = xn;
                    }
                }
                return x;
            }
            a = b;
            fa = fb;
        }
        (void)fa;
    }
    if (!ok) return r;              // "thetas.empty()" fallback
    return t < 0.0 ? 0.0 : t;
}

__device__ inline void lift_projective(const CameraDev& c, double u, double v, double& X, double& Y, double& Z) {
    if (c.model == 3) {   // SCARAMUZZA
        const double xc0 = u - c.cx, xc1 = v - c.cy;
        const double xa0 = dm(c.inv_scale, ds(xc0, dm(c.aff_d, xc1)));
        const double xa1 = dm(c.inv_scale, da(dm(-c.aff_e, xc0), dm(c.aff_c, xc1)));
        const double phi = sqrt(da(dm(xa0, xa0), dm(xa1, xa1)));
        double phi_i = 1.0, z = 0.0;
#pragma unroll
        for (int i = 0; i < 5; ++i) {
            z = da(z, dm(phi_i, c.poly[i]));
            phi_i = dm(phi_i, phi);
        }
        X = xc0; Y = xc1; Z = -z;
        return;
    }
    const double mx = da(dm(c.inv_K11, u), c.inv_K13);
    const double my = da(dm(c.inv_K22, v), c.inv_K23);
    if (c.model == 2) {   // PINHOLE
        radtan_undistort(c, mx, my, X, Y);
        Z = 1.0;
        return;
    }
    if (c.model == 1) {   // MEI
        radtan_undistort(c, mx, my, X, Y);
        const double xi = c.xi;
        if (xi == 1.0) {
            Z = ds(ds(1.0, dm(X, X)), dm(Y, Y)) / 2.0;
        } else {
            const double rho2 = da(dm(X, X), dm(Y, Y));
            const double den = da(xi, sqrt(da(1.0, dm(ds(1.0, dm(xi, xi)), rho2))));
            Z = ds(1.0, dm(xi, da(rho2, 1.0)) / den);
        }
        return;
    }
    // KANNALA_BRANDT
    const double r = sqrt(da(dm(mx, mx), dm(my, my)));
    const double phi = r < 1e-10 ? 0.0 : atan2(my, mx);
    const double theta = kb_theta(c, r);
    double st, ct, sp, cp;
    sincos(theta, &st, &ct);
    sincos(phi, &sp, &cp);
    X = dm(st, cp);
    Y = dm(st, sp);
    Z = ct;
}

}  // namespace vs
