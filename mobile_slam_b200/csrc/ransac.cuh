// (device code; kernels in ransac.cu and glue.cu)
// F-matrix RANSAC on the device: the outlier rejection of
// FeatureTracker::rejectWithFundamentalMatrix, /root/reference/src/frontend/feature_tracker.cpp:203-307
//   cv::findFundamentalMat(un_cur, un_next, FM_RANSAC, F_threshold, 0.99, status)   (:228-229)
//   + distance-aware edge recovery                                                     (:236-285)
// restated from OpenCV's published algorithm (calib3d fundam.cpp / ptsetreg.cpp): minimal 7-point samples, up to
// three real models per sample from det(l*F1 + (1-l)*F2) = 0, inlier test max(d1^2, d2^2) <= thr^2 on the
// point-to-epipolar-line distances in both images, iteration count adapted to the best inlier ratio at
// confidence 0.99 (cap 1000), output mask = inliers of the best model.
// OpenCV's sample order comes from its internal RNG, so masks are compared with cv2 statistically
// (tests/test_gpu_stages.py::test_ransac_*), not bitwise; the bit-for-bit path is VS_RANSAC_CALLBACK.
//
// One CTA per camera stream. A hypothesis is solved by a GROUP of 8 lanes (4 per warp, 32 per CTA round): lane j
// holds column j of the 7x9 system and every lane a copy of column 8, so the fp64 full-pivot reduction is seven
// rounds of {group arg-max, broadcast of the pivot column, one FMA per held element} instead of a 63-element
// serial elimination per thread.  The up-to-three models of a sample go to shared memory, each warp scores its
// own twelve model slots with its 32 lanes striding the stream's points (fp32), and the final mask is evaluated
// in fp64.  The iteration count follows OpenCV's rule after every round of 32 samples.
#pragma once
#include "glue.cuh"


namespace vs {

constexpr int kRansacThreads = 256;
constexpr int kRansacMaxIters = 1000;      // cv::findFundamentalMat default maxIters
constexpr int kRansacGroup = 8;            // lanes per hypothesis
constexpr int kRansacHyps = kRansacThreads / kRansacGroup;   // hypotheses per round
constexpr int kRansacWarps = kRansacThreads / 32;
constexpr int kRansacSlotsPerWarp = (32 / kRansacGroup) * 3;

__device__ __forceinline__ unsigned hash32(unsigned x) {
    x ^= x >> 16; x *= 0x7feb352du;
    x ^= x >> 15; x *= 0x846ca68bu;
    return x ^ (x >> 16);
}

__device__ __forceinline__ double det3(const double* f) {
    return f[0] * (f[4] * f[8] - f[5] * f[7]) - f[1] * (f[3] * f[8] - f[5] * f[6]) + f[2] * (f[3] * f[7] - f[4] * f[6]);
}

// real roots of c3 x^3 + c2 x^2 + c1 x + c0 (returns count, polished with two Newton steps)
__device__ inline int solve_cubic(double c3, double c2, double c1, double c0, double* r) {
    int n = 0;
    const double scale = fabs(c3) + fabs(c2) + fabs(c1) + fabs(c0);
    if (!(scale > 0.0)) return 0;
    if (fabs(c3) < 1e-12 * scale) {
        if (fabs(c2) < 1e-12 * scale) {
            if (fabs(c1) < 1e-12 * scale) return 0;
            r[n++] = -c0 / c1;
        } else {
            const double d = c1 * c1 - 4.0 * c2 * c0;
            if (d < 0.0) return 0;
            const double sq = sqrt(d), q = -0.5 * (c1 + (c1 >= 0.0 ? sq : -sq));
            r[n++] = q / c2;
            if (q != 0.0) r[n++] = c0 / q;
        }
    } else {
        const double a = c2 / c3, b = c1 / c3, c = c0 / c3;
        const double Q = (a * a - 3.0 * b) / 9.0, R = (2.0 * a * a * a - 9.0 * a * b + 27.0 * c) / 54.0;
        const double Q3 = Q * Q * Q, D = R * R - Q3;
        if (D < 0.0) {
            const double th = acos(fmax(-1.0, fmin(1.0, R / sqrt(Q3)))), sq = -2.0 * sqrt(Q);
            r[n++] = sq * cos(th / 3.0) - a / 3.0;
            r[n++] = sq * cos((th + 2.0 * 3.14159265358979323846) / 3.0) - a / 3.0;
            r[n++] = sq * cos((th - 2.0 * 3.14159265358979323846) / 3.0) - a / 3.0;
        } else {
            const double A = -copysign(cbrt(fabs(R) + sqrt(D)), R);
            const double B = A != 0.0 ? Q / A : 0.0;
            r[n++] = (A + B) - a / 3.0;
        }
    }
    for (int i = 0; i < n; ++i) {
        double x = r[i];
        for (int k = 0; k < 2; ++k) {
            const double f = ((c3 * x + c2) * x + c1) * x + c0, df = (3.0 * c3 * x + 2.0 * c2) * x + c1;
            if (fabs(df) > 1e-300) x -= f / df;
        }
        r[i] = x;
    }
    return n;
}

// Hartley normalisation of one image's points: x' = s (x - m)
struct Norm2 { double mx, my, s; };

// 7-point solver on normalised points, cooperative over a group of 8 lanes (j = lane within the group, gbase =
// the group's first lane; every lane of the warp must call it).  idx = the 7 sampled points (the same in all lanes
// of a group).  Writes up to three 3x3 models (row major, in the centred pixel frame the caller scores in) to
// slotF[0..2] with slot_valid[m] = 1 for the slots that hold a model.  a = points in image 1 (cur), b = image 2.
__device__ __forceinline__ void seven_point_group(const float2* pa, const float2* pb, const int* idx, const Norm2& na,
                                                  const Norm2& nb, int j, int gbase, double (*slotF)[9],
                                                  unsigned char* slot_valid) {
    constexpr unsigned FULL = 0xffffffffu;
    double A[7], B[7];                     // my column (j = 0..7) and column 8 of the system
    const int jd = j / 3, jm = j - 3 * jd;
#pragma unroll
    for (int r = 0; r < 7; ++r) {
        const float2 p = pa[idx[r]], q = pb[idx[r]];
        const double x1 = na.s * ((double)p.x - na.mx), y1 = na.s * ((double)p.y - na.my);
        const double x2 = nb.s * ((double)q.x - nb.mx), y2 = nb.s * ((double)q.y - nb.my);
        // row = [x2x1 x2y1 x2 | y2x1 y2y1 y2 | x1 y1 1]
        const double u = jd == 0 ? x2 : (jd == 1 ? y2 : 1.0);
        const double v = jm == 0 ? x1 : (jm == 1 ? y1 : 1.0);
        A[r] = u * v;
        B[r] = 1.0;
    }
    // reduced row echelon form with full pivoting.  Rows and columns stay where they are; used pivot rows / columns
    // are remembered in bit masks, so the step body does not depend on the step number and stays a rolled loop
    // (small code: the solver runs once per round, straight-line code would be fetched cold every time)
    int myrow = -1, brow = -1;             // row in which my column / column 8 is the pivot
    unsigned pmask = 0, rmask = 0;         // columns / rows used as pivots
    bool ok = true;
    const unsigned tiny = (unsigned)__double2hiint(1e-12);
#pragma unroll 1
    for (int k = 0; k < 7; ++k) {
        // arg-max |a| over the not yet used rows and columns; key = high word of |a| with (column, row) in
        // the low 7 bits (13 mantissa bits are plenty to pick a pivot)
        unsigned key = 0;
#pragma unroll
        for (int r = 0; r < 7; ++r) {
            const bool rfree = ((rmask >> r) & 1u) == 0;
            const unsigned ca = ((unsigned)__double2hiint(fabs(A[r])) & ~0x7Fu) | (unsigned)(j << 3 | r);
            const unsigned cb = ((unsigned)__double2hiint(fabs(B[r])) & ~0x7Fu) | (unsigned)(8 << 3 | r);
            key = max(key, (rfree && myrow < 0) ? ca : 0u);
            key = max(key, (rfree && brow < 0) ? cb : 0u);
        }
        key = max(key, __shfl_xor_sync(FULL, key, 1));
        key = max(key, __shfl_xor_sync(FULL, key, 2));
        key = max(key, __shfl_xor_sync(FULL, key, 4));
        const int pr = key & 7, pc = (key >> 3) & 15;
        ok = ok && (key & ~0x7Fu) >= tiny;                  // rank deficient sample otherwise
        double ak = A[0], bk = B[0];                        // my elements of the pivot row
#pragma unroll
        for (int r = 1; r < 7; ++r)
            if (pr == r) { ak = A[r]; bk = B[r]; }
        const int src = gbase + (pc & 7);
        double piv = __shfl_sync(FULL, ak, src);
        if (pc == 8) piv = bk;
        const double inv = 1.0 / piv;
        ak *= inv;
        bk *= inv;
#pragma unroll
        for (int r = 0; r < 7; ++r) {
            double f = __shfl_sync(FULL, A[r], src);
            if (pc == 8) f = B[r];
            if (r == pr) {
                A[r] = ak;
                B[r] = bk;
            } else {
                A[r] -= f * ak;
                B[r] -= f * bk;
            }
        }
        if (pc == j) myrow = pr;
        if (pc == 8) brow = pr;
        pmask |= 1u << pc;
        rmask |= 1u << pr;
    }
    // null space basis: the two columns that never became pivots are the free variables
    const unsigned fm = ~pmask & 0x1FFu;
    const int free1 = __ffs(fm) - 1, free2 = 31 - __clz(fm);        // free1 < free2, free1 <= 7
    double e1 = 0.0, e2 = 0.0, e81 = 0.0, e82 = 0.0;                // entries j and 8 of the two basis vectors
#pragma unroll
    for (int r = 0; r < 7; ++r) {
        const double v1 = __shfl_sync(FULL, A[r], gbase + free1);
        double v2 = __shfl_sync(FULL, A[r], gbase + (free2 & 7));
        if (free2 == 8) v2 = B[r];
        if (myrow == r) { e1 = -v1; e2 = -v2; }
        if (brow == r) { e81 = -v1; e82 = -v2; }
    }
    if (j == free1) { e1 = 1.0; e2 = 0.0; }
    if (j == free2) { e1 = 0.0; e2 = 1.0; }
    if (free2 == 8) { e81 = 0.0; e82 = 1.0; }
    double f1[9], f2[9];
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        f1[c] = __shfl_sync(FULL, e1, gbase + c);
        f2[c] = __shfl_sync(FULL, e2, gbase + c);
    }
    f1[8] = e81;
    f2[8] = e82;
    // det(l f1 + (1 - l) f2) = c3 l^3 + c2 l^2 + c1 l + c0 from four evaluations
    double m[9];
    const double d0 = det3(f2), d1 = det3(f1);
#pragma unroll
    for (int c = 0; c < 9; ++c) m[c] = 2.0 * f2[c] - f1[c];          // l = -1
    const double dm1 = det3(m);
#pragma unroll
    for (int c = 0; c < 9; ++c) m[c] = 2.0 * f1[c] - f2[c];          // l = 2
    const double d2 = det3(m);
    const double c0 = d0, c2 = 0.5 * (d1 + dm1) - c0, e = 0.5 * (d1 - dm1);
    const double c3 = (d2 - c0 - 4.0 * c2 - 2.0 * e) / 6.0, c1 = e - c3;
    double roots[3] = {0.0, 0.0, 0.0};
    const int nr = ok ? solve_cubic(c3, c2, c1, c0, roots) : 0;
    if (j < 3) {                                                     // lane m of the group finishes root m
        const double l = j == 0 ? roots[0] : (j == 1 ? roots[1] : roots[2]);
        bool have = j < nr && isfinite(l);
        double Fh[9];
#pragma unroll
        for (int c = 0; c < 9; ++c) Fh[c] = l * f1[c] + (1.0 - l) * f2[c];
        // F = Tb^T Fh Ta with T = [s 0 -s mx; 0 s -s my; 0 0 1]
        double G[9], F[9];
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            G[3 * r + 0] = na.s * Fh[3 * r + 0];
            G[3 * r + 1] = na.s * Fh[3 * r + 1];
            G[3 * r + 2] = Fh[3 * r + 2] - na.s * na.mx * Fh[3 * r + 0] - na.s * na.my * Fh[3 * r + 1];
        }
        double nrm = 0.0;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            F[c] = nb.s * G[c];
            F[3 + c] = nb.s * G[3 + c];
            F[6 + c] = G[6 + c] - nb.s * nb.mx * G[c] - nb.s * nb.my * G[3 + c];
        }
#pragma unroll
        for (int c = 0; c < 9; ++c) nrm += F[c] * F[c];
        have = have && nrm > 0.0 && isfinite(nrm);
        const double inv = have ? 1.0 / sqrt(nrm) : 0.0;
#pragma unroll
        for (int c = 0; c < 9; ++c) slotF[j][c] = F[c] * inv;
        slot_valid[j] = have ? 1 : 0;
    }
}

// cv::RANSACUpdateNumIters (ptsetreg.cpp): iterations needed at confidence p for outlier ratio ep
__device__ inline int ransac_update_num_iters(double p, double ep, int model_points, int max_iters) {
    p = fmax(p, 0.0); p = fmin(p, 1.0);
    ep = fmax(ep, 0.0); ep = fmin(ep, 1.0);
    const double num = fmax(1.0 - p, 2.2250738585072014e-308);
    const double denom = 1.0 - pow(1.0 - ep, (double)model_points);
    if (denom < 2.2250738585072014e-308) return 0;
    const double ln = log(num), ld = log(denom);
    return (ld >= 0.0 || -ln >= (double)max_iters * (-ld)) ? max_iters : (int)rint(ln / ld);
}

// sums NV doubles per thread over the CTA (result in every thread)
template <int NV>
__device__ __forceinline__ void block_sums(double* v, double (*s_red)[4], int tid) {
#pragma unroll
    for (int q = 0; q < NV; ++q)
#pragma unroll
        for (int k = 16; k > 0; k >>= 1) v[q] += __shfl_xor_sync(0xffffffffu, v[q], k);
    __syncthreads();
    if ((tid & 31) == 0)
#pragma unroll
        for (int q = 0; q < NV; ++q) s_red[tid >> 5][q] = v[q];
    __syncthreads();
#pragma unroll
    for (int q = 0; q < NV; ++q) {
        double t = 0.0;
#pragma unroll
        for (int k = 0; k < kRansacWarps; ++k) t += s_red[k][q];
        v[q] = t;
    }
}

// Final flags for a given model F (row-major, x2^T F x1 = 0): inlier iff the larger of the two point-to-epipolar-line
// distances is within f_threshold (what cv::findFundamentalMat's RANSAC reports), then the reference's distance-aware edge
// recovery (feature_tracker.cpp:236-285) for the rejected points.  keep_base = true takes the inlier flags already in
// w_rstatus (a teacher's, e.g. cv2's own mask) and applies only the recovery rule.
__device__ __forceinline__ void ransac_mark(const GlueParams& g, int64_t o, int n, const double* F, const float2* pa,
                                            const float2* pb, bool keep_base, int tid, int nthreads) {
    const double t2 = g.f_threshold * g.f_threshold;
    const double ef = g.f_edge_factor;
    const double dcx = (double)g.W / 2.0, dcy = (double)g.H / 2.0;
    const double r_max = sqrt(dcx * dcx + dcy * dcy), r_max_inv = r_max > 1e-6 ? 1.0 / r_max : 0.0;
    for (int i = tid; i < n; i += nthreads) {
        const double x1 = pa[i].x, y1 = pa[i].y, x2 = pb[i].x, y2 = pb[i].y;
        const double a = F[0] * x1 + F[1] * y1 + F[2], b = F[3] * x1 + F[4] * y1 + F[5], c = F[6] * x1 + F[7] * y1 + F[8];
        const double d = x2 * a + y2 * b + c;
        const double a2 = F[0] * x2 + F[3] * y2 + F[6], b2 = F[1] * x2 + F[4] * y2 + F[7];
        const double smin = fmin(a * a + b * b, a2 * a2 + b2 * b2);
        bool in = keep_base ? (g.w_rstatus[o + i] != 0) : (d * d <= t2 * smin);
        if (!in && ef > 0.0) {                   // edge recovery, feature_tracker.cpp:236-285
            const float2 px = g.w_next[o + i];
            const double dx = (double)px.x - dcx, dy = (double)px.y - dcy;
            const double rr = sqrt(dx * dx + dy * dy) * r_max_inv;
            const double nab = sqrt(a * a + b * b);
            if (rr >= 0.3 && nab >= 1e-12) in = fabs(d) / nab < g.f_threshold * (1.0 + ef * rr * rr);
        }
        g.w_rstatus[o + i] = in ? 1 : 0;
    }
}

// One stream's RANSAC; every thread of a 256-thread CTA calls it (dsm = 16 * cap bytes of dynamic shared memory).
__device__ __forceinline__ void ransac_body(const GlueParams& g, const unsigned int seed, const int s, const int tid, uint8_t* dsm) {
    float2* pa = reinterpret_cast<float2*>(dsm);          // [cap] image-1 (cur) points, centred
    float2* pb = pa + g.cap;                              // [cap] image-2 (next)
    __shared__ double s_red[kRansacWarps][4];
    __shared__ double s_F[kRansacWarps * kRansacSlotsPerWarp][9];   // this round's models (warp-private slots)
    __shared__ unsigned char s_valid[kRansacWarps * kRansacSlotsPerWarp];
    __shared__ double s_bestF[kRansacWarps][9];           // each warp's best model so far
    __shared__ unsigned long long s_key[2][kRansacWarps];
    const int lane = tid & 31, warp = tid >> 5;
    if (!g.ransac_flag[s]) return;
    const int n = g.wn[s];
    const int64_t o = (int64_t)s * g.cap;
    const float cx = (float)((double)g.W / 2.0), cy = (float)((double)g.H / 2.0);
    double sm[4] = {0.0, 0.0, 0.0, 0.0};
    for (int i = tid; i < n; i += kRansacThreads) {
        const float2 a = g.w_lcur[o + i], b = g.w_lnext[o + i];
        pa[i] = make_float2(a.x - cx, a.y - cy);
        pb[i] = make_float2(b.x - cx, b.y - cy);
        sm[0] += pa[i].x; sm[1] += pa[i].y; sm[2] += pb[i].x; sm[3] += pb[i].y;
    }
    block_sums<4>(sm, s_red, tid);
    Norm2 na, nb;
    na.mx = sm[0] / n; na.my = sm[1] / n;
    nb.mx = sm[2] / n; nb.my = sm[3] / n;
    double dd[2] = {0.0, 0.0};
    for (int i = tid; i < n; i += kRansacThreads) {
        dd[0] += sqrt((pa[i].x - na.mx) * (pa[i].x - na.mx) + (pa[i].y - na.my) * (pa[i].y - na.my));
        dd[1] += sqrt((pb[i].x - nb.mx) * (pb[i].x - nb.mx) + (pb[i].y - nb.my) * (pb[i].y - nb.my));
    }
    block_sums<2>(dd, s_red, tid);
    const double d1 = dd[0] / n, d2 = dd[1] / n;
    na.s = d1 > 1e-12 ? 1.4142135623730951 / d1 : 1.0;
    nb.s = d2 > 1e-12 ? 1.4142135623730951 / d2 : 1.0;

    const float thr2 = (float)(g.f_threshold * g.f_threshold);
    // The sample sequence is seeded from the stream's own points, not from its position in a batch, a lane or a GPU:
    // a stream's mask is the same wherever and with whatever neighbours it is tracked (`seed` only salts it).
    const unsigned seed0 = hash32((unsigned)n * 0x9E3779B1u ^ __float_as_uint(pa[0].x) ^ hash32(__float_as_uint(pb[0].y)) ^
                                  hash32(__float_as_uint(pa[n - 1].y) + 0x85ebca6bu) ^ hash32(__float_as_uint(pb[n >> 1].x) + seed));
    const int j = lane & (kRansacGroup - 1), gbase = lane & ~(kRansacGroup - 1), grp = lane / kRansacGroup;
    const int slot0 = warp * kRansacSlotsPerWarp;
    unsigned long long wbest = 0ull, best = 0ull;         // warp's / CTA's best (inliers << 32 | ~model id)
    int need = kRansacMaxIters, done = 0, par = 0;
    while (done < need) {
        const int h = done + warp * (32 / kRansacGroup) + grp;
        unsigned st = hash32(seed0 + (unsigned)h * 0x9E3779B1u);
        int idx[7];
#pragma unroll
        for (int k = 0; k < 7; ++k) {
            for (;;) {
                st ^= st << 13; st ^= st >> 17; st ^= st << 5;               // xorshift32
                const int c = (int)__umulhi(st, (unsigned)n);
                bool dup = false;
#pragma unroll
                for (int q = 0; q < k; ++q) dup |= idx[q] == c;
                if (!dup) { idx[k] = c; break; }
            }
        }
        seven_point_group(pa, pb, idx, na, nb, j, gbase, &s_F[slot0 + grp * 3], &s_valid[slot0 + grp * 3]);
        __syncwarp();
        for (int sl = 0; sl < kRansacSlotsPerWarp; ++sl) {
            if (!s_valid[slot0 + sl]) continue;                              // warp-uniform
            float f[9];
#pragma unroll
            for (int c = 0; c < 9; ++c) f[c] = (float)s_F[slot0 + sl][c];
            int cnt = 0;
            for (int i = lane; i < n; i += 32) {
                const float2 p = pa[i], q = pb[i];
                const float a = f[0] * p.x + f[1] * p.y + f[2], b = f[3] * p.x + f[4] * p.y + f[5];
                const float c = f[6] * p.x + f[7] * p.y + f[8];
                const float d = q.x * a + q.y * b + c;               // x2^T F x1 (the same scalar for both lines)
                const float a2 = f[0] * q.x + f[3] * q.y + f[6], b2 = f[1] * q.x + f[4] * q.y + f[7];
                const float smin = fminf(a * a + b * b, a2 * a2 + b2 * b2);
                cnt += (d * d <= thr2 * smin) ? 1 : 0;               // max(d1^2, d2^2) <= thr^2
            }
            cnt = __reduce_add_sync(0xffffffffu, cnt);
            const unsigned id = (unsigned)((done + warp * (32 / kRansacGroup) + sl / 3) * 4 + sl % 3);
            const unsigned long long key = ((unsigned long long)cnt << 32) | (0xFFFFFFFFu - id);
            if (key > wbest) {
                wbest = key;
                if (lane < 9) s_bestF[warp][lane] = s_F[slot0 + sl][lane];
            }
        }
        if (lane == 0) s_key[par][warp] = wbest;
        __syncthreads();
        unsigned long long b = 0ull;
#pragma unroll
        for (int w = 0; w < kRansacWarps; ++w) b = s_key[par][w] > b ? s_key[par][w] : b;
        par ^= 1;
        done += kRansacHyps;
        if (b != best) {
            best = b;
            const int best_cnt = (int)(best >> 32);
            need = ransac_update_num_iters(0.99, (double)(n - best_cnt) / (double)n, 7, need);
        }
    }
    const int best_cnt = (int)(best >> 32);
    if (best_cnt < 7) return;                    // no model: every point stays (w_rstatus was preset to 1)
    int wb = 0;
#pragma unroll
    for (int w = 1; w < kRansacWarps; ++w) wb = s_key[par ^ 1][w] == best ? w : wb;
    double F[9];
    for (int c = 0; c < 9; ++c) F[c] = s_bestF[wb][c];
    ransac_mark(g, o, n, F, pa, pb, false, tid, kRansacThreads);
}


}  // namespace vs
