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
// One CTA per camera stream, one hypothesis per thread per round (256 per round): the 7x9 system is reduced in
// fp64 on Hartley-normalised points, candidate models are scored in fp32 over the stream's points broadcast
// from shared memory, and the final mask is evaluated in fp64.
#include "tracker_state.h"

namespace vs {

constexpr int kRansacThreads = 256;
constexpr int kRansacMaxIters = 1000;      // cv::findFundamentalMat default maxIters

__device__ __forceinline__ unsigned long long splitmix64(unsigned long long& x) {
    x += 0x9E3779B97F4A7C15ull;
    unsigned long long z = x;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

__device__ __forceinline__ double det3(const double* f) {
    return f[0] * (f[4] * f[8] - f[5] * f[7]) - f[1] * (f[3] * f[8] - f[5] * f[6]) + f[2] * (f[3] * f[7] - f[4] * f[6]);
}

// real roots of c3 x^3 + c2 x^2 + c1 x + c0 (returns count, polished with two Newton steps)
__device__ int solve_cubic(double c3, double c2, double c1, double c0, double* r) {
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

// 7-point solver on normalised points: fills up to three 3x3 models (row major, in the centred pixel frame the
// caller scores in) and returns their number.  a = points in image 1 (cur), b = image 2 (next).
__device__ int seven_point(const float2* pa, const float2* pb, const int* idx, const Norm2& na, const Norm2& nb,
                           double (*Fout)[9]) {
    double A[7][9];
    for (int k = 0; k < 7; ++k) {
        const float2 p = pa[idx[k]], q = pb[idx[k]];
        const double x1 = na.s * ((double)p.x - na.mx), y1 = na.s * ((double)p.y - na.my);
        const double x2 = nb.s * ((double)q.x - nb.mx), y2 = nb.s * ((double)q.y - nb.my);
        A[k][0] = x2 * x1; A[k][1] = x2 * y1; A[k][2] = x2;
        A[k][3] = y2 * x1; A[k][4] = y2 * y1; A[k][5] = y2;
        A[k][6] = x1;      A[k][7] = y1;      A[k][8] = 1.0;
    }
    // reduced row echelon form with full pivoting (column order tracked in `col`)
    int col[9];
    for (int j = 0; j < 9; ++j) col[j] = j;
    for (int k = 0; k < 7; ++k) {
        int pr = k, pc = k;
        double best = 0.0;
        for (int r = k; r < 7; ++r)
            for (int c = k; c < 9; ++c) {
                const double v = fabs(A[r][c]);
                if (v > best) { best = v; pr = r; pc = c; }
            }
        if (best < 1e-12) return 0;                       // rank deficient sample
        if (pr != k) for (int c = 0; c < 9; ++c) { const double t = A[k][c]; A[k][c] = A[pr][c]; A[pr][c] = t; }
        if (pc != k) {
            for (int r = 0; r < 7; ++r) { const double t = A[r][k]; A[r][k] = A[r][pc]; A[r][pc] = t; }
            const int t = col[k]; col[k] = col[pc]; col[pc] = t;
        }
        const double inv = 1.0 / A[k][k];
        for (int c = k; c < 9; ++c) A[k][c] *= inv;
        for (int r = 0; r < 7; ++r) {
            if (r == k) continue;
            const double f = A[r][k];
            if (f != 0.0) for (int c = k; c < 9; ++c) A[r][c] -= f * A[k][c];
        }
    }
    // null space basis: free variables are the (permuted) columns 7 and 8
    double f1[9], f2[9];
    for (int k = 0; k < 7; ++k) { f1[col[k]] = -A[k][7]; f2[col[k]] = -A[k][8]; }
    f1[col[7]] = 1.0; f1[col[8]] = 0.0;
    f2[col[7]] = 0.0; f2[col[8]] = 1.0;
    // det(l f1 + (1 - l) f2) = c3 l^3 + c2 l^2 + c1 l + c0 from four evaluations
    double m[9];
    const double d0 = det3(f2), d1 = det3(f1);
    for (int j = 0; j < 9; ++j) m[j] = 2.0 * f2[j] - f1[j];          // l = -1
    const double dm1 = det3(m);
    for (int j = 0; j < 9; ++j) m[j] = 2.0 * f1[j] - f2[j];          // l = 2
    const double d2 = det3(m);
    const double c0 = d0, c2 = 0.5 * (d1 + dm1) - c0, e = 0.5 * (d1 - dm1);
    const double c3 = (d2 - c0 - 4.0 * c2 - 2.0 * e) / 6.0, c1 = e - c3;
    double roots[3];
    const int nr = solve_cubic(c3, c2, c1, c0, roots);
    int nm = 0;
    for (int i = 0; i < nr; ++i) {
        const double l = roots[i];
        if (!isfinite(l)) continue;
        double Fh[9];
        for (int j = 0; j < 9; ++j) Fh[j] = l * f1[j] + (1.0 - l) * f2[j];
        // F = Tb^T Fh Ta with T = [s 0 -s mx; 0 s -s my; 0 0 1]
        double G[9];
        for (int r = 0; r < 3; ++r) {
            G[3 * r + 0] = na.s * Fh[3 * r + 0];
            G[3 * r + 1] = na.s * Fh[3 * r + 1];
            G[3 * r + 2] = Fh[3 * r + 2] - na.s * na.mx * Fh[3 * r + 0] - na.s * na.my * Fh[3 * r + 1];
        }
        double* F = Fout[nm];
        double nrm = 0.0;
        for (int c = 0; c < 3; ++c) {
            F[c] = nb.s * G[c];
            F[3 + c] = nb.s * G[3 + c];
            F[6 + c] = G[6 + c] - nb.s * nb.mx * G[c] - nb.s * nb.my * G[3 + c];
        }
        for (int j = 0; j < 9; ++j) nrm += F[j] * F[j];
        if (!(nrm > 0.0) || !isfinite(nrm)) continue;
        const double inv = 1.0 / sqrt(nrm);
        for (int j = 0; j < 9; ++j) F[j] *= inv;
        ++nm;
    }
    return nm;
}


// cv::RANSACUpdateNumIters (ptsetreg.cpp): iterations needed at confidence p for outlier ratio ep
__device__ int ransac_update_num_iters(double p, double ep, int model_points, int max_iters) {
    p = fmax(p, 0.0); p = fmin(p, 1.0);
    ep = fmax(ep, 0.0); ep = fmin(ep, 1.0);
    const double num = fmax(1.0 - p, 2.2250738585072014e-308);
    const double denom = 1.0 - pow(1.0 - ep, (double)model_points);
    if (denom < 2.2250738585072014e-308) return 0;
    const double ln = log(num), ld = log(denom);
    return (ld >= 0.0 || -ln >= (double)max_iters * (-ld)) ? max_iters : (int)rint(ln / ld);
}

__device__ __forceinline__ double block_sum(double v, double* s_red, int tid) {
#pragma unroll
    for (int k = 16; k > 0; k >>= 1) v += __shfl_xor_sync(0xffffffffu, v, k);
    __syncthreads();
    if ((tid & 31) == 0) s_red[tid >> 5] = v;
    __syncthreads();
    double t = 0.0;
    for (int k = 0; k < kRansacThreads / 32; ++k) t += s_red[k];
    return t;
}

__global__ void __launch_bounds__(kRansacThreads) ransac_kernel(GlueParams g, unsigned int seed) {
    extern __shared__ __align__(16) uint8_t dsm[];
    float2* pa = reinterpret_cast<float2*>(dsm);          // [cap] image-1 (cur) points, centred
    float2* pb = pa + g.cap;                              // [cap] image-2 (next)
    __shared__ double s_red[kRansacThreads / 32];
    __shared__ unsigned long long s_key[kRansacThreads / 32];
    __shared__ double s_F[9];
    __shared__ unsigned long long s_best;
    const int s = blockIdx.x, tid = threadIdx.x;
    if (!g.ransac_flag[s]) return;
    const int n = g.wn[s];
    const int64_t o = (int64_t)s * g.cap;
    const float cx = (float)((double)g.W / 2.0), cy = (float)((double)g.H / 2.0);
    double sx1 = 0, sy1 = 0, sx2 = 0, sy2 = 0;
    for (int i = tid; i < n; i += kRansacThreads) {
        const float2 a = g.w_lcur[o + i], b = g.w_lnext[o + i];
        pa[i] = make_float2(a.x - cx, a.y - cy);
        pb[i] = make_float2(b.x - cx, b.y - cy);
        sx1 += pa[i].x; sy1 += pa[i].y; sx2 += pb[i].x; sy2 += pb[i].y;
    }
    Norm2 na, nb;
    na.mx = block_sum(sx1, s_red, tid) / n; na.my = block_sum(sy1, s_red, tid) / n;
    nb.mx = block_sum(sx2, s_red, tid) / n; nb.my = block_sum(sy2, s_red, tid) / n;
    double d1 = 0, d2 = 0;
    for (int i = tid; i < n; i += kRansacThreads) {
        d1 += sqrt((pa[i].x - na.mx) * (pa[i].x - na.mx) + (pa[i].y - na.my) * (pa[i].y - na.my));
        d2 += sqrt((pb[i].x - nb.mx) * (pb[i].x - nb.mx) + (pb[i].y - nb.my) * (pb[i].y - nb.my));
    }
    d1 = block_sum(d1, s_red, tid) / n;
    d2 = block_sum(d2, s_red, tid) / n;
    na.s = d1 > 1e-12 ? 1.4142135623730951 / d1 : 1.0;
    nb.s = d2 > 1e-12 ? 1.4142135623730951 / d2 : 1.0;

    const float thr2 = (float)(g.f_threshold * g.f_threshold);
    unsigned long long my_key = 0ull;
    double my_F[9];
    for (int j = 0; j < 9; ++j) my_F[j] = 0.0;
    int need = kRansacMaxIters, done = 0;
    while (done < need) {
        const int h = done + tid;
        unsigned long long st = ((unsigned long long)seed << 32) ^ ((unsigned long long)(unsigned)s * 0x9E3779B1ull) ^
                                ((unsigned long long)(unsigned)h << 20);
        int idx[7];
        for (int k = 0; k < 7; ++k) {
            for (;;) {
                const int c = (int)(((splitmix64(st) >> 32) * (unsigned long long)n) >> 32);
                bool dup = false;
                for (int q = 0; q < k; ++q) dup |= idx[q] == c;
                if (!dup) { idx[k] = c; break; }
            }
        }
        double F3[3][9];
        const int nm = seven_point(pa, pb, idx, na, nb, F3);
        for (int m = 0; m < nm; ++m) {
            float f[9];
            for (int j = 0; j < 9; ++j) f[j] = (float)F3[m][j];
            int cnt = 0;
            for (int i = 0; i < n; ++i) {
                const float2 p = pa[i], q = pb[i];                  // broadcast reads
                const float a = f[0] * p.x + f[1] * p.y + f[2], b = f[3] * p.x + f[4] * p.y + f[5];
                const float c = f[6] * p.x + f[7] * p.y + f[8];
                const float d = q.x * a + q.y * b + c;               // x2^T F x1 (the same scalar for both lines)
                const float a2 = f[0] * q.x + f[3] * q.y + f[6], b2 = f[1] * q.x + f[4] * q.y + f[7];
                const float smin = fminf(a * a + b * b, a2 * a2 + b2 * b2);
                cnt += (d * d <= thr2 * smin) ? 1 : 0;               // max(d1^2, d2^2) <= thr^2
            }
            const unsigned long long key = ((unsigned long long)cnt << 32) | (0xFFFFFFFFu - (unsigned)(h * 4 + m));
            if (key > my_key) {
                my_key = key;
                for (int j = 0; j < 9; ++j) my_F[j] = F3[m][j];
            }
        }
        unsigned long long k = my_key;
#pragma unroll
        for (int sft = 16; sft > 0; sft >>= 1) {
            const unsigned long long other = __shfl_xor_sync(0xffffffffu, k, sft);
            k = other > k ? other : k;
        }
        __syncthreads();
        if ((tid & 31) == 0) s_key[tid >> 5] = k;
        __syncthreads();
        if (tid == 0) {
            unsigned long long b = 0ull;
            for (int w = 0; w < kRansacThreads / 32; ++w) b = s_key[w] > b ? s_key[w] : b;
            s_best = b;
        }
        __syncthreads();
        done += kRansacThreads;
        const int best_cnt = (int)(s_best >> 32);
        need = ransac_update_num_iters(0.99, (double)(n - best_cnt) / (double)n, 7, need);
    }
    const unsigned long long best = s_best;
    const int best_cnt = (int)(best >> 32);
    if (best_cnt < 7) return;                    // no model: every point stays (w_rstatus was preset to 1)
    if (my_key == best)
        for (int j = 0; j < 9; ++j) s_F[j] = my_F[j];
    __syncthreads();
    double F[9];
    for (int j = 0; j < 9; ++j) F[j] = s_F[j];
    const double t2 = g.f_threshold * g.f_threshold;
    const double ef = g.f_edge_factor;
    const double dcx = (double)g.W / 2.0, dcy = (double)g.H / 2.0;
    const double r_max = sqrt(dcx * dcx + dcy * dcy), r_max_inv = r_max > 1e-6 ? 1.0 / r_max : 0.0;
    for (int i = tid; i < n; i += kRansacThreads) {
        const double x1 = pa[i].x, y1 = pa[i].y, x2 = pb[i].x, y2 = pb[i].y;
        const double a = F[0] * x1 + F[1] * y1 + F[2], b = F[3] * x1 + F[4] * y1 + F[5], c = F[6] * x1 + F[7] * y1 + F[8];
        const double d = x2 * a + y2 * b + c;
        const double a2 = F[0] * x2 + F[3] * y2 + F[6], b2 = F[1] * x2 + F[4] * y2 + F[7];
        const double smin = fmin(a * a + b * b, a2 * a2 + b2 * b2);
        bool in = d * d <= t2 * smin;
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

static int ransac_smem(int cap) { return cap * 16; }

int launch_ransac(vs_tracker* t) {
    static int configured = 0;
    const int smem = ransac_smem(t->cap);
    if (smem > 48 * 1024 && configured < smem) {
        cudaFuncSetAttribute(ransac_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        configured = smem;
    }
    ransac_kernel<<<t->B, kRansacThreads, smem, t->stream>>>(t->g, (unsigned)(t->lk_seq * 2654435761u + 12345u));
    t->launches += 1;
    return VS_OK;
}

}  // namespace vs

extern "C" int vs_stage_ransac(vs_tracker* t, const float* a_xy, const float* b_xy, int32_t n, uint8_t* status) {
    using namespace vs;
    if (!t || !a_xy || !b_xy || !status || n < 8) { set_error("vs_stage_ransac: bad argument (need n >= 8)"); return VS_ERR_INVALID; }
    if (!t->lanes.empty()) t = t->lanes[0];
    if (n > t->cap) { set_error("n %d above capacity %d", n, t->cap); return VS_ERR_CAPACITY; }
    VS_CUDA_TRY(cudaSetDevice(t->cfg.device));
    GlueParams& g = t->g;
    g.f_threshold = t->cfg.f_threshold;
    g.f_edge_factor = 0.0;                       // the stage takes lifted points only (no pixel positions)
    const int one = 1;
    VS_CUDA_TRY(cudaMemcpyAsync(g.w_lcur, a_xy, sizeof(float) * 2 * n, cudaMemcpyHostToDevice, t->stream));
    VS_CUDA_TRY(cudaMemcpyAsync(g.w_lnext, b_xy, sizeof(float) * 2 * n, cudaMemcpyHostToDevice, t->stream));
    VS_CUDA_TRY(cudaMemcpyAsync(g.wn, &n, sizeof(int), cudaMemcpyHostToDevice, t->stream));
    VS_CUDA_TRY(cudaMemcpyAsync(g.ransac_flag, &one, sizeof(int), cudaMemcpyHostToDevice, t->stream));
    VS_CUDA_TRY(cudaMemsetAsync(g.w_rstatus, 1, (size_t)n, t->stream));
    GlueParams g1 = g;
    g1.n_streams = 1;
    static int configured = 0;
    const int smem = ransac_smem(t->cap);
    if (smem > 48 * 1024 && configured < smem) {
        cudaFuncSetAttribute(ransac_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        configured = smem;
    }
    ransac_kernel<<<1, kRansacThreads, smem, t->stream>>>(g1, 777u);
    t->launches += 1;
    VS_CUDA_TRY(cudaMemcpyAsync(status, g.w_rstatus, (size_t)n, cudaMemcpyDeviceToHost, t->stream));
    VS_CUDA_TRY(cudaMemsetAsync(g.n, 0, sizeof(int) * t->B, t->stream));
    VS_CUDA_TRY(cudaMemsetAsync(g.ransac_flag, 0, sizeof(int), t->stream));
    VS_CUDA_TRY(cudaStreamSynchronize(t->stream));
    VS_CUDA_TRY(cudaGetLastError());
    t->has_prev = false;
    return VS_OK;
}
