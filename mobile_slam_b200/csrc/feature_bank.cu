// Device feature bank: frontend::FeatureManager (/root/reference/src/frontend/feature_manager.cpp,
// include/frontend/feature_manager.h) for B independent camera streams — SURVEY.md §8 f4, the per-feature-parallel
// consumers downstream of the tracker: addFeatureAndCheckParallax (fm:27-73) with compensatedParallax2 (fm:264-294),
// triangulateAcrossAllViews (fm:142-194) and the window maintenance they cannot run without (removeBack fm:237-249,
// removeFront fm:251-266, removeBackShiftDepth fm:208-235, removeFailures fm:112-118, setDepth / clearDepth /
// getDepthVector / getFeatureCount fm:15-25,94-140, getCorresponding fm:75-92).
//
// Layout (per stream): the reference's std::list<FeaturePerId> is an ORDER array of slot numbers (list position -> slot)
// plus struct-of-arrays slot storage; a feature's observations (<= WINDOW_SIZE + 1 rows of [ray xyz, observation uv,
// velocity xy], fm.h:25-47) sit in a ring inside its slot, so erasing the front observation is an index bump and erasing
// a feature never moves observation data: only the order array is compacted (order-preserving, like list::erase).
// One CTA per stream; every decision the reference takes in double is taken here in double with the same operation order
// and no FMA contraction (__dmul_rn / __dadd_rn), so key-frame decisions, parallax sums and shifted depths are bit-equal
// to the compiled reference.  The one exception is the triangulation's SVD (Eigen::JacobiSVD, fm:182): a one-sided Jacobi
// iteration per feature here — depths agree to rounding (tests: 1e-9 relative), not bit for bit.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "kernels.h"
#include "tracker_state.h"

namespace vs {

constexpr int kBankWin = 10;             // WINDOW_SIZE, include/utility/config.h:11
constexpr int kBankObs = kBankWin + 1;   // frames of the sliding window = observations one feature can hold
constexpr int kBankRec = 7;              // [x, y, z, u, v, vx, vy], common/image_frame.h:14
constexpr int kBankThreads = 256;

struct BankDev {
    int B, cap;
    int* count;        // [B] features in the list
    int* order;        // [B][cap] list position -> slot
    int* nfree;        // [B]
    int* free_slots;   // [B][cap] stack of unused slots
    int* id;           // [B][cap] per slot ...
    int* start;        // start_frame
    int* nobs;         // feature_per_frame.size()
    int* first;        // ring index of observation 0
    int* flag;         // solve_flag
    double* depth;     // estimated_depth
    double* obs;       // [B][cap][kBankObs][kBankRec]
    int* status;       // [B] sticky: 1 = capacity exceeded, 2 = more than kBankObs observations
    double* pval;      // [B][cap] scratch: per list position parallax (or -1)
};

// records of one frame: either host-style arrays (uploaded) or a tracker handle's result block
struct BankFrame {
    const int* frame_count;    // [B]
    // mode 0
    const int* n;              // [B]
    const int* ids;            // [B][max_n]
    const double* rec;         // [B][max_n][7]
    int max_n;
    // mode 1: tracker results (hdr | ids | cnt | pts | un | vel), records = entries with track_cnt > 1
    int from_tracker;
    const int* r_hdr;
    const int* r_ids;
    const int* r_cnt;
    const float2* r_pts;
    const float2* r_un;
    const float2* r_vel;
    int t_cap;
    int s0;                    // bank stream of the source's stream 0
    int rec_cap;               // records a stream can bring (max_n / the tracker's capacity): size of the match array
    double parallax_threshold; // min_parallax / focal_length
    int* o_key;                // [B]
    int* o_track;              // [B]
    double* o_psum;            // [B]
    int* o_pnum;               // [B]
};

__device__ __forceinline__ double* bank_obs(const BankDev& b, int s, int slot, int k) {
    const int ring = (b.first[(int64_t)s * b.cap + slot] + k) % kBankObs;
    return b.obs + (((int64_t)s * b.cap + slot) * kBankObs + ring) * kBankRec;
}

// compensatedParallax2 (fm:264-294): observation i = frame_count - 2, j = frame_count - 1
__device__ __forceinline__ double bank_parallax(const double* fi, const double* fj) {
    const double u_j = fj[0], v_j = fj[1];
    const double dep_i = fi[2];
    const double u_i = __ddiv_rn(fi[0], dep_i), v_i = __ddiv_rn(fi[1], dep_i);
    const double du = __dsub_rn(u_i, u_j), dv = __dsub_rn(v_i, v_j);
    const double d2 = __dadd_rn(__dmul_rn(du, du), __dmul_rn(dv, dv));      // the "compensated" term is the same value
    return fmax(0.0, __dsqrt_rn(d2));
}

__global__ void __launch_bounds__(kBankThreads) bank_add_kernel(BankDev b, BankFrame f, int n_src_streams) {
    extern __shared__ int smem[];
    const int ss = blockIdx.x;                 // stream of the source
    if (ss >= n_src_streams) return;
    const int s = f.s0 + ss;                   // stream of the bank
    const int tid = threadIdx.x;
    const int64_t so = (int64_t)s * b.cap;
    int* s_bid = smem;                         // [cap] ids in list order
    int* s_match = smem + b.cap;               // [max records] list position, -1 new, -2 not a record
    int* s_hash = s_match + f.rec_cap;         // [pow2 >= 2 * cap] open-addressing table: list position or -1
    __shared__ int s_tracked, s_new, s_new_ok;
    const int n_rec = f.from_tracker ? min(f.r_hdr[(int64_t)ss * kHdrInts + H_N], f.t_cap) : min(f.n[ss], f.max_n);
    const int cnt = b.count[s];
    const int fc = f.frame_count[ss];
    if (tid == 0) { s_tracked = 0; s_new = 0; s_new_ok = 0; }
    for (int i = tid; i < cnt; i += kBankThreads) s_bid[i] = b.id[so + b.order[so + i]];
    // the reference's find_if is a linear search (ids are unique in the list, so "first match" is "the match"): a hash
    // table over the list, sized to at most half full, gives the same answer in O(1) per record
    int hbits = 6;
    while ((1 << hbits) < 2 * cnt) ++hbits;
    const unsigned hmask = (1u << hbits) - 1u;
    for (int i = tid; i <= (int)hmask; i += kBankThreads) s_hash[i] = -1;
    __syncthreads();
    auto hash_of = [&](int id) { return ((unsigned)id * 2654435761u) >> (32 - hbits); };
    for (int i = tid; i < cnt; i += kBankThreads) {
        unsigned h = hash_of(s_bid[i]);
        while (atomicCAS(&s_hash[h], -1, i) != -1) h = (h + 1) & hmask;
    }
    __syncthreads();
    auto rec_valid = [&](int r) { return f.from_tracker ? f.r_cnt[(int64_t)ss * f.t_cap + r] > 1 : true; };
    auto rec_id = [&](int r) { return f.from_tracker ? f.r_ids[(int64_t)ss * f.t_cap + r] : f.ids[(int64_t)ss * f.max_n + r]; };
    auto rec_store = [&](int r, double* o) {
        if (f.from_tracker) {      // the publish loop of measurement_processor.cpp:210-231 / vio_engine.cpp:219-239
            const int64_t q = (int64_t)ss * f.t_cap + r;
            const float2 un = f.r_un[q], px = f.r_pts[q], ve = f.r_vel[q];
            o[0] = (double)un.x; o[1] = (double)un.y; o[2] = 1.0;
            o[3] = (double)px.x; o[4] = (double)px.y; o[5] = (double)ve.x; o[6] = (double)ve.y;
        } else {
            const double* src = f.rec + ((int64_t)ss * f.max_n + r) * kBankRec;
#pragma unroll
            for (int k = 0; k < kBankRec; ++k) o[k] = src[k];
        }
    };
    // ---- find_if over the bank for every record (fm:37-38): the first list position with that id ----
    for (int r = tid; r < n_rec; r += kBankThreads) {
        int m = -2;
        if (rec_valid(r)) {
            const int rid = rec_id(r);
            m = -1;
            for (unsigned h = hash_of(rid);; h = (h + 1) & hmask) {
                const int pos = s_hash[h];
                if (pos < 0) break;
                if (s_bid[pos] == rid) { m = pos; break; }
            }
            atomicAdd(m >= 0 ? &s_tracked : &s_new, 1);
        }
        s_match[r] = m;
    }
    __syncthreads();
    const int n_new = s_new;
    const int nfree = b.nfree[s];
    for (int r = tid; r < n_rec; r += kBankThreads) {
        const int m = s_match[r];
        if (m == -2) continue;
        if (m >= 0) {                          // tracked feature: push_back one observation (fm:45-48)
            const int slot = b.order[so + m];
            const int k = b.nobs[so + slot];
            if (k >= kBankObs) { atomicOr(b.status + s, 2); continue; }
            b.nobs[so + slot] = k + 1;
            rec_store(r, bank_obs(b, s, slot, k));
        } else {                               // new feature (fm:40-44): appended in the std::map's order = ascending id
            const int rid = rec_id(r);
            int rank = 0;
            for (int q = 0; q < n_rec; ++q)
                if (s_match[q] == -1 && rec_id(q) < rid) ++rank;
            if (cnt + rank >= b.cap || rank >= nfree) { atomicOr(b.status + s, 1); continue; }
            const int slot = b.free_slots[so + nfree - 1 - rank];
            b.order[so + cnt + rank] = slot;
            b.id[so + slot] = rid;
            b.start[so + slot] = fc;
            b.nobs[so + slot] = 1;
            b.first[so + slot] = 0;
            b.flag[so + slot] = 0;
            b.depth[so + slot] = -1.0;
            rec_store(r, bank_obs(b, s, slot, 0));
            atomicAdd(&s_new_ok, 1);
        }
    }
    __syncthreads();
    // ranks are dense from 0, so the accepted new features are exactly ranks 0 .. s_new_ok - 1
    const int cnt2 = cnt + min(s_new_ok, n_new);
    if (tid == 0) {
        b.count[s] = cnt2;
        b.nfree[s] = nfree - (cnt2 - cnt);
    }
    const int tracked = s_tracked;
    // ---- parallax (fm:51-72) ----
    if (fc < 2 || tracked < 20) {
        if (tid == 0) { f.o_key[s] = 1; f.o_track[s] = tracked; f.o_psum[s] = 0.0; f.o_pnum[s] = 0; }
        return;
    }
    __threadfence_block();
    for (int i = tid; i < cnt2; i += kBankThreads) {
        const int slot = b.order[so + i];
        const int st = b.start[so + slot], no = b.nobs[so + slot];
        double v = -1.0;
        if (st <= fc - 2 && st + no - 1 >= fc - 1) v = bank_parallax(bank_obs(b, s, slot, fc - 2 - st), bank_obs(b, s, slot, fc - 1 - st));
        b.pval[so + i] = v;
    }
    __syncthreads();
    if (tid < 32) {                            // the reference adds in list order: so does warp 0, 32 values per trip
        double sum = 0.0;
        int num = 0;
        for (int i0 = 0; i0 < cnt2; i0 += 32) {
            const double v = (i0 + tid < cnt2) ? b.pval[so + i0 + tid] : -1.0;
            for (int j = 0; j < 32; ++j) {
                const double vj = __shfl_sync(0xffffffffu, v, j);
                if (vj >= 0.0) { sum = __dadd_rn(sum, vj); ++num; }
            }
        }
        if (tid == 0) {
            f.o_track[s] = tracked;
            f.o_psum[s] = sum;
            f.o_pnum[s] = num;
            f.o_key[s] = (num == 0) ? 1 : (__ddiv_rn(sum, (double)num) >= f.parallax_threshold ? 1 : 0);
        }
    }
}

// ---- window slide / erase operations -----------------------------------------------------------------------------
enum BankOp { BANK_NONE = 0, BANK_REMOVE_BACK = 1, BANK_REMOVE_FRONT = 2, BANK_REMOVE_BACK_SHIFT_DEPTH = 3, BANK_REMOVE_FAILURES = 4 };

__device__ __forceinline__ void mat3_vec(const double* m, const double* v, double* o) {   // row-major m; ((a*b + c*d) + e*f)
#pragma unroll
    for (int i = 0; i < 3; ++i)
        o[i] = __dadd_rn(__dadd_rn(__dmul_rn(m[3 * i], v[0]), __dmul_rn(m[3 * i + 1], v[1])), __dmul_rn(m[3 * i + 2], v[2]));
}
__device__ __forceinline__ void mat3t_vec(const double* m, const double* v, double* o) {  // m^T v
#pragma unroll
    for (int i = 0; i < 3; ++i)
        o[i] = __dadd_rn(__dadd_rn(__dmul_rn(m[i], v[0]), __dmul_rn(m[3 + i], v[1])), __dmul_rn(m[6 + i], v[2]));
}

__global__ void __launch_bounds__(kBankThreads) bank_slide_kernel(BankDev b, const int* op, const int* frame_count,
                                                                  const double* rp, double init_depth) {
    const int s = blockIdx.x, tid = threadIdx.x;
    const int o = op[s];
    if (o == BANK_NONE) return;
    const int64_t so = (int64_t)s * b.cap;
    const int cnt = b.count[s];
    const int fc = frame_count ? frame_count[s] : 0;
    __shared__ int s_scan[kBankThreads];
    __shared__ int s_base, s_freed;
    if (tid == 0) { s_base = 0; s_freed = b.nfree[s]; }
    __syncthreads();
    for (int i0 = 0; i0 < cnt; i0 += kBankThreads) {
        const int i = i0 + tid;
        int slot = -1, keep = 0;
        if (i < cnt) {
            slot = b.order[so + i];
            keep = 1;
            int st = b.start[so + slot], no = b.nobs[so + slot];
            if (o == BANK_REMOVE_BACK) {                           // fm:237-249
                if (st != 0) b.start[so + slot] = st - 1;
                else {
                    b.first[so + slot] = (b.first[so + slot] + 1) % kBankObs;
                    b.nobs[so + slot] = --no;
                    if (no == 0) keep = 0;
                }
            } else if (o == BANK_REMOVE_FRONT) {                   // fm:251-266
                if (st == fc) b.start[so + slot] = st - 1;
                else {
                    const int j = kBankWin - 1 - st;
                    if (!(st + no - 1 < fc - 1) && j >= 0 && j < no) {
                        for (int k = j; k + 1 < no; ++k) {
                            double* dst = bank_obs(b, s, slot, k);
                            const double* src = bank_obs(b, s, slot, k + 1);
                            for (int q = 0; q < kBankRec; ++q) dst[q] = src[q];
                        }
                        b.nobs[so + slot] = --no;
                        if (no == 0) keep = 0;
                    }
                }
            } else if (o == BANK_REMOVE_BACK_SHIFT_DEPTH) {        // fm:208-235
                if (st != 0) b.start[so + slot] = st - 1;
                else {
                    const double* o0 = bank_obs(b, s, slot, 0);
                    const double uv[3] = {o0[0], o0[1], o0[2]};
                    b.first[so + slot] = (b.first[so + slot] + 1) % kBankObs;
                    b.nobs[so + slot] = --no;
                    if (no < 2) keep = 0;
                    else {
                        const double* q = rp + (int64_t)s * 24;    // marg_R (9) marg_P (3) new_R (9) new_P (3), row-major
                        const double dep = b.depth[so + slot];
                        const double pi[3] = {__dmul_rn(uv[0], dep), __dmul_rn(uv[1], dep), __dmul_rn(uv[2], dep)};
                        double rpv[3], w[3], pj[3];
                        mat3_vec(q, pi, rpv);
                        for (int k = 0; k < 3; ++k) w[k] = __dsub_rn(__dadd_rn(rpv[k], q[9 + k]), q[21 + k]);
                        mat3t_vec(q + 12, w, pj);
                        b.depth[so + slot] = pj[2] > 0 ? pj[2] : init_depth;
                    }
                }
            } else if (o == BANK_REMOVE_FAILURES) {                // fm:112-118
                if (b.flag[so + slot] == 2) keep = 0;
            }
        }
        // order-preserving compaction of this chunk (every position of the chunk has been read before any is written,
        // and writes only go to positions at or before the chunk)
        s_scan[tid] = keep;
        __syncthreads();
        for (int d = 1; d < kBankThreads; d <<= 1) {
            const int v = tid >= d ? s_scan[tid - d] : 0;
            __syncthreads();
            s_scan[tid] += v;
            __syncthreads();
        }
        const int incl = s_scan[tid], total = s_scan[kBankThreads - 1];
        const int base = s_base, freed = s_freed;
        const int n_here = min(kBankThreads, cnt - i0);
        if (i < cnt) {
            if (keep) b.order[so + base + incl - 1] = slot;
            else b.free_slots[so + freed + (tid + 1 - incl) - 1] = slot;
        }
        __syncthreads();
        if (tid == 0) { s_base = base + total; s_freed = freed + (n_here - total); }
        __syncthreads();
    }
    if (tid == 0) { b.count[s] = s_base; b.nfree[s] = s_freed; }
}

// ---- triangulateAcrossAllViews (fm:142-194) ----------------------------------------------------------------------
__device__ __forceinline__ void mat3_mul(const double* a, const double* bm, double* o) {   // row-major, left-to-right sums
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            o[3 * i + j] = __dadd_rn(__dadd_rn(__dmul_rn(a[3 * i], bm[j]), __dmul_rn(a[3 * i + 1], bm[3 + j])), __dmul_rn(a[3 * i + 2], bm[6 + j]));
}

// right singular vector of the smallest singular value of A (m x 4): one-sided Jacobi (Hestenes) on the columns
__device__ void smallest_right_singular_vector(double (*A)[4], int m, double* v_out) {
    double V[4][4];
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) V[i][j] = (i == j) ? 1.0 : 0.0;
    for (int sweep = 0; sweep < 40; ++sweep) {
        bool rotated = false;
        for (int p = 0; p < 3; ++p)
            for (int q = p + 1; q < 4; ++q) {
                double al = 0.0, be = 0.0, ga = 0.0;
                for (int r = 0; r < m; ++r) {
                    al += A[r][p] * A[r][p];
                    be += A[r][q] * A[r][q];
                    ga += A[r][p] * A[r][q];
                }
                if (ga == 0.0 || fabs(ga) <= 1e-16 * sqrt(al * be)) continue;
                rotated = true;
                const double zeta = (be - al) / (2.0 * ga);
                const double t = copysign(1.0, zeta) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                const double c = 1.0 / sqrt(1.0 + t * t), sn = c * t;
                for (int r = 0; r < m; ++r) {
                    const double ap = A[r][p], aq = A[r][q];
                    A[r][p] = c * ap - sn * aq;
                    A[r][q] = sn * ap + c * aq;
                }
                for (int r = 0; r < 4; ++r) {
                    const double vp = V[r][p], vq = V[r][q];
                    V[r][p] = c * vp - sn * vq;
                    V[r][q] = sn * vp + c * vq;
                }
            }
        if (!rotated) break;
    }
    int jmin = 0;
    double nmin = 0.0;
    for (int j = 0; j < 4; ++j) {
        double nn = 0.0;
        for (int r = 0; r < m; ++r) nn += A[r][j] * A[r][j];
        if (j == 0 || nn < nmin) { nmin = nn; jmin = j; }
    }
    for (int i = 0; i < 4; ++i) v_out[i] = V[i][jmin];
}

__global__ void __launch_bounds__(128) bank_triangulate_kernel(BankDev b, const double* poses, const double* t_ic,
                                                               const double* r_ic, double init_depth, const int* active) {
    const int s = blockIdx.x, tid = threadIdx.x;
    if (active && !active[s]) return;
    const int64_t so = (int64_t)s * b.cap;
    __shared__ double s_R[kBankObs][9], s_t[kBankObs][3];          // camera poses: R_i r_ic, P_i + R_i t_ic
    if (tid < kBankObs) {
        const double* ps = poses + ((int64_t)s * kBankObs + tid) * 12;
        double rt[3];
        mat3_vec(ps, t_ic, rt);
        for (int k = 0; k < 3; ++k) s_t[tid][k] = __dadd_rn(ps[9 + k], rt[k]);
        mat3_mul(ps, r_ic, s_R[tid]);
    }
    __syncthreads();
    const int cnt = b.count[s];
    for (int i = tid; i < cnt; i += blockDim.x) {
        const int slot = b.order[so + i];
        const int no = b.nobs[so + slot], st = b.start[so + slot];
        if (!(no >= 2 && st < kBankWin - 2)) continue;
        if (b.depth[so + slot] > 0) continue;
        if (st < 0 || st + no > kBankObs) continue;                 // outside the window: nothing the reference could index
        double A[2 * kBankObs][4];
        const double* R0 = s_R[st];
        const double* t0 = s_t[st];
        for (int k = 0; k < no; ++k) {
            const int j = st + k;
            const double* R1 = s_R[j];
            const double dt[3] = {__dsub_rn(s_t[j][0], t0[0]), __dsub_rn(s_t[j][1], t0[1]), __dsub_rn(s_t[j][2], t0[2])};
            double rel_t[3], rel_R[9];
            mat3t_vec(R0, dt, rel_t);                               // R0^T (t1 - t0)
            {                                                       // R0^T R1
                double R0t[9];
                for (int a = 0; a < 3; ++a)
                    for (int c = 0; c < 3; ++c) R0t[3 * a + c] = R0[3 * c + a];
                mat3_mul(R0t, R1, rel_R);
            }
            // P = [rel_R^T | (-rel_R^T) rel_t]
            double P[3][4];
            for (int a = 0; a < 3; ++a) {
                for (int c = 0; c < 3; ++c) P[a][c] = rel_R[3 * c + a];
                P[a][3] = __dadd_rn(__dadd_rn(__dmul_rn(-rel_R[a], rel_t[0]), __dmul_rn(-rel_R[3 + a], rel_t[1])), __dmul_rn(-rel_R[6 + a], rel_t[2]));
            }
            const double* ob = bank_obs(b, s, slot, k);
            double fx = ob[0], fy = ob[1], fz = ob[2];
            const double z2 = __dadd_rn(__dadd_rn(__dmul_rn(fx, fx), __dmul_rn(fy, fy)), __dmul_rn(fz, fz));
            if (z2 > 0) {
                const double nrm = __dsqrt_rn(z2);
                fx = __ddiv_rn(fx, nrm); fy = __ddiv_rn(fy, nrm); fz = __ddiv_rn(fz, nrm);
            }
            for (int c = 0; c < 4; ++c) {
                A[2 * k][c] = __dsub_rn(__dmul_rn(fx, P[2][c]), __dmul_rn(fz, P[0][c]));
                A[2 * k + 1][c] = __dsub_rn(__dmul_rn(fy, P[2][c]), __dmul_rn(fz, P[1][c]));
            }
        }
        double v[4];
        smallest_right_singular_vector(A, 2 * no, v);
        double dep = v[2] / v[3];
        if (!(dep >= 0.1)) dep = init_depth;                        // fm:190-192 (`< 0.1`; a NaN ratio also falls back)
        b.depth[so + slot] = dep;
    }
}

}  // namespace vs

// ---- host side / C ABI ---------------------------------------------------------------------------------------------
struct vs_bank {
    int device = 0;
    vs::BankDev d{};
    vs::DeviceArena arena;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev = nullptr;
    // per-call staging (device) and pinned host mirrors
    int* d_iparams = nullptr;     // [4][B]: frame_count | n / op | key | track
    int* d_pnum = nullptr;
    double* d_psum = nullptr;
    int* d_ids = nullptr;         // [B][max_n] grown on demand
    double* d_rec = nullptr;
    int staged_max_n = 0;
    double* d_poses = nullptr;    // [B][11][12]
    double* d_rp = nullptr;       // [B][24]
    double* d_ext = nullptr;      // t_ic (3) r_ic (9)
    int* h_i = nullptr;           // pinned [4][B]
    double* h_psum = nullptr;
    int* h_pnum = nullptr;
};

using vs::set_error;

#define VS_BANK_GUARD(bk)                                                   \
    if (!(bk)) { set_error("null bank handle"); return VS_ERR_INVALID; } \
    VS_CUDA_TRY(cudaSetDevice((bk)->device));

static int bank_add_smem(const vs_bank* b, int rec_cap) {
    int hsize = 64;
    while (hsize < 2 * b->d.cap) hsize <<= 1;
    return (b->d.cap + rec_cap + hsize) * (int)sizeof(int);
}

extern "C" {

int vs_bank_create(int32_t device, int32_t n_streams, int32_t capacity, vs_bank** out) {
    if (!out) { set_error("null out"); return VS_ERR_INVALID; }
    *out = nullptr;
    if (n_streams < 1 || capacity < 1 || capacity > 8192) { set_error("bank: n_streams >= 1 and 1 <= capacity <= 8192"); return VS_ERR_INVALID; }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
        set_error("no CUDA device %d (there is no CPU fallback)", device);
        return VS_ERR_CUDA;
    }
    VS_CUDA_TRY(cudaSetDevice(device));
    vs_bank* b = new vs_bank();
    b->device = device;
    b->d.B = n_streams;
    b->d.cap = capacity;
    {   // VS_DEBUG_GUARD=1: every device block sits between two bands of 0xA5 that nothing may write (vs_bank_debug_guard_check)
        const char* g = getenv("VS_DEBUG_GUARD");
        b->arena.guarded = g && atoi(g) != 0;
    }
    const size_t B = n_streams, C = capacity;
    auto fail = [&](cudaError_t e) { set_error("bank allocation: %s", cudaGetErrorString(e)); vs_bank_destroy(b); return VS_ERR_NOMEM; };
    cudaError_t e;
    if ((e = cudaStreamCreateWithFlags(&b->stream, cudaStreamNonBlocking)) != cudaSuccess) return fail(e);
    if ((e = cudaEventCreateWithFlags(&b->ev, cudaEventDisableTiming)) != cudaSuccess) return fail(e);
#define BANK_ALLOC(p, n) if ((e = b->arena.alloc(&(p), (n))) != cudaSuccess) return fail(e);
    BANK_ALLOC(b->d.count, B) BANK_ALLOC(b->d.order, B * C) BANK_ALLOC(b->d.nfree, B) BANK_ALLOC(b->d.free_slots, B * C)
    BANK_ALLOC(b->d.id, B * C) BANK_ALLOC(b->d.start, B * C) BANK_ALLOC(b->d.nobs, B * C) BANK_ALLOC(b->d.first, B * C)
    BANK_ALLOC(b->d.flag, B * C) BANK_ALLOC(b->d.depth, B * C) BANK_ALLOC(b->d.obs, B * C * vs::kBankObs * vs::kBankRec)
    BANK_ALLOC(b->d.status, B) BANK_ALLOC(b->d.pval, B * C)
    BANK_ALLOC(b->d_iparams, 4 * B) BANK_ALLOC(b->d_pnum, B) BANK_ALLOC(b->d_psum, B)
    BANK_ALLOC(b->d_poses, B * vs::kBankObs * 12) BANK_ALLOC(b->d_rp, B * 24) BANK_ALLOC(b->d_ext, 12)
#undef BANK_ALLOC
    if ((e = cudaMallocHost(&b->h_i, sizeof(int) * 4 * B)) != cudaSuccess) return fail(e);
    if ((e = cudaMallocHost(&b->h_psum, sizeof(double) * B)) != cudaSuccess) return fail(e);
    if ((e = cudaMallocHost(&b->h_pnum, sizeof(int) * B)) != cudaSuccess) return fail(e);
    *out = b;
    return vs_bank_clear(b, -1);
}

void vs_bank_destroy(vs_bank* b) {
    if (!b) return;
    cudaSetDevice(b->device);
    if (b->stream) cudaStreamSynchronize(b->stream);
    b->arena.release();
    if (b->d_ids) cudaFree(b->d_ids);
    if (b->d_rec) cudaFree(b->d_rec);
    if (b->h_i) cudaFreeHost(b->h_i);
    if (b->h_psum) cudaFreeHost(b->h_psum);
    if (b->h_pnum) cudaFreeHost(b->h_pnum);
    if (b->ev) cudaEventDestroy(b->ev);
    if (b->stream) cudaStreamDestroy(b->stream);
    delete b;
}

// bytes of the guard bands around the bank's device blocks that changed (-1: handle not created with VS_DEBUG_GUARD=1)
int64_t vs_bank_debug_guard_check(vs_bank* b) {
    if (!b) return -2;
    cudaSetDevice(b->device);
    cudaStreamSynchronize(b->stream);
    return b->arena.check();
}

// clearState (fm:11-13) for one stream or all (-1)
int vs_bank_clear(vs_bank* b, int32_t stream) {
    VS_BANK_GUARD(b);
    if (stream < -1 || stream >= b->d.B) { set_error("bank: stream %d out of range", stream); return VS_ERR_INVALID; }
    const int C = b->d.cap;
    std::vector<int> slots(C);
    for (int i = 0; i < C; ++i) slots[i] = C - 1 - i;          // the stack hands out slot 0 first
    const int zero = 0;
    for (int s = (stream < 0 ? 0 : stream); s < (stream < 0 ? b->d.B : stream + 1); ++s) {
        VS_CUDA_TRY(cudaMemcpyAsync(b->d.free_slots + (size_t)s * C, slots.data(), sizeof(int) * C, cudaMemcpyHostToDevice, b->stream));
        VS_CUDA_TRY(cudaMemcpyAsync(b->d.nfree + s, &C, sizeof(int), cudaMemcpyHostToDevice, b->stream));
        VS_CUDA_TRY(cudaMemcpyAsync(b->d.count + s, &zero, sizeof(int), cudaMemcpyHostToDevice, b->stream));
        VS_CUDA_TRY(cudaMemcpyAsync(b->d.status + s, &zero, sizeof(int), cudaMemcpyHostToDevice, b->stream));
    }
    VS_CUDA_TRY(cudaStreamSynchronize(b->stream));             // `slots`, `zero` and `C` leave scope
    return VS_OK;
}

static int bank_finish_add(vs_bank* b, int s0, int nB, int32_t* is_key, int32_t* last_track_num, double* parallax_sum,
                           int32_t* parallax_num) {
    const size_t B = b->d.B;
    VS_CUDA_TRY(cudaMemcpyAsync(b->h_i + 2 * B, b->d_iparams + 2 * B, sizeof(int) * 2 * B, cudaMemcpyDeviceToHost, b->stream));
    VS_CUDA_TRY(cudaMemcpyAsync(b->h_psum, b->d_psum, sizeof(double) * B, cudaMemcpyDeviceToHost, b->stream));
    VS_CUDA_TRY(cudaMemcpyAsync(b->h_pnum, b->d_pnum, sizeof(int) * B, cudaMemcpyDeviceToHost, b->stream));
    std::vector<int> st(B);
    VS_CUDA_TRY(cudaMemcpyAsync(st.data(), b->d.status, sizeof(int) * B, cudaMemcpyDeviceToHost, b->stream));
    VS_CUDA_TRY(cudaStreamSynchronize(b->stream));
    VS_CUDA_TRY(cudaGetLastError());
    for (int s = s0; s < s0 + nB; ++s) {
        if (is_key) is_key[s] = b->h_i[2 * B + s];
        if (last_track_num) last_track_num[s] = b->h_i[3 * B + s];
        if (parallax_sum) parallax_sum[s] = b->h_psum[s];
        if (parallax_num) parallax_num[s] = b->h_pnum[s];
    }
    for (int s = s0; s < s0 + nB; ++s)
        if (st[s]) {
            set_error("bank: stream %d %s", s, (st[s] & 1) ? "is full (capacity)" : "holds a feature with more than WINDOW_SIZE + 1 observations");
            return VS_ERR_CAPACITY;
        }
    return VS_OK;
}

int vs_bank_add_frame(vs_bank* b, const int32_t* frame_count, const int32_t* n, const int32_t* ids, const double* rec,
                      int32_t max_n, double min_parallax, double focal_length, int32_t* is_key, int32_t* last_track_num,
                      double* parallax_sum, int32_t* parallax_num) {
    VS_BANK_GUARD(b);
    if (!frame_count || !n || max_n < 0 || (max_n > 0 && (!ids || !rec)) || !is_key) { set_error("bank: bad argument"); return VS_ERR_INVALID; }
    const size_t B = b->d.B;
    for (size_t s = 0; s < B; ++s)
        if (n[s] < 0 || n[s] > max_n) { set_error("bank: n[%d] = %d outside 0..max_n", (int)s, n[s]); return VS_ERR_INVALID; }
    const int mn = max_n > 0 ? max_n : 1;
    if (mn > b->staged_max_n) {
        if (b->d_ids) cudaFree(b->d_ids);
        if (b->d_rec) cudaFree(b->d_rec);
        b->d_ids = nullptr; b->d_rec = nullptr; b->staged_max_n = 0;
        VS_CUDA_TRY(cudaMalloc(&b->d_ids, sizeof(int) * B * mn));
        VS_CUDA_TRY(cudaMalloc(&b->d_rec, sizeof(double) * B * mn * vs::kBankRec));
        b->staged_max_n = mn;
    }
    memcpy(b->h_i, frame_count, sizeof(int) * B);
    memcpy(b->h_i + B, n, sizeof(int) * B);
    VS_CUDA_TRY(cudaMemcpyAsync(b->d_iparams, b->h_i, sizeof(int) * 2 * B, cudaMemcpyHostToDevice, b->stream));
    if (max_n > 0) {
        VS_CUDA_TRY(cudaMemcpyAsync(b->d_ids, ids, sizeof(int) * B * max_n, cudaMemcpyHostToDevice, b->stream));
        VS_CUDA_TRY(cudaMemcpyAsync(b->d_rec, rec, sizeof(double) * B * max_n * vs::kBankRec, cudaMemcpyHostToDevice, b->stream));
    }
    vs::BankFrame f{};
    f.frame_count = b->d_iparams;
    f.n = b->d_iparams + B;
    f.ids = b->d_ids;
    f.rec = b->d_rec;
    f.max_n = max_n;
    f.from_tracker = 0;
    f.s0 = 0;
    f.rec_cap = mn;
    f.parallax_threshold = min_parallax / focal_length;
    f.o_key = b->d_iparams + 2 * B;
    f.o_track = b->d_iparams + 3 * B;
    f.o_psum = b->d_psum;
    f.o_pnum = b->d_pnum;
    const int smem = bank_add_smem(b, mn);
    if (smem > 200 * 1024) { set_error("bank: capacity + max_n = %d exceeds the shared memory of one CTA", b->d.cap + mn); return VS_ERR_CAPACITY; }
    VS_CUDA_TRY(vs::ensure_dyn_smem(reinterpret_cast<const void*>(vs::bank_add_kernel), smem));
    vs::bank_add_kernel<<<(unsigned)B, vs::kBankThreads, smem, b->stream>>>(b->d, f, (int)B);
    return bank_finish_add(b, 0, (int)B, is_key, last_track_num, parallax_sum, parallax_num);
}

int vs_bank_add_frame_from_tracker(vs_bank* b, vs_tracker* t, const int32_t* frame_count, double min_parallax,
                                   double focal_length, int32_t* is_key, int32_t* last_track_num) {
    VS_BANK_GUARD(b);
    if (!t || !frame_count || !is_key) { set_error("bank: bad argument"); return VS_ERR_INVALID; }
    if (t->cfg.device != b->device) { set_error("bank: tracker lives on device %d, the bank on %d", t->cfg.device, b->device); return VS_ERR_INVALID; }
    if (t->B != b->d.B) { set_error("bank: tracker has %d streams, the bank %d", t->B, b->d.B); return VS_ERR_INVALID; }
    const size_t B = b->d.B;
    memcpy(b->h_i, frame_count, sizeof(int) * B);
    VS_CUDA_TRY(cudaMemcpyAsync(b->d_iparams, b->h_i, sizeof(int) * B, cudaMemcpyHostToDevice, b->stream));
    std::vector<vs_tracker*> parts;
    std::vector<int> first;
    if (t->lanes.empty()) { parts.push_back(t); first.push_back(0); }
    else for (size_t l = 0; l < t->lanes.size(); ++l) { parts.push_back(t->lanes[l]); first.push_back(t->lane_s0[l]); }
    for (size_t l = 0; l < parts.size(); ++l) {
        vs_tracker* q = parts[l];
        if (!q->last_results) { set_error("bank: the tracker has not tracked a frame yet"); return VS_ERR_STATE; }
        VS_CUDA_TRY(cudaEventRecord(b->ev, q->stream));          // the frame's kernels wrote the result block on q->stream
        VS_CUDA_TRY(cudaStreamWaitEvent(b->stream, b->ev, 0));
        const int64_t qB = q->B, qC = q->cap;
        const uint8_t* p = q->last_results;
        vs::BankFrame f{};
        f.frame_count = b->d_iparams + first[l];
        f.from_tracker = 1;
        f.r_hdr = reinterpret_cast<const int*>(p); p += qB * vs::kHdrInts * 4;
        f.r_ids = reinterpret_cast<const int*>(p); p += qB * qC * 4;
        f.r_cnt = reinterpret_cast<const int*>(p); p += qB * qC * 4;
        f.r_pts = reinterpret_cast<const float2*>(p); p += qB * qC * 8;
        f.r_un = reinterpret_cast<const float2*>(p); p += qB * qC * 8;
        f.r_vel = reinterpret_cast<const float2*>(p);
        f.t_cap = (int)qC;
        f.s0 = first[l];
        f.rec_cap = (int)qC;
        f.parallax_threshold = min_parallax / focal_length;
        f.o_key = b->d_iparams + 2 * B;
        f.o_track = b->d_iparams + 3 * B;
        f.o_psum = b->d_psum;
        f.o_pnum = b->d_pnum;
        const int smem = bank_add_smem(b, (int)qC);
        if (smem > 200 * 1024) { set_error("bank: capacity %d + tracker capacity %d exceed the shared memory of one CTA", b->d.cap, (int)qC); return VS_ERR_CAPACITY; }
        VS_CUDA_TRY(vs::ensure_dyn_smem(reinterpret_cast<const void*>(vs::bank_add_kernel), smem));
        vs::bank_add_kernel<<<(unsigned)qB, vs::kBankThreads, smem, b->stream>>>(b->d, f, (int)qB);
    }
    return bank_finish_add(b, 0, (int)B, is_key, last_track_num, nullptr, nullptr);
}

int vs_bank_slide(vs_bank* b, const int32_t* op, const int32_t* frame_count, const double* rp, double init_depth) {
    VS_BANK_GUARD(b);
    if (!op) { set_error("bank: null op"); return VS_ERR_INVALID; }
    const size_t B = b->d.B;
    bool need_rp = false;
    for (size_t s = 0; s < B; ++s) {
        if (op[s] < 0 || op[s] > 4) { set_error("bank: op[%d] = %d", (int)s, op[s]); return VS_ERR_INVALID; }
        if (op[s] == vs::BANK_REMOVE_FRONT && !frame_count) { set_error("bank: removeFront needs frame_count"); return VS_ERR_INVALID; }
        need_rp |= op[s] == vs::BANK_REMOVE_BACK_SHIFT_DEPTH;
    }
    if (need_rp && !rp) { set_error("bank: removeBackShiftDepth needs the two poses"); return VS_ERR_INVALID; }
    memcpy(b->h_i, op, sizeof(int) * B);
    if (frame_count) memcpy(b->h_i + B, frame_count, sizeof(int) * B);
    else memset(b->h_i + B, 0, sizeof(int) * B);
    VS_CUDA_TRY(cudaMemcpyAsync(b->d_iparams, b->h_i, sizeof(int) * 2 * B, cudaMemcpyHostToDevice, b->stream));
    if (need_rp) VS_CUDA_TRY(cudaMemcpyAsync(b->d_rp, rp, sizeof(double) * B * 24, cudaMemcpyHostToDevice, b->stream));
    vs::bank_slide_kernel<<<(unsigned)B, vs::kBankThreads, 0, b->stream>>>(b->d, b->d_iparams, b->d_iparams + B, b->d_rp, init_depth);
    VS_CUDA_TRY(cudaStreamSynchronize(b->stream));
    VS_CUDA_TRY(cudaGetLastError());
    return VS_OK;
}

int vs_bank_triangulate(vs_bank* b, const double* poses, const double* t_ic, const double* r_ic, double init_depth,
                        const int32_t* active) {
    VS_BANK_GUARD(b);
    if (!poses || !t_ic || !r_ic) { set_error("bank: null argument"); return VS_ERR_INVALID; }
    const size_t B = b->d.B;
    if (active) {
        memcpy(b->h_i, active, sizeof(int) * B);
        VS_CUDA_TRY(cudaMemcpyAsync(b->d_iparams, b->h_i, sizeof(int) * B, cudaMemcpyHostToDevice, b->stream));
    }
    double ext[12];
    memcpy(ext, t_ic, sizeof(double) * 3);
    memcpy(ext + 3, r_ic, sizeof(double) * 9);
    VS_CUDA_TRY(cudaMemcpyAsync(b->d_poses, poses, sizeof(double) * B * vs::kBankObs * 12, cudaMemcpyHostToDevice, b->stream));
    VS_CUDA_TRY(cudaMemcpyAsync(b->d_ext, ext, sizeof(ext), cudaMemcpyHostToDevice, b->stream));
    vs::bank_triangulate_kernel<<<(unsigned)B, 128, 0, b->stream>>>(b->d, b->d_poses, b->d_ext, b->d_ext + 3, init_depth,
                                                                    active ? b->d_iparams : nullptr);
    VS_CUDA_TRY(cudaStreamSynchronize(b->stream));
    VS_CUDA_TRY(cudaGetLastError());
    return VS_OK;
}

// the list of one stream on the host (metadata only unless obs != null)
struct BankHost {
    int n = 0;
    std::vector<int> order, id, start, nobs, first, flag;
    std::vector<double> depth;
};
static int bank_fetch(vs_bank* b, int s, BankHost& h) {
    const size_t C = b->d.cap, so = (size_t)s * C;
    VS_CUDA_TRY(cudaMemcpyAsync(&h.n, b->d.count + s, sizeof(int), cudaMemcpyDeviceToHost, b->stream));
    h.order.resize(C); h.id.resize(C); h.start.resize(C); h.nobs.resize(C); h.first.resize(C); h.flag.resize(C); h.depth.resize(C);
    VS_CUDA_TRY(cudaMemcpyAsync(h.order.data(), b->d.order + so, sizeof(int) * C, cudaMemcpyDeviceToHost, b->stream));
    VS_CUDA_TRY(cudaMemcpyAsync(h.id.data(), b->d.id + so, sizeof(int) * C, cudaMemcpyDeviceToHost, b->stream));
    VS_CUDA_TRY(cudaMemcpyAsync(h.start.data(), b->d.start + so, sizeof(int) * C, cudaMemcpyDeviceToHost, b->stream));
    VS_CUDA_TRY(cudaMemcpyAsync(h.nobs.data(), b->d.nobs + so, sizeof(int) * C, cudaMemcpyDeviceToHost, b->stream));
    VS_CUDA_TRY(cudaMemcpyAsync(h.first.data(), b->d.first + so, sizeof(int) * C, cudaMemcpyDeviceToHost, b->stream));
    VS_CUDA_TRY(cudaMemcpyAsync(h.flag.data(), b->d.flag + so, sizeof(int) * C, cudaMemcpyDeviceToHost, b->stream));
    VS_CUDA_TRY(cudaMemcpyAsync(h.depth.data(), b->d.depth + so, sizeof(double) * C, cudaMemcpyDeviceToHost, b->stream));
    VS_CUDA_TRY(cudaStreamSynchronize(b->stream));
    return VS_OK;
}
static inline bool bank_countable(const BankHost& h, int slot) {      // used_num >= 2 && start_frame < WINDOW_SIZE - 2
    return h.nobs[slot] >= 2 && h.start[slot] < vs::kBankWin - 2;
}
#define VS_BANK_STREAM(bk, s) \
    if ((s) < 0 || (s) >= (bk)->d.B) { set_error("bank: stream %d out of range", (int)(s)); return VS_ERR_INVALID; }

// getDepthVector (fm:130-140); *n = getFeatureCount (fm:15-25)
int vs_bank_get_depth_vector(vs_bank* b, int32_t stream, int32_t max_n, double* out, int32_t* n) {
    VS_BANK_GUARD(b);
    VS_BANK_STREAM(b, stream);
    if (!n) { set_error("bank: null n"); return VS_ERR_INVALID; }
    BankHost h;
    const int rc = bank_fetch(b, stream, h);
    if (rc != VS_OK) return rc;
    int k = 0;
    for (int i = 0; i < h.n; ++i) {
        const int slot = h.order[i];
        if (!bank_countable(h, slot)) continue;
        if (out && k < max_n) out[k] = 1.0 / h.depth[slot];
        ++k;
    }
    *n = k;
    if (out && k > max_n) { set_error("bank: %d depths, room for %d", k, max_n); return VS_ERR_CAPACITY; }
    return VS_OK;
}

// setDepth (fm:94-110; clear = 0) / clearDepth (fm:120-128; clear = 1)
int vs_bank_set_depth(vs_bank* b, int32_t stream, const double* x, int32_t n, int32_t clear) {
    VS_BANK_GUARD(b);
    VS_BANK_STREAM(b, stream);
    if (!x && n > 0) { set_error("bank: null x"); return VS_ERR_INVALID; }
    BankHost h;
    const int rc = bank_fetch(b, stream, h);
    if (rc != VS_OK) return rc;
    int k = 0;
    for (int i = 0; i < h.n; ++i) {
        const int slot = h.order[i];
        if (!bank_countable(h, slot)) continue;
        if (k >= n) { set_error("bank: depth vector has %d entries, the bank more countable features", n); return VS_ERR_INVALID; }
        const double depth = 1.0 / x[k++];
        if (clear) h.depth[slot] = depth;
        else if (depth < 0) h.flag[slot] = 2;
        else { h.depth[slot] = depth; h.flag[slot] = 1; }
    }
    const size_t C = b->d.cap, so = (size_t)stream * C;
    VS_CUDA_TRY(cudaMemcpyAsync(b->d.depth + so, h.depth.data(), sizeof(double) * C, cudaMemcpyHostToDevice, b->stream));
    VS_CUDA_TRY(cudaMemcpyAsync(b->d.flag + so, h.flag.data(), sizeof(int) * C, cudaMemcpyHostToDevice, b->stream));
    VS_CUDA_TRY(cudaStreamSynchronize(b->stream));
    return VS_OK;
}

static int bank_obs_row(vs_bank* b, int s, const BankHost& h, int slot, int k, double* out) {
    const int ring = (h.first[slot] + k) % vs::kBankObs;
    const double* src = b->d.obs + (((size_t)s * b->d.cap + slot) * vs::kBankObs + ring) * vs::kBankRec;
    VS_CUDA_TRY(cudaMemcpyAsync(out, src, sizeof(double) * vs::kBankRec, cudaMemcpyDeviceToHost, b->stream));
    return VS_OK;
}

// getCorresponding (fm:75-92): pairs[k] = [ray in frame_l (3), ray in frame_r (3)], list order
int vs_bank_get_corresponding(vs_bank* b, int32_t stream, int32_t frame_l, int32_t frame_r, int32_t max_n, double* pairs,
                              int32_t* n) {
    VS_BANK_GUARD(b);
    VS_BANK_STREAM(b, stream);
    if (!n) { set_error("bank: null n"); return VS_ERR_INVALID; }
    BankHost h;
    int rc = bank_fetch(b, stream, h);
    if (rc != VS_OK) return rc;
    int k = 0;
    for (int i = 0; i < h.n; ++i) {
        const int slot = h.order[i];
        if (h.start[slot] <= frame_l && h.start[slot] + h.nobs[slot] - 1 >= frame_r) ++k;
    }
    *n = k;
    if (!pairs) return VS_OK;
    if (k > max_n) { set_error("bank: %d correspondences, room for %d", k, max_n); return VS_ERR_CAPACITY; }
    std::vector<double> tmp((size_t)k * 2 * vs::kBankRec);
    int q = 0;
    for (int i = 0; i < h.n; ++i) {
        const int slot = h.order[i];
        if (!(h.start[slot] <= frame_l && h.start[slot] + h.nobs[slot] - 1 >= frame_r)) continue;
        if ((rc = bank_obs_row(b, stream, h, slot, frame_l - h.start[slot], tmp.data() + (size_t)q * 14)) != VS_OK) return rc;
        if ((rc = bank_obs_row(b, stream, h, slot, frame_r - h.start[slot], tmp.data() + (size_t)q * 14 + 7)) != VS_OK) return rc;
        ++q;
    }
    VS_CUDA_TRY(cudaStreamSynchronize(b->stream));
    for (int i = 0; i < k; ++i)
        for (int c = 0; c < 3; ++c) {
            pairs[6 * i + c] = tmp[(size_t)i * 14 + c];
            pairs[6 * i + 3 + c] = tmp[(size_t)i * 14 + 7 + c];
        }
    return VS_OK;
}

// the list of one stream: ids, start_frame, observation counts, depths, solve flags and (optional) observations
// obs[max_n][WINDOW_SIZE + 1][7]; *n = list length (VS_ERR_CAPACITY if it exceeds max_n)
int vs_bank_export(vs_bank* b, int32_t stream, int32_t max_n, int32_t* ids, int32_t* start_frame, int32_t* n_obs,
                   double* depth, int32_t* solve_flag, double* obs, int32_t* n) {
    VS_BANK_GUARD(b);
    VS_BANK_STREAM(b, stream);
    if (!n) { set_error("bank: null n"); return VS_ERR_INVALID; }
    BankHost h;
    int rc = bank_fetch(b, stream, h);
    if (rc != VS_OK) return rc;
    *n = h.n;
    if (h.n > max_n) { set_error("bank: %d features, room for %d", h.n, max_n); return VS_ERR_CAPACITY; }
    std::vector<double> slab;
    if (obs) {
        slab.resize((size_t)b->d.cap * vs::kBankObs * vs::kBankRec);
        VS_CUDA_TRY(cudaMemcpyAsync(slab.data(), b->d.obs + (size_t)stream * slab.size(), sizeof(double) * slab.size(),
                                    cudaMemcpyDeviceToHost, b->stream));
        VS_CUDA_TRY(cudaStreamSynchronize(b->stream));
        memset(obs, 0, sizeof(double) * (size_t)max_n * vs::kBankObs * vs::kBankRec);
    }
    for (int i = 0; i < h.n; ++i) {
        const int slot = h.order[i];
        if (ids) ids[i] = h.id[slot];
        if (start_frame) start_frame[i] = h.start[slot];
        if (n_obs) n_obs[i] = h.nobs[slot];
        if (depth) depth[i] = h.depth[slot];
        if (solve_flag) solve_flag[i] = h.flag[slot];
        if (obs)
            for (int k = 0; k < h.nobs[slot] && k < vs::kBankObs; ++k) {
                const int ring = (h.first[slot] + k) % vs::kBankObs;
                memcpy(obs + ((size_t)i * vs::kBankObs + k) * vs::kBankRec,
                       slab.data() + ((size_t)slot * vs::kBankObs + ring) * vs::kBankRec, sizeof(double) * vs::kBankRec);
            }
    }
    return VS_OK;
}

}  // extern "C"
