// K10 / K8 / K7 / K11 — the per-stream glue of FeatureTracker::detectAndTrack on the device.
//   post_lk_kernel       status & inBorder filter + 6-array compaction, track_cnt++, RMS-displacement
//                        gate, virtual-pinhole lift for the F-matrix RANSAC   (feature_tracker.cpp:138-173, 211-225)
//   set_mask_kernel      RANSAC-status compaction, std::sort order, greedy keep with min_dist discs
//                        (feature_tracker.cpp:62-86, 300-305)
//   fill/paint_mask      the mask image goodFeaturesToTrack receives          (feature_tracker.cpp:40-60, 84)
//   select_finish_kernel candidate sort + greedy min-distance selection (second half of
//                        cv::goodFeaturesToTrack), addPoints, updateID, undistortedPoints + velocity
//                        (feature_tracker.cpp:89-95, 309-316, 338-380)
#include "camera.cuh"
#include "glue.cuh"
#include "introsort.cuh"
#include "ransac.cuh"

namespace vs {

__device__ __forceinline__ bool in_border(float2 p, int W, int H) {
    const int x = cv_round(p.x), y = cv_round(p.y);
    return 1 <= x && x < W - 1 && 1 <= y && y < H - 1;
}

// ordered block compaction helper: returns this thread's output slot (or -1) and advances *base
__device__ __forceinline__ int block_compact_slot(bool ok, int tid, int* s_warp_cnt, int& base) {
    const unsigned int bal = __ballot_sync(0xffffffffu, ok);
    const int lane = tid & 31, warp = tid >> 5;
    if (lane == 0) s_warp_cnt[warp] = __popc(bal);
    __syncthreads();
    int off = base, total = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const int c = s_warp_cnt[k];
        if (k < warp) off += c;
        total += c;
    }
    __syncthreads();
    const int slot = ok ? off + __popc(bal & ((1u << lane) - 1u)) : -1;
    base += total;
    return slot;
}

__device__ __forceinline__ void post_lk_body(const GlueParams& g, const int s, const int tid) {
    __shared__ int s_warp_cnt[8];
    __shared__ double s_red[8];
    __shared__ int s_flag;
    const int64_t o = (int64_t)s * g.cap;
    const int n = g.n[s];
    int base = 0;
    double acc = 0.0;
    for (int chunk = 0; chunk < n; chunk += 256) {
        const int i = chunk + tid;
        bool ok = false;
        float2 nx = make_float2(0.f, 0.f), cur = nx;
        if (i < n) {
            nx = g.lk_next[o + i];
            cur = g.pts[o + i];
            ok = g.lk_status[o + i] && in_border(nx, g.W, g.H);
            if (ok && g.fb_check) {
                const float2 b = g.lk_back[o + i];
                const float dx = b.x - cur.x, dy = b.y - cur.y;
                ok = g.lk_back_status[o + i] && (dx * dx + dy * dy <= g.fb_threshold * g.fb_threshold);
            }
        }
        const int slot = block_compact_slot(ok, tid, s_warp_cnt, base);
        if (slot >= 0) {
            g.w_cur[o + slot] = cur;
            g.w_next[o + slot] = nx;
            g.w_ids[o + slot] = g.ids[o + i];
            g.w_cnt[o + slot] = g.cnt[o + i] + 1;
            g.w_un[o + slot] = g.un[o + i];
            g.w_rstatus[o + slot] = 1;
            // float - float, then widened: `double dx = next_pts[i].x - cur_pts[i].x;` (feature_tracker.cpp:161-162)
            const double dx = (double)__fsub_rn(nx.x, cur.x), dy = (double)__fsub_rn(nx.y, cur.y);
            acc += dx * dx + dy * dy;
        }
    }
    const int wn = base;
#pragma unroll
    for (int k = 16; k > 0; k >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, k);
    if ((tid & 31) == 0) s_red[tid >> 5] = acc;
    __syncthreads();
    if (tid == 0) {
        double t = 0.0;
        for (int k = 0; k < 8; ++k) t += s_red[k];
        const double rms = wn > 0 ? sqrt(t / (double)wn) : 0.0;
        bool should_reject = true;
        if (wn > 0 && rms < 2.0) should_reject = false;
        const int flag = (g.ransac_enabled && should_reject && wn >= 30) ? 1 : 0;
        s_flag = flag;
        g.wn[s] = wn;
        g.ransac_flag[s] = flag;
        int* h = g.r_hdr + (int64_t)s * kHdrInts;
        h[H_LK_IN] = n;
        h[H_LK_OK] = wn;
        h[H_RANSAC_RAN] = flag;
        h[H_RANSAC_IN] = wn;
        h[H_RMS_BITS] = __float_as_int((float)rms);
    }
    __syncthreads();
    if (s_flag) {
        const double cx = (double)g.W / 2.0, cy = (double)g.H / 2.0;
        for (int i = tid; i < 2 * wn; i += 256) {
            const int j = i >> 1;
            const float2 p = (i & 1) ? g.w_next[o + j] : g.w_cur[o + j];
            double X, Y, Z;
            lift_projective(g.cam, (double)p.x, (double)p.y, X, Y, Z);
            const float2 r = make_float2((float)(g.focal_length * X / Z + cx), (float)(g.focal_length * Y / Z + cy));
            if (i & 1) g.w_lnext[o + j] = r; else g.w_lcur[o + j] = r;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// setMask: order by track_cnt (exact libstdc++ std::sort moves), greedy keep outside painted discs
// ---------------------------------------------------------------------------------------------
static __host__ __device__ inline int pow2_at_least(int n) {
    int p = 1;
    while (p < n) p <<= 1;
    return p;
}
constexpr int kFixedPointMax = 768;   // up to here the greedy keep uses a conflict bit matrix, above it neighbour lists
constexpr int kNbrMax = 12;           // neighbour-list slots per point (overflow -> warp-serial greedy)
constexpr int kGridCellsMax = 4096;   // cells of the neighbour-search grid
int set_mask_smem_bytes(int cap) {
    const int words = (cap + 31) / 32;
    // conflict bit matrix | neighbour lists + search grid (cell starts, fill counters, members)
    const int conf = cap <= kFixedPointMax ? cap * words * 4 : cap * kNbrMax * 2 + (2 * kGridCellsMax + 1) * 4 + cap * 2;
    return pow2_at_least(cap) * 8 + cap * (4 + 4 + 4 + 4 + 4 + 4 + 4) + conf + 3 * words * 4 + 16;
}

__device__ __forceinline__ void bitonic_sort_desc_u64(unsigned long long* v, int N, int tid, int nthreads) {
    for (int k = 2; k <= N; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int t = tid; t < (N >> 1); t += nthreads) {
                const int i = 2 * j * (t / j) + (t % j);
                const int l = i + j;
                const unsigned long long a = v[i], b = v[l];
                const bool desc = (i & k) == 0;
                if ((a < b) == desc) {
                    v[i] = b;
                    v[l] = a;
                }
            }
            __syncthreads();
        }
    }
}

__global__ void __launch_bounds__(256) post_lk_kernel(GlueParams g) {
    pdl_entry(); post_lk_body(g, blockIdx.x, threadIdx.x); }

__device__ __forceinline__ void set_mask_body(const GlueParams& g, const int s, const int tid, uint8_t* dsm) {
    unsigned long long* ck = reinterpret_cast<unsigned long long*>(dsm);   // (key, position) sort keys
    int* src = reinterpret_cast<int*>(ck + pow2_at_least(g.cap));          // surviving work indices (order preserved)
    int* key = src + g.cap;                          // track_cnt of src[j]
    int* perm = key + g.cap;                         // sorted order (indices into src)
    int* kidx = perm + g.cap;                        // kept work indices
    int* lpos = kidx + g.cap;                        // partition stop lists
    int* rpos = lpos + g.cap;
    short2* acc = reinterpret_cast<short2*>(rpos + g.cap);   // kept rounded centres
    __shared__ int s_warp_cnt[8];
    __shared__ int s_kn, s_overflow;
    __shared__ int s_isort[2];
    const int lane = tid & 31;
    const int64_t o = (int64_t)s * g.cap;
    const int wn = g.wn[s];
    const bool filt = g.ransac_flag[s] != 0;
    int n2 = 0;
    for (int chunk = 0; chunk < wn; chunk += 256) {
        const int i = chunk + tid;
        const bool ok = i < wn && (!filt || g.w_rstatus[o + i]);
        const int slot = block_compact_slot(ok, tid, s_warp_cnt, n2);
        if (slot >= 0) {
            src[slot] = i;
            key[slot] = g.w_cnt[o + i];
            perm[slot] = slot;
        }
    }
    __syncthreads();
    // std::sort(track_cnt desc) with libstdc++'s exact tie order: introsort partition phase on one warp,
    // then its final insertion sort == a stable sort of what the partitions left (introsort.cuh)
    block_introsort_loop_desc(perm, key, n2, lpos, rpos, kidx, kidx + g.cap / 2, s_isort, tid, 8);
    __syncthreads();
    {
        const int N = pow2_at_least(n2);
        for (int i = tid; i < N; i += 256)
            ck[i] = i < n2 ? ((unsigned long long)(unsigned int)key[perm[i]] << 32) |
                                 ((unsigned long long)(0xFFFFu - (unsigned int)i) << 16) | (unsigned int)perm[i]
                           : 0ull;
        __syncthreads();
        if (n2 <= 256) {
            // one key per thread: its rank among the (unique) keys is its sorted position
            if (tid < n2) {
                const unsigned long long mine = ck[tid];
                int rank = 0;
                for (int i = 0; i < n2; ++i) rank += ck[i] > mine ? 1 : 0;
                perm[rank] = (int)(mine & 0xFFFFull);
            }
        } else {
            bitonic_sort_desc_u64(ck, N, tid, 256);
            for (int i = tid; i < n2; i += 256) perm[i] = (int)(ck[i] & 0xFFFFull);
        }
    }
    __syncthreads();
    if (n2 <= kFixedPointMax && g.cap <= kFixedPointMax) {
        // Greedy keep (feature_tracker.cpp:79-86) as a fixed point: kept[q] = base(q) && no kept a < q whose
        // disc covers q.  The recurrence is triangular, so Jacobi sweeps from "all base-ok points kept"
        // reach its unique solution — the sequential result — after (longest rejection chain + 1) sweeps.
        // The pair tests are hoisted out of the sweeps into a conflict bit matrix (row q, bit a: a < q and
        // the disc of a covers q) built with all threads in parallel; a sweep is then nw ANDs per point.
        const int R = g.min_dist;
        const int nw = (n2 + 31) >> 5;
        uint32_t* conf = reinterpret_cast<uint32_t*>(acc + g.cap);      // [n2][nw]
        uint32_t* kb0 = conf + (size_t)g.cap * ((g.cap + 31) >> 5);     // kept bit masks, double buffered
        uint32_t* kb1 = kb0 + ((g.cap + 31) >> 5);
        uint32_t* baseb = kb1 + ((g.cap + 31) >> 5);
        for (int q0 = 0; q0 < n2; q0 += 256) {
            const int q = q0 + tid;
            bool b = false;
            if (q < n2) {
                const float2 p = g.w_next[o + src[perm[q]]];
                const int rx = cv_round(p.x), ry = cv_round(p.y);
                acc[q] = make_short2((short)rx, (short)ry);
                b = g.base_mask[(int64_t)ry * g.mask_pitch + rx] == 255;
            }
            const unsigned bal = __ballot_sync(0xffffffffu, b);
            if (lane == 0 && q0 + (tid & ~31) < n2) {
                baseb[(q0 + tid) >> 5] = bal;
                kb0[(q0 + tid) >> 5] = bal;
            }
        }
        __syncthreads();
        // a warp takes point q, its lanes the 32 earlier points of one word: one pair test per lane, ballot = the word
        // (words above q >> 5 are never read by the sweeps)
        for (int q = tid >> 5; q < n2; q += 8) {
            const int rx = acc[q].x, ry = acc[q].y;
            for (int w = 0; w <= (q >> 5); ++w) {
                const int a = 32 * w + lane;
                bool hit = false;
                if (a < q) {
                    const short2 c = acc[a];
                    const int dy = abs(ry - (int)c.y), dx = abs(rx - (int)c.x);
                    hit = dy <= R && dx <= __ldg(g.disc_hw + dy);
                }
                const unsigned bal = __ballot_sync(0xffffffffu, hit);
                if (lane == 0) conf[q * nw + w] = bal;
            }
        }
        __syncthreads();
        uint32_t* cur = kb0;
        uint32_t* nxt = kb1;
        for (int sweep = 0; sweep <= n2; ++sweep) {
            int changed = 0;
            for (int q0 = 0; q0 < n2; q0 += 256) {
                const int q = q0 + tid;
                bool k = false;
                if (q < n2) {
                    k = (baseb[q >> 5] >> (q & 31)) & 1u;
                    const uint32_t* row = conf + q * nw;
                    for (int w = 0; k && w <= (q >> 5); ++w)
                        if (row[w] & cur[w]) k = false;
                }
                const unsigned bal = __ballot_sync(0xffffffffu, k);
                if (lane == 0 && q0 + (tid & ~31) < n2) {
                    const int wi = (q0 + tid) >> 5;
                    nxt[wi] = bal;
                    changed |= (bal != cur[wi]);
                }
            }
            const int any = __syncthreads_or(changed);
            uint32_t* t = cur;
            cur = nxt;
            nxt = t;
            if (!any) break;
        }
        // ordered compaction of the kept points straight from the bit mask
        for (int q = tid; q < n2; q += 256) {
            if ((cur[q >> 5] >> (q & 31)) & 1u) {
                int slot = __popc(cur[q >> 5] & ((1u << (q & 31)) - 1u));
                for (int w = 0; w < (q >> 5); ++w) slot += __popc(cur[w]);
                kidx[slot] = src[perm[q]];
            }
        }
        if (tid == 0) {
            int tot = 0;
            for (int w = 0; w < nw; ++w) tot += __popc(cur[w]);
            s_kn = tot;
        }
    } else {
        // Large feature sets (cap > kFixedPointMax, e.g. 2000 features): the conflict relation is sparse (kept points
        // were >= min_dist apart one frame ago), so each point gets a short list of earlier points whose disc covers
        // it — n^2/2 pair tests spread over the CTA, same-address shared-memory broadcasts — and the greedy is the
        // same Jacobi fixed point over those lists.  A list overflow falls back to the warp-serial greedy below.
        const int R = g.min_dist;
        short* nbr = reinterpret_cast<short*>(acc + g.cap);        // [cap][kNbrMax]
        int* ncount = key;                                         // track_cnt keys are dead after the sort
        int* cur = lpos;
        int* nxt = rpos;
        if (tid == 0) s_overflow = 0;
        for (int q = tid; q < n2; q += 256) {
            const float2 p = g.w_next[o + src[perm[q]]];
            const int rx = cv_round(p.x), ry = cv_round(p.y);
            acc[q] = make_short2((short)rx, (short)ry);
            cur[q] = g.base_mask[(int64_t)ry * g.mask_pitch + rx] == 255;
            kidx[q] = cur[q];                        // base-ok flag, fixed
        }
        __syncthreads();
        // uniform grid over the rounded centres, cell edge >= R: every disc that can cover a point lies in the 3x3
        // cells around it, so a point is tested against a handful of earlier points instead of all of them
        const int cell = R > 16 ? R : 16;
        const int gx = (g.W + cell - 1) / cell, gy = (g.H + cell - 1) / cell, ncell = gx * gy;
        int* cstart = reinterpret_cast<int*>(nbr + (size_t)g.cap * kNbrMax);      // [ncell + 1]
        int* cfill = cstart + kGridCellsMax + 1;                                  // [ncell]
        short* members = reinterpret_cast<short*>(cfill + kGridCellsMax);         // [cap]
        const bool use_grid = ncell <= kGridCellsMax;
        if (use_grid) {
            for (int c = tid; c <= ncell; c += 256) cstart[c] = 0;
            for (int c = tid; c < ncell; c += 256) cfill[c] = 0;
            __syncthreads();
            for (int q = tid; q < n2; q += 256) {
                const int cx = min(max((int)acc[q].x, 0) / cell, gx - 1), cy = min(max((int)acc[q].y, 0) / cell, gy - 1);
                atomicAdd(&cstart[cy * gx + cx + 1], 1);
            }
            __syncthreads();
            if (tid < 32) {                       // exclusive prefix over <= 4096 cells on one warp
                int carry = 0;
                for (int c0 = 1; c0 <= ncell; c0 += 32) {
                    const int c = c0 + lane;
                    int v = c <= ncell ? cstart[c] : 0;
#pragma unroll
                    for (int o2 = 1; o2 < 32; o2 <<= 1) {
                        const int t = __shfl_up_sync(0xffffffffu, v, o2);
                        if (lane >= o2) v += t;
                    }
                    if (c <= ncell) cstart[c] = v + carry;
                    carry += __shfl_sync(0xffffffffu, v, 31);
                }
            }
            __syncthreads();
            for (int q = tid; q < n2; q += 256) {
                const int cx = min(max((int)acc[q].x, 0) / cell, gx - 1), cy = min(max((int)acc[q].y, 0) / cell, gy - 1);
                const int c = cy * gx + cx;
                members[cstart[c] + atomicAdd(&cfill[c], 1)] = (short)q;
            }
            __syncthreads();
        }
        for (int q = tid; q < n2; q += 256) {
            const int rx = acc[q].x, ry = acc[q].y;
            int cnt = 0;
            auto exact = [&](int a) {
                const short2 c = acc[a];
                const int dy = abs(ry - (int)c.y), dx = abs(rx - (int)c.x);
                if (dy <= R && dx <= __ldg(g.disc_hw + dy)) {
                    if (cnt < kNbrMax) nbr[q * kNbrMax + cnt] = (short)a;
                    ++cnt;
                }
            };
            if (use_grid) {
                const int cx = min(max(rx, 0) / cell, gx - 1), cy = min(max(ry, 0) / cell, gy - 1);
                for (int yy = max(cy - 1, 0); yy <= min(cy + 1, gy - 1); ++yy) {
                    const int c0 = yy * gx + max(cx - 1, 0), c1 = yy * gx + min(cx + 1, gx - 1);
                    for (int j = cstart[c0]; j < cstart[c1 + 1]; ++j) {      // the row's 2-3 cells are contiguous
                        const int a = members[j];
                        if (a < q) exact(a);
                    }
                }
            } else {
                for (int a = 0; a < q; ++a) exact(a);
            }
            ncount[q] = cnt;
            if (cnt > kNbrMax) s_overflow = 1;
        }
        __syncthreads();
        if (!s_overflow) {
            for (int sweep = 0; sweep <= n2; ++sweep) {
                int changed = 0;
                for (int q = tid; q < n2; q += 256) {
                    int k = kidx[q];
                    if (k) {
                        const int c = ncount[q];
                        for (int j = 0; j < c; ++j)
                            if (cur[nbr[q * kNbrMax + j]]) { k = 0; break; }
                    }
                    nxt[q] = k;
                    changed |= (k != cur[q]);
                }
                const int any = __syncthreads_or(changed);
                int* t = cur;
                cur = nxt;
                nxt = t;
                if (!any) break;
            }
            int base = 0;
            for (int chunk = 0; chunk < n2; chunk += 256) {
                const int q = chunk + tid;
                const bool ok = q < n2 && cur[q];
                const int widx = q < n2 ? src[perm[q]] : 0;
                __syncthreads();                     // kidx is about to be reused as the kept list
                const int slot = block_compact_slot(ok, tid, s_warp_cnt, base);
                if (slot >= 0) nxt[slot] = widx;
            }
            __syncthreads();
            for (int j = tid; j < base; j += 256) kidx[j] = nxt[j];
            if (tid == 0) s_kn = base;
        }
    }
    __syncthreads();
    if (!(n2 <= kFixedPointMax && g.cap <= kFixedPointMax) && s_overflow && tid < 32) {
        const int R = g.min_dist;
        int A = 0;
        for (int q0 = 0; q0 < n2; q0 += 32) {
            const int q = q0 + lane;
            bool alive = false;
            int rx = 0, ry = 0, widx = -1;
            if (q < n2) {
                widx = src[perm[q]];
                const float2 p = g.w_next[o + widx];
                rx = cv_round(p.x);
                ry = cv_round(p.y);
                alive = g.base_mask[(int64_t)ry * g.mask_pitch + rx] == 255;
                for (int a = 0; a < A && alive; ++a) {
                    const int dy = abs(ry - (int)acc[a].y), dx = abs(rx - (int)acc[a].x);
                    if (dy <= R && dx <= g.disc_hw[dy]) alive = false;
                }
            }
            unsigned int m = __ballot_sync(0xffffffffu, alive);
            while (m) {
                const int l = __ffs(m) - 1;
                const int ax = __shfl_sync(0xffffffffu, rx, l), ay = __shfl_sync(0xffffffffu, ry, l);
                if (lane == l) {
                    acc[A] = make_short2((short)rx, (short)ry);
                    kidx[A] = widx;
                    alive = false;
                } else if (alive && lane > l) {
                    const int dy = abs(ry - ay), dx = abs(rx - ax);
                    if (dy <= R && dx <= g.disc_hw[dy]) alive = false;
                }
                ++A;
                m = __ballot_sync(0xffffffffu, alive);
            }
            __syncwarp();
        }
        if (lane == 0) s_kn = A;
    }
    __syncthreads();
    const int kn = s_kn;
    for (int j = tid; j < kn; j += 256) {
        const int i = kidx[j];
        g.k_pts[o + j] = g.w_next[o + i];
        g.k_ids[o + j] = g.w_ids[o + i];
        g.k_cnt[o + j] = g.w_cnt[o + i];
        g.k_un[o + j] = g.w_un[o + i];
    }
    if (tid == 0) {
        g.kn[s] = kn;
        const int want = g.max_cnt - kn;
        g.want[s] = want;
        g.cand_count[s] = 0;
        g.max_enc[s] = 0;
        g.raw_count[s] = 0;
        g.l_enc[s] = 0;
        int* h = g.r_hdr + (int64_t)s * kHdrInts;
        h[H_RANSAC_OUT] = n2;
        h[H_MASK_KEPT] = kn;
        h[H_WANT] = want;
    }
}

__global__ void __launch_bounds__(256) set_mask_kernel(GlueParams g) {
    pdl_entry();
    extern __shared__ __align__(16) uint8_t dsm[];
    set_mask_body(g, blockIdx.x, threadIdx.x, dsm);
}

// post_lk + set_mask in one launch (no RANSAC stage between them): both are one CTA per stream, and a CTA only
// consumes what it wrote itself, so a block barrier is all the ordering needed
__global__ void __launch_bounds__(256) post_lk_set_mask_kernel(GlueParams g) {
    pdl_entry();
    extern __shared__ __align__(16) uint8_t dsm[];
    post_lk_body(g, blockIdx.x, threadIdx.x);
    __syncthreads();
    set_mask_body(g, blockIdx.x, threadIdx.x, dsm);
}

// ... and with the device F-matrix RANSAC between them (VS_RANSAC_DEVICE): still one CTA per stream end to end
__global__ void __launch_bounds__(256) post_lk_ransac_set_mask_kernel(GlueParams g, unsigned int seed) {
    pdl_entry();
    extern __shared__ __align__(16) uint8_t dsm[];
    static_assert(kRansacThreads == 256, "glue kernels run 256 threads per stream");
    post_lk_body(g, blockIdx.x, threadIdx.x);
    __syncthreads();
    ransac_body(g, seed, blockIdx.x, threadIdx.x, dsm);
    __syncthreads();
    set_mask_body(g, blockIdx.x, threadIdx.x, dsm);
}

// ---------------------------------------------------------------------------------------------
// K7: second half of cv::goodFeaturesToTrack — threshold 0.01 * max, order (eig desc, address desc),
// greedy min-distance accept until `want` corners — then the rest of the frame: addPoints, updateID,
// undistortedPoints + velocity (feature_tracker.cpp:89-95, 309-316, 338-380).
//
// The greedy consumes candidates in descending key order and stops at `want`, so only a prefix of the
// order is ever looked at.  A 1024-bin histogram over the key range (threshold, max] finds the bin
// boundary above which the M = max(64, 4*want) best keys live; only those are sorted (bitonic, shared
// memory) and fed to the greedy.  If fewer than `want` are accepted the prefix is widened 4x and the
// selection repeated, so the result always equals a full sort + greedy.
// ---------------------------------------------------------------------------------------------
constexpr int kSortSmemKeys = 4096;
constexpr int kFinishThreads = 256;
constexpr int kSelBins = 1024;

__device__ __forceinline__ void bitonic_sort_desc(unsigned long long* v, int N, int tid, int nthreads) {
    for (int k = 2; k <= N; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int t = tid; t < (N >> 1); t += nthreads) {
                const int i = 2 * j * (t / j) + (t % j);
                const int l = i + j;
                const unsigned long long a = v[i], b = v[l];
                const bool desc = (i & k) == 0;
                if ((a < b) == desc) {
                    v[i] = b;
                    v[l] = a;
                }
            }
            __syncthreads();
        }
    }
}

// greedy min-distance accept over keys sorted descending; returns the number accepted (<= want)
__device__ __forceinline__ int greedy_accept(const unsigned long long* v, int nv, int want, int W, int md2, short2* newp,
                                             int lane) {
    int A = 0;
    for (int q0 = 0; q0 < nv && A < want; q0 += 32) {
        const int q = q0 + lane;
        bool alive = q < nv;
        int x = 0, y = 0;
        if (alive) {
            const unsigned int lin = (unsigned int)(v[q] & 0xffffffffull);
            y = lin / W;
            x = lin - y * W;
            for (int a = 0; a < A && alive; ++a) {
                const int dx = x - (int)newp[a].x, dy = y - (int)newp[a].y;
                if (dx * dx + dy * dy < md2) alive = false;
            }
        }
        unsigned int m = __ballot_sync(0xffffffffu, alive);
        while (m && A < want) {
            const int l = __ffs(m) - 1;
            const int ax = __shfl_sync(0xffffffffu, x, l), ay = __shfl_sync(0xffffffffu, y, l);
            if (lane == l) {
                newp[A] = make_short2((short)x, (short)y);
                alive = false;
            } else if (alive && lane > l) {
                const int dx = x - ax, dy = y - ay;
                if (dx * dx + dy * dy < md2) alive = false;
            }
            ++A;
            m = __ballot_sync(0xffffffffu, alive);
        }
        __syncwarp();
    }
    return A;
}

__global__ void __launch_bounds__(kFinishThreads) select_finish_kernel(GlueParams g) {
    pdl_entry();
    extern __shared__ __align__(16) uint8_t dsm[];
    unsigned long long* skeys = reinterpret_cast<unsigned long long*>(dsm);       // [kSortSmemKeys]
    int* hist = reinterpret_cast<int*>(skeys + kSortSmemKeys);                     // [kSelBins]
    short2* newp = reinterpret_cast<short2*>(hist + kSelBins);                     // [cap]
    __shared__ int s_nnew, s_cnt, s_bin, s_total;
    const int s = blockIdx.x, tid = threadIdx.x, lane = tid & 31;
    const int64_t o = (int64_t)s * g.cap;
    const int kn = g.kn[s];
    int want = g.want[s];
    if (want > g.cap - kn) want = g.cap - kn;
    if (tid == 0) s_nnew = 0;
    int nc = 0;
    if (want > 0) {
        nc = g.cand_count[s];
        if (nc > g.cand_cap) nc = g.cand_cap;
    }
    __syncthreads();
    if (nc > 0) {
        unsigned long long* gk = g.cand + (int64_t)s * g.cand_cap;
        const unsigned int me = g.max_enc[s];
        const float max_val = me ? float_order_dec(me) : 0.f;       // minMaxLoc on an empty mask gives 0
        const float thr = (float)((double)max_val * 0.01);          // qualityLevel 0.01
        // keys are (float bits << 32 | address); eig > thr >= 0, so the float bits order like integers
        const unsigned int tb = __float_as_uint(thr > 0.f ? thr : 0.f);
        const unsigned int mb = __float_as_uint(max_val > 0.f ? max_val : 0.f);
        int shift = 0;
        while (mb > tb && ((mb - tb) >> shift) >= (unsigned)kSelBins) ++shift;
        for (int i = tid; i < kSelBins; i += kFinishThreads) hist[i] = 0;
        if (tid == 0) s_total = 0;
        __syncthreads();
        int mine = 0;
        for (int i = tid; i < nc; i += kFinishThreads) {
            const unsigned int eb = (unsigned int)(gk[i] >> 32);
            if (__uint_as_float(eb) > thr) {
                atomicAdd(&hist[min((eb - tb) >> shift, (unsigned)kSelBins - 1)], 1);
                ++mine;
            }
        }
        if (mine) atomicAdd(&s_total, mine);
        __syncthreads();
        const int n_valid = s_total;
        const int md2 = g.min_dist * g.min_dist;
        int target = max(64, 4 * want);
        while (n_valid > 0) {
            __syncthreads();
            // highest bin boundary b such that bins >= b hold at least `target` keys (or everything)
            if (tid < 32) {
                // lane L owns bins [32 L, 32 L + 32); suffix sums over the lanes find the lane where the running
                // total (from the top bin down) reaches `target`, that lane then walks its own 32 bins
                static_assert(kSelBins == 1024, "one warp x 32 bins");
                int part = 0;
#pragma unroll 8
                for (int b = 0; b < 32; ++b) part += hist[32 * lane + ((b + lane) & 31)];   // rotated: no bank conflicts
                int suf = part;
#pragma unroll
                for (int o2 = 1; o2 < 32; o2 <<= 1) {
                    const int t = __shfl_down_sync(0xffffffffu, suf, o2);
                    if (lane + o2 < 32) suf += t;
                }
                const unsigned m = __ballot_sync(0xffffffffu, suf >= target);
                if (m == 0) {                        // fewer than `target` keys in total: take everything
                    if (lane == 0) { s_bin = 0; s_cnt = 0; s_total = suf; }
                } else if (lane == 31 - __clz(m)) {
                    int acc = suf - part, b = 32 * lane + 31;
                    for (;; --b) {
                        acc += hist[b];
                        if (acc >= target || b == 32 * lane) break;
                    }
                    s_bin = b;
                    s_cnt = 0;
                    s_total = acc;
                }
            }
            __syncthreads();
            const int bsel = s_bin, nsel = s_total;
            const bool all = nsel >= n_valid;
            unsigned long long* v;
            int N = 1;
            while (N < nsel) N <<= 1;
            if (N <= kSortSmemKeys) {
                v = skeys;
                for (int i = tid; i < nc; i += kFinishThreads) {
                    const unsigned long long key = gk[i];
                    const unsigned int eb = (unsigned int)(key >> 32);
                    if (__uint_as_float(eb) > thr && (int)min((eb - tb) >> shift, (unsigned)kSelBins - 1) >= bsel)
                        skeys[atomicAdd(&s_cnt, 1)] = key;
                }
                for (int i = nsel + tid; i < N; i += kFinishThreads) skeys[i] = 0ull;
            } else {
                // too many for shared memory (first frames of feature-heavy configs): sort the whole list in place
                N = 1;
                while (N < nc) N <<= 1;
                for (int i = tid; i < N; i += kFinishThreads)
                    if (i >= nc || !(__uint_as_float((unsigned int)(gk[i] >> 32)) > thr)) gk[i] = 0ull;
                v = gk;
            }
            __syncthreads();
            if (v == skeys && nsel <= kFinishThreads) {
                // the usual case (a few dozen keys): rank sort, one key per thread (keys are unique: the low word is
                // the pixel address), instead of log^2 N barrier-separated bitonic stages
                unsigned long long mine = 0ull;
                int rank = 0;
                if (tid < nsel) {
                    mine = skeys[tid];
                    for (int i = 0; i < nsel; ++i) rank += skeys[i] > mine ? 1 : 0;
                }
                __syncthreads();
                if (tid < nsel) skeys[rank] = mine;
                __syncthreads();
            } else {
                bitonic_sort_desc(v, N, tid, kFinishThreads);
            }
            const int nv = (v == gk) ? n_valid : nsel;
            if (tid < 32) {
                const int A = greedy_accept(v, nv, want, g.W, md2, newp, lane);
                if (lane == 0) s_nnew = A;
            }
            __syncthreads();
            if (s_nnew >= want || all || v == gk) break;
            target *= 4;
        }
    }
    __syncthreads();
    const int n_new = s_nnew;
    const int total = kn + n_new;
    const int id0 = g.next_id[s];
    const double t_cur = g.cur_time[s];
    const double dt = t_cur - g.prev_time[s];
    for (int i = tid; i < total; i += kFinishThreads) {
        float2 p, un_prev = make_float2(0.f, 0.f);
        int id, c;
        if (i < kn) {
            p = g.k_pts[o + i];
            id = g.k_ids[o + i];
            c = g.k_cnt[o + i];
            un_prev = g.k_un[o + i];
        } else {
            p = make_float2((float)newp[i - kn].x, (float)newp[i - kn].y);
            id = id0 + (i - kn);
            c = 1;
        }
        double X, Y, Z;
        lift_projective(g.cam, (double)p.x, (double)p.y, X, Y, Z);
        const float2 un = make_float2((float)(X / Z), (float)(Y / Z));
        // velocity: only features whose id was already assigned when the previous frame's map was
        // built (track_cnt >= 3 now) are found in prev_undistorted_pts_map (feature_tracker.cpp:351-378)
        float2 vel = make_float2(0.f, 0.f);
        if (i < kn && c >= 3 && dt > 1e-4) {
            vel.x = (float)((double)__fsub_rn(un.x, un_prev.x) / dt);
            vel.y = (float)((double)__fsub_rn(un.y, un_prev.y) / dt);
        }
        g.pts[o + i] = p;
        g.ids[o + i] = id;
        g.cnt[o + i] = c;
        g.un[o + i] = un;
        g.r_pts[o + i] = p;
        g.r_ids[o + i] = id;
        g.r_cnt[o + i] = c;
        g.r_un[o + i] = un;
        g.r_vel[o + i] = vel;
    }
    if (tid == 0) {
        g.n[s] = total;
        g.next_id[s] = id0 + n_new;
        g.prev_time[s] = t_cur;
        int* h = g.r_hdr + (int64_t)s * kHdrInts;
        h[H_N] = total;
        h[H_NNEW] = n_new;
        h[H_NCAND] = nc;
    }
}

__global__ void undistort_kernel(CameraDev cam, const float2* __restrict__ pts, int n, float2* __restrict__ out) {
    pdl_entry();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double X, Y, Z;
    lift_projective(cam, (double)pts[i].x, (double)pts[i].y, X, Y, Z);
    out[i] = make_float2((float)(X / Z), (float)(Y / Z));
}

void launch_post_lk(const GlueParams& g, cudaStream_t st, int64_t* launches) {
    launch_pdl(post_lk_kernel, dim3(g.n_streams), dim3(256), 0, st, g);
    if (launches) *launches += 1;
}

void launch_set_mask(const GlueParams& g, cudaStream_t st, int64_t* launches, bool with_post_lk, bool with_ransac,
                     unsigned int seed) {
    int smem = set_mask_smem_bytes(g.cap);
    if (with_ransac && smem < g.cap * 16) smem = g.cap * 16;
    const void* k = with_post_lk ? (with_ransac ? reinterpret_cast<const void*>(post_lk_ransac_set_mask_kernel)
                                                : reinterpret_cast<const void*>(post_lk_set_mask_kernel))
                                 : reinterpret_cast<const void*>(set_mask_kernel);
    ensure_dyn_smem(k, smem);
    if (with_post_lk && with_ransac) launch_pdl(post_lk_ransac_set_mask_kernel, dim3(g.n_streams), dim3(256), smem, st, g, seed);
    else if (with_post_lk) launch_pdl(post_lk_set_mask_kernel, dim3(g.n_streams), dim3(256), smem, st, g);
    else launch_pdl(set_mask_kernel, dim3(g.n_streams), dim3(256), smem, st, g);
    if (launches) *launches += 1;
}

int select_finish_smem_bytes(int cap) { return kSortSmemKeys * 8 + kSelBins * 4 + cap * 4; }

// largest dynamic shared-memory request any glue launch makes for a handle of this capacity (vs_tracker_create checks
// it against the device's opt-in limit, so an oversized capacity fails at creation, not at the first track call)
int glue_max_smem_bytes(int cap) {
    int m = set_mask_smem_bytes(cap);
    if (m < cap * 16) m = cap * 16;                 // fused / stand-alone device RANSAC
    const int sf = select_finish_smem_bytes(cap);
    return m > sf ? m : sf;
}

void launch_select_finish(const GlueParams& g, cudaStream_t st, int64_t* launches) {
    const int smem = select_finish_smem_bytes(g.cap);
    ensure_dyn_smem(reinterpret_cast<const void*>(select_finish_kernel), smem);
    launch_pdl(select_finish_kernel, dim3(g.n_streams), dim3(kFinishThreads), smem, st, g);
    if (launches) *launches += 1;
}

void launch_undistort(const CameraDev& cam, const float2* pts, int n, float2* out, cudaStream_t st) {
    if (n <= 0) return;
    undistort_kernel<<<(n + 127) / 128, 128, 0, st>>>(cam, pts, n, out);
}

}  // namespace vs
