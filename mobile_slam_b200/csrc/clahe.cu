// K1/K2 — CLAHE(clipLimit 3.0, 8x8 tiles), bit-exact to cv::CLAHE::apply on CV_8UC1.
// Replaces the clahe_->apply call at /root/reference/src/frontend/feature_tracker.cpp:102-106.
// Arithmetic spec: SURVEY.md S1 (integer histogram/clip/redistribute, float32 LUT scale and
// bilinear LUT blend with no FMA contraction, round-half-even).
#include <cuda_fp16.h>

#include <cmath>
#include <cstdlib>
#include <type_traits>

#include "kernels.h"

namespace vs {

// ---------------------------------------------------------------------------------------------
// K1: one CTA per (tile, stream): 256-bin histogram -> clip -> redistribute -> scan -> LUT.
//
// Histogram without atomics: every thread owns a private set of 256 byte counters, packed four to a
// 32-bit word (bins v, v+64, v+128, v+192 share word row v & 63) and laid out so that thread t only
// ever touches shared-memory bank t % 32 — the read-modify-write of a pixel is conflict free no
// matter which bins the 32 lanes of a warp hit.  A thread sees at most 255 pixels (the launcher
// picks the CTA size from the tile area), so the byte counters cannot overflow.  The private
// counters are then summed with packed 16-bit adds.
// ---------------------------------------------------------------------------------------------
// v >> s (0 < s < 32) as mul.hi(v, 2^(32 - s)): same value, issued to the multiplier pipe instead of the shifter
__device__ __forceinline__ uint32_t shr_mul(uint32_t v, int s) {
    if (s == 0) return v;
    uint32_t d;
    asm("mul.hi.u32 %0, %1, %2;" : "=r"(d) : "r"(v), "r"(1u << (32 - s)));
    return d;
}

template <int NT, bool FULLW>
__global__ void __launch_bounds__(NT) clahe_lut_kernel(const uint8_t* __restrict__ src, int pitch, int64_t img_stride,
                                                       int tw, int th, int clip, float lut_scale,
                                                       uint8_t* __restrict__ luts) {
    pdl_entry();
    // [64 word rows][NT threads]; static up to 32 KB so its address is an immediate of the shared atomics below
    extern __shared__ __align__(16) uint32_t s_cnt_dyn[];
    __shared__ __align__(16) uint32_t s_cnt_static[NT <= 128 ? 64 * NT : 4];
    uint32_t* const s_cnt = NT <= 128 ? s_cnt_static : s_cnt_dyn;
    __shared__ uint32_t s_part[NT >= 64 ? NT / 64 : 1][64][2];
    __shared__ int s_hist[256];
    __shared__ unsigned int s_warp_tot[NT / 32];
    __shared__ unsigned int s_clipped;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.x, s = blockIdx.y;
    const int tx = tile % kClaheTiles, ty = tile / kClaheTiles;
    const uint8_t* img = src + (int64_t)s * img_stride;
#pragma unroll 4
    for (int r = 0; r < 64; ++r) s_cnt[r * NT + tid] = 0;
    if (tid == 0) s_clipped = 0;
    // no barrier needed before the pixel loop: a thread only touches its own column of s_cnt

    // aligned words covering the tile's columns [x0, x0 + tw); the first / last word may be partial
    const int x0 = tx * tw, y0 = ty * th;
    const int xa = x0 & ~3, nw = ((x0 + tw - 1) >> 2) - (xa >> 2) + 1;
    const uint32_t tid4 = 4u * (uint32_t)tid;
    // thread walks the tile's words in row-major order, NT apart: (y, q) advance incrementally
    const int dq = NT % nw, dy = NT / nw;
    const int yt = tid / nw;
    int q = tid - yt * nw;
    const uint8_t* rowp = img + (int64_t)(y0 + yt) * pitch + xa;
    // a thread's words come in trips of four (four independent loads in flight before the dependent shared-memory
    // updates); only the last trip can hold past-the-end slots, so the full trips carry no validity test at all
    const int total_words = nw * th;
    const int n_mine = tid < total_words ? (total_words - tid + NT - 1) / NT : 0;
    auto trip = [&](auto tail_tag, int n_valid) {
        constexpr bool TAIL = decltype(tail_tag)::value;
        uint32_t v[4];
        int lo[4], hi[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const bool ok = !TAIL || u < n_valid;
            v[u] = ok ? __ldg(reinterpret_cast<const uint32_t*>(rowp) + q) : 0u;
            const int xw = xa + 4 * q;
            lo[u] = ok ? max(x0 - xw, 0) : 4;              // valid bytes of this word: lo .. hi-1
            hi[u] = min(x0 + tw - xw, 4);
            q += dq;
            rowp += (int64_t)dy * pitch;
            if (q >= nw) {
                q -= nw;
                rowp += pitch;
            }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                // FULLW: every tile starts on a word boundary and is a whole number of words wide: every byte of a valid
                // word counts
                const bool valid = FULLW ? (!TAIL || lo[u] == 0) : (k >= lo[u] && k < hi[u]);
                // The loop is bound by the integer-ALU pipe (shifts, logic), the multiplier pipe idles: right shifts are
                // therefore taken as the high half of a multiplication by a power of two (mul.hi on the FMA pipe).
                // byte lane of bin b = b >> 6: the counter increment is 1 << 8 * (b >> 6)
                const uint32_t inc = valid ? (1u << (shr_mul(v[u], 8 * k + 3) & 0x18u)) : 0u;
                // byte offset of word row (b & 63) in this thread's column, taken straight from the pixel word: one
                // shift + one mask-or per pixel (NT * 4 is a power of two)
                constexpr int SH = (NT == 32 ? 7 : NT == 64 ? 8 : NT == 128 ? 9 : 10);     // log2(NT * 4)
                const uint32_t off = ((8 * k >= SH ? shr_mul(v[u], 8 * k - SH) : v[u] << (SH - 8 * k)) & (63u << SH)) | tid4;
                // ATOMS without a result: one instruction, no load-add-store chain
                atomicAdd(reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(s_cnt) + off), inc);
            }
        }
    };
#pragma unroll 1
    for (int t4 = 0; t4 < (n_mine >> 2); ++t4) trip(std::false_type{}, 4);
    if (n_mine & 3) trip(std::true_type{}, n_mine & 3);
    __syncthreads();
    // sum the NT private copies (packed 16-bit sums of bytes (0, 2) and (1, 3)), rotated so lanes hit distinct banks
    if constexpr (NT == 32) {
        // one warp per tile: lane sums word rows lane and lane + 32 over the 32 copies (<= 32 * 255 per 16-bit lane)
#pragma unroll
        for (int rr = 0; rr < 2; ++rr) {
            const int r = tid + 32 * rr;
            uint32_t a = 0, b = 0;
            const uint32_t* row = s_cnt + r * NT;
#pragma unroll 8
            for (int j = 0; j < 32; ++j) {
                const uint32_t w = row[(j + lane) & 31];
                a += w & 0x00FF00FFu;
                b += (w >> 8) & 0x00FF00FFu;
            }
            s_hist[r] = (int)(a & 0xFFFFu);
            s_hist[r + 64] = (int)(b & 0xFFFFu);
            s_hist[r + 128] = (int)(a >> 16);
            s_hist[r + 192] = (int)(b >> 16);
        }
    } else {
        // thread -> (word row r, 64-column group cg)
        {
            const int r = tid & 63, cg = tid >> 6;
            uint32_t a = 0, b = 0;
            const uint32_t* row = s_cnt + r * NT + cg * 64;
#pragma unroll 8
            for (int j = 0; j < 64; ++j) {
                const uint32_t w = row[(j + lane) & 63];
                a += w & 0x00FF00FFu;
                b += (w >> 8) & 0x00FF00FFu;
            }
            s_part[cg][r][0] = a;
            s_part[cg][r][1] = b;
        }
        __syncthreads();
        if (tid < 64) {
            uint32_t a = 0, b = 0;
#pragma unroll
            for (int cg = 0; cg < NT / 64; ++cg) {     // per-16-bit-lane totals stay below 65536 (tile area checked by the launcher)
                a += s_part[cg][tid][0];
                b += s_part[cg][tid][1];
            }
            s_hist[tid] = (int)(a & 0xFFFFu);            // byte 0 -> bin r
            s_hist[tid + 64] = (int)(b & 0xFFFFu);       // byte 1 -> bin r + 64
            s_hist[tid + 128] = (int)(a >> 16);          // byte 2 -> bin r + 128
            s_hist[tid + 192] = (int)(b >> 16);          // byte 3 -> bin r + 192
        }
    }
    __syncthreads();
    // clip, count the excess; thread t owns bins t*BPT .. t*BPT+BPT-1
    constexpr int BPT = 256 / NT;
    int h[BPT];
    unsigned int e = 0;
#pragma unroll
    for (int k = 0; k < BPT; ++k) {
        h[k] = s_hist[tid * BPT + k];
        if (h[k] > clip) {
            e += (unsigned int)(h[k] - clip);
            h[k] = clip;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) e += __shfl_xor_sync(0xffffffffu, e, o);
    if (lane == 0 && e) atomicAdd(&s_clipped, e);
    __syncthreads();
    const int clipped = (int)s_clipped;
    const int residual = clipped & 255;    // clipped % 256
    int step = residual ? 256 / residual : 1;
    if (step < 1) step = 1;
    unsigned int run = 0;
#pragma unroll
    for (int k = 0; k < BPT; ++k) {
        const int bin = tid * BPT + k;
        h[k] += clipped >> 8;              // clipped / 256
        // serial loop "for (i = 0; i < 256 && residual > 0; i += step, residual--) hist[i]++"
        if (residual && (bin % step) == 0 && (bin / step) < residual) h[k] += 1;
        run += (unsigned int)h[k];
        h[k] = (int)run;                   // inclusive prefix inside the thread
    }
    unsigned int v = run;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned int n = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += n;
    }
    if (lane == 31) s_warp_tot[warp] = v;
    __syncthreads();
    unsigned int base = v - run;
#pragma unroll
    for (int k = 0; k < NT / 32; ++k)
        if (k < warp) base += s_warp_tot[k];
    uint32_t packed = 0;
#pragma unroll
    for (int k = 0; k < BPT; ++k) {
        int q = __float2int_rn(__fmul_rn((float)(int)(base + (unsigned int)h[k]), lut_scale));
        q = q < 0 ? 0 : (q > 255 ? 255 : q);
        packed |= (uint32_t)q << (8 * (k & 3));
        if ((k & 3) == 3 || k == BPT - 1) {
            uint8_t* d = luts + ((int64_t)s * (kClaheTiles * kClaheTiles) + tile) * 256 + tid * BPT + (k & ~3);
            if (BPT >= 4) *reinterpret_cast<uint32_t*>(d) = packed;
            else
                for (int j = 0; j <= (k & 3); ++j) d[j] = (uint8_t)(packed >> (8 * j));
            packed = 0;
        }
    }
}

// Generic K1 (images whose size is not a multiple of 8, or tiles above 65 280 pixels): shared-memory
// atomics on warp-private sub-histograms, pixels fetched one by one with the REFLECT_101 padding
// cv::CLAHE applies to the bottom/right.
__global__ void __launch_bounds__(256) clahe_lut_generic_kernel(const uint8_t* __restrict__ src, int pitch,
                                                                int64_t img_stride, int W, int H, int tw, int th,
                                                                int clip, float lut_scale,
                                                                uint8_t* __restrict__ luts) {
    pdl_entry();
    __shared__ unsigned int hist[8][256];
    __shared__ unsigned int warp_tot[8];
    __shared__ unsigned int s_clipped;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int tile = blockIdx.x, s = blockIdx.y;
    const int tx = tile % kClaheTiles, ty = tile / kClaheTiles;
    const uint8_t* img = src + (int64_t)s * img_stride;
#pragma unroll
    for (int k = 0; k < 8; ++k) hist[k][tid] = 0;
    if (tid == 0) s_clipped = 0;
    __syncthreads();
    const int x0 = tx * tw, y0 = ty * th;
    for (int idx = tid; idx < tw * th; idx += 256) {
        const int y = idx / tw, x = idx - y * tw;
        const int gx = reflect101(x0 + x, W), gy = reflect101(y0 + y, H);
        atomicAdd(&hist[warp][img[(int64_t)gy * pitch + gx]], 1u);
    }
    __syncthreads();
    int h = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) h += (int)hist[k][tid];
    int excess = h > clip ? h - clip : 0;
    h = h > clip ? clip : h;
    unsigned int e = (unsigned int)excess;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) e += __shfl_xor_sync(0xffffffffu, e, o);
    if (lane == 0 && e) atomicAdd(&s_clipped, e);
    __syncthreads();
    const int clipped = (int)s_clipped;
    h += clipped >> 8;
    const int residual = clipped & 255;
    if (residual) {
        int step = 256 / residual;
        if (step < 1) step = 1;
        if ((tid % step) == 0 && (tid / step) < residual) h += 1;
    }
    unsigned int v = (unsigned int)h;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        unsigned int n = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += n;
    }
    if (lane == 31) warp_tot[warp] = v;
    __syncthreads();
    unsigned int base = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k)
        if (k < warp) base += warp_tot[k];
    int q = __float2int_rn(__fmul_rn((float)(int)(v + base), lut_scale));
    q = q < 0 ? 0 : (q > 255 ? 255 : q);
    luts[((int64_t)s * (kClaheTiles * kClaheTiles) + tile) * 256 + tid] = (uint8_t)q;
}

// ---------------------------------------------------------------------------------------------
// K2: bilinear blend of the four neighbouring tile LUTs.
//
// One CTA = one interpolation band (the rows that share the same pair of tile rows) x one chunk of
// 128 columns.  The CTA first gathers, for every interpolation cell its chunk touches, a table of
// 256 entries holding the four LUT values a pixel of that cell needs (as 4 x fp16, exact for 0..255),
// so a pixel costs ONE 8-byte shared-memory load instead of four byte loads; then each warp walks
// down the band, one aligned 32-bit word (4 pixels) per lane and row, with the per-column weights
// held in registers.
// ---------------------------------------------------------------------------------------------
constexpr int kApplyCells = kClaheTiles + 1;   // interpolation cells per row: cx = -1 .. 7
// CTAs per band (each rebuilds the band's cell tables): chosen by the launcher — 1 when the batch alone fills the GPU
// (the table build is ~10 % of a CTA's instructions), more with few streams so that enough CTAs exist

struct ClaheBands {
    int start[kClaheTiles + 2];   // band b (cell row b-1) covers rows start[b] .. start[b+1]-1
};

__device__ __forceinline__ float round_magic(float r) {   // round-half-even of 0 <= r <= 255 into the low byte
    return __fadd_rn(r, 12582912.0f);
}

// blockDim.x = 32 * (number of 256-column chunks): warp w owns columns 256w .. 256w+255 of the band and every lane
// 8 consecutive pixels of a row (one 64-bit load / store).  The four LUT values a pixel blends are kept as 4 x fp16 per
// (cell column, pixel value) in shared memory — one 64-bit load per pixel (a float4 table halves the instruction count
// of the unpack but doubles the shared-memory wavefronts and was measured slower) — the horizontal weights are lane
// constants, the vertical weights of the band's rows sit in a small shared table.
constexpr int kApplyMaxBandRows = 512;
__global__ void __launch_bounds__(1024) clahe_apply_kernel(const uint8_t* __restrict__ src, int src_pitch,
                                                           int64_t src_stride, uint8_t* __restrict__ dst,
                                                           int dst_pitch, int64_t dst_stride, int W, int H,
                                                           float inv_tw, float inv_th, ClaheBands bands,
                                                           const uint8_t* __restrict__ luts, int row_split) {
    pdl_entry();
    __shared__ __align__(16) uint2 s_cell[kApplyCells][256];         // (p1, p2 | q1, q2) as 4 x fp16 (bytes are exact in fp16)
    __shared__ float2 s_ya[kApplyMaxBandRows];                        // (ya1, ya) of the CTA's rows
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int s = blockIdx.y;
    const int band = blockIdx.x / row_split, part = blockIdx.x - band * row_split;
    const int b0 = bands.start[band], b1 = bands.start[band + 1];
    const int per = (b1 - b0 + row_split - 1) / row_split;
    const int y_begin = b0 + part * per, y_end = min(y_begin + per, b1);
    if (y_begin >= y_end) return;
    const int cy = band - 1;                                         // cell row: tile rows (cy, cy+1) clamped
    const int ty1 = max(cy, 0), ty2 = min(cy + 1, kClaheTiles - 1);
    const uint8_t* L = luts + (int64_t)s * (kClaheTiles * kClaheTiles * 256);
    for (int idx = tid; idx < kApplyCells * 256; idx += blockDim.x) {
        const int c = idx >> 8, v = idx & 255;
        const int cx = c - 1;
        const int tx1 = max(cx, 0), tx2 = min(cx + 1, kClaheTiles - 1);
        const __half2 top = __halves2half2(__ushort2half_rn(L[(ty1 * kClaheTiles + tx1) * 256 + v]),
                                           __ushort2half_rn(L[(ty1 * kClaheTiles + tx2) * 256 + v]));
        const __half2 bot = __halves2half2(__ushort2half_rn(L[(ty2 * kClaheTiles + tx1) * 256 + v]),
                                           __ushort2half_rn(L[(ty2 * kClaheTiles + tx2) * 256 + v]));
        s_cell[c][v] = make_uint2(*reinterpret_cast<const uint32_t*>(&top), *reinterpret_cast<const uint32_t*>(&bot));
    }
    for (int r = tid; r < y_end - y_begin && r < kApplyMaxBandRows; r += blockDim.x) {
        const float tyf = __fsub_rn(__fmul_rn((float)(y_begin + r), inv_th), 0.5f);
        const float ya = __fsub_rn(tyf, floorf(tyf));
        s_ya[r] = make_float2(__fsub_rn(1.0f, ya), ya);
    }
    __syncthreads();
    const bool ya_tab = (y_end - y_begin) <= kApplyMaxBandRows;
    const int nchunks = (W + 255) >> 8, nwarps = blockDim.x >> 5;
    const uint32_t cells = (uint32_t)__cvta_generic_to_shared(&s_cell[0][0]);
    // work unit = (256-column chunk, row group): a CTA carries >= 8 warps to amortise its table build; with few chunks
    // per row the warps interleave over row pairs
    const int rgroups = max(1, nwarps / nchunks);
    for (int unit = warp; unit < nchunks * rgroups; unit += nwarps) {
        const int chunk = unit % nchunks, rg = unit / nchunks;
        const int x = 256 * chunk + 8 * lane;
        if (x >= W) continue;
        float xa[8], xa1[8];
        uint32_t cell_addr[8];                                       // shared-memory address of the pixel's cell table
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const float txf = __fsub_rn(__fmul_rn((float)(x + k), inv_tw), 0.5f);
            const int tx1r = (int)floorf(txf);
            xa[k] = __fsub_rn(txf, (float)tx1r);
            xa1[k] = __fsub_rn(1.0f, xa[k]);
            cell_addr[k] = cells + (uint32_t)min(max(tx1r + 1, 0), kApplyCells - 1) * (256u * 8u);
        }
        const uint8_t* img = src + (int64_t)s * src_stride + x;
        uint8_t* out = dst + (int64_t)s * dst_stride + x;
        const bool full = (x + 7 < W);
        auto fetch = [&](int y) -> uint2 {
            if (y >= y_end) return make_uint2(0u, 0u);
            const uint8_t* sp = img + (int64_t)y * src_pitch;
            if (full) return __ldg(reinterpret_cast<const uint2*>(sp));
            uint2 v = make_uint2(0u, 0u);
            for (int k = 0; k < 8 && x + k < W; ++k) {
                if (k < 4) v.x |= (uint32_t)sp[k] << (8 * k);
                else v.y |= (uint32_t)sp[k] << (8 * (k - 4));
            }
            return v;
        };
        auto process = [&](int y, uint2 v8) {
            if (y >= y_end) return;
            float2 yw;
            if (ya_tab) yw = s_ya[y - y_begin];
            else {                                       // bands taller than the table (images above ~4K rows)
                const float tyf = __fsub_rn(__fmul_rn((float)y, inv_th), 0.5f);
                const float ya = __fsub_rn(tyf, floorf(tyf));
                yw = make_float2(__fsub_rn(1.0f, ya), ya);
            }
            uint32_t o[2] = {0u, 0u};
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const uint32_t word = k < 4 ? v8.x : v8.y;
                const uint32_t byte = __byte_perm(word, 0u, 0x4440u | (uint32_t)(k & 3));
                uint2 ent;
                asm("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(ent.x), "=r"(ent.y) : "r"(cell_addr[k] + byte * 8u));
                const float2 pt = __half22float2(*reinterpret_cast<const __half2*>(&ent.x));
                const float2 pb = __half22float2(*reinterpret_cast<const __half2*>(&ent.y));
                const float top = __fadd_rn(__fmul_rn(pt.x, xa1[k]), __fmul_rn(pt.y, xa[k]));
                const float bot = __fadd_rn(__fmul_rn(pb.x, xa1[k]), __fmul_rn(pb.y, xa[k]));
                const float r = __fadd_rn(__fmul_rn(top, yw.x), __fmul_rn(bot, yw.y));
                // 0 <= r <= 255 (convex blend of bytes): the low byte of the magic-rounded float is the result
                const uint32_t rb = __float_as_uint(round_magic(r));
                o[k >> 2] = __byte_perm(o[k >> 2], rb, (k & 3) == 0 ? 0x3214u : (k & 3) == 1 ? 0x3240u : (k & 3) == 2 ? 0x3410u : 0x4210u);
            }
            uint8_t* dp = out + (int64_t)y * dst_pitch;
            if (full) *reinterpret_cast<uint2*>(dp) = make_uint2(o[0], o[1]);
            else
                for (int k = 0; k < 8 && x + k < W; ++k) dp[k] = (uint8_t)(o[k >> 2] >> (8 * (k & 3)));
        };
        // two rows (16 pixels) per trip; the NEXT trip's rows are requested before this trip's table lookups, so the
        // global-load latency is covered by a whole trip of arithmetic instead of being waited out at its start
        int y = y_begin + 2 * rg;
        uint2 a0 = fetch(y), a1 = fetch(y + 1);
        for (; y < y_end; y += 2 * rgroups) {
            const uint2 n0 = fetch(y + 2 * rgroups), n1 = fetch(y + 2 * rgroups + 1);
            process(y, a0);
            process(y + 1, a1);
            a0 = n0;
            a1 = n1;
        }
    }   // chunk
}

// plain strided copy used when equalize == 0 (level 0 = the raw frame)
__global__ void __launch_bounds__(256) copy_image_kernel(const uint8_t* __restrict__ src, int src_pitch,
                                                         int64_t src_stride, uint8_t* __restrict__ dst,
                                                         int dst_pitch, int64_t dst_stride, int W, int H) {
    pdl_entry();
    const int s = blockIdx.y;
    const int wq = (W + 3) >> 2;
    for (int idx = blockIdx.x * 256 + threadIdx.x; idx < wq * H; idx += gridDim.x * 256) {
        const int y = idx / wq, x = (idx - y * wq) << 2;
        const uint8_t* sp = src + (int64_t)s * src_stride + (int64_t)y * src_pitch + x;
        uint8_t* dp = dst + (int64_t)s * dst_stride + (int64_t)y * dst_pitch + x;
        if (x + 3 < W && ((src_pitch & 3) == 0) && ((reinterpret_cast<uintptr_t>(sp) & 3) == 0))
            *reinterpret_cast<uint32_t*>(dp) = *reinterpret_cast<const uint32_t*>(sp);
        else
            for (int k = 0; k < 4 && x + k < W; ++k) dp[k] = sp[k];
    }
}

struct ClaheGeom {
    int tw, th, clip;
    float lut_scale;
};
static ClaheGeom clahe_geom(int W, int H) {
    // cv::CLAHE::apply pads right / bottom by tiles - (size % tiles) as soon as EITHER side is not a multiple of the
    // tile grid, so a side that already is a multiple grows by a full 8 (checked against cv2 4.13: 120 x 90 -> 128 x 96)
    const bool pad = (W % kClaheTiles) != 0 || (H % kClaheTiles) != 0;
    const int We = pad ? W + kClaheTiles - W % kClaheTiles : W;
    const int He = pad ? H + kClaheTiles - H % kClaheTiles : H;
    ClaheGeom g;
    g.tw = We / kClaheTiles;
    g.th = He / kClaheTiles;
    const int area = g.tw * g.th;
    g.clip = (int)(3.0 * area / 256);   // clipLimit * tileSizeTotal / histSize
    if (g.clip < 1) g.clip = 1;
    g.lut_scale = 255.0f / (float)area;
    return g;
}

template <int NT, bool FULLW>
static void launch_lut_nt2(const uint8_t* src, int src_pitch, int64_t src_stride, const ClaheGeom& g, int n_streams,
                           uint8_t* luts, cudaStream_t st) {
    const int smem = NT <= 128 ? 0 : 64 * NT * 4;
    ensure_dyn_smem(reinterpret_cast<const void*>(clahe_lut_kernel<NT, FULLW>), smem);
    launch_pdl(clahe_lut_kernel<NT, FULLW>, dim3(kClaheTiles * kClaheTiles, n_streams), dim3(NT), smem, st,
        src, src_pitch, src_stride, g.tw, g.th, g.clip, g.lut_scale, luts);
}
template <int NT>
static void launch_lut_nt(const uint8_t* src, int src_pitch, int64_t src_stride, const ClaheGeom& g, int n_streams,
                          uint8_t* luts, cudaStream_t st) {
    if ((g.tw & 3) == 0) launch_lut_nt2<NT, true>(src, src_pitch, src_stride, g, n_streams, luts, st);
    else launch_lut_nt2<NT, false>(src, src_pitch, src_stride, g, n_streams, luts, st);
}

void launch_clahe_lut_only(const uint8_t* src, int src_pitch, int64_t src_stride, int W, int H, int n_streams,
                           uint8_t* luts, cudaStream_t st) {
    const ClaheGeom g = clahe_geom(W, H);
    const int area = g.tw * g.th;
    const bool exact_tiles = (W % kClaheTiles == 0) && (H % kClaheTiles == 0) && (src_pitch % 4 == 0);
    // a thread of the private-counter kernel may see up to 4 * ceil(words/NT) pixels; keep that below 256
    const int words = ((g.tw + 6) / 4 + 1) * g.th;
    // fewest private copies that fit the byte counters: the merge costs 64 shared-memory reads per copy
    // (with only a few streams the launch is latency-bound: spread a tile over more threads instead)
    if (exact_tiles && n_streams >= 8 && 4 * ((words + 31) / 32) <= 255) launch_lut_nt<32>(src, src_pitch, src_stride, g, n_streams, luts, st);
    else if (exact_tiles && n_streams < 8 && 4 * ((words + 127) / 128) <= 255) launch_lut_nt<128>(src, src_pitch, src_stride, g, n_streams, luts, st);
    else if (exact_tiles && 4 * ((words + 63) / 64) <= 255) launch_lut_nt<64>(src, src_pitch, src_stride, g, n_streams, luts, st);
    else if (exact_tiles && 4 * ((words + 127) / 128) <= 255) launch_lut_nt<128>(src, src_pitch, src_stride, g, n_streams, luts, st);
    else if (exact_tiles && 4 * ((words + 255) / 256) <= 255 && area < 65536) launch_lut_nt<256>(src, src_pitch, src_stride, g, n_streams, luts, st);
    else
        launch_pdl(clahe_lut_generic_kernel, dim3(kClaheTiles * kClaheTiles, n_streams), dim3(256), 0, st,
            src, src_pitch, src_stride, W, H, g.tw, g.th, g.clip, g.lut_scale, luts);
}

void launch_clahe_apply_only(const uint8_t* src, int src_pitch, int64_t src_stride, uint8_t* dst, int dst_pitch,
                             int64_t dst_stride, int W, int H, int n_streams, const uint8_t* luts,
                             cudaStream_t st) {
    const ClaheGeom g = clahe_geom(W, H);
    const float inv_tw = 1.0f / (float)g.tw, inv_th = 1.0f / (float)g.th;
    // interpolation bands: rows grouped by floor(y * inv_th - 0.5f), evaluated in float32 exactly like the kernel
    ClaheBands bands;
    int b = 0;
    bands.start[0] = 0;
    for (int y = 0; y < H; ++y) {
        volatile float prod = (float)y * inv_th;     // two separately rounded float32 operations, like the kernel
        volatile float tyf = prod - 0.5f;
        const int cy = (int)std::floor(tyf);
        while (b < cy + 1 && b < kClaheTiles) bands.start[++b] = y;
    }
    while (b < kClaheTiles + 1) bands.start[++b] = H;
    const int nchunks = (W + 255) / 256;
    int nwarps = nchunks < 32 ? nchunks : 32;
    if (nwarps < 8) nwarps = (8 / nchunks) * nchunks;          // row groups fill the CTA up to 8 warps
    int row_split = n_streams >= 64 ? 1 : n_streams >= 8 ? 2 : 4;
    {
        static const char* e = getenv("VS_CLAHE_ROW_SPLIT");      // measurement knob
        if (e && atoi(e) > 0) row_split = atoi(e);
    }
    launch_pdl(clahe_apply_kernel, dim3((kClaheTiles + 1) * row_split, n_streams), dim3(32 * nwarps), 0, st,
        src, src_pitch, src_stride, dst, dst_pitch, dst_stride, W, H, inv_tw, inv_th, bands, luts, row_split);
}

void launch_clahe(const uint8_t* src, int src_pitch, int64_t src_stride, uint8_t* dst, int dst_pitch,
                  int64_t dst_stride, int W, int H, int n_streams, uint8_t* luts, cudaStream_t st,
                  int64_t* launches) {
    launch_clahe_lut_only(src, src_pitch, src_stride, W, H, n_streams, luts, st);
    launch_clahe_apply_only(src, src_pitch, src_stride, dst, dst_pitch, dst_stride, W, H, n_streams, luts, st);
    if (launches) *launches += 2;
}

void launch_copy_image(const uint8_t* src, int src_pitch, int64_t src_stride, uint8_t* dst, int dst_pitch,
                       int64_t dst_stride, int W, int H, int n_streams, cudaStream_t st, int64_t* launches) {
    const int wq = (W + 3) >> 2;
    int blocks = (wq * H + 255) / 256;
    if (blocks > 1024) blocks = 1024;
    launch_pdl(copy_image_kernel, dim3(blocks, n_streams), dim3(256), 0, st, src, src_pitch, src_stride, dst, dst_pitch,
                                                               dst_stride, W, H);
    if (launches) *launches += 1;
}

}  // namespace vs
