// Frame ingest — the step BEFORE the hot path, as a pipeline that keeps the GPU fed.
//
// Reference: MeasurementProcessor::extractImageFeatures (/root/reference/src/utility/measurement_processor.cpp:178-231)
// does `cv::imread(full_path, cv::IMREAD_GRAYSCALE)`, hands `rowRange(0, g_config.camera.row)` to detectAndTrack, runs
// the updateID loop and publishes the track_cnt > 1 features; VIOEngine::processFrame (src/vio_engine.cpp:180) wraps a
// raw grey pointer instead of reading a file.  Here:
//   vs_image_decode_gray / vs_image_read_gray   PNG (every colour type / bit depth libpng hands cv::imread, converted to
//                                               8-bit grey the way cv::imread(IMREAD_GRAYSCALE) does) and binary PGM
//   vs_ingest_*                                 decode threads -> ring of pinned host slots -> vs_tracker_submit_batch /
//                                               vs_tracker_wait on one coordinator thread: batch k+2 is being decoded while
//                                               batch k+1 is uploading and batch k is in the kernels
// Host code only (zlib's inflate is the one outside dependency); the tracker calls go through the public C ABI.
#include <zlib.h>

#include <atomic>
#include <condition_variable>
#include <cstdio>
#include <cstring>
#include <deque>
#include <exception>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "tracker_state.h"

namespace vs {
void set_error(const char* fmt, ...);

namespace {

inline uint32_t be32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

// libpng's png_set_rgb_to_gray(png, 1, 0.299, 0.587) as cv::imread(IMREAD_GRAYSCALE) sets it up: 15-bit fixed-point
// weights red = 29900 * 32768 / 100000, green = 58700 * 32768 / 100000 (truncated), blue = the rest; pixels with
// r == g == b pass through unchanged; 8-bit samples are truncated ("the historical approach"), 16-bit ones rounded
// (pngrtran.c, png_do_rgb_to_gray) — both checked against cv2.imread in tests/test_ingest.py
constexpr uint32_t kRC = 9797, kGC = 19234, kBC = 32768 - 9797 - 19234;
inline uint32_t rgb_to_gray8(uint32_t r, uint32_t g, uint32_t b) {
    if (r == g && r == b) return r;
    return (kRC * r + kGC * g + kBC * b) >> 15;
}
inline uint32_t rgb_to_gray16(uint32_t r, uint32_t g, uint32_t b) {
    if (r == g && r == b) return r;
    return (kRC * r + kGC * g + kBC * b + 16384) >> 15;
}

inline int paeth(int a, int b, int c) {
    const int p = a + b - c, pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

// PNG -> 8-bit grey.  Writes min(height, max_h) rows (the caller's rowRange(0, row)); width must be <= max_w.
int decode_png(const uint8_t* d, int64_t n, uint8_t* out, int out_stride, int max_w, int max_h, int* w_out, int* h_out) {
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    if (n < 8 + 25 || memcmp(d, sig, 8) != 0) { set_error("not a PNG stream"); return VS_ERR_INVALID; }
    int64_t pos = 8;
    uint32_t W = 0, H = 0;
    int depth = 0, ctype = -1, interlace = 0;
    uint8_t pal[256][3];
    memset(pal, 0, sizeof(pal));
    std::vector<uint8_t> idat;
    bool end = false;
    while (!end && pos + 12 <= n) {
        const uint32_t len = be32(d + pos);
        const uint8_t* type = d + pos + 4;
        const uint8_t* body = d + pos + 8;
        if (pos + 12 + (int64_t)len > n) { set_error("truncated PNG chunk"); return VS_ERR_INVALID; }
        if (!memcmp(type, "IHDR", 4)) {
            if (len < 13) { set_error("bad IHDR"); return VS_ERR_INVALID; }
            W = be32(body);
            H = be32(body + 4);
            depth = body[8];
            ctype = body[9];
            interlace = body[12];
        } else if (!memcmp(type, "PLTE", 4)) {
            for (uint32_t i = 0; i < len / 3 && i < 256; ++i) { pal[i][0] = body[3 * i]; pal[i][1] = body[3 * i + 1]; pal[i][2] = body[3 * i + 2]; }
        } else if (!memcmp(type, "IDAT", 4)) {
            idat.insert(idat.end(), body, body + len);
        } else if (!memcmp(type, "IEND", 4)) {
            end = true;
        }
        pos += 12 + (int64_t)len;
    }
    int channels = 0;
    switch (ctype) {
        case 0: channels = 1; break;
        case 2: channels = 3; break;
        case 3: channels = 1; break;
        case 4: channels = 2; break;
        case 6: channels = 4; break;
        default: set_error("PNG: unsupported colour type %d", ctype); return VS_ERR_INVALID;
    }
    const bool depth_ok = (ctype == 0 && (depth == 1 || depth == 2 || depth == 4 || depth == 8 || depth == 16)) ||
                          (ctype == 3 && (depth == 1 || depth == 2 || depth == 4 || depth == 8)) ||
                          ((ctype == 2 || ctype == 4 || ctype == 6) && (depth == 8 || depth == 16));
    if (!depth_ok || W == 0 || H == 0 || W > 0x7fffffffu || H > 0x7fffffffu) { set_error("PNG: bad header (%ux%u, depth %d, type %d)", W, H, depth, ctype); return VS_ERR_INVALID; }
    if (interlace) { set_error("PNG: Adam7 interlacing is not supported"); return VS_ERR_INVALID; }
    if (W > (uint32_t)max_w) { set_error("PNG is %u wide, the tracker expects at most %d", W, max_w); return VS_ERR_INVALID; }
    *w_out = (int)W;
    *h_out = (int)H;
    const int rows = H < (uint32_t)max_h ? (int)H : max_h;
    if (idat.size() > 0xffffffffull) { set_error("PNG: more than 4 GiB of image data"); return VS_ERR_INVALID; }
    const size_t rowbytes = ((size_t)W * channels * depth + 7) / 8;
    const int bpp = channels * depth >= 8 ? channels * depth / 8 : 1;
    std::vector<uint8_t> raw((rowbytes + 1) * (size_t)rows);
    {
        z_stream zs;
        memset(&zs, 0, sizeof(zs));
        if (inflateInit(&zs) != Z_OK) { set_error("zlib inflateInit failed"); return VS_ERR_NOMEM; }
        zs.next_in = idat.data();
        zs.avail_in = (uInt)idat.size();
        zs.next_out = raw.data();
        zs.avail_out = (uInt)raw.size();
        const int zr = inflate(&zs, Z_FINISH);            // Z_BUF_ERROR = output full: the rows asked for are there
        const size_t got = raw.size() - zs.avail_out;
        inflateEnd(&zs);
        if ((zr != Z_STREAM_END && zr != Z_BUF_ERROR && zr != Z_OK) || got < raw.size()) {
            set_error("PNG: corrupt or short image data (zlib %d, %zu of %zu bytes)", zr, got, raw.size());
            return VS_ERR_INVALID;
        }
    }
    std::vector<uint8_t> prev(rowbytes, 0), cur(rowbytes);
    for (int y = 0; y < rows; ++y) {
        const uint8_t* src = raw.data() + (rowbytes + 1) * (size_t)y;
        const int ft = src[0];
        ++src;
        for (size_t i = 0; i < rowbytes; ++i) {
            const int a = i >= (size_t)bpp ? cur[i - bpp] : 0, b = prev[i], c = i >= (size_t)bpp ? prev[i - bpp] : 0;
            int v = src[i];
            switch (ft) {
                case 0: break;
                case 1: v += a; break;
                case 2: v += b; break;
                case 3: v += (a + b) >> 1; break;
                case 4: v += paeth(a, b, c); break;
                default: set_error("PNG: bad filter type %d", ft); return VS_ERR_INVALID;
            }
            cur[i] = (uint8_t)v;
        }
        uint8_t* o = out + (size_t)y * out_stride;
        if (ctype == 0 && depth == 8) {
            memcpy(o, cur.data(), W);
        } else if (ctype == 0 && depth == 16) {
            for (uint32_t x = 0; x < W; ++x) o[x] = cur[2 * x];                   // png_set_strip_16: the high byte
        } else if (ctype == 0) {                                                   // 1/2/4-bit grey, expanded to 0..255
            const int maxv = (1 << depth) - 1, per = 8 / depth;
            for (uint32_t x = 0; x < W; ++x) {
                const int v = (cur[x / per] >> (8 - depth * (1 + (int)(x % per)))) & maxv;
                o[x] = (uint8_t)(v * 255 / maxv);
            }
        } else if (ctype == 3) {
            const int maxv = (1 << depth) - 1, per = 8 / depth;
            for (uint32_t x = 0; x < W; ++x) {
                const int idx = depth == 8 ? cur[x] : ((cur[x / per] >> (8 - depth * (1 + (int)(x % per)))) & maxv);
                o[x] = (uint8_t)rgb_to_gray8(pal[idx][0], pal[idx][1], pal[idx][2]);
            }
        } else if (ctype == 4) {
            const int step = depth == 8 ? 2 : 4;
            for (uint32_t x = 0; x < W; ++x) o[x] = cur[(size_t)step * x];        // alpha dropped, 16-bit -> high byte
        } else {                                                                   // RGB / RGBA, alpha dropped
            if (depth == 8) {
                for (uint32_t x = 0; x < W; ++x) {
                    const uint8_t* p = &cur[(size_t)channels * x];
                    o[x] = (uint8_t)rgb_to_gray8(p[0], p[1], p[2]);
                }
            } else {                                                               // grey in 16 bits, then the high byte
                for (uint32_t x = 0; x < W; ++x) {
                    const uint8_t* p = &cur[(size_t)channels * 2 * x];
                    const uint32_t r = (p[0] << 8) | p[1], g = (p[2] << 8) | p[3], b = (p[4] << 8) | p[5];
                    o[x] = (uint8_t)(rgb_to_gray16(r, g, b) >> 8);
                }
            }
        }
        prev.swap(cur);
    }
    return VS_OK;
}

// binary PGM (P5), maxval <= 255
int decode_pgm(const uint8_t* d, int64_t n, uint8_t* out, int out_stride, int max_w, int max_h, int* w_out, int* h_out) {
    int64_t pos = 2;
    int vals[3], k = 0;
    while (k < 3 && pos < n) {
        if (d[pos] == '#') { while (pos < n && d[pos] != '\n') ++pos; continue; }
        if (d[pos] == ' ' || d[pos] == '\t' || d[pos] == '\n' || d[pos] == '\r') { ++pos; continue; }
        int v = 0;
        bool any = false;
        while (pos < n && d[pos] >= '0' && d[pos] <= '9' && v < 100000000) { v = v * 10 + (d[pos] - '0'); ++pos; any = true; }
        if (!any || v >= 100000000) { set_error("PGM: bad header"); return VS_ERR_INVALID; }
        vals[k++] = v;
    }
    if (k < 3 || pos >= n) { set_error("PGM: truncated header"); return VS_ERR_INVALID; }
    ++pos;                                                  // the single whitespace byte after maxval
    const int W = vals[0], H = vals[1];
    if (vals[2] > 255 || W <= 0 || H <= 0) { set_error("PGM: only 8-bit images are supported"); return VS_ERR_INVALID; }
    if (W > max_w) { set_error("PGM is %d wide, the tracker expects at most %d", W, max_w); return VS_ERR_INVALID; }
    const int rows = H < max_h ? H : max_h;
    if (pos + (int64_t)W * rows > n) { set_error("PGM: short pixel data"); return VS_ERR_INVALID; }
    for (int y = 0; y < rows; ++y) memcpy(out + (size_t)y * out_stride, d + pos + (int64_t)y * W, W);
    *w_out = W;
    *h_out = H;
    return VS_OK;
}

int read_file(const char* path, std::vector<uint8_t>* buf) {
    FILE* f = fopen(path, "rb");
    if (!f) { set_error("Cannot load image: %s", path); return VS_ERR_INVALID; }     // measurement_processor.cpp:186-189
    fseek(f, 0, SEEK_END);
    const long sz = ftell(f);
    fseek(f, 0, SEEK_SET);
    if (sz <= 0) { fclose(f); set_error("Cannot load image: %s (empty)", path); return VS_ERR_INVALID; }
    try {
        buf->resize((size_t)sz);
    } catch (const std::exception&) {
        fclose(f);
        set_error("Cannot load image: %s (%ld bytes do not fit in memory)", path, sz);
        return VS_ERR_NOMEM;
    }
    const size_t got = fread(buf->data(), 1, (size_t)sz, f);
    fclose(f);
    if (got != (size_t)sz) { set_error("Cannot load image: %s (short read)", path); return VS_ERR_INVALID; }
    return VS_OK;
}

}  // namespace
}  // namespace vs

using namespace vs;

// =================================================================================================
// ingest pipeline
// =================================================================================================
struct vs_ingest {
    static constexpr int kSlots = 3;     // one batch decoding + two in flight in the tracker
    vs_tracker* t = nullptr;
    int B = 0, W = 0, H = 0, cap = 0;
    size_t img_bytes = 0;
    uint8_t* ring = nullptr;             // pinned: kSlots x B x H x W
    struct Job {
        int slot = 0;
        std::vector<std::string> paths;             // empty = raw frames were already copied into the slot
        std::vector<double> times;
        std::atomic<int> pending{0};                // streams still being decoded
        int rc = VS_OK;
        std::string err;
    };
    struct Done {
        int rc = VS_OK;
        std::string err;
        std::vector<vs_track_result> res;           // pointers into `store`
        std::vector<uint8_t> store;
    };
    std::mutex mu;
    std::condition_variable cv_work, cv_done, cv_dec;
    std::deque<Job*> decoding;           // submitted by the caller, in order
    std::deque<std::pair<Job*, int>> tasks;     // (job, stream) decode tasks
    std::deque<Done*> done;
    Done* last_returned = nullptr;
    int64_t submitted = 0, returned = 0;
    bool stop = false;
    std::vector<std::thread> decoders;
    std::thread coordinator;
};

namespace {

void decode_worker(vs_ingest* g) {
    std::vector<uint8_t> file;
    for (;;) {
        std::pair<vs_ingest::Job*, int> task;
        {
            std::unique_lock<std::mutex> lk(g->mu);
            g->cv_dec.wait(lk, [&] { return g->stop || !g->tasks.empty(); });
            if (g->stop && g->tasks.empty()) return;
            task = g->tasks.front();
            g->tasks.pop_front();
        }
        vs_ingest::Job* j = task.first;
        const int s = task.second;
        uint8_t* dst = g->ring + ((size_t)j->slot * g->B + s) * g->img_bytes;
        int w = 0, h = 0;
        int rc = read_file(j->paths[s].c_str(), &file);
        if (rc == VS_OK) rc = vs_image_decode_gray(file.data(), (int64_t)file.size(), dst, g->W, g->W, g->H, &w, &h);
        if (rc == VS_OK && (w != g->W || h < g->H)) {
            set_error("%s is %dx%d, the tracker needs %d columns and at least %d rows", j->paths[s].c_str(), w, h, g->W, g->H);
            rc = VS_ERR_INVALID;
        }
        {
            std::lock_guard<std::mutex> lk(g->mu);
            if (rc != VS_OK && j->rc == VS_OK) { j->rc = rc; j->err = vs_last_error(); }
            if (--j->pending == 0) g->cv_work.notify_all();
        }
    }
}

// the only thread that touches the tracker handle: submits decoded batches in order (at most two in flight) and waits
// for the oldest one whenever nothing can be submitted
void coordinator_loop(vs_ingest* g) {
    std::deque<vs_ingest::Job*> flying;
    std::vector<const uint8_t*> ptrs(g->B);
    std::vector<vs_track_result> res(g->B);
    for (;;) {
        vs_ingest::Job* ready = nullptr;
        bool stopping = false;
        {
            std::unique_lock<std::mutex> lk(g->mu);
            g->cv_work.wait(lk, [&] {
                return g->stop || !flying.empty() || (!g->decoding.empty() && g->decoding.front()->pending == 0);
            });
            stopping = g->stop;
            if (stopping) {                          // shutting down: undecoded work is dropped, batches in flight are retired
                for (auto* j : g->decoding) delete j;
                g->decoding.clear();
            } else if (!g->decoding.empty() && g->decoding.front()->pending == 0 && (int)flying.size() < 2) {
                ready = g->decoding.front();
                g->decoding.pop_front();
            }
        }
        if (ready) {
            if (ready->rc == VS_OK) {
                for (int s = 0; s < g->B; ++s) ptrs[s] = g->ring + ((size_t)ready->slot * g->B + s) * g->img_bytes;
                ready->rc = vs_tracker_submit_batch(g->t, ptrs.data(), g->W, ready->times.data());
                if (ready->rc != VS_OK) ready->err = vs_last_error();
            }
            flying.push_back(ready);                 // a failed batch rides along, so results keep the submission order
            continue;
        }
        if (flying.empty()) {
            if (stopping) return;
            continue;
        }
        // nothing to submit: retire the oldest batch
        vs_ingest::Job* j = flying.front();
        flying.pop_front();
        auto* d = new vs_ingest::Done();
        if (j->rc != VS_OK) {
            d->rc = j->rc;
            d->err = j->err;
        } else {
            d->rc = vs_tracker_wait(g->t, res.data());
            if (d->rc != VS_OK) d->err = vs_last_error();
            else {
                // the tracker's result block is reused two waits later: hand the caller a copy it owns until its next wait
                size_t bytes = 0;
                for (int s = 0; s < g->B; ++s) bytes += (size_t)res[s].n * 32;
                d->store.resize(bytes + 64);
                d->res = res;
                uint8_t* p = d->store.data();
                for (int s = 0; s < g->B; ++s) {
                    const size_t n = (size_t)res[s].n;
                    vs_track_result& r = d->res[s];
                    memcpy(p, res[s].ids, n * 4); r.ids = reinterpret_cast<const int32_t*>(p); p += n * 4;
                    memcpy(p, res[s].track_cnt, n * 4); r.track_cnt = reinterpret_cast<const int32_t*>(p); p += n * 4;
                    memcpy(p, res[s].cur_pts, n * 8); r.cur_pts = reinterpret_cast<const float*>(p); p += n * 8;
                    memcpy(p, res[s].undist, n * 8); r.undist = reinterpret_cast<const float*>(p); p += n * 8;
                    memcpy(p, res[s].velocity, n * 8); r.velocity = reinterpret_cast<const float*>(p); p += n * 8;
                }
            }
        }
        delete j;
        std::lock_guard<std::mutex> lk(g->mu);
        g->done.push_back(d);
        g->cv_done.notify_all();
    }
}

int ingest_enqueue(vs_ingest* g, vs_ingest::Job* j) {
    std::lock_guard<std::mutex> lk(g->mu);
    g->decoding.push_back(j);
    if (!j->paths.empty()) {
        for (int s = 0; s < g->B; ++s) g->tasks.emplace_back(j, s);
        g->cv_dec.notify_all();
    }
    ++g->submitted;
    g->cv_work.notify_all();
    return VS_OK;
}

}  // namespace

extern "C" {

int vs_image_decode_gray(const uint8_t* bytes, int64_t n, uint8_t* out, int32_t out_stride, int32_t max_w, int32_t max_h,
                         int32_t* w, int32_t* h) {
    if (!bytes || !out || !w || !h || n < 4 || out_stride < max_w || max_w <= 0 || max_h <= 0) { set_error("vs_image_decode_gray: bad argument"); return VS_ERR_INVALID; }
    try {                                       // nothing may unwind through the C boundary (std::vector growth)
        if (bytes[0] == 0x89 && bytes[1] == 'P') return decode_png(bytes, n, out, out_stride, max_w, max_h, w, h);
        if (bytes[0] == 'P' && bytes[1] == '5') return decode_pgm(bytes, n, out, out_stride, max_w, max_h, w, h);
    } catch (const std::exception& e) {
        set_error("image decode: %s", e.what());
        return VS_ERR_NOMEM;
    }
    set_error("unsupported image format (PNG and binary PGM are)");
    return VS_ERR_INVALID;
}

int vs_image_read_gray(const char* path, uint8_t* out, int32_t out_stride, int32_t max_w, int32_t max_h, int32_t* w, int32_t* h) {
    if (!path) { set_error("vs_image_read_gray: null path"); return VS_ERR_INVALID; }
    std::vector<uint8_t> buf;
    int rc;
    try {
        rc = read_file(path, &buf);
    } catch (const std::exception& e) {
        set_error("Cannot load image: %s (%s)", path, e.what());
        return VS_ERR_NOMEM;
    }
    if (rc != VS_OK) return rc;
    return vs_image_decode_gray(buf.data(), (int64_t)buf.size(), out, out_stride, max_w, max_h, w, h);
}

int vs_ingest_create(vs_tracker* t, int32_t decode_threads, vs_ingest** out) {
    if (!t || !out) { set_error("vs_ingest_create: null argument"); return VS_ERR_INVALID; }
    *out = nullptr;
    if (t->cfg.ransac_mode == VS_RANSAC_CALLBACK) { set_error("the ingest pipeline uses submit/wait: not available with VS_RANSAC_CALLBACK"); return VS_ERR_STATE; }
    vs_ingest* g = new (std::nothrow) vs_ingest();
    if (!g) { set_error("out of host memory"); return VS_ERR_NOMEM; }
    g->t = t;
    g->B = t->B;
    g->W = t->W;
    g->H = t->H;
    g->cap = t->cap;
    g->img_bytes = (size_t)t->W * t->H;
    VS_CUDA_TRY(cudaSetDevice(t->cfg.device));
    if (cudaMallocHost(&g->ring, (size_t)vs_ingest::kSlots * g->B * g->img_bytes) != cudaSuccess) {
        delete g;
        set_error("cannot pin %zu bytes for the ingest ring", (size_t)vs_ingest::kSlots * t->B * (size_t)t->W * t->H);
        return VS_ERR_NOMEM;
    }
    const int nt = decode_threads > 0 ? decode_threads : 4;
    for (int i = 0; i < nt; ++i) g->decoders.emplace_back(decode_worker, g);
    g->coordinator = std::thread(coordinator_loop, g);
    *out = g;
    return VS_OK;
}

static int ingest_slot_check(vs_ingest* g) {
    // a ring slot is free again once the batch that used it (kSlots submissions ago) has been handed back
    std::unique_lock<std::mutex> lk(g->mu);
    if (g->submitted - g->returned >= vs_ingest::kSlots) {
        set_error("vs_ingest: %d batches outstanding, call vs_ingest_wait first", vs_ingest::kSlots);
        return VS_ERR_STATE;
    }
    return VS_OK;
}

int vs_ingest_submit_files(vs_ingest* g, const char* const* paths, const double* time_s) {
    if (!g || !paths || !time_s) { set_error("vs_ingest_submit_files: null argument"); return VS_ERR_INVALID; }
    const int rc = ingest_slot_check(g);
    if (rc != VS_OK) return rc;
    auto* j = new vs_ingest::Job();
    j->slot = (int)(g->submitted % vs_ingest::kSlots);
    j->paths.resize(g->B);
    for (int s = 0; s < g->B; ++s) {
        if (!paths[s]) { delete j; set_error("paths[%d] is null", s); return VS_ERR_INVALID; }
        j->paths[s] = paths[s];
    }
    j->times.assign(time_s, time_s + g->B);
    j->pending = g->B;
    return ingest_enqueue(g, j);
}

int vs_ingest_submit_raw(vs_ingest* g, const uint8_t* const* frames, int32_t stride, const double* time_s) {
    if (!g || !frames || !time_s || stride < g->W) { set_error("vs_ingest_submit_raw: bad argument"); return VS_ERR_INVALID; }
    const int rc = ingest_slot_check(g);
    if (rc != VS_OK) return rc;
    auto* j = new vs_ingest::Job();
    j->slot = (int)(g->submitted % vs_ingest::kSlots);
    for (int s = 0; s < g->B; ++s) {                     // the caller's buffer is only valid during this call (vio_engine.cpp:180)
        if (!frames[s]) { delete j; set_error("frames[%d] is null", s); return VS_ERR_INVALID; }
        uint8_t* dst = g->ring + ((size_t)j->slot * g->B + s) * g->img_bytes;
        for (int y = 0; y < g->H; ++y) memcpy(dst + (size_t)y * g->W, frames[s] + (size_t)y * stride, g->W);
    }
    j->times.assign(time_s, time_s + g->B);
    j->pending = 0;
    return ingest_enqueue(g, j);
}

int vs_ingest_wait(vs_ingest* g, vs_track_result* out) {
    if (!g || !out) { set_error("vs_ingest_wait: null argument"); return VS_ERR_INVALID; }
    vs_ingest::Done* d = nullptr;
    {
        std::unique_lock<std::mutex> lk(g->mu);
        if (g->submitted == g->returned) { set_error("vs_ingest_wait: nothing submitted"); return VS_ERR_STATE; }
        g->cv_done.wait(lk, [&] { return !g->done.empty(); });
        d = g->done.front();
        g->done.pop_front();
        ++g->returned;
    }
    delete g->last_returned;
    g->last_returned = d;
    if (d->rc != VS_OK) { set_error("%s", d->err.c_str()); return d->rc; }
    for (int s = 0; s < g->B; ++s) out[s] = d->res[s];
    return VS_OK;
}

void vs_ingest_destroy(vs_ingest* g) {
    if (!g) return;
    {
        std::lock_guard<std::mutex> lk(g->mu);
        g->stop = true;
        g->tasks.clear();
        for (auto* j : g->decoding) j->pending = 0;
    }
    g->cv_dec.notify_all();
    g->cv_work.notify_all();
    for (auto& th : g->decoders) th.join();
    if (g->coordinator.joinable()) g->coordinator.join();
    for (auto* d : g->done) delete d;
    delete g->last_returned;
    if (g->ring) cudaFreeHost(g->ring);
    delete g;
}

}  // extern "C"
