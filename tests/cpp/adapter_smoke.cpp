// Exercises the C++ host layer (include/vs_feature_tracker.hpp) the way the reference's caller does
// (src/utility/measurement_processor.cpp:191-231): detectAndTrack, the updateID loop, then read the public vectors.
// usage: adapter_smoke <frames.bin> <W> <H> <n_frames> <calib.yaml|-> [n_gpus n_streams]
// frames.bin = n_frames tightly packed W x H u8 images.  Prints one line per frame: "k n id0 x0 y0 ... " (first 4 features)
// and a final "published <count>".  Exit code 3 = the library refused to run without CUDA (expected on a CPU-only host).
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "vs_feature_tracker.hpp"

int main(int argc, char** argv) {
    if (argc < 6) { std::fprintf(stderr, "usage\n"); return 2; }
    const int W = std::atoi(argv[2]), H = std::atoi(argv[3]), N = std::atoi(argv[4]);
    std::vector<uint8_t> buf((size_t)W * H * N);
    FILE* f = std::fopen(argv[1], "rb");
    if (!f || std::fread(buf.data(), 1, buf.size(), f) != buf.size()) { std::fprintf(stderr, "cannot read frames\n"); return 2; }
    std::fclose(f);
    try {
        if (argc >= 8) {     // stream-sharded multi-GPU host layer: every stream gets the same frames
            const int G = std::atoi(argv[6]), S = std::atoi(argv[7]);
            vs_tracker_config tc;
            vs_tracker_default_config(&tc);
            tc.width = W; tc.height = H; tc.max_cnt = 60; tc.min_dist = 20; tc.ransac_mode = VS_RANSAC_OFF;
            tc.camera.width = W; tc.camera.height = H; tc.camera.fx = tc.camera.fy = 300.0; tc.camera.cx = W / 2.0; tc.camera.cy = H / 2.0;
            vs::ShardedTracker st(tc, S, G, 2);
            // persistent per-GPU workers, two batches in flight: submit frame k+1 before waiting for frame k
            std::vector<std::vector<const uint8_t*>> fr(N, std::vector<const uint8_t*>(S));
            std::vector<std::vector<double>> ts(N, std::vector<double>(S));
            for (int k = 0; k < N; ++k)
                for (int s = 0; s < S; ++s) { fr[k][s] = buf.data() + (size_t)k * W * H; ts[k][s] = 0.05 * k; }
            st.submit(fr[0].data(), W, ts[0].data());
            for (int k = 0; k < N; ++k) {
                if (k + 1 < N) st.submit(fr[k + 1].data(), W, ts[k + 1].data());
                const auto& r = st.wait();
                for (int s = 1; s < S; ++s)
                    if (r[s].n != r[0].n || std::memcmp(r[s].cur_pts, r[0].cur_pts, sizeof(float) * 2 * r[0].n)) { std::printf("stream %d differs\n", s); return 1; }
                std::printf("%d %d\n", k, r[0].n);
            }
            // and the synchronous form on top of the same workers
            const auto& r2 = st.track(fr[N - 1].data(), W, ts[N - 1].data());
            if (r2[0].n <= 0) { std::printf("track() returned nothing\n"); return 1; }
            std::printf("sharded ok %d gpus %d streams\n", st.n_gpus(), st.n_streams());
            return 0;
        }
        frontend::FeatureTracker ft;
        ft.config.row = H; ft.config.col = W; ft.config.max_cnt = 60; ft.config.min_dist = 20;
        ft.config.ransac_mode = VS_RANSAC_OFF;
        if (argv[5][0] == '-') ft.setIntrinsicParameter(VS_CAM_PINHOLE, W, H, 300.0, 300.0, W / 2.0, H / 2.0, -0.1, 0.01, 1e-4, -1e-4);
        else ft.readIntrinsicParameter(argv[5]);
        size_t published = 0;
        for (int k = 0; k < N; ++k) {
            frontend::FeatureTracker::Image img(buf.data() + (size_t)k * W * H, W, H);
            ft.detectAndTrack(img, 0.05 * k);
            for (unsigned i = 0;; ++i) if (!ft.updateID(i)) break;
            std::printf("%d %zu", k, ft.ids.size());
            for (size_t j = 0; j < ft.ids.size() && j < 4; ++j)
                std::printf(" %d %.6f %.6f %.9f %.9f %d", ft.ids[j], ft.cur_pts[j].x, ft.cur_pts[j].y, ft.cur_undistorted_pts[j].x,
                            ft.cur_undistorted_pts[j].y, ft.track_cnt[j]);
            std::printf("\n");
            published = ft.imageData().size();
        }
        std::printf("published %zu n_id %d\n", published, frontend::FeatureTracker::n_id);
    } catch (const vs::Error& e) {
        std::fprintf(stderr, "%s\n", e.what());
        return 3;
    }
    return 0;
}
