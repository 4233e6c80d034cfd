// include/vs_feature_tracker.hpp in the configuration INTEGRATION.md §2 tells a maintainer to build:
//   -DVS_WITH_OPENCV -DVS_WITH_REFERENCE_CONFIG
// OpenCV's SDK is not installed, so <opencv2/opencv.hpp> resolves to the stand-in of oracle/ref_shim/ (same public names
// and signatures for the slice used; its compute entry points are callbacks, unset here) and "utility/config.h" to the
// reference's own header when /root/reference exists, else to tests/cpp/stubs/.  The program drives the adapter the way
// VIOEngine does: cv::Mat frames wrapping raw pointers, g_config mutated between frames (setMobileParams /
// setTrackingParams / equalize), updateID loop, imageData().
// usage: adapter_opencv_probe <frames.bin> <W> <H> <n_frames>     exit 3 = no CUDA device (expected on a CPU-only host)
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "vs_feature_tracker.hpp"

namespace utility { Config g_config; }
extern "C" {
vs_ref_callbacks vs_ref_cb = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
void (*vs_ref_eigvals_cb)(int, const double*, double*, double*) = nullptr;
}

int main(int argc, char** argv) {
    if (argc < 5) { std::fprintf(stderr, "usage\n"); return 2; }
    const int W = std::atoi(argv[2]), H = std::atoi(argv[3]), N = std::atoi(argv[4]);
    std::vector<uint8_t> buf((size_t)W * H * N);
    FILE* f = std::fopen(argv[1], "rb");
    if (!f || std::fread(buf.data(), 1, buf.size(), f) != buf.size()) { std::fprintf(stderr, "cannot read frames\n"); return 2; }
    std::fclose(f);
    auto& g = utility::g_config;
    g.camera.row = H; g.camera.col = W; g.camera.focal_length = 460.0;
    g.feature_tracker.max_cnt = 60; g.feature_tracker.min_dist = 20;
    try {
        frontend::FeatureTracker ft;
        ft.config.ransac_mode = VS_RANSAC_OFF;     // the callback mode would call cv::findFundamentalMat, which is a stand-in here
        ft.setIntrinsicParameter(VS_CAM_PINHOLE, W, H, 300.0, 300.0, W / 2.0, H / 2.0, -0.1, 0.01, 1e-4, -1e-4);
        for (int k = 0; k < N; ++k) {
            if (k == 2) g.feature_tracker.max_cnt = 80;                       // VIOEngine::setMobileParams
            if (k == 3) { g.feature_tracker.lk_window_size = 15; g.feature_tracker.lk_iterations = 20; g.feature_tracker.lk_eps = 0.03; }
            if (k == 4) g.feature_tracker.equalize = 0;                       // re-read on every call (feature_tracker.cpp:102)
            cv::Mat img(H, W, CV_8UC1, buf.data() + (size_t)k * W * H);      // vio_engine.cpp:180
            const auto data = vs::processFrameFrontend(ft, img.data, W, H, 0.05 * k);
            std::printf("%d %zu %zu", k, ft.ids.size(), data.size());
            for (size_t j = 0; j < ft.ids.size() && j < 4; ++j)
                std::printf(" %d %.6f %.6f %d", ft.ids[j], ft.cur_pts[j].x, ft.cur_pts[j].y, ft.track_cnt[j]);
            std::printf("\n");
        }
        // a frame of another size must be refused, not read out of bounds
        try {
            cv::Mat small(H / 2, W, CV_8UC1, buf.data());
            ft.detectAndTrack(small, 1.0);
            std::printf("size change accepted\n");
            return 1;
        } catch (const std::runtime_error& e) {
            std::printf("size change refused\n");
        }
    } catch (const vs::Error& e) {
        std::fprintf(stderr, "%s\n", e.what());
        return 3;
    }
    return 0;
}
