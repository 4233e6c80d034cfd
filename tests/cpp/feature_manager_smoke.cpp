// Drives frontend::FeatureManager (include/vs_feature_manager.hpp) through a deterministic estimator-style loop and prints
// every decision and the final bank; tests/test_cpp_feature_manager.py replays the same loop on the FeatureManager oracle.
// usage: feature_manager_smoke <seed> <n_frames> <step>     exit code 3 = no CUDA device (there is no CPU fallback)
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "vs_feature_manager.hpp"

static unsigned int g_state;
static unsigned int rnd() {
    g_state = (g_state * 1103515245u + 12345u) & 0x7fffffffu;
    return g_state;
}

int main(int argc, char** argv) {
    if (argc < 4) return 2;
    g_state = (unsigned int)atoi(argv[1]);
    const int n_frames = atoi(argv[2]);
    const double step = atof(argv[3]);
    const int N = 120, WIN = VS_BANK_WINDOW_SIZE;
    double px[N], py[N], pd[N];
    int gen[N];
    for (int i = 0; i < N; ++i) {
        px[i] = (double)(rnd() % 4096) / 4096.0 - 0.5;
        py[i] = (double)(rnd() % 4096) / 8192.0 - 0.25;
        pd[i] = 3.0 + (double)(rnd() % 1024) / 128.0;
        gen[i] = 0;
    }
    try {
        frontend::FeatureManager fm(1024);
        vs::Poses window;
        for (auto& w : window) { w.first = {1, 0, 0, 0, 1, 0, 0, 0, 1}; w.second = {0, 0, 0}; }
        const vs::Vec3 t_ic = {0, 0, 0};
        const vs::Mat3 r_ic = {1, 0, 0, 0, 1, 0, 0, 0, 1};
        int frame_count = 0;
        bool nonlinear = false;
        for (int k = 0; k < n_frames; ++k) {
            const double cam_x = step * k;
            vs::ImageData image;
            for (int i = 0; i < N; ++i) {
                if (rnd() % 37 == 0) ++gen[i];                       // the track breaks: the point comes back under a new id
                const double x = px[i] - cam_x / pd[i], y = py[i];
                if (x < -0.6 || x > 0.6) continue;
                image[i + 1000 * gen[i]] = {x, y, 1.0, x * 460.0 + 376.0, y * 460.0 + 240.0, 0.0, 0.0};
            }
            window[frame_count].second = {cam_x, 0, 0};
            const bool key = fm.addFeatureAndCheckParallax(frame_count, image);
            std::printf("frame %d fc %d key %d tracked %d bank %d count %d psum %.17g pnum %d\n", k, frame_count, key ? 1 : 0,
                        fm.last_track_num_, (int)fm.feature_bank_.size(), fm.getFeatureCount(), fm.last_parallax_sum,
                        (int)fm.last_parallax_num);
            if (frame_count < WIN) { ++frame_count; continue; }
            if (nonlinear) fm.triangulateAcrossAllViews(window, t_ic, r_ic);
            if (key) {
                const vs::Mat3 R0 = window[0].first;
                const vs::Vec3 P0 = window[0].second;
                for (int i = 0; i < WIN; ++i) window[i] = window[i + 1];
                if (nonlinear) fm.removeBackShiftDepth(R0, P0, window[0].first, window[0].second);
                else fm.removeBack();
            } else {
                window[WIN - 1] = window[WIN];
                fm.removeFront(frame_count);
            }
            nonlinear = true;
        }
        for (const auto& f : fm.feature_bank_)
            std::printf("feature %d start %d nobs %d flag %d depth %.17g x0 %.17g\n", f.feature_id, f.start_frame,
                        (int)f.feature_per_frame.size(), f.solve_flag, f.estimated_depth, f.feature_per_frame[0].ray_vector[0]);
        const auto c = fm.getCorresponding(WIN - 2, WIN);
        std::printf("corresponding %d\n", (int)c.size());
        const auto d = fm.getDepthVector();
        std::printf("depth_vector %d\n", (int)d.size());
    } catch (const vs::BankError& e) {
        std::fprintf(stderr, "%s\n", e.what());
        return (e.code == VS_ERR_CUDA) ? 3 : 1;
    }
    return 0;
}
