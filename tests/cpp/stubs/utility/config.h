// TEST STUB — the fields of the reference's utility::g_config that the tracker reads
// (/root/reference/include/utility/config.h:18-56, 118), so that include/vs_feature_tracker.hpp can be compiled with
// -DVS_WITH_REFERENCE_CONFIG where the reference tree is absent (the GPU box).  Same names, types and defaults.
#ifndef VS_TEST_STUB_UTILITY_CONFIG_H
#define VS_TEST_STUB_UTILITY_CONFIG_H
#include <string>
namespace utility {
struct CameraConfig {
    double focal_length = 460.0;
    double row = 480.0;
    double col = 752.0;
};
struct FeatureTrackerConfig {
    int max_cnt = 150;
    int min_dist = 30;
    int window_size = 20;
    double f_threshold = 1.0;
    int show_track = 1;
    int equalize = 1;
    int fisheye = 0;
    int lk_window_size = 21;
    int lk_pyramid_levels = 3;
    int lk_iterations = 30;
    double lk_eps = 0.01;
    double f_threshold_edge_factor = 0.0;
    std::string fisheye_mask;
};
struct Config {
    CameraConfig camera;
    FeatureTrackerConfig feature_tracker;
};
extern Config g_config;
}  // namespace utility
#endif
