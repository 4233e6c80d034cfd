// Device-resident tracker state and the per-stream "glue" kernels' parameter block.
// Everything detectAndTrack does between its OpenCV calls (feature_tracker.cpp:138-199) runs on the
// device, one CTA per camera stream, so a frame needs no host round trip (except the optional host
// RANSAC callback).
#pragma once
#include "kernels.h"

namespace vs {

constexpr int kHdrInts = 16;   // per-stream result header
enum Hdr { H_N = 0, H_NNEW, H_LK_IN, H_LK_OK, H_RANSAC_IN, H_RANSAC_OUT, H_MASK_KEPT, H_NCAND, H_RANSAC_RAN,
           H_RMS_BITS, H_WANT };

struct GlueParams {
    int n_streams, cap;
    int W, H;
    int max_cnt, min_dist;
    int ransac_enabled;        // mode != OFF
    double focal_length;
    double f_threshold, f_edge_factor;
    int fb_check;
    float fb_threshold;
    CameraDev cam;

    // persistent per-stream feature state (what the reference keeps in cur_pts/ids/track_cnt/...)
    int* n;                    // [B]
    float2* pts;               // [B][cap] cur_pts
    int* ids;                  // [B][cap]
    int* cnt;                  // [B][cap]
    float2* un;                // [B][cap] cur_undistorted_pts
    int* next_id;              // [B]     per-stream n_id
    double* prev_time;         // [B]
    const double* cur_time;    // [B]     (device copy of this frame's timestamps)

    // LK outputs
    float2* lk_next;           // [B][cap]
    uint8_t* lk_status;        // [B][cap]
    float2* lk_back;           // [B][cap] (fb_check)
    uint8_t* lk_back_status;

    // compacted work set (after status/inBorder filtering)
    int* wn;                   // [B]
    float2* w_cur;             // previous-frame position
    float2* w_next;
    int* w_ids;
    int* w_cnt;
    float2* w_un;
    float2* w_lcur;            // virtual-pinhole lifts handed to the F-matrix RANSAC
    float2* w_lnext;
    uint8_t* w_rstatus;        // RANSAC inlier flags
    int* ransac_flag;          // [B]

    // kept set (after setMask)
    int* kn;                   // [B]
    float2* k_pts;
    int* k_ids;
    int* k_cnt;
    float2* k_un;
    int* want;                 // [B] max_cnt - kn

    // mask
    const uint8_t* base_mask;  // [H][mask_pitch] fisheye / file mask / all 255 (shared by all streams)
    uint8_t* mask;             // [B][H][mask_pitch]
    int mask_pitch;
    int64_t mask_stride;
    const int* disc_hw;        // [min_dist + 1] half widths of cv::circle(r = min_dist) rows

    // Shi-Tomasi candidates
    unsigned long long* cand;  // [B][cand_cap]
    int* cand_count;           // [B]
    int cand_cap;
    unsigned int* max_enc;     // [B]
    unsigned long long* cand_raw;   // [B][cand_cap] two-tier Shi-Tomasi: records awaiting exact evaluation
    int* raw_count;            // [B]
    unsigned int* l_enc;       // [B] lower bound of the masked max (fp32 scan)

    // results (device mirror of the pinned host block)
    int* r_hdr;                // [B][kHdrInts]
    int* r_ids;                // [B][cap]
    int* r_cnt;
    float2* r_pts;
    float2* r_un;
    float2* r_vel;
};

void launch_post_lk(const GlueParams& g, cudaStream_t st, int64_t* launches);
// with_post_lk: post_lk fused in front; with_ransac (needs with_post_lk): device RANSAC fused between them
void launch_set_mask(const GlueParams& g, cudaStream_t st, int64_t* launches, bool with_post_lk = false,
                     bool with_ransac = false, unsigned int seed = 0);
void launch_select_finish(const GlueParams& g, cudaStream_t st, int64_t* launches);
void launch_undistort(const CameraDev& cam, const float2* pts, int n, float2* out, cudaStream_t st);
int set_mask_smem_bytes(int cap);
int select_finish_smem_bytes(int cap);
int glue_max_smem_bytes(int cap);

}  // namespace vs
