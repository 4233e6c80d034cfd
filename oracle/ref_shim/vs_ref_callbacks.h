/* TEST INFRASTRUCTURE ONLY — the OpenCV entry points the reference's feature_tracker.cpp calls, as C
 * callbacks.  oracle/ref_binding.py fills them with the real cv2 (opencv-python 4.13.0) functions, so the
 * reference's own compiled glue runs on top of the real OpenCV arithmetic although OpenCV's C++ SDK is absent. */
#ifndef VS_REF_CALLBACKS_H
#define VS_REF_CALLBACKS_H
#ifdef __cplusplus
extern "C" {
#endif
typedef struct vs_ref_callbacks {
    /* cv::CLAHE::apply (createCLAHE(clip, Size(tx,ty))) */
    void (*clahe_apply)(double clip, int tx, int ty, const unsigned char* src, int rows, int cols, long src_step,
                        unsigned char* dst /* rows*cols, packed */);
    /* cv::calcOpticalFlowPyrLK on the two level-0 images (flags = 0, minEigThreshold default) */
    void (*lk)(const unsigned char* prev, long prev_step, const unsigned char* next, long next_step, int rows, int cols,
               const float* prev_pts, int n, float* next_pts, unsigned char* status, float* err, int win_w, int win_h,
               int max_level, int crit_type, int crit_count, double crit_eps);
    /* cv::goodFeaturesToTrack(img, corners, max_corners, quality, min_dist, mask); returns the corner count */
    int (*gftt)(const unsigned char* img, long step, const unsigned char* mask, long mask_step, int rows, int cols,
                int max_corners, double quality, double min_dist, float* out_xy);
    /* cv::findFundamentalMat(p1, p2, method, thr, conf, status); returns 1 when F (3x3, row-major doubles) is valid,
     * 0 when OpenCV returned an empty matrix; has_status tells whether OpenCV filled the mask */
    int (*find_f)(const float* p1, const float* p2, int n, int method, double thr, double conf, double* f9,
                  unsigned char* status, int* has_status);
    /* cv::circle(img, Point(cx,cy), radius, Scalar(color), thickness) */
    void (*circle)(unsigned char* img, long step, int rows, int cols, int cx, int cy, int radius, int color, int thickness);
    /* cv::imread(path, IMREAD_GRAYSCALE): call with out == NULL to get the size (0x0 when unreadable), then again */
    void (*imread_gray)(const char* path, unsigned char* out, int* rows, int* cols);
    /* cv::FileStorage(path, READ)[k1][k2] (k2 = "" for a top-level key) through cv2.FileStorage: returns -1 when the file
     * cannot be opened, 0 for a missing node, 1 for an integer node, 2 for a real node (value in *num), 3 for a string
     * node (copied into str) */
    int (*yaml_get)(const char* path, const char* k1, const char* k2, double* num, char* str, int str_cap);
    /* Eigen::EigenSolver stand-in: eigenvalues of an n x n column-major matrix (numpy.linalg.eigvals) */
    void (*eigvals)(int n, const double* a_colmajor, double* wr, double* wi);
} vs_ref_callbacks;
extern vs_ref_callbacks vs_ref_cb;
#ifdef __cplusplus
}
#endif
#endif
