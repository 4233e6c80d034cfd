// TEST INFRASTRUCTURE ONLY — exposes libstdc++'s real std::sort with the comparator of
// /root/reference/src/frontend/feature_tracker.cpp:70-73 (sort by track_cnt descending, UNSTABLE),
// so the Python oracle reproduces the reference's tie order exactly instead of guessing it.
// Build: g++ -O2 -shared -fPIC -o oracle/_build/libstdsort.so oracle/stdsort.cpp   (oracle/Makefile)
#include <algorithm>
#include <utility>
#include <vector>

extern "C" void oracle_setmask_order(const int* track_cnt, int n, int* perm_out) {
    // element type mirrors pair<int, pair<Point2f,int>>: only .first takes part in the comparison,
    // the payload here is the original index.
    std::vector<std::pair<int, std::pair<std::pair<float, float>, int>>> v;
    v.reserve(n);
    for (int i = 0; i < n; ++i) v.push_back({track_cnt[i], {{0.f, 0.f}, i}});
    std::sort(v.begin(), v.end(),
              [](const std::pair<int, std::pair<std::pair<float, float>, int>>& a,
                 const std::pair<int, std::pair<std::pair<float, float>, int>>& b) { return a.first > b.first; });
    for (int i = 0; i < n; ++i) perm_out[i] = v[i].second.second;
}
