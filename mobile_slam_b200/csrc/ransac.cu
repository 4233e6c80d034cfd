// F-matrix RANSAC on the device — placeholder until the 7-point kernel lands.
#include "tracker_state.h"

namespace vs {
int launch_ransac(vs_tracker*) {
    set_error("VS_RANSAC_DEVICE is not built yet: use VS_RANSAC_OFF or VS_RANSAC_CALLBACK");
    return VS_ERR_STATE;
}
}  // namespace vs

extern "C" int vs_stage_ransac(vs_tracker*, const float*, const float*, int32_t, uint8_t*) {
    vs::set_error("vs_stage_ransac: device RANSAC is not built yet");
    return VS_ERR_STATE;
}
