// dh_k1.cuh — K1 arithmetic: box -> (integer centre, radius), one rounding per reference operation.
// Restates get_centers_and_radii (/root/reference/packages/draw_heatmap/tests/_test_helpers.py:20-28):
//   tl = c - bbox[:2]; br = bbox[2:] - c; radius = int32(ceil(min(tl, br) / osf)), radius < 1 -> 1;
//   center_int = int32(c / osf)  (fp32 divide, truncation toward zero).
#pragma once

namespace dh {

__device__ __forceinline__ void radius_center_one(float cx, float cy, float x1, float y1, float x2, float y2,
                                                  float osf, int& xi, int& yi, int& r) {
    const float m = fminf(fminf(__fsub_rn(cx, x1), __fsub_rn(cy, y1)), fminf(__fsub_rn(x2, cx), __fsub_rn(y2, cy)));
    r = (int)ceilf(__fdiv_rn(m, osf));
    if (r < 1) r = 1;
    xi = (int)__fdiv_rn(cx, osf);
    yi = (int)__fdiv_rn(cy, osf);
}

}  // namespace dh
