// TEST INFRASTRUCTURE (oracle/). C driver around the reference's UNMODIFIED DALI operator
//   /root/reference/packages/dali_pipeline_framework/ext_impl/DrawGaussians.cc:88-157 (RunImpl) / :31-75 (draw_gaussian)
// compiled against oracle/dali_stub (DALI is not installed here). Built by oracle/build_oracle.py into
// oracle/_ref/libdg_ref.so; used to generate tests/golden/ref_dali.npz and to pin the C restatement
// (dho_draw_gaussians in dh_oracle.c). Never part of the product.
#include <cstdint>
#include <cstring>
#include <vector>

#include "DrawGaussians.h"

extern "C" __attribute__((visibility("default"))) int dgref_draw_gaussians(
    const float* heat_in, float* heat_out, int B, int C /* 0: 2-D heatmaps [H,W] */, int H, int W, const uint8_t* active,
    const int32_t* slice_ids, const int32_t* centers_xy, const float* radii, const int64_t* offsets /* [B+1] */,
    const float* k_for_classes, int num_k, float radius_to_sigma_factor) {
    try {
        dali::OpSpec spec;
        spec.float_vecs["k_for_classes"] = std::vector<float>(k_for_classes, k_for_classes + num_k);
        spec.floats["radius_to_sigma_factor"] = radius_to_sigma_factor;
        custom_operators::DrawGaussians<dali::CPUBackend> op(spec);
        dali::Workspace ws;
        ws.inputs.resize(5);
        ws.outputs.resize(1);
        const dali::DALIDataType types[5] = {DALI_FLOAT, DALI_BOOL, DALI_INT32, DALI_INT32, DALI_FLOAT};
        const size_t sample_elems = (size_t)(C > 0 ? C : 1) * H * W;
        static_assert(sizeof(bool) == 1, "active is passed as bytes");
        for (int i = 0; i < 5; ++i) ws.inputs[i].dtype = types[i];
        for (int s = 0; s < B; ++s) {
            const int64_t o0 = offsets[s], n = offsets[s + 1] - offsets[s];
            const dali::TensorShape<> hs = C > 0 ? dali::TensorShape<>(C, H, W) : dali::TensorShape<>(H, W);
            ws.inputs[0].shapes.push_back(hs);
            ws.inputs[0].data.push_back(const_cast<float*>(heat_in) + s * sample_elems);
            ws.inputs[1].shapes.push_back(dali::TensorShape<>(n));
            ws.inputs[1].data.push_back(const_cast<uint8_t*>(active) + o0);
            ws.inputs[2].shapes.push_back(dali::TensorShape<>(n));
            ws.inputs[2].data.push_back(const_cast<int32_t*>(slice_ids) + o0);
            ws.inputs[3].shapes.push_back(dali::TensorShape<>(n, (int64_t)2));
            ws.inputs[3].data.push_back(const_cast<int32_t*>(centers_xy) + 2 * o0);
            ws.inputs[4].shapes.push_back(dali::TensorShape<>(n));
            ws.inputs[4].data.push_back(const_cast<float*>(radii) + o0);
            ws.outputs[0].shapes.push_back(hs);
            ws.outputs[0].data.push_back(heat_out + s * sample_elems);
        }
        std::vector<dali::OutputDesc> desc;
        if (!op.Setup(desc, ws)) return 2;
        op.Run(ws);
        return 0;
    } catch (const std::exception&) {
        return 1;
    }
}
