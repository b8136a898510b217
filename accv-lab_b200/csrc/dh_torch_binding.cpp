// dh_torch_binding.cpp — thin ATen/pybind layer over the C ABI (include/dh_b200.h).
//
// Mirrors the reference extension module `accvlab.draw_heatmap.draw_heatmap_ext`
// (/root/reference/packages/draw_heatmap/accvlab/draw_heatmap/csrc/draw_heatmap.cpp:131-145): same function names,
// argument names, order and defaults; same host-side validation as draw_heatmap_cuda.cu:62-165 (is_cuda /
// is_contiguous / shape checks -> RuntimeError), plus explicit dtype checks. Differences, all documented in
// nums_targets may be int32 OR int64 (no per-call cast); launch failures raise instead of printf.
// Everything named *_b200 / radius_center / render_* is an addition (FRESH mode, fused K1), not in the reference.
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <torch/extension.h>

#include <algorithm>
#include <tuple>
#include <vector>

#include "dh_b200.h"

namespace {

#define CHECK_CUDA(x) TORCH_CHECK(x.is_cuda(), #x " must be a CUDA tensor")
#define CHECK_CONTIGUOUS(x) TORCH_CHECK(x.is_contiguous(), #x " must be contiguous")
#define CHECK_INPUT(x) \
    CHECK_CUDA(x);     \
    CHECK_CONTIGUOUS(x)
#define CHECK_DTYPE(x, dt) TORCH_CHECK(x.scalar_type() == dt, #x " must have dtype " #dt ", got ", x.scalar_type())
#define CHECK_SAME_DEVICE(x, ref) TORCH_CHECK(x.device() == ref.device(), #x " must be on the same device as " #ref)

void check_status(int st, const char* what) {
    TORCH_CHECK(st == DH_OK, what, " failed: ", dh_status_string(st), " (status ", st, ")");
}

// The C ABI reads "labels == NULL" as "one plane per sample". A class-wise call with max_num_targets == 0 has an empty
// labels tensor whose data pointer is NULL too: hand over a non-NULL stand-in (never dereferenced: there are no slots).
const int32_t* labels_ptr(const at::Tensor& labels) {
    static const int32_t dummy = 0;
    const int32_t* p = labels.data_ptr<int>();
    return p != nullptr ? p : &dummy;
}

const void* sizes_ptr(const at::Tensor& nums_targets, int& tag) {
    TORCH_CHECK(nums_targets.scalar_type() == at::kInt || nums_targets.scalar_type() == at::kLong,
                "nums_targets must have dtype int32 or int64, got ", nums_targets.scalar_type());
    tag = nums_targets.scalar_type() == at::kLong ? DH_SIZES_I64 : DH_SIZES_I32;
    return nums_targets.data_ptr();
}

// Bin-list workspace of the batched entries (K2 pre-pass, dh_b200.h dh_batched_workspace_bytes): from torch's caching
// allocator on the current stream — stream-ordered, no device synchronisation, re-used across calls by the allocator.
at::Tensor bin_workspace(const at::Tensor& like, int64_t B, int64_t T, int64_t C, size_t& bytes) {
    bytes = dh_batched_workspace_bytes(B, T, C);
    return at::empty({(int64_t)std::max<size_t>(bytes, 16)}, like.options().dtype(at::kByte));
}

// ---- flat entry (draw_heatmap_cuda.cu:62-89) ---------------------------------------------------------------
void draw_heatmap_mode(at::Tensor& heatmap, const at::Tensor& centers, const at::Tensor& radii,
                       const at::Tensor& heatmap_idxes, float diameter_to_sigma_factor, float k_scale, int mode) {
    CHECK_INPUT(heatmap);
    CHECK_INPUT(centers);
    CHECK_INPUT(radii);
    CHECK_INPUT(heatmap_idxes);
    c10::cuda::CUDAGuard guard(heatmap.device());
    auto stream = at::cuda::getCurrentCUDAStream();
    TORCH_CHECK(centers.size(0) == radii.size(0), "centers and radii must have the same size at dim0");
    TORCH_CHECK(centers.size(0) == heatmap_idxes.size(0), "centers and heatmap_idxes must have the same size at dim0");
    TORCH_CHECK(heatmap.dim() == 3, "heatmap must be of shape [num_heatmaps, height, width]");
    TORCH_CHECK(centers.dim() == 2 && centers.size(1) == 2, "centers must be of shape [num_targets, 2]");
    TORCH_CHECK(radii.dim() == 1 && heatmap_idxes.dim() == 1, "radii and heatmap_idxes must be of shape [num_targets]");
    CHECK_DTYPE(heatmap, at::kFloat);
    CHECK_DTYPE(centers, at::kInt);
    CHECK_DTYPE(radii, at::kInt);
    CHECK_DTYPE(heatmap_idxes, at::kInt);
    CHECK_SAME_DEVICE(centers, heatmap);
    CHECK_SAME_DEVICE(radii, heatmap);
    CHECK_SAME_DEVICE(heatmap_idxes, heatmap);
    const int64_t n = centers.size(0), m = heatmap.size(0);
    const size_t ws_bytes = dh_draw_heatmap_workspace_bytes(n, m);
    at::Tensor ws = at::empty({(int64_t)ws_bytes}, heatmap.options().dtype(at::kByte));
    check_status(dh_draw_heatmap(stream.stream(), heatmap.data_ptr<float>(), centers.data_ptr<int>(), radii.data_ptr<int>(),
                                 heatmap_idxes.data_ptr<int>(), n, m, heatmap.size(1), heatmap.size(2),
                                 diameter_to_sigma_factor, k_scale, mode, ws.data_ptr(), ws_bytes, nullptr),
                 "draw_heatmap");
}

// `assume_zero_init` (keyword-only, default false; NOT in the reference): the caller promises that the heatmap holds
// zeros or nothing it wants to keep — every pixel is then written once (DH_MODE_FRESH, zero-init fused) instead of
// read-max-written, which also makes the caller-side torch.zeros (e.g. DH/tests/test_draw_heatmap.py:70) unnecessary.
// The positional signature is the reference's (draw_heatmap.cpp:132-144).
void draw_heatmap(at::Tensor& heatmaps, const at::Tensor& centers, const at::Tensor& radii,
                  const at::Tensor& heatmap_idxes, float diameter_to_sigma_factor, float k_scale, bool assume_zero_init) {
    draw_heatmap_mode(heatmaps, centers, radii, heatmap_idxes, diameter_to_sigma_factor, k_scale,
                      assume_zero_init ? DH_MODE_FRESH : DH_MODE_ACCUMULATE);
}

// ---- batched entries (draw_heatmap_cuda.cu:91-165) ---------------------------------------------------------
// 16-bit planes (b200 additions only): 0 for float32, else the DH_OUT_* tag
int out16_tag(const at::Tensor& heatmap) {
    if (heatmap.scalar_type() == at::kHalf) return DH_OUT_F16;
    if (heatmap.scalar_type() == at::kBFloat16) return DH_OUT_BF16;
    return 0;
}

void batched_checks(const at::Tensor& heatmap, const at::Tensor& centers, const at::Tensor& radii,
                    const at::Tensor& nums_targets, const at::Tensor* labels, bool allow16 = false) {
    CHECK_INPUT(heatmap);
    CHECK_INPUT(centers);
    CHECK_INPUT(radii);
    CHECK_INPUT(nums_targets);
    TORCH_CHECK(radii.dim() == 2, "radii must be of shape [batch_size, num_targets]");
    TORCH_CHECK(centers.dim() == 3 && centers.size(2) == 2, "centers must be of shape [batch_size, num_targets, 2]");
    const int64_t batch_size = heatmap.size(0);
    const int64_t num_targets = radii.size(1);
    TORCH_CHECK(nums_targets.dim() == 1, "nums_targets must be of shape [batch_size]");
    TORCH_CHECK(batch_size == radii.size(0) && batch_size == centers.size(0) && batch_size == nums_targets.size(0),
                "batch_size (dim 0) need to be the same for all inputs");
    TORCH_CHECK(num_targets == centers.size(1), "maximum number of targets (dim 1) need to be the same centers and radii");
    if (labels == nullptr) {
        TORCH_CHECK(heatmap.dim() == 3, "heatmap must be of shape [batch_size, height, width]");
    } else {
        const at::Tensor& labels_ = *labels;
        CHECK_INPUT(labels_);
        TORCH_CHECK(heatmap.dim() == 4, "heatmap must be of shape [batch_size, max_num_classes, height, width]");
        TORCH_CHECK(labels_.dim() == 2, "labels must be of shape [batch_size, radii.size(1)]");
        TORCH_CHECK(labels_.size(0) == batch_size && labels_.size(1) == num_targets,
                    "labels shape must be [batch_size, radii.size(1)]");
        CHECK_DTYPE(labels_, at::kInt);
        CHECK_SAME_DEVICE(labels_, heatmap);
    }
    if (!(allow16 && out16_tag(heatmap) != 0)) CHECK_DTYPE(heatmap, at::kFloat);
    CHECK_SAME_DEVICE(centers, heatmap);
    CHECK_SAME_DEVICE(radii, heatmap);
    CHECK_SAME_DEVICE(nums_targets, heatmap);
}

void batched_mode(at::Tensor& heatmap, const at::Tensor& centers, const at::Tensor& radii, const at::Tensor& nums_targets,
                  const at::Tensor* labels, float factor, float k, int mode, bool use_workspace = false) {
    batched_checks(heatmap, centers, radii, nums_targets, labels, mode == DH_MODE_FRESH);
    CHECK_DTYPE(centers, at::kInt);
    CHECK_DTYPE(radii, at::kInt);
    c10::cuda::CUDAGuard guard(heatmap.device());
    auto stream = at::cuda::getCurrentCUDAStream();
    int tag = 0;
    const void* sizes = sizes_ptr(nums_targets, tag);
    const bool cw = labels != nullptr;
    // workspace-less entries = the single-launch renderer, which measured faster than the two-launch (K2 pre-pass) path
    // of the *_ws entries on every BASELINE config (profiles/r02/experiments.md); pass `use_workspace` through
    // b200.render_* to try the other one
    size_t wsb = 0;
    at::Tensor ws;
    if (use_workspace) ws = bin_workspace(heatmap, heatmap.size(0), radii.size(1), cw ? heatmap.size(1) : 1, wsb);
    void* const wsp = use_workspace ? ws.data_ptr() : nullptr;
    if (const int o16 = out16_tag(heatmap)) {  // fp16 / bf16 planes: FRESH only
        check_status(dh_render_heatmap_batched_16_ws(stream.stream(), heatmap.data_ptr(), o16, centers.data_ptr<int>(),
                                                     radii.data_ptr<int>(), nullptr, nullptr, 1.0f, sizes, tag,
                                                     cw ? labels_ptr(*labels) : nullptr, heatmap.size(0), radii.size(1),
                                                     cw ? heatmap.size(1) : 1, heatmap.size(cw ? 2 : 1),
                                                     heatmap.size(cw ? 3 : 2), factor, k, nullptr, wsp, wsb),
                     "render_heatmap_batched (16-bit planes)");
        return;
    }
    check_status(dh_draw_heatmap_batched_ws(stream.stream(), heatmap.data_ptr<float>(), centers.data_ptr<int>(),
                                            radii.data_ptr<int>(), sizes, tag, cw ? labels_ptr(*labels) : nullptr,
                                            heatmap.size(0), radii.size(1), cw ? heatmap.size(1) : 1,
                                            heatmap.size(cw ? 2 : 1), heatmap.size(cw ? 3 : 2), factor, k, mode, nullptr,
                                            wsp, wsb),
                 "draw_heatmap_batched");
}

void draw_heatmap_batched(at::Tensor& heatmaps, const at::Tensor& centers, const at::Tensor& radii,
                          const at::Tensor& nums_targets, float diameter_to_sigma_factor, float k_scale,
                          bool assume_zero_init) {
    if (assume_zero_init) CHECK_DTYPE(heatmaps, at::kFloat);  // the reference-named entries stay float32-only
    batched_mode(heatmaps, centers, radii, nums_targets, nullptr, diameter_to_sigma_factor, k_scale,
                 assume_zero_init ? DH_MODE_FRESH : DH_MODE_ACCUMULATE);
}

void draw_heatmap_batched_classwise(at::Tensor& heatmaps, const at::Tensor& centers, const at::Tensor& radii,
                                    const at::Tensor& nums_targets, const at::Tensor& labels,
                                    float diameter_to_sigma_factor, float k_scale, bool assume_zero_init) {
    if (assume_zero_init) CHECK_DTYPE(heatmaps, at::kFloat);
    batched_mode(heatmaps, centers, radii, nums_targets, &labels, diameter_to_sigma_factor, k_scale,
                 assume_zero_init ? DH_MODE_FRESH : DH_MODE_ACCUMULATE);
}

// ---- additions -----------------------------------------------------------------------------------------------
void render_heatmap_batched(at::Tensor& heatmaps, const at::Tensor& centers, const at::Tensor& radii,
                            const at::Tensor& nums_targets, const c10::optional<at::Tensor>& labels,
                            float diameter_to_sigma_factor, float k_scale, bool fresh, bool use_workspace) {
    batched_mode(heatmaps, centers, radii, nums_targets, labels.has_value() ? &labels.value() : nullptr,
                 diameter_to_sigma_factor, k_scale, fresh ? DH_MODE_FRESH : DH_MODE_ACCUMULATE, use_workspace);
}

void render_heatmap_flat(at::Tensor& heatmaps, const at::Tensor& centers, const at::Tensor& radii,
                         const at::Tensor& heatmap_idxes, float diameter_to_sigma_factor, float k_scale, bool fresh) {
    draw_heatmap_mode(heatmaps, centers, radii, heatmap_idxes, diameter_to_sigma_factor, k_scale,
                      fresh ? DH_MODE_FRESH : DH_MODE_ACCUMULATE);
}

std::tuple<at::Tensor, at::Tensor> radius_center(const at::Tensor& centers, const at::Tensor& bboxes,
                                                 double out_size_factor) {
    CHECK_INPUT(centers);
    CHECK_INPUT(bboxes);
    CHECK_DTYPE(centers, at::kFloat);
    CHECK_DTYPE(bboxes, at::kFloat);
    CHECK_SAME_DEVICE(bboxes, centers);
    TORCH_CHECK(centers.dim() >= 1 && centers.size(-1) == 2, "centers must be of shape [..., 2]");
    TORCH_CHECK(bboxes.dim() == centers.dim() && bboxes.size(-1) == 4, "bboxes must be of shape [..., 4]");
    for (int64_t d = 0; d + 1 < centers.dim(); ++d)
        TORCH_CHECK(centers.size(d) == bboxes.size(d), "centers and bboxes must agree in all leading dimensions");
    c10::cuda::CUDAGuard guard(centers.device());
    auto stream = at::cuda::getCurrentCUDAStream();
    at::Tensor centers_out = at::empty(centers.sizes(), centers.options().dtype(at::kInt));
    auto lead = centers.sizes().vec();
    lead.pop_back();
    at::Tensor radii_out = at::empty(lead, centers.options().dtype(at::kInt));
    check_status(dh_radius_center(stream.stream(), centers.data_ptr<float>(), bboxes.data_ptr<float>(),
                                  (float)out_size_factor, radii_out.numel(), centers_out.data_ptr<int>(),
                                  radii_out.data_ptr<int>()),
                 "radius_center");
    return {centers_out, radii_out};
}

void render_heatmap_from_boxes(at::Tensor& heatmaps, const at::Tensor& centers, const at::Tensor& bboxes,
                               double out_size_factor, const at::Tensor& nums_targets,
                               const c10::optional<at::Tensor>& labels, float diameter_to_sigma_factor, float k_scale,
                               bool fresh, bool use_workspace) {
    CHECK_INPUT(bboxes);
    CHECK_DTYPE(centers, at::kFloat);
    CHECK_DTYPE(bboxes, at::kFloat);
    TORCH_CHECK(bboxes.dim() == 3 && bboxes.size(2) == 4, "bboxes must be of shape [batch_size, num_targets, 4]");
    CHECK_SAME_DEVICE(bboxes, heatmaps);
    TORCH_CHECK(centers.dim() == 3 && centers.size(2) == 2, "centers must be of shape [batch_size, num_targets, 2]");
    TORCH_CHECK(centers.size(0) == bboxes.size(0) && centers.size(1) == bboxes.size(1),
                "centers and bboxes must agree in batch_size and num_targets");
    const at::Tensor* lab = labels.has_value() ? &labels.value() : nullptr;
    CHECK_INPUT(heatmaps);
    CHECK_INPUT(centers);
    CHECK_INPUT(nums_targets);
    const int o16 = fresh ? out16_tag(heatmaps) : 0;
    if (o16 == 0) CHECK_DTYPE(heatmaps, at::kFloat);
    TORCH_CHECK(nums_targets.dim() == 1 && nums_targets.size(0) == heatmaps.size(0) && centers.size(0) == heatmaps.size(0),
                "batch_size (dim 0) need to be the same for all inputs");
    CHECK_SAME_DEVICE(centers, heatmaps);
    CHECK_SAME_DEVICE(nums_targets, heatmaps);
    const bool cw = lab != nullptr;
    if (cw) {
        const at::Tensor& labels_ = *lab;
        CHECK_INPUT(labels_);
        CHECK_DTYPE(labels_, at::kInt);
        CHECK_SAME_DEVICE(labels_, heatmaps);
        TORCH_CHECK(heatmaps.dim() == 4, "heatmap must be of shape [batch_size, max_num_classes, height, width]");
        TORCH_CHECK(labels_.dim() == 2 && labels_.size(0) == centers.size(0) && labels_.size(1) == centers.size(1),
                    "labels shape must be [batch_size, num_targets]");
    } else {
        TORCH_CHECK(heatmaps.dim() == 3, "heatmap must be of shape [batch_size, height, width]");
    }
    c10::cuda::CUDAGuard guard(heatmaps.device());
    auto stream = at::cuda::getCurrentCUDAStream();
    int tag = 0;
    const void* sizes = sizes_ptr(nums_targets, tag);
    size_t wsb = 0;
    at::Tensor ws;
    if (use_workspace) ws = bin_workspace(heatmaps, heatmaps.size(0), centers.size(1), cw ? heatmaps.size(1) : 1, wsb);
    void* const wsp = use_workspace ? ws.data_ptr() : nullptr;
    if (o16 != 0) {  // fp16 / bf16 planes: FRESH only
        check_status(dh_render_heatmap_batched_16_ws(stream.stream(), heatmaps.data_ptr(), o16, nullptr, nullptr,
                                                     centers.data_ptr<float>(), bboxes.data_ptr<float>(),
                                                     (float)out_size_factor, sizes, tag, cw ? labels_ptr(*lab) : nullptr,
                                                     heatmaps.size(0), centers.size(1), cw ? heatmaps.size(1) : 1,
                                                     heatmaps.size(cw ? 2 : 1), heatmaps.size(cw ? 3 : 2),
                                                     diameter_to_sigma_factor, k_scale, nullptr, wsp, wsb),
                     "render_heatmap_from_boxes (16-bit planes)");
        return;
    }
    check_status(dh_draw_heatmap_batched_from_boxes_ws(
                     stream.stream(), heatmaps.data_ptr<float>(), centers.data_ptr<float>(), bboxes.data_ptr<float>(),
                     (float)out_size_factor, sizes, tag, cw ? labels_ptr(*lab) : nullptr, heatmaps.size(0),
                     centers.size(1), cw ? heatmaps.size(1) : 1, heatmaps.size(cw ? 2 : 1), heatmaps.size(cw ? 3 : 2),
                     diameter_to_sigma_factor, k_scale, fresh ? DH_MODE_FRESH : DH_MODE_ACCUMULATE, nullptr, nullptr,
                     nullptr, wsp, wsb),
                 "render_heatmap_from_boxes");
}

// digest += checksum(values); digest is a CUDA int64[3] tensor (wrapping uint64 arithmetic, see dh_b200.h)
void checksum_(const at::Tensor& values, at::Tensor& digest, int64_t index_base) {
    CHECK_INPUT(values);
    CHECK_INPUT(digest);
    CHECK_DTYPE(values, at::kFloat);
    CHECK_DTYPE(digest, at::kLong);
    CHECK_SAME_DEVICE(digest, values);
    TORCH_CHECK(digest.numel() == 3, "digest must have 3 int64 elements");
    c10::cuda::CUDAGuard guard(values.device());
    check_status(dh_checksum(at::cuda::getCurrentCUDAStream().stream(), values.data_ptr<float>(), values.numel(),
                             index_base, reinterpret_cast<uint64_t*>(digest.data_ptr<int64_t>())),
                 "checksum");
}

int64_t host_scratch_bytes(int64_t B, int64_t T, int64_t C, int64_t H, int64_t W, bool has_labels, int64_t chunk_samples,
                           bool staging) {
    return (int64_t)dh_host_scratch_bytes(B, T, C, H, W, has_labels, chunk_samples, staging);
}

// Host-buffer entry: CPU (ideally pinned) inputs -> heatmap on the host and/or the device (+ optional digest).
void render_heatmap_from_boxes_host(const c10::optional<at::Tensor>& heatmap_host,
                                    const c10::optional<at::Tensor>& heatmap_device,
                                    const c10::optional<at::Tensor>& checksum_host, const at::Tensor& centers,
                                    const at::Tensor& bboxes, double out_size_factor, const at::Tensor& nums_targets,
                                    const c10::optional<at::Tensor>& labels, int64_t num_classes, int64_t height,
                                    int64_t width, float diameter_to_sigma_factor, float k_scale, at::Tensor& scratch,
                                    int64_t chunk_samples) {
    auto check_cpu = [](const at::Tensor& t, const char* name) {
        TORCH_CHECK(t.device().is_cpu(), name, " must be a CPU (host) tensor");
        TORCH_CHECK(t.is_contiguous(), name, " must be contiguous");
    };
    check_cpu(centers, "centers");
    check_cpu(bboxes, "bboxes");
    check_cpu(nums_targets, "nums_targets");
    CHECK_DTYPE(centers, at::kFloat);
    CHECK_DTYPE(bboxes, at::kFloat);
    CHECK_DTYPE(nums_targets, at::kLong);
    TORCH_CHECK(centers.dim() == 3 && centers.size(2) == 2, "centers must be of shape [batch_size, num_targets, 2]");
    TORCH_CHECK(bboxes.dim() == 3 && bboxes.size(2) == 4, "bboxes must be of shape [batch_size, num_targets, 4]");
    const int64_t B = centers.size(0), T = centers.size(1);
    TORCH_CHECK(bboxes.size(0) == B && bboxes.size(1) == T && nums_targets.dim() == 1 && nums_targets.size(0) == B,
                "batch_size (dim 0) and num_targets (dim 1) need to be the same for all inputs");
    const bool cw = labels.has_value();
    if (cw) {
        check_cpu(*labels, "labels");
        CHECK_DTYPE((*labels), at::kInt);
        TORCH_CHECK(labels->dim() == 2 && labels->size(0) == B && labels->size(1) == T,
                    "labels shape must be [batch_size, num_targets]");
    }
    const int64_t C = cw ? num_classes : 1;
    const int64_t numel = B * C * height * width;
    float* hh = nullptr;
    float* hd = nullptr;
    if (heatmap_host.has_value()) {
        check_cpu(*heatmap_host, "heatmap_host");
        CHECK_DTYPE((*heatmap_host), at::kFloat);
        TORCH_CHECK(heatmap_host->numel() == numel, "heatmap_host has the wrong number of elements");
        hh = heatmap_host->data_ptr<float>();
    }
    CHECK_INPUT(scratch);
    if (heatmap_device.has_value()) {
        const at::Tensor& heatmap_device_ = *heatmap_device;
        CHECK_INPUT(heatmap_device_);
        CHECK_DTYPE(heatmap_device_, at::kFloat);
        CHECK_SAME_DEVICE(heatmap_device_, scratch);
        TORCH_CHECK(heatmap_device->numel() == numel, "heatmap_device has the wrong number of elements");
        hd = heatmap_device->data_ptr<float>();
    }
    TORCH_CHECK(hh || hd, "at least one of heatmap_host / heatmap_device must be given");
    uint64_t* ck = nullptr;
    if (checksum_host.has_value()) {
        check_cpu(*checksum_host, "checksum_host");
        CHECK_DTYPE((*checksum_host), at::kLong);
        TORCH_CHECK(checksum_host->numel() == 3, "checksum_host must have 3 int64 elements");
        ck = reinterpret_cast<uint64_t*>(checksum_host->data_ptr<int64_t>());
    }
    c10::cuda::CUDAGuard guard(scratch.device());
    check_status(dh_draw_heatmap_batched_from_boxes_host(
                     at::cuda::getCurrentCUDAStream().stream(), hh, hd, ck, centers.data_ptr<float>(),
                     bboxes.data_ptr<float>(), (float)out_size_factor, nums_targets.data_ptr<int64_t>(),
                     cw ? labels_ptr(*labels) : nullptr, B, T, C, height, width, diameter_to_sigma_factor, k_scale,
                     scratch.data_ptr(), (size_t)scratch.nbytes(), chunk_samples),
                 "render_heatmap_from_boxes_host");
}

// ---- N1: list of per-sample tensors -> padded RaggedBatch tensor (+ sample sizes), one launch -----------------
// Host side of dh_pack_ragged / dh_pack_radius_center: the B source pointers and B sizes go into ONE pinned host
// table that crosses PCIe in one asynchronous copy (stream-ordered; the caching host allocator keeps the block alive
// until the copy has run). Replaces the per-sample slice-assign loop of combine_data
// (batching_helpers/batched_processing_py.py:411-423) and of DH/examples/draw_heatmap_batched.py:42-50.
struct PackTable {
    at::Tensor dev;  // int64 [n_arrays*B + B]: pointer columns, then sizes
    int64_t B = 0, T = 0;
};

PackTable make_pack_table(const std::vector<const std::vector<at::Tensor>*>& lists, int64_t max_sample_size) {
    const auto& first = *lists[0];
    const int64_t B = (int64_t)first.size();
    TORCH_CHECK(B > 0, "`data_list` must not be empty");
    const int64_t na = (int64_t)lists.size();
    at::Tensor host = at::empty({(na + 1) * B}, at::TensorOptions().dtype(at::kLong).pinned_memory(true));
    int64_t* h = host.data_ptr<int64_t>();
    int64_t T = 0;
    const at::Device device = first[0].device();
    for (int64_t i = 0; i < B; ++i) {
        const at::Tensor& el = first[i];
        const int64_t n = el.dim() == 0 ? 0 : std::min<int64_t>(el.size(0), el.numel());
        for (int64_t a = 0; a < na; ++a) {
            const at::Tensor& t = (*lists[a])[i];
            TORCH_CHECK(t.is_cuda() && t.device() == device, "all list elements must be CUDA tensors on one device");
            TORCH_CHECK(t.is_contiguous(), "list elements must be contiguous");
            const int64_t nt = t.dim() == 0 ? 0 : std::min<int64_t>(t.size(0), t.numel());
            TORCH_CHECK(nt == n, "lists disagree on the size of sample ", i);
            h[a * B + i] = (int64_t) reinterpret_cast<uintptr_t>(t.data_ptr());
        }
        h[na * B + i] = n;
        T = std::max(T, n);
    }
    if (max_sample_size >= 0) {
        TORCH_CHECK(max_sample_size >= T, "max_sample_size (", max_sample_size, ") is smaller than the largest sample (", T, ")");
        T = max_sample_size;
    }
    PackTable r;
    r.dev = host.to(device, /*non_blocking=*/true);
    r.B = B;
    r.T = T;
    return r;
}

std::tuple<at::Tensor, at::Tensor> combine_data(const std::vector<at::Tensor>& data_list, int64_t max_sample_size) {
    TORCH_CHECK(!data_list.empty(), "`data_list` must not be empty");
    const at::Tensor* sample = nullptr;  // first non-empty element defines dtype and trailing dims
    for (const auto& t : data_list)
        if (t.numel() > 0) {
            sample = &t;
            break;
        }
    if (!sample) sample = &data_list[0];
    std::vector<int64_t> trailing;
    for (int64_t d = 1; d < sample->dim(); ++d) trailing.push_back(sample->size(d));
    int64_t row = 1;
    for (auto v : trailing) row *= v;
    for (const auto& t : data_list) {
        TORCH_CHECK(t.scalar_type() == sample->scalar_type(), "all list elements must have the same dtype");
        if (t.numel() > 0) {
            TORCH_CHECK(t.dim() == sample->dim(), "all non-empty list elements must have the same number of dimensions");
            for (int64_t d = 1; d < t.dim(); ++d)
                TORCH_CHECK(t.size(d) == sample->size(d), "list elements must match in every dimension but the first");
        }
    }
    c10::cuda::CUDAGuard guard(sample->device());
    PackTable tab = make_pack_table({&data_list}, max_sample_size);
    std::vector<int64_t> shape = {tab.B, tab.T};
    shape.insert(shape.end(), trailing.begin(), trailing.end());
    at::Tensor out = at::empty(shape, sample->options());
    const int64_t* d = tab.dev.data_ptr<int64_t>();
    check_status(dh_pack_ragged(at::cuda::getCurrentCUDAStream().stream(), out.data_ptr(),
                                reinterpret_cast<const void* const*>(d), d + tab.B, tab.B, tab.T, row,
                                (int)sample->element_size()),
                 "combine_data");
    return {out, tab.dev.slice(0, tab.B, 2 * tab.B)};
}

std::tuple<at::Tensor, at::Tensor, at::Tensor> pack_radius_center(const std::vector<at::Tensor>& centers_list,
                                                                  const std::vector<at::Tensor>& bboxes_list,
                                                                  double out_size_factor, int64_t max_sample_size) {
    TORCH_CHECK(!centers_list.empty(), "`centers_list` must not be empty");
    TORCH_CHECK(centers_list.size() == bboxes_list.size(), "centers_list and bboxes_list must have the same length");
    for (size_t i = 0; i < centers_list.size(); ++i) {
        const at::Tensor &c = centers_list[i], &b = bboxes_list[i];
        TORCH_CHECK(c.scalar_type() == at::kFloat && b.scalar_type() == at::kFloat, "centres and boxes must be float32");
        TORCH_CHECK(c.numel() == 0 || (c.dim() == 2 && c.size(1) == 2), "centres must be of shape [n, 2]");
        TORCH_CHECK(b.numel() == 0 || (b.dim() == 2 && b.size(1) == 4), "bboxes must be of shape [n, 4]");
        TORCH_CHECK(reinterpret_cast<uintptr_t>(c.data_ptr()) % 8 == 0 && reinterpret_cast<uintptr_t>(b.data_ptr()) % 16 == 0,
                    "centres / boxes must be aligned to their row size (8 / 16 bytes)");
    }
    c10::cuda::CUDAGuard guard(centers_list[0].device());
    PackTable tab = make_pack_table({&centers_list, &bboxes_list}, max_sample_size);
    auto opt = centers_list[0].options().dtype(at::kInt);
    at::Tensor centers_out = at::empty({tab.B, tab.T, 2}, opt), radii_out = at::empty({tab.B, tab.T}, opt);
    const int64_t* d = tab.dev.data_ptr<int64_t>();
    check_status(dh_pack_radius_center(at::cuda::getCurrentCUDAStream().stream(),
                                       reinterpret_cast<const float* const*>(d),
                                       reinterpret_cast<const float* const*>(d + tab.B), d + 2 * tab.B,
                                       (float)out_size_factor, tab.B, tab.T, centers_out.data_ptr<int>(),
                                       radii_out.data_ptr<int>()),
                 "pack_radius_center");
    return {centers_out, radii_out, tab.dev.slice(0, 2 * tab.B, 3 * tab.B)};
}

// ---- N2: DALI draw_gaussians semantics (DrawGaussians.cc:88-157), RaggedBatch layout ---------------------------
at::Tensor draw_gaussians(const c10::optional<at::Tensor>& heatmap_in, at::Tensor& heatmap_out,
                          const c10::optional<at::Tensor>& active, const c10::optional<at::Tensor>& slice_ids,
                          const at::Tensor& centers, const at::Tensor& radii, const at::Tensor& nums_targets,
                          const at::Tensor& k_for_classes, double radius_to_sigma_factor, int64_t fast) {
    CHECK_INPUT(heatmap_out);
    CHECK_INPUT(centers);
    CHECK_INPUT(radii);
    CHECK_INPUT(nums_targets);
    CHECK_INPUT(k_for_classes);
    CHECK_DTYPE(heatmap_out, at::kFloat);
    CHECK_DTYPE(centers, at::kInt);
    CHECK_DTYPE(radii, at::kFloat);
    CHECK_DTYPE(k_for_classes, at::kFloat);
    TORCH_CHECK(heatmap_out.dim() == 3 || heatmap_out.dim() == 4, "Heatmap has to have 2 or 3 dimensions per sample");
    const int64_t B = heatmap_out.size(0), C = heatmap_out.dim() == 4 ? heatmap_out.size(1) : 1;
    const int64_t H = heatmap_out.size(-2), W = heatmap_out.size(-1);
    TORCH_CHECK(radii.dim() == 2 && radii.size(0) == B, "radii must be of shape [batch_size, max_num_targets]");
    const int64_t T = radii.size(1);
    TORCH_CHECK(centers.dim() == 3 && centers.size(0) == B && centers.size(1) == T && centers.size(2) == 2,
                "centers must be of shape [batch_size, max_num_targets, 2]");
    TORCH_CHECK(nums_targets.dim() == 1 && nums_targets.size(0) == B, "nums_targets must be of shape [batch_size]");
    TORCH_CHECK(k_for_classes.dim() == 1, "k_for_classes must be 1-D");
    CHECK_SAME_DEVICE(centers, heatmap_out);
    CHECK_SAME_DEVICE(radii, heatmap_out);
    CHECK_SAME_DEVICE(nums_targets, heatmap_out);
    CHECK_SAME_DEVICE(k_for_classes, heatmap_out);
    const float* in_ptr = nullptr;
    if (heatmap_in.has_value()) {
        const at::Tensor& hin = *heatmap_in;
        CHECK_INPUT(hin);
        CHECK_DTYPE(hin, at::kFloat);
        CHECK_SAME_DEVICE(hin, heatmap_out);
        TORCH_CHECK(hin.sizes() == heatmap_out.sizes(), "heatmap_in and heatmap_out must have the same shape");
        in_ptr = hin.data_ptr<float>();
    }
    const uint8_t* act_ptr = nullptr;
    if (active.has_value()) {
        const at::Tensor& a = *active;
        CHECK_INPUT(a);
        TORCH_CHECK(a.scalar_type() == at::kBool || a.scalar_type() == at::kByte, "active has to be of type BOOL");
        TORCH_CHECK(a.dim() == 2 && a.size(0) == B && a.size(1) == T, "active must be of shape [batch_size, max_num_targets]");
        CHECK_SAME_DEVICE(a, heatmap_out);
        act_ptr = static_cast<const uint8_t*>(a.data_ptr());
    }
    const int32_t* sl_ptr = nullptr;
    if (slice_ids.has_value()) {
        const at::Tensor& sl = *slice_ids;
        CHECK_INPUT(sl);
        CHECK_DTYPE(sl, at::kInt);
        TORCH_CHECK(sl.dim() == 2 && sl.size(0) == B && sl.size(1) == T, "class_ids must be of shape [batch_size, max_num_targets]");
        CHECK_SAME_DEVICE(sl, heatmap_out);
        sl_ptr = sl.data_ptr<int>();
    }
    int tag = 0;
    const void* sizes = sizes_ptr(nums_targets, tag);
    c10::cuda::CUDAGuard guard(heatmap_out.device());
    // FRESH fast path (classify -> main renderer -> exact fix-up): three launches instead of one, selected by handing the C
    // ABI a workspace. fast = 1 / 0 forces / forbids it; -1 (default) takes it from 256 MiB of output on (measured
    // break-even ~200 MB on B200).
    at::Tensor ws;
    size_t wsb = 0;
    const bool want_fast = fast > 0 || (fast < 0 && heatmap_out.numel() * 4 >= ((int64_t)256 << 20));
    if (!heatmap_in.has_value() && T > 0 && want_fast) {
        wsb = dh_draw_gaussians_workspace_bytes(B, T);
        ws = at::empty({(int64_t)wsb}, heatmap_out.options().dtype(at::kByte));
    }
    check_status(dh_draw_gaussians(at::cuda::getCurrentCUDAStream().stream(), in_ptr, heatmap_out.data_ptr<float>(),
                                   centers.data_ptr<int>(), radii.data_ptr<float>(), sl_ptr, act_ptr, sizes, tag,
                                   k_for_classes.data_ptr<float>(), k_for_classes.numel(), B, T, C, H, W,
                                   (float)radius_to_sigma_factor, nullptr, ws.defined() ? ws.data_ptr() : nullptr, wsb),
                 "draw_gaussians");
    return heatmap_out;
}

// ---- front half of BoundingBoxToHeatmapConverter (bounding_box_to_heatmap_converter.py:321-376) ---------------
std::vector<at::Tensor> boxes_to_gaussian_params(const at::Tensor& bboxes, const c10::optional<at::Tensor>& centers,
                                                 const c10::optional<at::Tensor>& classes,
                                                 const c10::optional<at::Tensor>& is_valid, const at::Tensor& nums_targets,
                                                 const at::Tensor& scale_xy, const c10::optional<at::Tensor>& min_sizes_hw,
                                                 int64_t num_classes, int64_t heatmap_h, int64_t heatmap_w,
                                                 double radius_scaling_factor, double min_radius, double max_radius,
                                                 double min_fraction_area) {
    CHECK_INPUT(bboxes);
    CHECK_INPUT(nums_targets);
    CHECK_INPUT(scale_xy);
    CHECK_DTYPE(bboxes, at::kFloat);
    CHECK_DTYPE(scale_xy, at::kFloat);
    TORCH_CHECK(bboxes.dim() == 3 && bboxes.size(2) == 4, "bboxes must be of shape [batch_size, max_num_targets, 4]");
    const int64_t B = bboxes.size(0), T = bboxes.size(1);
    TORCH_CHECK(nums_targets.dim() == 1 && nums_targets.size(0) == B, "nums_targets must be of shape [batch_size]");
    TORCH_CHECK(scale_xy.dim() == 2 && scale_xy.size(0) == B && scale_xy.size(1) == 2, "scale_xy must be of shape [batch_size, 2]");
    CHECK_SAME_DEVICE(nums_targets, bboxes);
    CHECK_SAME_DEVICE(scale_xy, bboxes);
    const float* c_ptr = nullptr;
    const int32_t* k_ptr = nullptr;
    const uint8_t* v_ptr = nullptr;
    const double* m_ptr = nullptr;
    int64_t nm = 0;
    if (centers.has_value()) {
        const at::Tensor& c = *centers;
        CHECK_INPUT(c);
        CHECK_DTYPE(c, at::kFloat);
        CHECK_SAME_DEVICE(c, bboxes);
        TORCH_CHECK(c.dim() == 3 && c.size(0) == B && c.size(1) == T && c.size(2) == 2, "centers must be of shape [batch_size, max_num_targets, 2]");
        c_ptr = c.data_ptr<float>();
    }
    if (classes.has_value()) {
        const at::Tensor& k = *classes;
        CHECK_INPUT(k);
        CHECK_DTYPE(k, at::kInt);
        CHECK_SAME_DEVICE(k, bboxes);
        TORCH_CHECK(k.dim() == 2 && k.size(0) == B && k.size(1) == T, "classes must be of shape [batch_size, max_num_targets]");
        k_ptr = k.data_ptr<int>();
    }
    if (is_valid.has_value()) {
        const at::Tensor& v = *is_valid;
        CHECK_INPUT(v);
        TORCH_CHECK(v.scalar_type() == at::kBool || v.scalar_type() == at::kByte, "is_valid must be bool");
        CHECK_SAME_DEVICE(v, bboxes);
        TORCH_CHECK(v.dim() == 2 && v.size(0) == B && v.size(1) == T, "is_valid must be of shape [batch_size, max_num_targets]");
        v_ptr = static_cast<const uint8_t*>(v.data_ptr());
    }
    if (min_sizes_hw.has_value()) {
        const at::Tensor& m = *min_sizes_hw;
        CHECK_INPUT(m);
        CHECK_DTYPE(m, at::kDouble);
        CHECK_SAME_DEVICE(m, bboxes);
        TORCH_CHECK(m.dim() == 2 && m.size(1) == 2, "min_sizes_hw must be of shape [n, 2]");
        m_ptr = m.data_ptr<double>();
        nm = m.size(0);
    }
    int tag = 0;
    const void* sizes = sizes_ptr(nums_targets, tag);
    c10::cuda::CUDAGuard guard(bboxes.device());
    auto f = bboxes.options();
    at::Tensor bc = at::empty({B, T, 4}, f), ci = at::empty({B, T, 2}, f.dtype(at::kInt)), off = at::empty({B, T, 2}, f),
               rad = at::empty({B, T}, f), act = at::empty({B, T}, f.dtype(at::kBool)), hw = at::empty({B, T, 2}, f),
               fr = at::empty({B, T}, f);
    check_status(dh_boxes_to_gaussian_params(at::cuda::getCurrentCUDAStream().stream(), bboxes.data_ptr<float>(), c_ptr, k_ptr,
                                             v_ptr, sizes, tag, scale_xy.data_ptr<float>(), m_ptr, nm, num_classes, B, T,
                                             heatmap_h, heatmap_w, radius_scaling_factor, (float)min_radius,
                                             (float)max_radius, min_fraction_area, bc.data_ptr<float>(), ci.data_ptr<int>(),
                                             off.data_ptr<float>(), rad.data_ptr<float>(),
                                             static_cast<uint8_t*>(act.data_ptr()), hw.data_ptr<float>(), fr.data_ptr<float>()),
                 "boxes_to_gaussian_params");
    return {bc, ci, off, rad, act, hw, fr};
}

// ---- N4: CenterNet / CenterPoint gaussian_radius (not in the reference; parity unpinned) -------------------------
std::tuple<at::Tensor, at::Tensor> gaussian_radius(const at::Tensor& det_hw, double min_overlap, int64_t variant,
                                                   int64_t min_radius) {
    CHECK_INPUT(det_hw);
    CHECK_DTYPE(det_hw, at::kFloat);
    TORCH_CHECK(det_hw.dim() >= 1 && det_hw.size(-1) == 2, "det_size must be of shape [..., 2] (height, width)");
    c10::cuda::CUDAGuard guard(det_hw.device());
    auto lead = det_hw.sizes().vec();
    lead.pop_back();
    at::Tensor ri = at::empty(lead, det_hw.options().dtype(at::kInt)), rf = at::empty(lead, det_hw.options());
    check_status(dh_gaussian_radius(at::cuda::getCurrentCUDAStream().stream(), det_hw.data_ptr<float>(), min_overlap,
                                    (int)variant, (int32_t)min_radius, ri.numel(), ri.data_ptr<int>(), rf.data_ptr<float>()),
                 "gaussian_radius");
    return {ri, rf};
}

// ---- N3: fused Gaussian focal loss (DH/docs/intro.rst:7-26; no reference implementation) -----------------------
std::tuple<at::Tensor, at::Tensor> gaussian_focal_loss(const at::Tensor& pred, const at::Tensor& centers,
                                                       const at::Tensor& radii, const at::Tensor& nums_targets,
                                                       const c10::optional<at::Tensor>& labels,
                                                       float diameter_to_sigma_factor, float k_scale, double alpha,
                                                       double gamma, double eps, bool want_grad, bool fast_math) {
    CHECK_INPUT(pred);
    CHECK_DTYPE(pred, at::kFloat);
    const bool classwise = labels.has_value();
    batched_checks(pred, centers, radii, nums_targets, classwise ? &*labels : nullptr);
    const int64_t B = pred.size(0), T = radii.size(1);
    const int64_t C = classwise ? pred.size(1) : 1, H = pred.size(-2), W = pred.size(-1);
    int tag = 0;
    const void* sizes = sizes_ptr(nums_targets, tag);
    c10::cuda::CUDAGuard guard(pred.device());
    at::Tensor out2 = at::empty({2}, pred.options().dtype(at::kDouble));
    at::Tensor grad = want_grad ? at::empty_like(pred) : at::Tensor();
    const size_t wsb = dh_focal_loss_workspace_bytes(B, C, H, W);
    at::Tensor ws = at::empty({(int64_t)wsb}, pred.options().dtype(at::kByte));
    check_status(dh_gaussian_focal_loss(at::cuda::getCurrentCUDAStream().stream(), pred.data_ptr<float>(),
                                        centers.data_ptr<int>(), radii.data_ptr<int>(), sizes, tag,
                                        classwise ? labels_ptr(*labels) : nullptr, B, T, C, H, W,
                                        diameter_to_sigma_factor, k_scale, (float)alpha, (float)gamma, (float)eps,
                                        fast_math ? 1 : 0, want_grad ? grad.data_ptr<float>() : nullptr, out2.data_ptr<double>(),
                                        ws.data_ptr(), wsb),
                 "gaussian_focal_loss");
    return {out2, grad};
}

// grad * g (g: 0-dim / 1-element float32 CUDA tensor), into a new tensor or in place
at::Tensor scale_by(const at::Tensor& grad, const at::Tensor& g, bool in_place) {
    CHECK_INPUT(grad);
    CHECK_DTYPE(grad, at::kFloat);
    TORCH_CHECK(g.is_cuda() && g.numel() == 1 && g.scalar_type() == at::kFloat, "scale must be a 1-element float32 CUDA tensor");
    CHECK_SAME_DEVICE(g, grad);
    c10::cuda::CUDAGuard guard(grad.device());
    at::Tensor out = in_place ? grad : at::empty_like(grad);
    check_status(dh_scale_by_device_scalar(at::cuda::getCurrentCUDAStream().stream(), out.data_ptr<float>(),
                                           grad.data_ptr<float>(), g.data_ptr<float>(), grad.numel()),
                 "scale_by");
    return out;
}

void fill_probe(at::Tensor& t) {
    CHECK_INPUT(t);
    c10::cuda::CUDAGuard guard(t.device());
    check_status(dh_fill_probe(at::cuda::getCurrentCUDAStream().stream(), t.data_ptr(), (size_t)t.nbytes()), "fill_probe");
}

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
    m.doc() = "B200-native (sm_100a) Gaussian heatmap drawing; drop-in for accvlab.draw_heatmap.draw_heatmap_ext";
    // --- the reference's surface (draw_heatmap.cpp:131-145) ---
    m.def("draw_heatmap", &draw_heatmap, "Draw Gaussians (max-blend, in place) from concatenated objects.",
          pybind11::arg("heatmaps"), pybind11::arg("centers"), pybind11::arg("radii"), pybind11::arg("heatmap_idxes"),
          pybind11::arg("diameter_to_sigma_factor") = 6.f, pybind11::arg("k_scale") = 1.f, pybind11::kw_only(),
          pybind11::arg("assume_zero_init") = false);
    m.def("draw_heatmap_batched_impl", &draw_heatmap_batched, "Draw Gaussians (max-blend, in place), one plane per sample.",
          pybind11::arg("heatmaps"), pybind11::arg("centers"), pybind11::arg("radii"), pybind11::arg("nums_targets"),
          pybind11::arg("diameter_to_sigma_factor") = 6.f, pybind11::arg("k_scale") = 1.f, pybind11::kw_only(),
          pybind11::arg("assume_zero_init") = false);
    m.def("draw_heatmap_batched_classwise_impl", &draw_heatmap_batched_classwise,
          "Draw Gaussians (max-blend, in place), one plane per (sample, class).", pybind11::arg("heatmaps"),
          pybind11::arg("centers"), pybind11::arg("radii"), pybind11::arg("nums_targets"), pybind11::arg("labels"),
          pybind11::arg("diameter_to_sigma_factor") = 6.f, pybind11::arg("k_scale") = 1.f, pybind11::kw_only(),
          pybind11::arg("assume_zero_init") = false);
    // --- additions ---
    m.def("render_heatmap_batched", &render_heatmap_batched,
          "Batched drawing with explicit mode: fresh=True overwrites every pixel (zero-init fused, nothing read).",
          pybind11::arg("heatmaps"), pybind11::arg("centers"), pybind11::arg("radii"), pybind11::arg("nums_targets"),
          pybind11::arg("labels") = pybind11::none(), pybind11::arg("diameter_to_sigma_factor") = 6.f,
          pybind11::arg("k_scale") = 1.f, pybind11::arg("fresh") = false, pybind11::arg("use_workspace") = false);
    m.def("render_heatmap_flat", &render_heatmap_flat, "Flat drawing with explicit mode.", pybind11::arg("heatmaps"),
          pybind11::arg("centers"), pybind11::arg("radii"), pybind11::arg("heatmap_idxes"),
          pybind11::arg("diameter_to_sigma_factor") = 6.f, pybind11::arg("k_scale") = 1.f, pybind11::arg("fresh") = false);
    m.def("radius_center", &radius_center, "bbox -> (int32 centres, int32 radii); tests/_test_helpers.py:20-28 semantics.",
          pybind11::arg("centers"), pybind11::arg("bboxes"), pybind11::arg("out_size_factor"));
    m.def("render_heatmap_from_boxes", &render_heatmap_from_boxes,
          "radius_center fused into the batched renderer (float boxes in, one launch).", pybind11::arg("heatmaps"),
          pybind11::arg("centers"), pybind11::arg("bboxes"), pybind11::arg("out_size_factor"),
          pybind11::arg("nums_targets"), pybind11::arg("labels") = pybind11::none(),
          pybind11::arg("diameter_to_sigma_factor") = 6.f, pybind11::arg("k_scale") = 1.f, pybind11::arg("fresh") = false,
          pybind11::arg("use_workspace") = false);
    m.def("fill_probe", &fill_probe, "Zero-fill a tensor with the renderer's store path (write-only ceiling).",
          pybind11::arg("tensor"));
    m.def("checksum_", &checksum_, "digest[3] (int64, CUDA) += order-independent digest of a float32 CUDA tensor.",
          pybind11::arg("values"), pybind11::arg("digest"), pybind11::arg("index_base") = 0);
    m.def("host_scratch_bytes", &host_scratch_bytes, pybind11::arg("batch_size"), pybind11::arg("max_num_targets"),
          pybind11::arg("num_classes"), pybind11::arg("height"), pybind11::arg("width"), pybind11::arg("has_labels"),
          pybind11::arg("chunk_samples") = 0, pybind11::arg("staging") = false);
    m.def("render_heatmap_from_boxes_host", &render_heatmap_from_boxes_host,
          "Host-buffer entry: stage inputs, render (K1 fused, fresh), optionally copy the heatmap / a digest back.",
          pybind11::arg("heatmap_host"), pybind11::arg("heatmap_device"), pybind11::arg("checksum_host"),
          pybind11::arg("centers"), pybind11::arg("bboxes"), pybind11::arg("out_size_factor"),
          pybind11::arg("nums_targets"), pybind11::arg("labels"), pybind11::arg("num_classes"), pybind11::arg("height"),
          pybind11::arg("width"), pybind11::arg("diameter_to_sigma_factor") = 6.f, pybind11::arg("k_scale") = 1.f,
          pybind11::arg("scratch"), pybind11::arg("chunk_samples") = 0);
    m.def("draw_gaussians", &draw_gaussians,
          "DALI draw_gaussians semantics (float radii, per-class k, active flags) on RaggedBatch-layout inputs; one launch.",
          pybind11::arg("heatmap_in"), pybind11::arg("heatmap_out"), pybind11::arg("active"), pybind11::arg("slice_ids"),
          pybind11::arg("centers"), pybind11::arg("radii"), pybind11::arg("nums_targets"), pybind11::arg("k_for_classes"),
          pybind11::arg("radius_to_sigma_factor"), pybind11::arg("fast") = -1);
    m.def("boxes_to_gaussian_params", &boxes_to_gaussian_params,
          "Boxes -> (clipped boxes, int centres, offsets, float radii, is_active, clipped hw, area fraction): the front half of the DALI heatmap converter, one launch.",
          pybind11::arg("bboxes"), pybind11::arg("centers"), pybind11::arg("classes"), pybind11::arg("is_valid"),
          pybind11::arg("nums_targets"), pybind11::arg("scale_xy"), pybind11::arg("min_sizes_hw"), pybind11::arg("num_classes"),
          pybind11::arg("heatmap_h"), pybind11::arg("heatmap_w"), pybind11::arg("radius_scaling_factor"),
          pybind11::arg("min_radius"), pybind11::arg("max_radius"), pybind11::arg("min_fraction_area"));
    m.def("gaussian_radius", &gaussian_radius, "CenterNet/CenterPoint gaussian_radius(det_size, min_overlap) -> (int32 radii, float32 radii).",
          pybind11::arg("det_hw"), pybind11::arg("min_overlap"), pybind11::arg("variant"), pybind11::arg("min_radius"));
    m.def("scale_by", &scale_by, "grad * g with g a 1-element float32 CUDA tensor (no host synchronisation), at copy rate.",
          pybind11::arg("grad"), pybind11::arg("g"), pybind11::arg("in_place") = false);
    m.def("gaussian_focal_loss", &gaussian_focal_loss,
          "Fused target rendering + Gaussian focal loss: (double[2] = {loss sum, #target==1}, d loss / d pred or undefined).",
          pybind11::arg("pred"), pybind11::arg("centers"), pybind11::arg("radii"), pybind11::arg("nums_targets"),
          pybind11::arg("labels"), pybind11::arg("diameter_to_sigma_factor") = 6.f, pybind11::arg("k_scale") = 1.f,
          pybind11::arg("alpha") = 2.0, pybind11::arg("gamma") = 4.0, pybind11::arg("eps") = 1e-12,
          pybind11::arg("want_grad") = true, pybind11::arg("fast_math") = false);
    m.def("combine_data", &combine_data,
          "List of per-sample CUDA tensors [n_i, *data] -> (zero-padded tensor [B, T, *data], int64 sample sizes [B]); one launch.",
          pybind11::arg("data_list"), pybind11::arg("max_sample_size") = -1);
    m.def("pack_radius_center", &pack_radius_center,
          "Per-sample float centres [n_i,2] / boxes [n_i,4] -> (int32 centres [B,T,2], int32 radii [B,T], int64 sizes [B]), K1 fused.",
          pybind11::arg("centers_list"), pybind11::arg("bboxes_list"), pybind11::arg("out_size_factor"),
          pybind11::arg("max_sample_size") = -1);
    m.def("launch_count", []() { return dh_launch_count(); }, "Kernels launched by this library in this process.");
    m.def("build_info", []() { return std::string(dh_build_info()); });
}
