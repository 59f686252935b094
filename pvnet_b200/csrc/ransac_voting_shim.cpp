// ransac_voting_shim.cpp -- the PyTorch C++ extension that REPLACES the reference's
//   lib/ransac_voting_gpu_layer/src/ransac_voting.cpp (+ ransac_voting_kernel.cu, setup.py)
// as a thin shim over the C ABI of libpvvote.so (include/pvvote.h).  It contains no kernel and no arithmetic: it
// checks tensors the way the reference's CHECK_INPUT does (ransac_voting.cpp:7-9), takes pointers, strides and the
// CURRENT stream/device from ATen, and forwards.  Module name and the two primitive signatures are the reference's
// (ransac_voting.cpp:102-107); the two fused entry points are what ransac_voting_gpu.py:520-604 / :333-406 collapse to.
#include <torch/extension.h>

#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>

#include <atomic>
#include <random>
#include <tuple>

#include "../../include/pvvote.h"

#define CHECK_CUDA(x) TORCH_CHECK((x).is_cuda(), #x " must be a CUDA tensor")
#define CHECK_CONTIGUOUS(x) TORCH_CHECK((x).is_contiguous(), #x " must be contiguous")
#define CHECK_INPUT(x) \
    CHECK_CUDA(x);     \
    CHECK_CONTIGUOUS(x)

namespace {

void check_status(int st, const char *what)
{
    TORCH_CHECK(st == PVV_OK, what, " failed with pvv_status ", st, st == PVV_ERR_CUDA ? " cudaError " : "",
                st == PVV_ERR_CUDA ? std::to_string(pvv_last_cuda_error()) : std::string());
}

// `seed=None` means a fresh draw stream per call, like the reference's .random_() (ransac_voting_gpu.py:553): a process-local
// counter, started from std::random_device, hashed (splitmix64).  The user's torch RNG state is not touched.
int64_t fresh_seed()
{
    static std::atomic<uint64_t> ctr{((uint64_t)std::random_device{}() << 32) ^ (uint64_t)std::random_device{}()};
    uint64_t z = ctr.fetch_add(0x9E3779B97F4A7C15ull) + 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return (int64_t)((z ^ (z >> 31)) >> 2);
}

void *cur_stream(const at::Tensor &t) { return (void *)c10::cuda::getCurrentCUDAStream(t.get_device()).stream(); }

int mask_dtype(const at::Tensor &m)
{
    switch (m.scalar_type()) {
        case at::kByte:
        case at::kBool: return PVV_MASK_U8;
        case at::kInt: return PVV_MASK_I32;
        case at::kLong: return PVV_MASK_I64;
        case at::kFloat: return PVV_MASK_F32;
        default: TORCH_CHECK(false, "mask dtype must be uint8/bool/int32/int64/float32");
    }
}

pvv_job make_job(const at::Tensor &mask, const at::Tensor &vertex, int64_t hn_total, double thresh, int64_t min_num,
                 int64_t max_num, int fg_rule, const c10::optional<at::Tensor> &idxs,
                 const c10::optional<at::Tensor> &selection, int64_t seed, int64_t image_id_base)
{
    CHECK_CUDA(mask);
    CHECK_CUDA(vertex);
    const bool logits = mask.dim() == 4;  // seg_pred [b,C,h,w]: argmax fused into the compaction pass
    TORCH_CHECK(vertex.dim() == 5 && vertex.size(4) == 2 && (mask.dim() == 3 || logits),
                "expected mask [b,h,w] (or seg_pred [b,C,h,w]), vertex [b,h,w,vn,2]");
    TORCH_CHECK(vertex.scalar_type() == at::kFloat, "vertex must be float32");
    TORCH_CHECK(mask.size(0) == vertex.size(0) && mask.size(-2) == vertex.size(1) && mask.size(-1) == vertex.size(2),
                "mask and vertex disagree on [b,h,w]");
    pvv_job j = {};
    j.mask = mask.data_ptr();
    j.fg_rule = fg_rule;
    if (logits) {
        TORCH_CHECK(mask.scalar_type() == at::kFloat, "seg_pred must be float32");
        j.mask_dtype = PVV_MASK_LOGITS_F32;
        j.mask_strides[0] = mask.stride(0); j.mask_strides[1] = mask.stride(2); j.mask_strides[2] = mask.stride(3);
        j.seg_classes = (int32_t)mask.size(1);
        j.seg_class_stride = mask.stride(1);
    } else {
        j.mask_dtype = mask_dtype(mask);
        for (int i = 0; i < 3; ++i) j.mask_strides[i] = mask.stride(i);
    }
    j.vertex = vertex.data_ptr<float>();
    for (int i = 0; i < 5; ++i) j.vertex_strides[i] = vertex.stride(i);
    j.b = (int32_t)vertex.size(0); j.h = (int32_t)vertex.size(1); j.w = (int32_t)vertex.size(2);
    j.vn = (int32_t)vertex.size(3);
    j.hn_total = (int32_t)hn_total;
    j.inlier_thresh = (float)thresh;
    j.min_num = (int32_t)min_num;
    j.max_num = (int32_t)std::min<int64_t>(max_num, INT32_MAX);
    if (idxs.has_value()) {
        CHECK_INPUT(*idxs);
        TORCH_CHECK(idxs->scalar_type() == at::kInt && idxs->numel() == (int64_t)j.b * j.hn_total * j.vn * 2,
                    "idxs must be int32 [b,hn,vn,2]");
        j.idxs = idxs->data_ptr<int32_t>();
    }
    if (selection.has_value()) {
        CHECK_INPUT(*selection);
        TORCH_CHECK(selection->scalar_type() == at::kFloat && selection->numel() == (int64_t)j.b * j.h * j.w,
                    "selection must be float32 [b,h,w]");
        j.selection = selection->data_ptr<float>();
    }
    j.seed = (uint64_t)seed;
    j.image_id_base = image_id_base;
    j.stream = cur_stream(mask);
    return j;
}

}  // namespace

// reference: at::Tensor generate_hypothesis(direct, coords, idxs)            ransac_voting.cpp:20-31
at::Tensor generate_hypothesis(at::Tensor direct, at::Tensor coords, at::Tensor idxs)
{
    CHECK_INPUT(direct);
    CHECK_INPUT(coords);
    CHECK_INPUT(idxs);
    TORCH_CHECK(direct.scalar_type() == at::kFloat && coords.scalar_type() == at::kFloat && idxs.scalar_type() == at::kInt,
                "direct/coords float32, idxs int32");
    const c10::cuda::CUDAGuard guard(direct.device());
    const int tn = (int)direct.size(0), vn = (int)direct.size(1), hn = (int)idxs.size(0);
    auto hyp = at::empty({hn, vn, 2}, direct.options());
    check_status(pvv_generate_hypothesis(direct.data_ptr<float>(), coords.data_ptr<float>(), idxs.data_ptr<int32_t>(), tn, vn,
                                         hn, hyp.data_ptr<float>(), cur_stream(direct)),
                 "pvv_generate_hypothesis");
    return hyp;
}

// reference: void voting_for_hypothesis(direct, coords, hypo_pts, inliers, inlier_thresh)   ransac_voting.cpp:41-55
void voting_for_hypothesis(at::Tensor direct, at::Tensor coords, at::Tensor hypo_pts, at::Tensor inliers,
                           float inlier_thresh)
{
    CHECK_INPUT(direct);
    CHECK_INPUT(coords);
    CHECK_INPUT(hypo_pts);
    CHECK_INPUT(inliers);
    TORCH_CHECK(inliers.scalar_type() == at::kByte, "inliers must be uint8");
    const c10::cuda::CUDAGuard guard(direct.device());
    const int tn = (int)direct.size(0), vn = (int)direct.size(1), hn = (int)hypo_pts.size(0);
    check_status(pvv_voting_for_hypothesis(direct.data_ptr<float>(), coords.data_ptr<float>(), hypo_pts.data_ptr<float>(),
                                           inliers.data_ptr<uint8_t>(), tn, vn, hn, inlier_thresh, cur_stream(direct)),
                 "pvv_voting_for_hypothesis");
}

// reference: at::Tensor generate_hypothesis_vanishing_point(direct, coords, idxs)       ransac_voting.cpp:64-75
at::Tensor generate_hypothesis_vanishing_point(at::Tensor direct, at::Tensor coords, at::Tensor idxs)
{
    CHECK_INPUT(direct);
    CHECK_INPUT(coords);
    CHECK_INPUT(idxs);
    TORCH_CHECK(direct.scalar_type() == at::kFloat && coords.scalar_type() == at::kFloat && idxs.scalar_type() == at::kInt,
                "direct/coords float32, idxs int32");
    const c10::cuda::CUDAGuard guard(direct.device());
    const int tn = (int)direct.size(0), vn = (int)direct.size(1), hn = (int)idxs.size(0);
    auto hyp = at::empty({hn, vn, 3}, direct.options());
    check_status(pvv_generate_hypothesis_vanishing_point(direct.data_ptr<float>(), coords.data_ptr<float>(),
                                                         idxs.data_ptr<int32_t>(), tn, vn, hn, hyp.data_ptr<float>(),
                                                         cur_stream(direct)),
                 "pvv_generate_hypothesis_vanishing_point");
    return hyp;
}

// reference: void voting_for_hypothesis_vanishing_point(direct, coords, hypo_pts, inliers, thresh)   ransac_voting.cpp:85-98
void voting_for_hypothesis_vanishing_point(at::Tensor direct, at::Tensor coords, at::Tensor hypo_pts, at::Tensor inliers,
                                           float inlier_thresh)
{
    CHECK_INPUT(direct);
    CHECK_INPUT(coords);
    CHECK_INPUT(hypo_pts);
    CHECK_INPUT(inliers);
    TORCH_CHECK(inliers.scalar_type() == at::kByte, "inliers must be uint8");
    const c10::cuda::CUDAGuard guard(direct.device());
    const int tn = (int)direct.size(0), vn = (int)direct.size(1), hn = (int)hypo_pts.size(0);
    check_status(pvv_voting_for_hypothesis_vanishing_point(direct.data_ptr<float>(), coords.data_ptr<float>(),
                                                           hypo_pts.data_ptr<float>(), inliers.data_ptr<uint8_t>(), tn, vn,
                                                           hn, inlier_thresh, cur_stream(direct)),
                 "pvv_voting_for_hypothesis_vanishing_point");
}

// ransac_voting_gpu.py:520-604 as one call
at::Tensor ransac_voting_layer_v3(at::Tensor mask, at::Tensor vertex, int64_t round_hyp_num, double inlier_thresh,
                                  double confidence, int64_t max_iter, int64_t min_num, int64_t max_num,
                                  c10::optional<at::Tensor> idxs, c10::optional<at::Tensor> selection,
                                  c10::optional<int64_t> seed, int64_t image_id_base)
{
    pvv_job j = make_job(mask, vertex, round_hyp_num, inlier_thresh, min_num, max_num, PVV_FG_BYTE, idxs, selection,
                         seed.has_value() ? *seed : fresh_seed(), image_id_base);
    const c10::cuda::CUDAGuard guard(mask.device());
    j.stream = cur_stream(mask);
    auto ws = at::empty({(int64_t)pvv_workspace_bytes(&j)}, mask.options().dtype(at::kByte));
    j.workspace = ws.data_ptr();
    j.workspace_bytes = (size_t)ws.numel();
    auto out = at::empty({j.b, j.vn, 2}, vertex.options());
    check_status(pvv_v3(&j, (float)confidence, (int)max_iter, out.data_ptr<float>(), nullptr), "pvv_v3");
    return out;
}

// ransac_voting_gpu.py:333-406 as one call
std::tuple<at::Tensor, at::Tensor> estimate_voting_distribution_with_mean(
    at::Tensor mask, at::Tensor vertex, at::Tensor mean, int64_t round_hyp_num, int64_t min_hyp_num, double inlier_thresh,
    int64_t min_num, int64_t max_num, c10::optional<at::Tensor> idxs, c10::optional<at::Tensor> selection,
    c10::optional<int64_t> seed, int64_t image_id_base)
{
    const int64_t rounds = (min_hyp_num + round_hyp_num - 1) / round_hyp_num;
    pvv_job j = make_job(mask, vertex, rounds * round_hyp_num, inlier_thresh, min_num, max_num, PVV_FG_EQ1, idxs, selection,
                         seed.has_value() ? *seed : fresh_seed(), image_id_base);
    const c10::cuda::CUDAGuard guard(mask.device());
    j.stream = cur_stream(mask);
    auto mean_c = mean.to(vertex.options()).contiguous();
    TORCH_CHECK(mean_c.numel() == (int64_t)j.b * j.vn * 2, "mean must be [b,vn,2]");
    auto ws = at::empty({(int64_t)pvv_workspace_bytes(&j)}, mask.options().dtype(at::kByte));
    j.workspace = ws.data_ptr();
    j.workspace_bytes = (size_t)ws.numel();
    auto cov = at::empty({j.b, j.vn, 2, 2}, vertex.options());
    check_status(pvv_cov(&j, mean_c.data_ptr<float>(), cov.data_ptr<float>(), nullptr), "pvv_cov");
    return std::make_tuple(mean, cov);
}

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m)
{
    m.def("generate_hypothesis", &generate_hypothesis, "generate hypothesis");
    m.def("voting_for_hypothesis", &voting_for_hypothesis, "voting for hypothesis");
    m.def("generate_hypothesis_vanishing_point", &generate_hypothesis_vanishing_point, "generate hypothesis vanishing point");
    m.def("voting_for_hypothesis_vanishing_point", &voting_for_hypothesis_vanishing_point,
          "voting for hypothesis vanishing point");
    m.def("ransac_voting_layer_v3", &ransac_voting_layer_v3, "fused v3", py::arg("mask"), py::arg("vertex"),
          py::arg("round_hyp_num"), py::arg("inlier_thresh") = 0.99, py::arg("confidence") = 0.99, py::arg("max_iter") = 100,
          py::arg("min_num") = 100, py::arg("max_num") = 30000, py::arg("idxs") = py::none(),
          py::arg("selection") = py::none(), py::arg("seed") = py::none(), py::arg("image_id_base") = 0);
    m.def("estimate_voting_distribution_with_mean", &estimate_voting_distribution_with_mean, "fused distribution",
          py::arg("mask"), py::arg("vertex"), py::arg("mean"), py::arg("round_hyp_num") = 256,
          py::arg("min_hyp_num") = 4096, py::arg("inlier_thresh") = 0.99, py::arg("min_num") = 20,
          py::arg("max_num") = 30000, py::arg("idxs") = py::none(), py::arg("selection") = py::none(),
          py::arg("seed") = py::none(), py::arg("image_id_base") = 0);
}
