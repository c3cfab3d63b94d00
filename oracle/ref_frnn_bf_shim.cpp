// Test infrastructure: C-callable shim around the REFERENCE's own CPU fixed-radius kNN
// (third_party/FRNN/frnn/csrc/bruteforce/bruteforce_cpu.cpp: FRNNBruteForceCPU), which is compiled
// from where it lies under /root/reference by oracle/Makefile into oracle/_ref/.  No reference source
// is copied here; this file only declares the function and forwards raw pointers to it.
#include <torch/torch.h>
#include <cstdint>
#include <cstring>
#include <tuple>

std::tuple<at::Tensor, at::Tensor> FRNNBruteForceCPU(const at::Tensor& p1, const at::Tensor& p2,
                                                     const at::Tensor& lengths1, const at::Tensor& lengths2, int K,
                                                     float r);

extern "C" int ref_frnn_bruteforce_cpu(const float* p1, const float* p2, const int64_t* len1, const int64_t* len2,
                                       int64_t N, int64_t P1, int64_t P2, int K, float r, int64_t* idx_out,
                                       float* dist_out) {
    auto f = torch::TensorOptions().dtype(torch::kFloat32);
    auto l = torch::TensorOptions().dtype(torch::kInt64);
    at::Tensor t1 = torch::from_blob(const_cast<float*>(p1), {N, P1, 3}, f);
    at::Tensor t2 = torch::from_blob(const_cast<float*>(p2), {N, P2, 3}, f);
    at::Tensor l1 = torch::from_blob(const_cast<int64_t*>(len1), {N}, l);
    at::Tensor l2 = torch::from_blob(const_cast<int64_t*>(len2), {N}, l);
    auto out = FRNNBruteForceCPU(t1, t2, l1, l2, K, r);
    at::Tensor idx = std::get<0>(out).contiguous(), d = std::get<1>(out).contiguous();
    std::memcpy(idx_out, idx.data_ptr<int64_t>(), sizeof(int64_t) * N * P1 * K);
    std::memcpy(dist_out, d.data_ptr<float>(), sizeof(float) * N * P1 * K);
    return 0;
}
