// Test infrastructure (not product code): compiles the reference's vendored rasteriser
// (third_party/minimal_pytorch_rasterizer/minimal_pytorch_rasterizer/cuda/rasterizer_kernel.cu, from where it lies under
// the reference tree) into oracle/_ref/ref_mpr_cuda.so so that tests can run the REFERENCE kernels beside ours on the GPU box.
// The reference dispatches over float / double / half; its half instantiation does not compile against this image's torch
// headers (c10::Half comparisons are ambiguous under nvcc 12.9), and the simulator only ever calls it in float32
// (scanning_simulator/utils/minimal_rasterizer.py:41-45), so the dispatch macro is narrowed to float / double here - the
// reference sources themselves are untouched.
#include <ATen/cuda/CUDAContext.h>
#include <torch/extension.h>
#undef AT_DISPATCH_FLOATING_TYPES_AND_HALF
#define AT_DISPATCH_FLOATING_TYPES_AND_HALF AT_DISPATCH_FLOATING_TYPES
#include REF_MPR_KERNEL
