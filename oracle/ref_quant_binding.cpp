// TEST INFRASTRUCTURE (oracle/_ref build only, see oracle/build_ref.py).
// Python binding for the reference's own fuse_sym_quant launcher: the function below is DEFINED in the reference's
// third-party/QuaRot/quarot/kernels/quant.cu (compiled from where it lies); the reference binds it in
// quarot/kernels/bindings.cpp:128-146 with block size 256 together with GEMM / flashinfer sources that are not on this path.
#include <torch/extension.h>

void rowAbsMaxQuantize(at::Tensor x, at::Tensor scale, at::Tensor q, float clip_ratio, int block_size);
void rowAbsMaxQuantize_i8(at::Tensor x, at::Tensor scale, at::Tensor q, float clip_ratio, int block_size);

static void fuse_sym_quant(at::Tensor x, at::Tensor scale, at::Tensor q, double input_clip_ratio) {
  TORCH_CHECK(x.is_cuda() && scale.is_cuda() && q.is_cuda(), "fuse_sym_quant: CUDA tensors expected");
  TORCH_CHECK(x.is_contiguous() && scale.is_contiguous() && q.is_contiguous(), "fuse_sym_quant: contiguous tensors expected");
  rowAbsMaxQuantize(x, scale, q, static_cast<float>(input_clip_ratio), 256);
}

// bindings.cpp:148-164 (W4A8 quantiser), same block size
static void fuse_sym_quant_i8(at::Tensor x, at::Tensor scale, at::Tensor q, double input_clip_ratio) {
  TORCH_CHECK(x.is_cuda() && scale.is_cuda() && q.is_cuda(), "fuse_sym_quant_i8: CUDA tensors expected");
  TORCH_CHECK(x.is_contiguous() && scale.is_contiguous() && q.is_contiguous(), "fuse_sym_quant_i8: contiguous tensors expected");
  rowAbsMaxQuantize_i8(x, scale, q, static_cast<float>(input_clip_ratio), 256);
}

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.def("fuse_sym_quant", &fuse_sym_quant);
  m.def("fuse_sym_quant_i8", &fuse_sym_quant_i8);
}
