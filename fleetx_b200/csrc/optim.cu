// optim.cu — FusedAdamW step with ClipGradByGlobalNorm: ppfleetx/optims/optimizer.py:29-50
// (paddle.optimizer.AdamW with apply_decay_param_fun, multi_precision master weights) and
// pretrain_gpt_base.yaml:69-84 (beta 0.9 / 0.999, eps 1e-8, weight_decay 0.01, clip_norm 1.0).
//
// All parameters of the model are updated by TWO launches: a table of tensors and a table of fixed-size
// chunks live in device memory; one kernel accumulates the squared gradient norm over every chunk, the
// second applies clip-scale, decoupled weight decay and the Adam update per chunk. Nothing synchronises with
// the host: the clip factor is computed on the device from the norm. HBM-bound: per parameter element the
// update reads grad (2 B) + master, m, v (12 B) and writes master, m, v, param (14 B).
#include "common.cuh"
#include "tc5.cuh"

namespace fx {
namespace optim {

__global__ void __launch_bounds__(256)
sq_norm_kernel(const fx_adam_tensor* __restrict__ tensors, const int2* __restrict__ chunks, int chunk_elems,
               float* __restrict__ out_sq) {
  __shared__ float red[8];
  const int2 ck = chunks[blockIdx.x];
  const fx_adam_tensor t = tensors[ck.x];
  const int64_t lo = static_cast<int64_t>(ck.y) * chunk_elems;
  const int64_t hi = lo + chunk_elems < t.n ? lo + chunk_elems : t.n;
  float acc = 0.f;
  if (t.grad_f32) {
    const float* g = static_cast<const float*>(t.grad);
    for (int64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) acc += g[i] * g[i];
  } else {
    const __nv_bfloat16* g = static_cast<const __nv_bfloat16*>(t.grad);
    for (int64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) {
      const float x = __bfloat162float(g[i]);
      acc += x * x;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int w = 0; w < 8; ++w) s += red[w];
    atomicAdd(out_sq + (t.is_distributed ? 0 : 1), s);
  }
}

// Paddle AdamW: p *= 1 - lr * coeff (decoupled, only where apply_decay_param_fun says so), then
//   m = b1 m + (1 - b1) g,  v = b2 v + (1 - b2) g^2,  p -= lr / (1 - b1^t) * m / (sqrt(v) / sqrt(1 - b2^t) + eps)
// with g already scaled by clip_norm / max(global_norm, clip_norm) (ClipGradByGlobalNorm).
__global__ void __launch_bounds__(256)
adamw_kernel(const fx_adam_tensor* __restrict__ tensors, const int2* __restrict__ chunks, int chunk_elems, float lr,
             float beta1, float beta2, float eps, float bc1, float bc2, const float* __restrict__ sq_norm,
             float clip_norm) {
  const int2 ck = chunks[blockIdx.x];
  const fx_adam_tensor t = tensors[ck.x];
  const int64_t lo = static_cast<int64_t>(ck.y) * chunk_elems;
  const int64_t hi = lo + chunk_elems < t.n ? lo + chunk_elems : t.n;
  float gscale = 1.f;
  if (sq_norm != nullptr && clip_norm > 0.f) {
    const float gn = sqrtf(sq_norm[0] + sq_norm[1]);
    gscale = clip_norm / fmaxf(gn, clip_norm);
  }
  const float decay = 1.f - lr * t.decay;
  const float step = lr / bc1, inv_sqrt_bc2 = rsqrtf(bc2);
  for (int64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) {
    const float g = gscale * (t.grad_f32 ? static_cast<const float*>(t.grad)[i]
                                         : __bfloat162float(static_cast<const __nv_bfloat16*>(t.grad)[i]));
    float p = t.master != nullptr ? t.master[i]
                                  : (t.param_f32 ? static_cast<float*>(t.param)[i]
                                                 : __bfloat162float(static_cast<__nv_bfloat16*>(t.param)[i]));
    const float m = beta1 * t.m[i] + (1.f - beta1) * g;
    const float v = beta2 * t.v[i] + (1.f - beta2) * g * g;
    p = p * decay - step * m / (sqrtf(v) * inv_sqrt_bc2 + eps);
    t.m[i] = m;
    t.v[i] = v;
    if (t.master != nullptr) t.master[i] = p;
    if (t.param_f32) static_cast<float*>(t.param)[i] = p;
    else static_cast<__nv_bfloat16*>(t.param)[i] = __float2bfloat16_rn(p);
  }
}

}  // namespace optim
}  // namespace fx

extern "C" int fx_multi_sq_norm(void* stream, const fx_adam_tensor* tensors_dev, const int32_t* chunks_dev, int n_chunks,
                                int chunk_elems, float* out_sq2) {
  using namespace fx;
  FX_CHECK_ARG(tensors_dev && chunks_dev && out_sq2 && n_chunks > 0 && chunk_elems > 0, "fx_multi_sq_norm: bad arguments");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  FX_CUDA(cudaMemsetAsync(out_sq2, 0, 2 * sizeof(float), st));
  optim::sq_norm_kernel<<<n_chunks, 256, 0, st>>>(tensors_dev, reinterpret_cast<const int2*>(chunks_dev), chunk_elems, out_sq2);
  FX_LAUNCH_CHECK();
  return FX_OK;
}

extern "C" int fx_multi_adamw(void* stream, const fx_adam_tensor* tensors_dev, const int32_t* chunks_dev, int n_chunks,
                              int chunk_elems, float lr, float beta1, float beta2, float eps, int64_t step,
                              const float* sq_norm2, float clip_norm) {
  using namespace fx;
  FX_CHECK_ARG(tensors_dev && chunks_dev && n_chunks > 0 && chunk_elems > 0, "fx_multi_adamw: bad arguments");
  FX_CHECK_ARG(step >= 1 && beta1 >= 0.f && beta1 < 1.f && beta2 >= 0.f && beta2 < 1.f, "fx_multi_adamw: bad step / betas");
  const float bc1 = 1.f - powf(beta1, static_cast<float>(step)), bc2 = 1.f - powf(beta2, static_cast<float>(step));
  optim::adamw_kernel<<<n_chunks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      tensors_dev, reinterpret_cast<const int2*>(chunks_dev), chunk_elems, lr, beta1, beta2, eps, bc1, bc2, sq_norm2, clip_norm);
  FX_LAUNCH_CHECK();
  return FX_OK;
}
