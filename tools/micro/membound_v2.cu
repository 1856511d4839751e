// membound_v2.cu — CANDIDATE rewrites of the layer's weakest memory-bound kernels, as a self-checking
// microbenchmark. NOT part of the library and not yet run on hardware: profiles/README.md (round 1) shows
// LayerNorm backward at 0.42, dropout backward at 0.55 and the attention row-dot at 0.37 of the measured HBM
// peak; each is latency-bound by construction (one row / one vector in flight per thread, a late load behind
// a block reduction). The candidates keep more loads in flight. This program
//   * dlopens libfleetx_b200.so and runs the shipped entry point as the baseline (LayerNorm / dropout backward),
//     or a one-thread-per-row statement of the same sum (row-dot),
//   * runs the candidate on the same buffers, compares (dx / dy bit for bit, fp32 sums to 1e-4 relative),
//   * times both with CUDA events, a 256 MiB L2 flush between repetitions, and prints GB/s.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/micro/membound_v2 \
//        tools/micro/membound_v2.cu -ldl
//   tools/micro/membound_v2 fleetx_b200/lib/libfleetx_b200.so [rows=16384] [h=2048]
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#define CK(x)                                                                                   \
  do {                                                                                          \
    cudaError_t e__ = (x);                                                                      \
    if (e__ != cudaSuccess) {                                                                   \
      fprintf(stderr, "%s:%d: %s\n", __FILE__, __LINE__, cudaGetErrorString(e__));              \
      exit(2);                                                                                  \
    }                                                                                           \
  } while (0)

__device__ __forceinline__ float lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float hi(uint32_t w) { return __uint_as_float(w & 0xFFFF0000u); }
__device__ __forceinline__ uint32_t pack(float a, float b) {
  __nv_bfloat162 t = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
  f[0] = lo(u.x); f[1] = hi(u.x); f[2] = lo(u.y); f[3] = hi(u.y);
  f[4] = lo(u.z); f[5] = hi(u.z); f[6] = lo(u.w); f[7] = hi(u.w);
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  return make_uint4(pack(f[0], f[1]), pack(f[2], f[3]), pack(f[4], f[5]), pack(f[6], f[7]));
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// =====================================================================================================
// 1. LayerNorm backward, candidate: the three row-sized loads (x, dy, skip-path gradient) of row i+1 are
//    issued BEFORE row i is reduced and stored, so each CTA always has one whole row in flight behind the
//    one it computes on. Same arithmetic, same partial-sum layout (partial[cta][{dgamma, dbeta}][h]) as the
//    shipped kernel, so the shipped finish kernel can follow it.
template <int VPT, int NT>
__global__ void __launch_bounds__(NT)
ln_bwd_v2(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x, const float* __restrict__ gamma,
          const float* __restrict__ mean, const float* __restrict__ rstd, const __nv_bfloat16* __restrict__ dres,
          __nv_bfloat16* __restrict__ dx, float* __restrict__ partial, int rows, int h) {
  __shared__ float2 red[NT / 32];
  const int nvec = h >> 3;
  const float inv_h = 1.0f / static_cast<float>(h);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  float g[VPT][8], dg[VPT][8], db[VPT][8];
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int v = threadIdx.x + i * NT;
#pragma unroll
    for (int j = 0; j < 8; ++j) dg[i][j] = db[i][j] = g[i][j] = 0.f;
    if (v < nvec) {
      const float4* gp = reinterpret_cast<const float4*>(gamma + v * 8);
      const float4 g0 = __ldg(gp), g1 = __ldg(gp + 1);
      g[i][0] = g0.x; g[i][1] = g0.y; g[i][2] = g0.z; g[i][3] = g0.w;
      g[i][4] = g1.x; g[i][5] = g1.y; g[i][6] = g1.z; g[i][7] = g1.w;
    }
  }

  uint4 nx[VPT], ndy[VPT], ndr[VPT];
  float nmu = 0.f, nrs = 0.f;
  auto fetch = [&](int row) {
    const uint4* xp = reinterpret_cast<const uint4*>(x + (size_t)row * h);
    const uint4* dyp = reinterpret_cast<const uint4*>(dy + (size_t)row * h);
    const uint4* drp = reinterpret_cast<const uint4*>(dres ? dres + (size_t)row * h : nullptr);
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int v = threadIdx.x + i * NT;
      if (v < nvec) {
        nx[i] = __ldg(xp + v);
        ndy[i] = __ldg(dyp + v);
        ndr[i] = dres ? __ldg(drp + v) : make_uint4(0, 0, 0, 0);
      }
    }
    nmu = __ldg(mean + row);
    nrs = __ldg(rstd + row);
  };

  int row = blockIdx.x;
  if (row < rows) fetch(row);
  for (; row < rows; row += gridDim.x) {
    uint4 cx[VPT], cdy[VPT], cdr[VPT];
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      cx[i] = nx[i];
      cdy[i] = ndy[i];
      cdr[i] = ndr[i];
    }
    const float mu = nmu, rs = nrs;
    if (row + (int)gridDim.x < rows) fetch(row + gridDim.x);  // next row's loads are in flight from here on

    float xh[VPT][8], gy[VPT][8];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int v = threadIdx.x + i * NT;
      if (v < nvec) {
        float xv[8], dv[8];
        unpack8(cx[i], xv);
        unpack8(cdy[i], dv);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          xh[i][j] = (xv[j] - mu) * rs;
          gy[i][j] = dv[j] * g[i][j];
          s1 += gy[i][j];
          s2 += gy[i][j] * xh[i][j];
          dg[i][j] += dv[j] * xh[i][j];
          db[i][j] += dv[j];
        }
      }
    }
    s1 = warp_sum(s1);
    s2 = warp_sum(s2);
    __syncthreads();  // readers of red[] from the previous row are done
    if (lane == 0) red[warp] = make_float2(s1, s2);
    __syncthreads();
    float2 s = red[0];
#pragma unroll
    for (int i = 1; i < NT / 32; ++i) {
      s.x += red[i].x;
      s.y += red[i].y;
    }
    const float c1 = s.x * inv_h, c2 = s.y * inv_h;
    uint4* dxp = reinterpret_cast<uint4*>(dx + (size_t)row * h);
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int v = threadIdx.x + i * NT;
      if (v < nvec) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = rs * (gy[i][j] - c1 - xh[i][j] * c2);
        if (dres != nullptr) {
          float r[8];
          unpack8(cdr[i], r);
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] += r[j];
        }
        dxp[v] = pack8(o);
      }
    }
  }

  float* pg = partial + (size_t)blockIdx.x * 2 * h;
  float* pb = pg + h;
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int v = threadIdx.x + i * NT;
    if (v < nvec) {
      float4* a = reinterpret_cast<float4*>(pg + v * 8);
      float4* b = reinterpret_cast<float4*>(pb + v * 8);
      a[0] = make_float4(dg[i][0], dg[i][1], dg[i][2], dg[i][3]);
      a[1] = make_float4(dg[i][4], dg[i][5], dg[i][6], dg[i][7]);
      b[0] = make_float4(db[i][0], db[i][1], db[i][2], db[i][3]);
      b[1] = make_float4(db[i][4], db[i][5], db[i][6], db[i][7]);
    }
  }
}

__global__ void ln_finish(const float* __restrict__ partial, float* __restrict__ dgamma, float* __restrict__ dbeta,
                          int nparts, int h) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= 2 * h) return;
  float s = 0.f;
  for (int r = 0; r < nparts; ++r) s += partial[(size_t)r * 2 * h + c];
  if (c < h) dgamma[c] = s;
  else dbeta[c - h] = s;
}

// =====================================================================================================
// 2. Dropout backward (p = 0 here: the Philox draw is the same code in both versions and not what limits the
//    kernel), candidate: four independent 16-byte vectors per thread per trip instead of one.
template <int U>
__global__ void __launch_bounds__(256)
scale_copy_v2(const uint4* __restrict__ in, uint4* __restrict__ out, size_t ngroups, float scale) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t g0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; g0 + (U - 1) * stride < ngroups; g0 += U * stride) {
    uint4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) v[u] = __ldg(in + g0 + u * stride);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      float f[8];
      unpack8(v[u], f);
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] *= scale;
      out[g0 + u * stride] = pack8(f);
    }
  }
  for (; g0 < ngroups; g0 += stride) {
    float f[8];
    unpack8(__ldg(in + g0), f);
#pragma unroll
    for (int j = 0; j < 8; ++j) f[j] *= scale;
    out[g0] = pack8(f);
  }
}

// =====================================================================================================
// 3. Attention row-dot delta[b, head, row] = sum_d dO * O over one head's hd columns of a [b, s, nh*hd] pair.
//    Baseline: one thread per (row, head), hd/8 sequential 16-byte loads at a 2*hd-byte stride between lanes.
//    Candidate: hd/8 adjacent lanes share a (row, head) — a warp's load is one contiguous 512-byte run — and
//    finish with shuffles.
__global__ void __launch_bounds__(256)
rowdot_v1(const __nv_bfloat16* __restrict__ a, const __nv_bfloat16* __restrict__ o, float* __restrict__ out, size_t n_rh,
          int hd) {
  const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n_rh) return;
  const uint4* pa = reinterpret_cast<const uint4*>(a + idx * hd);
  const uint4* po = reinterpret_cast<const uint4*>(o + idx * hd);
  float acc = 0.f;
  for (int v = 0; v < hd / 8; ++v) {
    float x[8], y[8];
    unpack8(__ldg(pa + v), x);
    unpack8(__ldg(po + v), y);
#pragma unroll
    for (int j = 0; j < 8; ++j) acc += x[j] * y[j];
  }
  out[idx] = acc;
}
template <int LANES>  // LANES = hd / 8
__global__ void __launch_bounds__(256)
rowdot_v2(const uint4* __restrict__ a, const uint4* __restrict__ o, float* __restrict__ out, size_t n_vec) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t v = (size_t)blockIdx.x * blockDim.x + threadIdx.x; v < n_vec; v += stride) {  // n_vec % 32 == 0
    float x[8], y[8];
    unpack8(__ldg(a + v), x);
    unpack8(__ldg(o + v), y);
    float acc = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) acc += x[j] * y[j];
#pragma unroll
    for (int m = LANES / 2; m > 0; m >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, m);
    if ((threadIdx.x & (LANES - 1)) == 0) out[v / LANES] = acc;
  }
}

// ----------------------------------------------------------------------------------------------------- host
typedef int (*ln_bwd_fn)(void*, const void*, const void*, const float*, const float*, const float*, const void*, void*,
                         float*, float*, void*, int, int);
typedef size_t (*ln_ws_fn)(int, int);
typedef int (*bdr_bwd_fn)(void*, const void*, void*, int, int, float, uint64_t, uint64_t, const uint8_t*);

static uint32_t lcg = 12345u;
static float frand() {
  lcg = lcg * 1664525u + 1013904223u;
  return ((lcg >> 8) & 0xFFFF) / 32768.0f - 1.0f;
}
static void fill_bf16(std::vector<__nv_bfloat16>& v) {
  for (auto& e : v) e = __float2bfloat16(frand());
}

template <class F>
static float time_ms(F&& f, void* flush, int reps = 20) {
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  float total = 0.f;
  for (int i = 0; i < 3; ++i) f();
  for (int i = 0; i < reps; ++i) {
    CK(cudaMemsetAsync(flush, i, 256u << 20));
    CK(cudaEventRecord(e0));
    f();
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    total += ms;
  }
  return total / reps;
}

int main(int argc, char** argv) {
  if (argc < 2) {
    fprintf(stderr, "usage: %s path/to/libfleetx_b200.so [rows] [h]\n", argv[0]);
    return 1;
  }
  const int rows = argc > 2 ? atoi(argv[2]) : 16384, h = argc > 3 ? atoi(argv[3]) : 2048;
  void* lib = dlopen(argv[1], RTLD_NOW);
  if (!lib) {
    fprintf(stderr, "dlopen: %s\n", dlerror());
    return 1;
  }
  auto ln_bwd = (ln_bwd_fn)dlsym(lib, "fx_layernorm_bwd");
  auto ln_ws = (ln_ws_fn)dlsym(lib, "fx_layernorm_bwd_workspace_bytes");
  auto bdr_bwd = (bdr_bwd_fn)dlsym(lib, "fx_bias_dropout_residual_bwd");
  if (!ln_bwd || !ln_ws || !bdr_bwd) return 1;
  if (h % 8 || h > 4096 || h < 1024) {
    fprintf(stderr, "this harness instantiates the 128-thread kernels only: 1024 <= h <= 4096\n");
    return 1;
  }

  const size_t n = (size_t)rows * h;
  std::vector<__nv_bfloat16> hx(n), hdy(n), hdr(n);
  fill_bf16(hx), fill_bf16(hdy), fill_bf16(hdr);
  std::vector<float> hg(h), hmean(rows), hrstd(rows);
  for (auto& e : hg) e = 1.0f + 0.1f * frand();
  for (auto& e : hmean) e = 0.05f * frand();
  for (auto& e : hrstd) e = 1.5f + 0.5f * frand();
  __nv_bfloat16 *x, *dy, *dr, *dx0, *dx1;
  float *g, *mean, *rstd, *dg0, *db0, *dg1, *db1, *ws0, *ws1;
  void* flush;
  CK(cudaMalloc(&x, n * 2)); CK(cudaMalloc(&dy, n * 2)); CK(cudaMalloc(&dr, n * 2));
  CK(cudaMalloc(&dx0, n * 2)); CK(cudaMalloc(&dx1, n * 2));
  CK(cudaMalloc(&g, h * 4)); CK(cudaMalloc(&mean, rows * 4)); CK(cudaMalloc(&rstd, rows * 4));
  CK(cudaMalloc(&dg0, h * 4)); CK(cudaMalloc(&db0, h * 4)); CK(cudaMalloc(&dg1, h * 4)); CK(cudaMalloc(&db1, h * 4));
  const size_t wsb = ln_ws(rows, h);
  CK(cudaMalloc(&ws0, wsb)); CK(cudaMalloc(&ws1, wsb));
  CK(cudaMalloc(&flush, 256u << 20));
  CK(cudaMemcpy(x, hx.data(), n * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dy, hdy.data(), n * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dr, hdr.data(), n * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(g, hg.data(), h * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(mean, hmean.data(), rows * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(rstd, hrstd.data(), rows * 4, cudaMemcpyHostToDevice));
  const double peak = 6545.6;

  // ---- 1. LayerNorm backward
  const int grid = rows < 592 ? rows : 592;
  auto base_ln = [&] { if (ln_bwd(nullptr, dy, x, g, mean, rstd, dr, dx0, dg0, db0, ws0, rows, h)) exit(3); };
  auto cand_ln = [&] {
    if (h <= 1024) ln_bwd_v2<1, 128><<<grid, 128>>>(dy, x, g, mean, rstd, dr, dx1, ws1, rows, h);
    else if (h <= 2048) ln_bwd_v2<2, 128><<<grid, 128>>>(dy, x, g, mean, rstd, dr, dx1, ws1, rows, h);
    else ln_bwd_v2<2, 256><<<grid, 256>>>(dy, x, g, mean, rstd, dr, dx1, ws1, rows, h);
    ln_finish<<<(2 * h + 255) / 256, 256>>>(ws1, dg1, db1, grid, h);
  };
  base_ln();
  cand_ln();
  CK(cudaDeviceSynchronize());
  {
    std::vector<__nv_bfloat16> a(n), b(n);
    std::vector<float> ga(h), gb(h), ba(h), bb(h);
    CK(cudaMemcpy(a.data(), dx0, n * 2, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(b.data(), dx1, n * 2, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(ga.data(), dg0, h * 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(gb.data(), dg1, h * 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(ba.data(), db0, h * 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(bb.data(), db1, h * 4, cudaMemcpyDeviceToHost));
    size_t bad = 0;
    for (size_t i = 0; i < n; ++i) bad += memcmp(&a[i], &b[i], 2) != 0;
    double worst = 0, scale = 0;
    for (int i = 0; i < h; ++i) {
      worst = fmax(worst, fmax(fabs(ga[i] - gb[i]), fabs(ba[i] - bb[i])));
      scale = fmax(scale, fmax(fabs(ga[i]), fabs(ba[i])));
    }
    printf("ln_bwd: dx differs in %zu of %zu elements; dgamma/dbeta max abs diff %.3e (scale %.3e) -> %s\n", bad, n, worst,
           scale, (bad == 0 && worst <= 1e-4 * scale) ? "MATCH" : "MISMATCH");
  }
  {
    const double bytes = 8.0 * h * rows;
    const float t0 = time_ms(base_ln, flush), t1 = time_ms(cand_ln, flush);
    printf("ln_bwd   rows %d h %d: shipped %.1f us = %.0f GB/s (%.2f of peak) | candidate %.1f us = %.0f GB/s (%.2f)\n", rows, h,
           t0 * 1e3, bytes / t0 / 1e6, bytes / t0 / 1e6 / peak, t1 * 1e3, bytes / t1 / 1e6, bytes / t1 / 1e6 / peak);
  }

  // ---- 2. dropout backward at p = 0 (dy = dout * 1): pure streaming
  const size_t ngroups = n / 8;
  auto base_bdr = [&] { if (bdr_bwd(nullptr, dy, dx0, rows, h, 0.0f, 0, 0, nullptr)) exit(3); };
  auto cand_bdr = [&] {
    scale_copy_v2<4><<<148 * 8, 256>>>(reinterpret_cast<const uint4*>(dy), reinterpret_cast<uint4*>(dx1), ngroups, 1.0f);
  };
  base_bdr();
  cand_bdr();
  CK(cudaDeviceSynchronize());
  {
    std::vector<__nv_bfloat16> a(n), b(n);
    CK(cudaMemcpy(a.data(), dx0, n * 2, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(b.data(), dx1, n * 2, cudaMemcpyDeviceToHost));
    printf("bdr_bwd: %s\n", memcmp(a.data(), b.data(), n * 2) == 0 ? "MATCH" : "MISMATCH");
    const double bytes = 4.0 * h * rows;
    const float t0 = time_ms(base_bdr, flush), t1 = time_ms(cand_bdr, flush);
    printf("bdr_bwd  rows %d h %d: shipped %.1f us = %.0f GB/s (%.2f of peak) | candidate %.1f us = %.0f GB/s (%.2f)\n", rows, h,
           t0 * 1e3, bytes / t0 / 1e6, bytes / t0 / 1e6 / peak, t1 * 1e3, bytes / t1 / 1e6, bytes / t1 / 1e6 / peak);
  }

  // ---- 3. attention row-dot, head_dim 64
  {
    const int hd = 64;
    const size_t n_rh = n / hd;
    float *d0, *d1;
    CK(cudaMalloc(&d0, n_rh * 4));
    CK(cudaMalloc(&d1, n_rh * 4));
    auto base = [&] { rowdot_v1<<<(unsigned)((n_rh + 255) / 256), 256>>>(dy, x, d0, n_rh, hd); };
    auto cand = [&] {
      rowdot_v2<8><<<148 * 8, 256>>>(reinterpret_cast<const uint4*>(dy), reinterpret_cast<const uint4*>(x), d1, n / 8);
    };
    base();
    cand();
    CK(cudaDeviceSynchronize());
    std::vector<float> a(n_rh), b(n_rh);
    CK(cudaMemcpy(a.data(), d0, n_rh * 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(b.data(), d1, n_rh * 4, cudaMemcpyDeviceToHost));
    double worst = 0;
    for (size_t i = 0; i < n_rh; ++i) worst = fmax(worst, fabs(a[i] - b[i]));
    printf("rowdot: max abs diff %.3e -> %s\n", worst, worst < 1e-3 ? "MATCH" : "MISMATCH");
    const double bytes = 4.0 * h * rows;
    const float t0 = time_ms(base, flush), t1 = time_ms(cand, flush);
    printf("rowdot   rows %d h %d hd %d: one thread per row %.1f us = %.0f GB/s (%.2f of peak) | candidate %.1f us = %.0f GB/s (%.2f)\n",
           rows, h, hd, t0 * 1e3, bytes / t0 / 1e6, bytes / t0 / 1e6 / peak, t1 * 1e3, bytes / t1 / 1e6, bytes / t1 / 1e6 / peak);
  }
  return 0;
}
