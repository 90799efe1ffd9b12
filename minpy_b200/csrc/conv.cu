// NCHW convolution and max-pooling: the MXNet `Convolution` / `Pooling` symbols the
// reference reaches through core.Function (minpy/core.py:148-190; examples/nn/cnn.py:32-40).
//
// Round-1 state: direct FP32 CUDA-core kernels (exact FP32 accumulation), shared-memory
// staging of the filter slice, coalesced along W.  They are the correctness baseline and
// the FP32-accumulate cross-check for the tcgen05 implicit-GEMM path ("next").
#include "common.cuh"

namespace mc {

constexpr int CV_THREADS = 256;

// y[n,f,ho,wo] = bias[f] + sum_{c,kh,kw} x[n,c,ho*s+kh-p,wo*s+kw-p] * w[f,c,kh,kw]
// One CTA = one (n, f) plane tile; the (C,KH,KW) filter for f sits in shared memory.
__global__ void __launch_bounds__(CV_THREADS) conv_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                              const float* __restrict__ bias, float* __restrict__ y, int C,
                                                              int H, int W, int F, int KH, int KW, int stride, int pad,
                                                              int OH, int OW, int relu) {
  mc::pdl_sync();
  extern __shared__ float wsm[];
  const int f = blockIdx.y, n = blockIdx.z;
  const int wsz = C * KH * KW;
  for (int i = threadIdx.x; i < wsz; i += CV_THREADS) wsm[i] = w[(int64_t)f * wsz + i];
  __syncthreads();
  const float b = bias ? bias[f] : 0.f;
  const float* __restrict__ xn = x + (int64_t)n * C * H * W;
  float* __restrict__ yp = y + ((int64_t)n * F + f) * OH * OW;
  for (int o = blockIdx.x * CV_THREADS + threadIdx.x; o < OH * OW; o += gridDim.x * CV_THREADS) {
    const int ho = o / OW, wo = o % OW;
    float acc = b;
    for (int c = 0; c < C; ++c) {
      const float* __restrict__ xc = xn + (int64_t)c * H * W;
      for (int kh = 0; kh < KH; ++kh) {
        const int h = ho * stride + kh - pad;
        if (h < 0 || h >= H) continue;
        for (int kw = 0; kw < KW; ++kw) {
          const int ww = wo * stride + kw - pad;
          if (ww < 0 || ww >= W) continue;
          acc = fmaf(xc[h * W + ww], wsm[(c * KH + kh) * KW + kw], acc);
        }
      }
    }
    if (relu) acc = fmaxf(acc, 0.f);
    yp[o] = acc;
  }
}

// dx[n,c,h,w] = sum_{f,kh,kw} dy[n,f,(h+p-kh)/s,(w+p-kw)/s] * w[f,c,kh,kw]   (when divisible)
__global__ void __launch_bounds__(CV_THREADS) conv_dgrad_kernel(const float* __restrict__ dy, const float* __restrict__ w,
                                                                float* __restrict__ dx, int C, int H, int W, int F, int KH,
                                                                int KW, int stride, int pad, int OH, int OW) {
  mc::pdl_sync();
  extern __shared__ float wsm[];  // w[:, c, :, :] -> (F, KH, KW)
  const int c = blockIdx.y, n = blockIdx.z;
  const int ksz = KH * KW;
  for (int i = threadIdx.x; i < F * ksz; i += CV_THREADS) {
    const int f = i / ksz, k = i % ksz;
    wsm[i] = w[((int64_t)f * C + c) * ksz + k];
  }
  __syncthreads();
  const float* __restrict__ dyn = dy + (int64_t)n * F * OH * OW;
  float* __restrict__ dxp = dx + ((int64_t)n * C + c) * H * W;
  for (int o = blockIdx.x * CV_THREADS + threadIdx.x; o < H * W; o += gridDim.x * CV_THREADS) {
    const int h = o / W, ww = o % W;
    float acc = 0.f;
    for (int f = 0; f < F; ++f) {
      const float* __restrict__ dyf = dyn + (int64_t)f * OH * OW;
      for (int kh = 0; kh < KH; ++kh) {
        const int hs = h + pad - kh;
        if (hs < 0 || hs % stride) continue;
        const int ho = hs / stride;
        if (ho >= OH) continue;
        for (int kw = 0; kw < KW; ++kw) {
          const int ws_ = ww + pad - kw;
          if (ws_ < 0 || ws_ % stride) continue;
          const int wo = ws_ / stride;
          if (wo >= OW) continue;
          acc = fmaf(dyf[ho * OW + wo], wsm[(f * KH + kh) * KW + kw], acc);
        }
      }
    }
    dxp[o] = acc;
  }
}

// dw[f,c,kh,kw] = sum_{n,ho,wo} dy[n,f,ho,wo] * x[n,c,ho*s+kh-p,wo*s+kw-p]
// One CTA per (f, c, kh, kw); the batch is split over grid.y into a workspace.
__global__ void __launch_bounds__(CV_THREADS) conv_wgrad_kernel(const float* __restrict__ x, const float* __restrict__ dy,
                                                                float* __restrict__ part, int N, int C, int H, int W, int F,
                                                                int KH, int KW, int stride, int pad, int OH, int OW,
                                                                int n_per_split) {
  mc::pdl_sync();
  __shared__ float smem[32];
  const int idx = blockIdx.x;  // ((f*C + c)*KH + kh)*KW + kw
  const int kw = idx % KW, kh = (idx / KW) % KH, c = (idx / (KW * KH)) % C, f = idx / (KW * KH * C);
  const int n0 = blockIdx.y * n_per_split, n1 = min(N, n0 + n_per_split);
  const int plane = OH * OW;
  float acc = 0.f;
  for (int n = n0; n < n1; ++n) {
    const float* __restrict__ dyp = dy + ((int64_t)n * F + f) * plane;
    const float* __restrict__ xp = x + ((int64_t)n * C + c) * H * W;
    for (int o = threadIdx.x; o < plane; o += CV_THREADS) {
      const int ho = o / OW, wo = o % OW;
      const int h = ho * stride + kh - pad, ww = wo * stride + kw - pad;
      if (h >= 0 && h < H && ww >= 0 && ww < W) acc = fmaf(dyp[o], xp[h * W + ww], acc);
    }
  }
  acc = block_reduce(acc, SumOp(), smem);
  if (threadIdx.x == 0) part[(int64_t)blockIdx.y * gridDim.x + idx] = acc;
}

__global__ void __launch_bounds__(CV_THREADS) maxpool_fwd_kernel(const float* __restrict__ x, float* __restrict__ y,
                                                                 int32_t* __restrict__ amax, int64_t NC, int H, int W, int KH,
                                                                 int KW, int stride, int OH, int OW) {
  mc::pdl_sync();
  const int64_t n = NC * OH * OW;
  const int64_t gs = (int64_t)gridDim.x * CV_THREADS;
  for (int64_t t = (int64_t)blockIdx.x * CV_THREADS + threadIdx.x; t < n; t += gs) {
    const int wo = (int)(t % OW), ho = (int)((t / OW) % OH);
    const int64_t p = t / ((int64_t)OW * OH);
    const float* __restrict__ xp = x + p * H * W;
    float best = -INFINITY;
    int bi = -1;
    for (int kh = 0; kh < KH; ++kh) {
      const int h = ho * stride + kh;
      if (h >= H) break;
      for (int kw = 0; kw < KW; ++kw) {
        const int ww = wo * stride + kw;
        if (ww >= W) break;
        const float v = xp[h * W + ww];
        if (bi < 0 || v > best || (v != v && best == best)) { best = v; bi = h * W + ww; }
      }
    }
    y[t] = best;
    if (amax) amax[t] = bi;
  }
}

__global__ void __launch_bounds__(CV_THREADS) maxpool_bwd_kernel(const float* __restrict__ dy, const int32_t* __restrict__ amax,
                                                                 float* __restrict__ dx, int64_t NC, int H, int W, int OH,
                                                                 int OW) {
  mc::pdl_sync();
  const int64_t n = NC * OH * OW;
  const int64_t gs = (int64_t)gridDim.x * CV_THREADS;
  for (int64_t t = (int64_t)blockIdx.x * CV_THREADS + threadIdx.x; t < n; t += gs) {
    const int64_t p = t / ((int64_t)OW * OH);
    atomicAdd(dx + p * H * W + amax[t], dy[t]);  // windows may overlap when stride < kernel
  }
}

// 'valid' average pooling over NCHW planes (MXNet Pooling pool_type='avg': the mean over the kh x kw window).
__global__ void __launch_bounds__(CV_THREADS) avgpool_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, int64_t NC,
                                                                 int H, int W, int KH, int KW, int stride, int OH, int OW) {
  mc::pdl_sync();
  const int64_t n = NC * OH * OW;
  const int64_t gs = (int64_t)gridDim.x * CV_THREADS;
  const float inv = 1.0f / (float)(KH * KW);
  for (int64_t t = (int64_t)blockIdx.x * CV_THREADS + threadIdx.x; t < n; t += gs) {
    const int wo = (int)(t % OW), ho = (int)((t / OW) % OH);
    const float* __restrict__ xp = x + (t / ((int64_t)OW * OH)) * H * W + (int64_t)ho * stride * W + wo * stride;
    float acc = 0.f;
    for (int kh = 0; kh < KH; ++kh)
      for (int kw = 0; kw < KW; ++kw) acc += xp[kh * W + kw];
    y[t] = acc * inv;
  }
}
// dx[h, w] = (1 / (kh kw)) * sum of dy over the windows that contain (h, w): a gather in window order
// (deterministic; overlapping windows need no atomics).
__global__ void __launch_bounds__(CV_THREADS) avgpool_bwd_kernel(const float* __restrict__ dy, float* __restrict__ dx, int64_t NC,
                                                                 int H, int W, int KH, int KW, int stride, int OH, int OW) {
  mc::pdl_sync();
  const int64_t n = NC * H * W;
  const int64_t gs = (int64_t)gridDim.x * CV_THREADS;
  const float inv = 1.0f / (float)(KH * KW);
  for (int64_t t = (int64_t)blockIdx.x * CV_THREADS + threadIdx.x; t < n; t += gs) {
    const int w = (int)(t % W), h = (int)((t / W) % H);
    const float* __restrict__ gp = dy + (t / ((int64_t)W * H)) * OH * OW;
    // windows ho with ho*stride <= h <= ho*stride + KH - 1
    const int ho_lo = h - KH + 1 > 0 ? (h - KH + 1 + stride - 1) / stride : 0, ho_hi = min(OH - 1, h / stride);
    const int wo_lo = w - KW + 1 > 0 ? (w - KW + 1 + stride - 1) / stride : 0, wo_hi = min(OW - 1, w / stride);
    float acc = 0.f;
    for (int ho = ho_lo; ho <= ho_hi; ++ho)
      for (int wo = wo_lo; wo <= wo_hi; ++wo) acc += gp[ho * OW + wo];
    dx[t] = acc * inv;
  }
}

int reduce_sum_internal(const float* x, float* out, int64_t outer, int64_t red, int64_t inner);

static int round4(int64_t v) { return (int)((v + 3) / 4 * 4); }

// Scratch device buffer freed (stream-ordered) when it goes out of scope.
struct Scratch {
  float* p = nullptr;
  int alloc(size_t n) {
    cudaError_t e = cudaMallocAsync((void**)&p, sizeof(float) * (n ? n : 1), stream());
    if (e != cudaSuccess) { set_error("cudaMallocAsync(conv scratch): %s", cudaGetErrorString(e)); p = nullptr; return 1; }
    return 0;
  }
  ~Scratch() { if (p) cudaFreeAsync(p, stream()); }
};

// The tensor-core path works on channels-last tensors (conv_tc.cu).  Stride-1 convolutions on
// NCHW tensors are served by converting at the boundary; callers that keep activations
// channels-last (the symbol executor) use the mc_conv2d_nhwc_* entry points directly.
static bool tc_eligible(int64_t stride, int64_t C, int64_t KH, int64_t KW) {
  return stride == 1 && round4(C) <= 256 && KH * KW <= 255;
}

static thread_local int g_conv_last_path = -1;

}  // namespace mc

using namespace mc;

extern "C" {

int mc_conv2d_fwd(const float* x, const float* w, const float* bias, float* y, int64_t N, int64_t C, int64_t H, int64_t W,
                  int64_t F, int64_t KH, int64_t KW, int64_t stride, int64_t pad, int relu, int precision) {
  MC_REQUIRE(stride >= 1 && pad >= 0, "bad stride/pad");
  const int64_t OH = (H + 2 * pad - KH) / stride + 1, OW = (W + 2 * pad - KW) / stride + 1;
  MC_REQUIRE(OH > 0 && OW > 0, "kernel larger than padded input");
  if (precision == MC_GEMM_TF32 && tc_eligible(stride, C, KH, KW) && N * C * H * W > 0) {
    g_conv_last_path = 1;
    const int Cp = round4(C), Fp = round4(F);
    const int64_t wpad = (KW * Cp <= 32) ? pad : 0;
    Scratch xh, yh;
    if (xh.alloc((size_t)N * H * (W + 2 * wpad) * Cp) || yh.alloc((size_t)N * OH * OW * Fp)) return 1;
    if (int rc = mc_nchw_to_nhwc(x, xh.p, N, C, H, W, Cp, wpad)) return rc;
    if (int rc = mc_conv2d_nhwc_fwd(xh.p, w, bias, yh.p, N, C, H, W, Cp, wpad, F, Fp, KH, KW, pad, relu)) return rc;
    return mc_nhwc_to_nchw(yh.p, y, N, F, OH, OW, Fp);
  }
  g_conv_last_path = 0;
  MC_REQUIRE(F <= 65535 && N <= 65535, "F and N must be <= 65535");
  const size_t smem = sizeof(float) * C * KH * KW;
  MC_REQUIRE(smem <= 48 * 1024, "filter slice exceeds 48 KB of shared memory");
  dim3 grid((unsigned)((OH * OW + CV_THREADS - 1) / CV_THREADS), (unsigned)F, (unsigned)N);
  mc::launch(conv_fwd_kernel, grid, CV_THREADS, smem, x, w, bias, y, (int)C, (int)H, (int)W, (int)F, (int)KH, (int)KW,
                                                        (int)stride, (int)pad, (int)OH, (int)OW, relu);
  MC_LAUNCHED();
  return 0;
}

int mc_conv2d_dgrad(const float* dy, const float* w, float* dx, int64_t N, int64_t C, int64_t H, int64_t W, int64_t F,
                    int64_t KH, int64_t KW, int64_t stride, int64_t pad, int precision) {
  const int64_t OH = (H + 2 * pad - KH) / stride + 1, OW = (W + 2 * pad - KW) / stride + 1;
  MC_REQUIRE(OH > 0 && OW > 0, "kernel larger than padded input");
  // data gradient = stride-1 correlation of dy with the flipped filter, pad' = K - 1 - pad
  if (precision == MC_GEMM_TF32 && KH == KW && KH - 1 - pad >= 0 && tc_eligible(stride, F, KH, KW) && N * C * H * W > 0) {
    g_conv_last_path = 1;
    const int Cp = round4(C), Fp = round4(F);
    Scratch dyh, dxh;
    if (dyh.alloc((size_t)N * OH * OW * Fp) || dxh.alloc((size_t)N * H * W * Cp)) return 1;
    if (int rc = mc_nchw_to_nhwc(dy, dyh.p, N, F, OH, OW, Fp, 0)) return rc;
    if (int rc = mc_conv2d_nhwc_dgrad(dyh.p, w, dxh.p, N, C, H, W, Cp, F, Fp, KH, KW, pad)) return rc;
    return mc_nhwc_to_nchw(dxh.p, dx, N, C, H, W, Cp);
  }
  g_conv_last_path = 0;
  MC_REQUIRE(C <= 65535 && N <= 65535, "C and N must be <= 65535");
  const size_t smem = sizeof(float) * F * KH * KW;
  MC_REQUIRE(smem <= 48 * 1024, "filter slice exceeds 48 KB of shared memory");
  dim3 grid((unsigned)((H * W + CV_THREADS - 1) / CV_THREADS), (unsigned)C, (unsigned)N);
  mc::launch(conv_dgrad_kernel, grid, CV_THREADS, smem, dy, w, dx, (int)C, (int)H, (int)W, (int)F, (int)KH, (int)KW,
                                                          (int)stride, (int)pad, (int)OH, (int)OW);
  MC_LAUNCHED();
  return 0;
}

int mc_conv2d_wgrad(const float* x, const float* dy, float* dw, float* dbias, int64_t N, int64_t C, int64_t H, int64_t W,
                    int64_t F, int64_t KH, int64_t KW, int64_t stride, int64_t pad, int precision) {
  const int64_t OH = (H + 2 * pad - KH) / stride + 1, OW = (W + 2 * pad - KW) / stride + 1;
  MC_REQUIRE(OH > 0 && OW > 0, "kernel larger than padded input");
  const int64_t nw = F * C * KH * KW;
  bool dw_done = false;
  if (dw && precision == MC_GEMM_TF32 && tc_eligible(stride, C, KH, KW) && N * C * H * W > 0) {
    g_conv_last_path = 1;
    const int Cp = round4(C), Fp = round4(F);
    const int64_t wpad = (KW * Cp <= 32) ? pad : 0;
    Scratch xh, dyh;
    if (xh.alloc((size_t)N * H * (W + 2 * wpad) * Cp) || dyh.alloc((size_t)N * OH * OW * Fp)) return 1;
    if (int rc = mc_nchw_to_nhwc(x, xh.p, N, C, H, W, Cp, wpad)) return rc;
    if (int rc = mc_nchw_to_nhwc(dy, dyh.p, N, F, OH, OW, Fp, 0)) return rc;
    if (int rc = mc_conv2d_nhwc_wgrad(xh.p, dyh.p, dw, dbias, N, C, H, W, Cp, wpad, F, Fp, KH, KW, pad)) return rc;
    dw_done = true;
    dbias = nullptr;   // computed by the same kernel
  } else if (dw) {
    g_conv_last_path = 0;
  }
  if (dw && !dw_done) {
    int splits = 1;
    if (nw < 4 * (int64_t)sm_count()) {
      splits = (int)((4 * (int64_t)sm_count() + nw - 1) / nw);
      if (splits > N) splits = (int)N;
    }
    int n_per = (int)((N + splits - 1) / splits);
    splits = (int)((N + n_per - 1) / n_per);
    float* part = dw;
    if (splits > 1) MC_CHECK(cudaMallocAsync((void**)&part, sizeof(float) * splits * nw, stream()));
    dim3 grid((unsigned)nw, (unsigned)splits);
    mc::launch(conv_wgrad_kernel, grid, CV_THREADS, 0, x, dy, part, (int)N, (int)C, (int)H, (int)W, (int)F, (int)KH, (int)KW,
                                                         (int)stride, (int)pad, (int)OH, (int)OW, n_per);
    MC_LAUNCHED();
    if (splits > 1) {
      int rc = reduce_sum_internal(part, dw, 1, splits, nw);
      MC_CHECK(cudaFreeAsync(part, stream()));
      if (rc) return rc;
    }
  }
  if (dbias) {
    // sum over (n, ho, wo): rows of the (N*F, OH*OW) view, then over n
    float* tmp = nullptr;
    MC_CHECK(cudaMallocAsync((void**)&tmp, sizeof(float) * N * F, stream()));
    int rc = reduce_sum_internal(dy, tmp, N * F, OH * OW, 1);
    if (rc == 0) rc = reduce_sum_internal(tmp, dbias, 1, N, F);
    MC_CHECK(cudaFreeAsync(tmp, stream()));
    if (rc) return rc;
  }
  return 0;
}

int mc_conv_last_path(int* path) {
  *path = g_conv_last_path;
  return 0;
}

int mc_maxpool2d_fwd(const float* x, float* y, int32_t* argmax, int64_t NC, int64_t H, int64_t W, int64_t KH, int64_t KW,
                     int64_t stride) {
  const int64_t OH = (H - KH) / stride + 1, OW = (W - KW) / stride + 1;
  MC_REQUIRE(OH > 0 && OW > 0, "pool kernel larger than input");
  const int64_t n = NC * OH * OW;
  if (n <= 0) return 0;
  mc::launch(maxpool_fwd_kernel, grid_for(n, CV_THREADS), CV_THREADS, 0, x, y, argmax, NC, (int)H, (int)W, (int)KH, (int)KW,
                                                                        (int)stride, (int)OH, (int)OW);
  MC_LAUNCHED();
  return 0;
}

int mc_avgpool2d_fwd(const float* x, float* y, int64_t NC, int64_t H, int64_t W, int64_t KH, int64_t KW, int64_t stride) {
  const int64_t OH = (H - KH) / stride + 1, OW = (W - KW) / stride + 1;
  MC_REQUIRE(OH > 0 && OW > 0 && stride > 0, "pool kernel larger than input");
  const int64_t n = NC * OH * OW;
  if (n <= 0) return 0;
  mc::launch(avgpool_fwd_kernel, grid_for(n, CV_THREADS), CV_THREADS, 0, x, y, NC, (int)H, (int)W, (int)KH, (int)KW, (int)stride,
             (int)OH, (int)OW);
  MC_LAUNCHED();
  return 0;
}

int mc_avgpool2d_bwd(const float* dy, float* dx, int64_t NC, int64_t H, int64_t W, int64_t KH, int64_t KW, int64_t stride) {
  const int64_t OH = (H - KH) / stride + 1, OW = (W - KW) / stride + 1;
  MC_REQUIRE(OH > 0 && OW > 0 && stride > 0, "pool kernel larger than input");
  const int64_t n = NC * H * W;
  if (n <= 0) return 0;
  mc::launch(avgpool_bwd_kernel, grid_for(n, CV_THREADS), CV_THREADS, 0, dy, dx, NC, (int)H, (int)W, (int)KH, (int)KW, (int)stride,
             (int)OH, (int)OW);
  MC_LAUNCHED();
  return 0;
}

int mc_maxpool2d_bwd(const float* dy, const int32_t* argmax, float* dx, int64_t NC, int64_t H, int64_t W, int64_t OH,
                     int64_t OW) {
  const int64_t n = NC * OH * OW;
  if (n <= 0) return 0;
  mc::launch(maxpool_bwd_kernel, grid_for(n, CV_THREADS), CV_THREADS, 0, dy, argmax, dx, NC, (int)H, (int)W, (int)OH, (int)OW);
  MC_LAUNCHED();
  return 0;
}

}  // extern "C"
