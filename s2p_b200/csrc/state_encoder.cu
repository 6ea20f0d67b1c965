// State encoder: nerf-style positional encoding and the
// fp32 Linear(+LeakyReLU) layers of the StyleGAN-style mapping MLP and of the
// per-MAT state heads (all heads of the generator are batched into ONE linear by the
// host module).  Tiny next to the convs (0.006 GMAC of 56.4 per frame), so plain
// fp32 CUDA-core SGEMM keeps the state path exact; latency at small batch is hidden
// by capturing the whole forward in a CUDA graph.
#include "ptx.cuh"
#include "s2p_internal.h"

namespace s2p {

__global__ void pos_encode_kernel(const float* __restrict__ state, float* __restrict__ pe, int pe_pitch,
                                  int B, int S, int L) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * S) return;
  const int b = idx / S, i = idx % S;
  const float s = state[idx];
  float* row = pe + static_cast<size_t>(b) * pe_pitch;
  row[i] = s;
  float f = 1.f;
  for (int k = 0; k < L; ++k) {
    const float a = s * f;                 // exact: f is a power of two
    row[S * (1 + 2 * k) + i] = sinf(a);    // accurate (non fast-math) range reduction
    row[S * (2 + 2 * k) + i] = cosf(a);
    f *= 2.f;
  }
}

// y[B,N] = act(x[B,K] @ w[N,K]^T + bias).  64x64 tile, BK=16, 256 threads, 4x4 per thread.
__global__ void __launch_bounds__(256)
linear_kernel(const float* __restrict__ x, int x_pitch, const float* __restrict__ w,
              const float* __restrict__ bias, float* __restrict__ y, int y_pitch, int B, int K, int N, int act,
              int k_per_split, float* __restrict__ partial) {
  __shared__ float xs[16][64 + 4];
  __shared__ float wsh[16][64 + 4];
  const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;
  const int m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  // split-K (gridDim.z > 1): this CTA sums k in [k_begin, k_end) into partial[z][B][N]; combined in a fixed
  // order by linear_splitk_finalize_kernel (bit-stable, and the split depends on (K, N) only, never on B)
  const int k_begin = blockIdx.z * k_per_split;
  const int k_end = min(K, k_begin + k_per_split);
  for (int k0 = k_begin; k0 < k_end; k0 += 16) {
    for (int i = threadIdx.x; i < 64 * 16; i += 256) {
      const int r = i / 16, kk = i % 16;
      const int k = k0 + kk;
      xs[kk][r] = (m0 + r < B && k < k_end) ? x[static_cast<size_t>(m0 + r) * x_pitch + k] : 0.f;
      wsh[kk][r] = (n0 + r < N && k < k_end) ? w[static_cast<size_t>(n0 + r) * K + k] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < 16; ++kk) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = xs[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = wsh[kk][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + ty * 4 + i;
    if (m >= B) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n >= N) continue;
      if (partial != nullptr) {
        partial[(static_cast<size_t>(blockIdx.z) * B + m) * N + n] = acc[i][j];
        continue;
      }
      float v = acc[i][j] + (bias ? bias[n] : 0.f);
      if (act == 2) v = v > 0.f ? v : 0.2f * v;
      else if (act == 1) v = fmaxf(v, 0.f);
      y[static_cast<size_t>(m) * y_pitch + n] = v;
    }
  }
}

__global__ void linear_splitk_finalize_kernel(const float* __restrict__ partial, int splits, const float* __restrict__ bias,
                                             float* __restrict__ y, int y_pitch, int B, int N, int act) {
  const long long total = static_cast<long long>(B) * N;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float v = 0.f;
    for (int z = 0; z < splits; ++z) v += partial[z * total + i];
    const int n = static_cast<int>(i % N);
    const long long m = i / N;
    v += bias ? bias[n] : 0.f;
    if (act == 2) v = v > 0.f ? v : 0.2f * v;
    else if (act == 1) v = fmaxf(v, 0.f);
    y[m * y_pitch + n] = v;
  }
}

// split count: a function of (K, N) only, so a row's summation order never depends on the batch
static inline int linear_splits(int K) { return K >= 1024 ? (K / 512 < 16 ? K / 512 : 16) : (K >= 256 ? 4 : 1); }

// OIHW fp32 -> [Cout][R][S][Cin] bf16
__global__ void pack_conv_weight_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ dst, int Cout,
                                        int Cin, int R, int S, float scale) {
  const long long total = static_cast<long long>(Cout) * Cin * R * S;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int ci = static_cast<int>(i % Cin);
    long long t = i / Cin;
    const int s = static_cast<int>(t % S); t /= S;
    const int r = static_cast<int>(t % R);
    const int co = static_cast<int>(t / R);
    dst[i] = __float2bfloat16_rn(w[((static_cast<long long>(co) * Cin + ci) * R + r) * S + s] * scale);
  }
}

// torch.optim.Adam update (no weight decay, no amsgrad), fp32 master parameters
__global__ void adam_step_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                                 float* __restrict__ v, long long n, float lr, float b1, float b2, float eps,
                                 float bc1, float bc2_sqrt) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float gi = g[i];
    const float mi = b1 * m[i] + (1.f - b1) * gi;
    const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    const float denom = sqrtf(vi) / bc2_sqrt + eps;
    p[i] -= (lr / bc1) * (mi / denom);
  }
}

}  // namespace s2p

extern "C" int s2p_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, long long n,
                             float lr, float beta1, float beta2, float eps, int step, void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(param && grad && exp_avg && exp_avg_sq && n > 0 && step >= 1, "adam_step: bad args");
  const float bc1 = 1.f - powf(beta1, static_cast<float>(step));
  const float bc2 = 1.f - powf(beta2, static_cast<float>(step));
  const int blocks = static_cast<int>((n + 255) / 256 < 1184 ? (n + 255) / 256 : 1184);
  adam_step_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(param, grad, exp_avg, exp_avg_sq, n, lr, beta1,
                                                                         beta2, eps, bc1, sqrtf(bc2));
  count_launch();
  return check_launch("adam_step_kernel");
}

extern "C" int s2p_pos_encode_fwd(const float* state, float* pe, int pe_pitch, int B, int S, int n_freqs,
                                  void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(state && pe, "pos_encode: null pointer");
  S2P_CHECK_ARG(B > 0 && S > 0 && n_freqs >= 0 && pe_pitch >= S * (1 + 2 * n_freqs), "pos_encode: bad dims");
  const int total = B * S;
  pos_encode_kernel<<<(total + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(state, pe, pe_pitch, B, S,
                                                                                     n_freqs);
  count_launch();
  return check_launch("pos_encode_kernel");
}

extern "C" long long s2p_linear_workspace_floats(int B, int K, int N) {
  if (B <= 0 || K <= 0 || N <= 0) return -1;
  const int splits = s2p::linear_splits(K);
  return splits > 1 ? static_cast<long long>(splits) * B * N : 0;
}

extern "C" int s2p_linear_fwd(const float* x, int x_pitch, const float* w, const float* bias, float* y,
                              int y_pitch, int B, int K, int N, int act, float* workspace, long long workspace_floats,
                              void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(x && w && y, "linear: null pointer");
  S2P_CHECK_ARG(B > 0 && K > 0 && N > 0 && x_pitch >= K && y_pitch >= N, "linear: bad dims");
  S2P_CHECK_ARG(act >= 0 && act <= 2, "linear: act=%d", act);
  const int splits = linear_splits(K);
  const long long need = s2p_linear_workspace_floats(B, K, N);
  S2P_CHECK_ARG(need == 0 || (workspace && workspace_floats >= need), "linear: workspace of %lld floats needed (K=%d)", need, K);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  dim3 grid((N + 63) / 64, (B + 63) / 64, splits);
  if (splits == 1) {
    linear_kernel<<<grid, 256, 0, st>>>(x, x_pitch, w, bias, y, y_pitch, B, K, N, act, K, nullptr);
    count_launch();
    return check_launch("linear_kernel");
  }
  const int kps = (((K + splits - 1) / splits) + 15) / 16 * 16;
  linear_kernel<<<grid, 256, 0, st>>>(x, x_pitch, w, bias, y, y_pitch, B, K, N, act, kps, workspace);
  count_launch();
  int rc = check_launch("linear_kernel");
  if (rc != S2P_OK) return rc;
  const long long total = static_cast<long long>(B) * N;
  const int blocks = static_cast<int>((total + 255) / 256 < 1184 ? (total + 255) / 256 : 1184);
  linear_splitk_finalize_kernel<<<blocks, 256, 0, st>>>(workspace, splits, bias, y, y_pitch, B, N, act);
  count_launch();
  return check_launch("linear_splitk_finalize_kernel");
}

extern "C" int s2p_pack_conv_weight(const float* w_oihw, void* dst, int Cout, int Cin, int R, int S, float scale,
                                    void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(w_oihw && dst && Cout > 0 && Cin > 0 && R > 0 && S > 0, "pack_conv_weight: bad args");
  const long long total = static_cast<long long>(Cout) * Cin * R * S;
  const int blocks = static_cast<int>((total + 255) / 256 < 4096 ? (total + 255) / 256 : 4096);
  pack_conv_weight_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      w_oihw, static_cast<__nv_bfloat16*>(dst), Cout, Cin, R, S, scale);
  count_launch();
  return check_launch("pack_conv_weight_kernel");
}
