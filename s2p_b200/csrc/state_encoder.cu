// State encoder: nerf-style positional encoding and the
// fp32 Linear(+LeakyReLU) layers of the StyleGAN-style mapping MLP and of the
// per-MAT state heads (all heads of the generator are batched into ONE linear by the
// host module).  Tiny next to the convs (0.006 GMAC of 56.4 per frame), so plain
// fp32 CUDA-core SGEMM keeps the state path exact; latency at small batch is hidden
// by capturing the whole forward in a CUDA graph.
#include <cuda_fp16.h>

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
__host__ __device__ static inline int linear_splits(int K) { return K >= 1024 ? (K / 512 < 16 ? K / 512 : 16) : (K >= 256 ? 4 : 1); }

// ---------------------------------------------------------------------------------------------------------------
// The whole state path in ONE launch (small batches: the configs[0]/[1] batch-16 latency case, where the nine
// GEMM launches + split-K finalizes above are pure launch / dependency latency): positional encoding -> the
// mapping MLP -> all MAT state heads, as a persistent cooperative kernel.  A phase = one Linear layer; between
// phases the grid meets at a global-memory barrier (one arrival counter, spun on with ld.acquire), activations
// ping-pong through an L2-resident workspace, and every CTA fetches the weight tile of its first task of the
// NEXT phase before it waits at the barrier (weights do not depend on it).  A task = 16 batch rows x NC output
// columns (NC = 4 for the 512-wide MLP layers -> 128 tasks per layer, one per SM at batch 16; NC = 16 for the
// wide head layer).
//
// Bit-compatible with linear_kernel + linear_splitk_finalize_kernel by construction: an output is the same
// `linear_splits(K)` partial sums over the same k slices, each one serial ascending chain of fmaf starting at 0,
// combined in the same order, then + bias, then the activation.  So the generator's bytes do not depend on which
// of the two state paths a batch size selects (shard / micro-batch invariance, SURVEY.md section 8 e).
namespace se {

constexpr int kRows = 16;         // batch rows per task
constexpr int kThreads = 256;     // 4 k-splits x 16 rows x 4 column slots
constexpr int kMaxPhases = 17;    // mapping layers + the head layer
constexpr int kMaxK = 512;
constexpr int kPitch = kMaxK + 4; // tile row pitch in floats: rows r, r+1 land 4 banks apart (8 rows conflict-free)

struct Params {
  const float* state; int B, S, L;       // L = n_freqs
  const float* w[kMaxPhases]; const float* b[kMaxPhases];
  int K[kMaxPhases], N[kMaxPhases], act[kMaxPhases];
  int n_phases;
  float* buf[2]; int width;              // ping-pong activations [B][width]
  float* out; int out_pitch;             // last phase's output
  unsigned* sync;                        // [0] barrier arrivals, [1] CTAs done (both zero between launches)
};

__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float ld_cg(const float* p) {   // L2 only: written by other SMs earlier in this launch
  float v;
  asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(v) : "l"(p));
  return v;
}

// element k of the positional encoding of s (same expressions as pos_encode_kernel: bit-identical)
__device__ __forceinline__ float pe_element(const float* __restrict__ srow, int S, int k) {
  if (k < S) return srow[k];
  const int j = (k - S) / S, i = (k - S) - j * S;
  const float a = srow[i] * static_cast<float>(1u << (j >> 1));   // exact: power-of-two frequency
  return (j & 1) ? cosf(a) : sinf(a);
}

template <int CPT>   // output columns per thread; a task covers NC = 4 * CPT columns
__device__ __forceinline__ void run_task(const Params& p, int phase, int rb, int cg, float* a_s, float* w_s, float* part,
                                         bool w_loaded) {
  constexpr int NC = 4 * CPT;
  const int t = threadIdx.x;
  const int K = p.K[phase], N = p.N[phase];
  const int splits = linear_splits(K);
  const int kps = splits == 1 ? K : (((K + splits - 1) / splits) + 15) / 16 * 16;
  const int row0 = rb * kRows, col0 = cg * NC;
  if (!w_loaded) {
    for (int i = t; i < NC * K; i += kThreads) {
      const int c = i / K, k = i - c * K;
      w_s[c * kPitch + k] = (col0 + c < N) ? __ldg(p.w[phase] + static_cast<size_t>(col0 + c) * K + k) : 0.f;
    }
  }
  if (phase == 0) {
    for (int i = t; i < kRows * K; i += kThreads) {
      const int r = i / K, k = i - r * K;
      a_s[r * kPitch + k] = (row0 + r < p.B) ? pe_element(p.state + static_cast<size_t>(row0 + r) * p.S, p.S, k) : 0.f;
    }
  } else {
    const float* in = p.buf[(phase - 1) & 1];
    for (int i = t; i < kRows * K; i += kThreads) {
      const int r = i / K, k = i - r * K;
      a_s[r * kPitch + k] = (row0 + r < p.B) ? ld_cg(in + static_cast<size_t>(row0 + r) * p.width + k) : 0.f;
    }
  }
  __syncthreads();
  const int split = t >> 6, row = (t >> 2) & 15, q = t & 3;
  float acc[CPT];
#pragma unroll
  for (int j = 0; j < CPT; ++j) acc[j] = 0.f;
  if (split < splits) {
    const int k_begin = split * kps, k_end = min(K, k_begin + kps);
    const float* ar = a_s + row * kPitch;
#pragma unroll 8
    for (int k = k_begin; k < k_end; ++k) {
      const float a = ar[k];
#pragma unroll
      for (int j = 0; j < CPT; ++j) acc[j] = fmaf(a, w_s[(q + 4 * j) * kPitch + k], acc[j]);
    }
  }
#pragma unroll
  for (int j = 0; j < CPT; ++j) part[(split * kRows + row) * NC + q + 4 * j] = acc[j];
  __syncthreads();
  if (t < kRows * NC) {
    const int r = t / NC, c = t - r * NC;
    if (row0 + r < p.B && col0 + c < N) {
      float v;
      if (splits == 1) {
        v = part[r * NC + c];
      } else {
        v = 0.f;
        for (int z = 0; z < splits; ++z) v += part[(z * kRows + r) * NC + c];
      }
      v += p.b[phase] ? __ldg(p.b[phase] + col0 + c) : 0.f;
      if (p.act[phase] == 2) v = v > 0.f ? v : 0.2f * v;
      else if (p.act[phase] == 1) v = fmaxf(v, 0.f);
      if (phase == p.n_phases - 1) p.out[static_cast<size_t>(row0 + r) * p.out_pitch + col0 + c] = v;
      else p.buf[phase & 1][static_cast<size_t>(row0 + r) * p.width + col0 + c] = v;
    }
  }
  __syncthreads();
}

__global__ void __launch_bounds__(kThreads, 1) state_encoder_fused_kernel(const Params p) {
  extern __shared__ __align__(16) float sm[];
  float* a_s = sm;                           // [16][kPitch]
  float* w_s = a_s + kRows * kPitch;         // [16][kPitch]
  float* part = w_s + 16 * kPitch;           // [4][16][16]
  const int t = threadIdx.x;
  const unsigned G = gridDim.x;
  const int rbs = (p.B + kRows - 1) / kRows;
  bool w_loaded = false;
  for (int phase = 0; phase < p.n_phases; ++phase) {
    const bool head = phase == p.n_phases - 1;
    const int NC = head ? 16 : 4;
    const int cgs = (p.N[phase] + NC - 1) / NC;
    const int ntasks = rbs * cgs;
    bool first = true;
    if (phase > 0) {
      // weights of this phase's first task were fetched before the barrier (see below); now wait for every CTA
      // to have finished the previous phase, then read its activations.  EVERY CTA waits here, also one without a
      // task in this phase (B <= 16: 128 tasks per MLP layer on 148 CTAs): the counter is monotonic, so a CTA that
      // ran ahead and arrived for later phases would release this barrier before the previous phase is complete
      // (seen as a rare mismatch on the first, cold-cache launch of a process).
      if (t == 0) {
        const unsigned target = static_cast<unsigned>(phase) * G;
        while (ld_acquire_u32(p.sync) < target) { }
      }
      __syncthreads();
    }
    for (int task = blockIdx.x; task < ntasks; task += G) {
      const int rb = task / cgs, cg = task - rb * cgs;
      if (head) run_task<4>(p, phase, rb, cg, a_s, w_s, part, first && w_loaded);
      else run_task<1>(p, phase, rb, cg, a_s, w_s, part, first && w_loaded);
      first = false;
    }
    w_loaded = false;
    if (phase + 1 < p.n_phases) {
      // arrive: this CTA's outputs of `phase` are complete and visible
      __syncthreads();
      if (t == 0) {
        __threadfence();
        atomicAdd(p.sync, 1u);
      }
      // prefetch the weight tile of this CTA's first task of the next phase while the others finish
      const int nph = phase + 1;
      const bool nhead = nph == p.n_phases - 1;
      const int nNC = nhead ? 16 : 4;
      const int ncgs = (p.N[nph] + nNC - 1) / nNC;
      if (static_cast<int>(blockIdx.x) < rbs * ncgs) {
        const int cg = static_cast<int>(blockIdx.x) % ncgs;
        const int K = p.K[nph], N = p.N[nph], col0 = cg * nNC;
        for (int i = t; i < nNC * K; i += kThreads) {
          const int c = i / K, k = i - c * K;
          w_s[c * kPitch + k] = (col0 + c < N) ? __ldg(p.w[nph] + static_cast<size_t>(col0 + c) * K + k) : 0.f;
        }
        w_loaded = true;
      }
    }
  }
  // leave the counters zeroed for the next launch: the last CTA to finish resets them (every CTA has passed its
  // last barrier wait by the time it gets here)
  __syncthreads();
  if (t == 0) {
    const unsigned done = atomicAdd(p.sync + 1, 1u);
    if (done == G - 1) {
      p.sync[0] = 0u;
      p.sync[1] = 0u;
      __threadfence();
    }
  }
}

constexpr size_t kSmemBytes = (2 * 16 * kPitch + 4 * 16 * 16) * sizeof(float);

}  // namespace se

// OIHW fp32 -> [Cout][R][S][Cin] bf16
__global__ void pack_conv_weight_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ dst, int Cout,
                                        int Cin, int R, int S, float scale, int f16) {
  const long long total = static_cast<long long>(Cout) * Cin * R * S;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int ci = static_cast<int>(i % Cin);
    long long t = i / Cin;
    const int s = static_cast<int>(t % S); t /= S;
    const int r = static_cast<int>(t % R);
    const int co = static_cast<int>(t / R);
    const float v = w[((static_cast<long long>(co) * Cin + ci) * R + r) * S + s] * scale;
    if (f16) reinterpret_cast<__half*>(dst)[i] = __float2half_rn(v);
    else dst[i] = __float2bfloat16_rn(v);
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
                                    int f16, void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(w_oihw && dst && Cout > 0 && Cin > 0 && R > 0 && S > 0, "pack_conv_weight: bad args");
  const long long total = static_cast<long long>(Cout) * Cin * R * S;
  const int blocks = static_cast<int>((total + 255) / 256 < 4096 ? (total + 255) / 256 : 4096);
  pack_conv_weight_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      w_oihw, static_cast<__nv_bfloat16*>(dst), Cout, Cin, R, S, scale, f16);
  count_launch();
  return check_launch("pack_conv_weight_kernel");
}

extern "C" long long s2p_state_encoder_workspace_bytes(int B, int width) {
  if (B <= 0 || width <= 0) return -1;
  return 256 + 2LL * B * width * 4;   // [sync counters, 256 B] [ping] [pong]
}

extern "C" int s2p_state_encoder_fwd(const float* state, int B, int S, int n_freqs, const float* const* layer_w,
                                     const float* const* layer_b, int n_layers, int width, const float* heads_w,
                                     const float* heads_b, int heads_n, float* gsbs, int gsbs_pitch, void* workspace,
                                     long long workspace_bytes, void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(state && layer_w && layer_b && heads_w && gsbs && workspace, "state_encoder: null pointer");
  S2P_CHECK_ARG(B > 0 && S > 0 && n_freqs >= 0 && n_freqs <= 16, "state_encoder: bad dims B=%d S=%d L=%d", B, S, n_freqs);
  S2P_CHECK_ARG(n_layers >= 1 && n_layers + 1 <= se::kMaxPhases, "state_encoder: n_layers=%d", n_layers);
  const int K0 = S * (1 + 2 * n_freqs);
  S2P_CHECK_ARG(K0 <= se::kMaxK && width > 0 && width <= se::kMaxK, "state_encoder: K0=%d width=%d exceed %d", K0, width,
                se::kMaxK);
  S2P_CHECK_ARG(heads_n > 0 && gsbs_pitch >= heads_n, "state_encoder: heads_n=%d pitch=%d", heads_n, gsbs_pitch);
  S2P_CHECK_ARG(workspace_bytes >= s2p_state_encoder_workspace_bytes(B, width) &&
                    (reinterpret_cast<uintptr_t>(workspace) & 255) == 0,
                "state_encoder: workspace of %lld bytes (256-B aligned, zero-initialised once) needed",
                s2p_state_encoder_workspace_bytes(B, width));
  se::Params p{};
  p.state = state; p.B = B; p.S = S; p.L = n_freqs;
  for (int l = 0; l < n_layers; ++l) {
    S2P_CHECK_ARG(layer_w[l] != nullptr, "state_encoder: layer %d weight is null", l);
    p.w[l] = layer_w[l]; p.b[l] = layer_b[l];
    p.K[l] = l == 0 ? K0 : width; p.N[l] = width; p.act[l] = S2P_ACT_LRELU;
  }
  p.w[n_layers] = heads_w; p.b[n_layers] = heads_b;
  p.K[n_layers] = width; p.N[n_layers] = heads_n; p.act[n_layers] = S2P_ACT_NONE;
  p.n_phases = n_layers + 1;
  p.sync = static_cast<unsigned*>(workspace);
  p.buf[0] = reinterpret_cast<float*>(static_cast<char*>(workspace) + 256);
  p.buf[1] = p.buf[0] + static_cast<size_t>(B) * width;
  p.width = width;
  p.out = gsbs; p.out_pitch = gsbs_pitch;
  static bool configured[kMaxDevices] = {};
  const int dev = current_device();
  if (!configured[dev]) {
    S2P_CHECK_CUDA(cudaFuncSetAttribute(se::state_encoder_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        static_cast<int>(se::kSmemBytes)));
    configured[dev] = true;
  }
  const int rbs = (B + se::kRows - 1) / se::kRows;
  long long max_tasks = static_cast<long long>(rbs) * ((heads_n + 15) / 16);
  const long long mlp_tasks = static_cast<long long>(rbs) * ((width + 3) / 4);
  if (mlp_tasks > max_tasks) max_tasks = mlp_tasks;
  const int sms = num_sms();
  const int grid = static_cast<int>(max_tasks < sms ? max_tasks : sms);
  // cooperative launch: the grid barrier needs every CTA resident (grid <= SM count, one CTA per SM)
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid, 1, 1);
  cfg.blockDim = dim3(se::kThreads, 1, 1);
  cfg.dynamicSmemBytes = se::kSmemBytes;
  cfg.stream = static_cast<cudaStream_t>(stream);
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeCooperative;
  attr[0].val.cooperative = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  S2P_CHECK_CUDA(cudaLaunchKernelEx(&cfg, se::state_encoder_fused_kernel, p));
  count_launch();
  return check_launch("state_encoder_fused_kernel");
}

// ---------------------------------------------------------------------------------------------------------------
// The same launch for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses
// that forbid cudaMalloc): the workspace of s2p_state_encoder_fwd lives in a file-scope device array instead of a
// caller-provided buffer.  Every GPU that loads the library holds S2P_STATE_ENCODER_STATIC_BYTES (256 B of barrier
// counters + ping/pong activations for up to 32768 floats of B * width, e.g. B = 64 at width 512) until exit.
// No counterpart in the reference.
namespace s2p {
namespace se {
constexpr long long kStaticFloats = 32768;                                  // max B * width
constexpr long long kStaticBytes = 256 + 2 * kStaticFloats * 4;             // 262 400 B
static __device__ __align__(256) unsigned char g_static_scratch[kStaticBytes];
}  // namespace se
}  // namespace s2p

extern "C" long long s2p_state_encoder_static_max_floats(void) { return s2p::se::kStaticFloats; }

extern "C" int s2p_state_encoder_fwd_static(const float* state, int B, int S, int n_freqs, const float* const* layer_w,
                                            const float* const* layer_b, int n_layers, int width, const float* heads_w,
                                            const float* heads_b, int heads_n, float* gsbs, int gsbs_pitch, void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(B > 0 && width > 0, "state_encoder_static: bad dims B=%d width=%d", B, width);
  if (static_cast<long long>(B) * width > se::kStaticFloats)
    return set_error(S2P_ERR_UNSUPPORTED,
                     "state_encoder_static: B * width = %lld floats exceeds the built-in scratch of %lld floats "
                     "(use s2p_state_encoder_fwd with a caller-provided workspace)",
                     static_cast<long long>(B) * width, se::kStaticFloats);
  // In host code the array's name is a host-side stand-in, not a device address: ask the runtime for this device's
  // copy once per device, with that device current.
  static void* scratch[kMaxDevices] = {};
  const int dev = current_device();
  if (dev < 0 || dev >= kMaxDevices) return set_error(S2P_ERR_CUDA, "state_encoder_static: bad device %d", dev);
  if (scratch[dev] == nullptr) {
    void* ptr = nullptr;
    S2P_CHECK_CUDA(cudaGetSymbolAddress(&ptr, se::g_static_scratch));
    if (ptr == nullptr) return set_error(S2P_ERR_CUDA, "state_encoder_static: cudaGetSymbolAddress returned a null address");
    scratch[dev] = ptr;
  }
  // The array is shared by every call and keeps the last call's contents.  Only the barrier counters must start at
  // zero (the activations are fully written before they are read); a finished launch leaves them zeroed, an aborted
  // one may not, so they are cleared on the stream before every launch (no allocation: a memset node / async fill).
  S2P_CHECK_CUDA(cudaMemsetAsync(scratch[dev], 0, 256, static_cast<cudaStream_t>(stream)));
  return s2p_state_encoder_fwd(state, B, S, n_freqs, layer_w, layer_b, n_layers, width, heads_w, heads_b, heads_n, gsbs,
                               gsbs_pitch, scratch[dev], se::kStaticBytes, stream);
}
