// Small training-path kernels: the pieces of the conv backward that are
// not GEMMs, and the optimiser update.
//   pack_conv_weight_dgrad : OIHW fp32 -> the packed weight of the data-gradient conv
//                            (filter flipped, in/out channels swapped) in one pass
//   act_bwd_bias_grad      : dy' = dy * act'(y) (ReLU / LeakyReLU(0.2), from the activated
//                            output) and db[c] = sum_pixels dy'[pix, c], one read of dy / y,
//                            fixed-order two-stage column reduction (bit-stable)
//   adam_step (device step): torch.optim.Adam's update with the step count read from device
//                            memory, so a captured CUDA graph of the whole step stays valid
#include "ptx.cuh"
#include "s2p_internal.h"

namespace s2p {

__global__ void pack_conv_weight_dgrad_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ dst, int Cout,
                                              int Cin, int R, int S) {
  // dst[ci][r'][s'][co] = w[co][ci][R-1-r'][S-1-s']
  const long long total = static_cast<long long>(Cout) * Cin * R * S;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int co = static_cast<int>(i % Cout);
    long long t = i / Cout;
    const int s = static_cast<int>(t % S); t /= S;
    const int r = static_cast<int>(t % R);
    const int ci = static_cast<int>(t / R);
    dst[i] = __float2bfloat16_rn(w[((static_cast<long long>(co) * Cin + ci) * R + (R - 1 - r)) * S + (S - 1 - s)]);
  }
}

// Both packed forms of up to kPackMulti weight tensors in ONE launch: a training step re-packs every conv weight
// after the optimiser step, and tensor by tensor that was ~105 launches of 7-8 us each (0.8 ms of a 14 ms batch-16
// step).  The pointer table travels in kernel parameter space (graph-capturable); blocks are dealt out by tensor size.
constexpr int kPackMulti = 48;
constexpr int kPackChunk = 4096;   // elements of ONE packed form per block
struct PackMultiArgs {
  const float* src[kPackMulti];
  __nv_bfloat16* fwd[kPackMulti];      // [Cout][R][S][Cin]
  __nv_bfloat16* dgrad[kPackMulti];    // [Cin][R][S][Cout], filter flipped; nullptr = not wanted
  int cout[kPackMulti], cin[kPackMulti], r[kPackMulti], s[kPackMulti];
  int first_block[kPackMulti + 1];
  int count;
};

__global__ void __launch_bounds__(256) pack_conv_weights_multi_kernel(const __grid_constant__ PackMultiArgs a) {
  int t = 0;
  while (t + 1 < a.count && static_cast<int>(blockIdx.x) >= a.first_block[t + 1]) ++t;   // block-uniform
  const float* __restrict__ w = a.src[t];
  const int Cout = a.cout[t], Cin = a.cin[t], R = a.r[t], S = a.s[t];
  const long long total = static_cast<long long>(Cout) * Cin * R * S;
  const long long i0 = static_cast<long long>(blockIdx.x - a.first_block[t]) * kPackChunk;
  const long long i1 = (i0 + kPackChunk < total) ? i0 + kPackChunk : total;
  __nv_bfloat16* __restrict__ fwd = a.fwd[t];
  for (long long i = i0 + threadIdx.x; i < i1; i += 256) {
    const int ci = static_cast<int>(i % Cin);
    long long q = i / Cin;
    const int s = static_cast<int>(q % S); q /= S;
    const int r = static_cast<int>(q % R);
    const int co = static_cast<int>(q / R);
    fwd[i] = __float2bfloat16_rn(__ldg(w + ((static_cast<long long>(co) * Cin + ci) * R + r) * S + s));
  }
  __nv_bfloat16* __restrict__ dg = a.dgrad[t];
  if (dg == nullptr) return;
  for (long long i = i0 + threadIdx.x; i < i1; i += 256) {
    const int co = static_cast<int>(i % Cout);
    long long q = i / Cout;
    const int s = static_cast<int>(q % S); q /= S;
    const int r = static_cast<int>(q % R);
    const int ci = static_cast<int>(q / R);
    dg[i] = __float2bfloat16_rn(__ldg(w + ((static_cast<long long>(co) * Cin + ci) * R + (R - 1 - r)) * S + (S - 1 - s)));
  }
}

__device__ __forceinline__ void unpack8f(const uint4& u, float (&f)[8]) {
  f[0] = bf16_lo(u.x); f[1] = bf16_hi(u.x); f[2] = bf16_lo(u.y); f[3] = bf16_hi(u.y);
  f[4] = bf16_lo(u.z); f[5] = bf16_hi(u.z); f[6] = bf16_lo(u.w); f[7] = bf16_hi(u.w);
}

// blockDim = (C/8, PY); grid = chunks of rows.  partial[chunk][C]
__global__ void __launch_bounds__(256)
act_bwd_colsum_kernel(const __nv_bfloat16* __restrict__ dy, int dy_pitch, const __nv_bfloat16* __restrict__ y,
                      int y_pitch, int act, __nv_bfloat16* __restrict__ out, int out_pitch,
                      float* __restrict__ partial, long long rows, int C, long long rows_per_cta) {
  extern __shared__ float red[];   // [PY][C]
  const int c0 = threadIdx.x * 8;
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  const long long r0 = blockIdx.x * rows_per_cta;
  const long long r1 = (r0 + rows_per_cta < rows) ? r0 + rows_per_cta : rows;
  // two rows per iteration with all loads issued first: the single-row loop kept one 16-B load pair in flight per
  // thread and ran at 48 % of HBM peak
  const long long stride = blockDim.y;
  for (long long r = r0 + threadIdx.y; r < r1; r += 2 * stride) {
    const long long rb = r + stride;
    const bool has_b = rb < r1;
    uint4 da = __ldg(reinterpret_cast<const uint4*>(dy + r * dy_pitch + c0)), db = make_uint4(0u, 0u, 0u, 0u);
    uint4 ya = make_uint4(0u, 0u, 0u, 0u), yb = ya;
    if (has_b) db = __ldg(reinterpret_cast<const uint4*>(dy + rb * dy_pitch + c0));
    if (act != 0) {
      ya = __ldg(reinterpret_cast<const uint4*>(y + r * y_pitch + c0));
      if (has_b) yb = __ldg(reinterpret_cast<const uint4*>(y + rb * y_pitch + c0));
    }
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      if (q == 1 && !has_b) break;
      float d[8];
      unpack8f(q ? db : da, d);
      if (act != 0) {
        float v[8];
        unpack8f(q ? yb : ya, v);
#pragma unroll
        for (int j = 0; j < 8; ++j) d[j] = v[j] > 0.f ? d[j] : (act == 2 ? 0.2f * d[j] : 0.f);
        *reinterpret_cast<uint4*>(out + (q ? rb : r) * out_pitch + c0) =
            make_uint4(pack_bf16x2(d[0], d[1]), pack_bf16x2(d[2], d[3]), pack_bf16x2(d[4], d[5]), pack_bf16x2(d[6], d[7]));
        // the sum must be over the values as stored (bf16), like the tensor the wgrad kernel reads
#pragma unroll
        for (int j = 0; j < 8; ++j) d[j] = __bfloat162float(__float2bfloat16_rn(d[j]));
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += d[j];
    }
  }
  if (partial == nullptr) return;
#pragma unroll
  for (int j = 0; j < 8; ++j) red[threadIdx.y * C + c0 + j] = acc[j];
  __syncthreads();
  for (int c = threadIdx.y * blockDim.x + threadIdx.x; c < C; c += blockDim.x * blockDim.y) {
    float s = 0.f;
    for (int k = 0; k < static_cast<int>(blockDim.y); ++k) s += red[k * C + c];
    partial[static_cast<size_t>(blockIdx.x) * C + c] = s;
  }
}

// blockDim = (32 channels, 32 slices): slice k sums chunks k, k+32, ... in fp64, the 32 slices are then combined in
// a fixed order (bit-stable).  The walk is a chain of dependent L2 round trips per thread: with 8 slices over up to
// 592 chunks a launch took 19 us for 1 MB of partials (28 launches per training step).
constexpr int kColsumSlices = 32;
__global__ void __launch_bounds__(32 * kColsumSlices)
colsum_finalize_kernel(const float* __restrict__ partial, int chunks, float* __restrict__ out, int C) {
  __shared__ double red[kColsumSlices][32];
  const int c = blockIdx.x * 32 + threadIdx.x;
  double s = 0.0;
  if (c < C)
    for (int k = threadIdx.y; k < chunks; k += kColsumSlices) s += static_cast<double>(partial[static_cast<size_t>(k) * C + c]);
  red[threadIdx.y][threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.y == 0 && c < C) {
    double t = 0.0;
#pragma unroll
    for (int k = 0; k < kColsumSlices; ++k) t += red[k][threadIdx.x];
    out[c] = static_cast<float>(t);
  }
}

// torch.optim.Adam's update for up to kAdamMulti tensors in one launch.  Blocks are dealt out in proportion to
// tensor size (one block per kAdamChunk elements; first_block[] is the prefix sum), so a 2.4 M-element conv weight
// and a 64-element bias in the same launch both run at full memory parallelism.
constexpr int kAdamMulti = 64;
constexpr int kAdamChunk = 8192;
struct AdamMultiArgs {
  float* p[kAdamMulti];
  const float* g[kAdamMulti];
  float* m[kAdamMulti];
  float* v[kAdamMulti];
  long long n[kAdamMulti];
  int first_block[kAdamMulti + 1];
  int count;
};

__global__ void __launch_bounds__(256)
adam_step_multi_kernel(const __grid_constant__ AdamMultiArgs a, float lr, float b1, float b2, float eps,
                       const int* __restrict__ step_dev) {
  int t = 0;
  while (t + 1 < a.count && static_cast<int>(blockIdx.x) >= a.first_block[t + 1]) ++t;   // <= 64 steps, block-uniform
  const float step = static_cast<float>(*step_dev);
  const float bc1 = 1.f - powf(b1, step);
  const float bc2_sqrt = sqrtf(1.f - powf(b2, step));
  float* __restrict__ p = a.p[t];
  const float* __restrict__ g = a.g[t];
  float* __restrict__ m = a.m[t];
  float* __restrict__ v = a.v[t];
  const long long i0 = static_cast<long long>(blockIdx.x - a.first_block[t]) * kAdamChunk;
  const long long i1 = (i0 + kAdamChunk < a.n[t]) ? i0 + kAdamChunk : a.n[t];
  for (long long i = i0 + threadIdx.x; i < i1; i += 256) {
    const float gi = g[i];
    const float mi = b1 * m[i] + (1.f - b1) * gi;
    const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    p[i] -= (lr / bc1) * (mi / (sqrtf(vi) / bc2_sqrt + eps));
  }
}

__global__ void adam_step_dev_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                                     float* __restrict__ v, long long n, float lr, float b1, float b2, float eps,
                                     const int* __restrict__ step_dev) {
  const float step = static_cast<float>(*step_dev);
  const float bc1 = 1.f - powf(b1, step);
  const float bc2_sqrt = sqrtf(1.f - powf(b2, step));
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float gi = g[i];
    const float mi = b1 * m[i] + (1.f - b1) * gi;
    const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    p[i] -= (lr / bc1) * (mi / (sqrtf(vi) / bc2_sqrt + eps));
  }
}

static const int kColsumMaxChunks = 1184;   // 8 CTAs of 256 threads per SM: full occupancy for the streaming pass

}  // namespace s2p

extern "C" int s2p_pack_conv_weight_dgrad(const float* w_oihw, void* dst, int Cout, int Cin, int R, int S,
                                          void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(w_oihw && dst && Cout > 0 && Cin > 0 && R > 0 && S > 0, "pack_conv_weight_dgrad: bad args");
  const long long total = static_cast<long long>(Cout) * Cin * R * S;
  const int blocks = static_cast<int>((total + 255) / 256 < 4096 ? (total + 255) / 256 : 4096);
  pack_conv_weight_dgrad_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      w_oihw, static_cast<__nv_bfloat16*>(dst), Cout, Cin, R, S);
  count_launch();
  return check_launch("pack_conv_weight_dgrad_kernel");
}

extern "C" int s2p_pack_conv_weights_multi(const float* const* w_oihw, void* const* dst_fwd, void* const* dst_dgrad,
                                           const int* Cout, const int* Cin, const int* R, const int* S, int count,
                                           void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(w_oihw && dst_fwd && dst_dgrad && Cout && Cin && R && S, "pack_conv_weights_multi: null table");
  S2P_CHECK_ARG(count > 0 && count <= kPackMulti, "pack_conv_weights_multi: count=%d (1..%d)", count, kPackMulti);
  PackMultiArgs a;
  int blocks = 0;
  for (int i = 0; i < count; ++i) {
    S2P_CHECK_ARG(w_oihw[i] && dst_fwd[i] && Cout[i] > 0 && Cin[i] > 0 && R[i] > 0 && S[i] > 0,
                  "pack_conv_weights_multi: tensor %d", i);
    a.src[i] = w_oihw[i];
    a.fwd[i] = static_cast<__nv_bfloat16*>(dst_fwd[i]);
    a.dgrad[i] = static_cast<__nv_bfloat16*>(dst_dgrad[i]);
    a.cout[i] = Cout[i]; a.cin[i] = Cin[i]; a.r[i] = R[i]; a.s[i] = S[i];
    a.first_block[i] = blocks;
    const long long total = static_cast<long long>(Cout[i]) * Cin[i] * R[i] * S[i];
    const long long nb = (total + kPackChunk - 1) / kPackChunk;
    S2P_CHECK_ARG(blocks + nb < (1ll << 30), "pack_conv_weights_multi: too many elements");
    blocks += static_cast<int>(nb);
  }
  for (int i = count; i < kPackMulti; ++i) {
    a.src[i] = nullptr; a.fwd[i] = nullptr; a.dgrad[i] = nullptr;
    a.cout[i] = a.cin[i] = a.r[i] = a.s[i] = 0;
    a.first_block[i] = blocks;
  }
  a.first_block[kPackMulti] = blocks;
  a.count = count;
  pack_conv_weights_multi_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(a);
  count_launch();
  return check_launch("pack_conv_weights_multi_kernel");
}

extern "C" int s2p_act_bwd_bias_grad_chunks(void) { return s2p::kColsumMaxChunks; }

extern "C" int s2p_act_bwd_bias_grad(const void* dy, int dy_pitch, const void* y, int y_pitch, int act,
                                     void* dy_out, int out_pitch, float* bias_grad, float* partial,
                                     long long rows, int C, void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(dy && rows > 0 && C > 0 && C % 8 == 0 && C <= 2048, "act_bwd_bias_grad: bad dims");
  S2P_CHECK_ARG(act >= 0 && act <= 2, "act_bwd_bias_grad: act=%d", act);
  S2P_CHECK_ARG(act == 0 || (y && dy_out), "act_bwd_bias_grad: y and dy_out are needed when act != 0");
  S2P_CHECK_ARG((bias_grad == nullptr) == (partial == nullptr), "act_bwd_bias_grad: bias_grad/partial together");
  S2P_CHECK_ARG(dy_pitch % 8 == 0 && y_pitch % 8 == 0 && out_pitch % 8 == 0, "act_bwd_bias_grad: pitch");
  const int groups = C / 8;
  int py = 256 / groups;
  if (py < 1) py = 1;
  long long chunks = (rows + py * 8 - 1) / (py * 8);
  if (chunks > kColsumMaxChunks) chunks = kColsumMaxChunks;
  const long long rpc = (rows + chunks - 1) / chunks;
  chunks = (rows + rpc - 1) / rpc;
  const size_t smem = static_cast<size_t>(py) * C * sizeof(float);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  act_bwd_colsum_kernel<<<static_cast<unsigned>(chunks), dim3(groups, py), smem, st>>>(
      static_cast<const __nv_bfloat16*>(dy), dy_pitch, static_cast<const __nv_bfloat16*>(y), y_pitch, act,
      static_cast<__nv_bfloat16*>(dy_out), out_pitch, partial, rows, C, rpc);
  count_launch();
  int rc = check_launch("act_bwd_colsum_kernel");
  if (rc != S2P_OK || bias_grad == nullptr) return rc;
  colsum_finalize_kernel<<<(C + 31) / 32, dim3(32, kColsumSlices), 0, st>>>(partial, static_cast<int>(chunks), bias_grad, C);
  count_launch();
  return check_launch("colsum_finalize_kernel");
}

extern "C" int s2p_adam_step_dev(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, long long n,
                                 float lr, float beta1, float beta2, float eps, const int* step_dev, void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(param && grad && exp_avg && exp_avg_sq && step_dev && n > 0, "adam_step_dev: bad args");
  const int blocks = static_cast<int>((n + 255) / 256 < 1184 ? (n + 255) / 256 : 1184);
  adam_step_dev_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(param, grad, exp_avg, exp_avg_sq, n, lr,
                                                                             beta1, beta2, eps, step_dev);
  count_launch();
  return check_launch("adam_step_dev_kernel");
}

// ---------------------------------------------------------------------------------------------------
// 2x2 stride-2 max pool (floor mode) on NHWC bf16 and its backward, for the VGG19 perceptual loss of the
// training step (rebuttal.md:71,131-135).  HBM-bound: one read of x, a quarter-size
// write.  The backward recomputes the argmax from x with nn.MaxPool2d's rule (first maximum in (dy, dx) scan
// order wins; after a ReLU ties at 0 are common) and also zero-fills the row / column an odd input size
// leaves uncovered, so dx needs no memset.
namespace s2p {

__device__ __forceinline__ void unpack8(const uint4& v, float* f) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}

__global__ void maxpool2_fwd_kernel(const __nv_bfloat16* __restrict__ x, int x_pitch, __nv_bfloat16* __restrict__ y,
                                    int y_pitch, int B, int H, int W, int C) {
  const int Ho = H / 2, Wo = W / 2, cv = C / 8;
  const long long total = static_cast<long long>(B) * Ho * Wo * cv;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c8 = static_cast<int>(i % cv);
    long long t = i / cv;
    const int wo = static_cast<int>(t % Wo); t /= Wo;
    const int ho = static_cast<int>(t % Ho);
    const long long b = t / Ho;
    const __nv_bfloat16* px = x + ((b * H + 2 * ho) * W + 2 * wo) * static_cast<long long>(x_pitch) + c8 * 8;
    const uint4 v00 = __ldg(reinterpret_cast<const uint4*>(px));
    const uint4 v01 = __ldg(reinterpret_cast<const uint4*>(px + x_pitch));
    const uint4 v10 = __ldg(reinterpret_cast<const uint4*>(px + static_cast<long long>(W) * x_pitch));
    const uint4 v11 = __ldg(reinterpret_cast<const uint4*>(px + static_cast<long long>(W + 1) * x_pitch));
    uint4 r;
    __nv_bfloat162* rh = reinterpret_cast<__nv_bfloat162*>(&r);
    const __nv_bfloat162* a = reinterpret_cast<const __nv_bfloat162*>(&v00);
    const __nv_bfloat162* bq = reinterpret_cast<const __nv_bfloat162*>(&v01);
    const __nv_bfloat162* c = reinterpret_cast<const __nv_bfloat162*>(&v10);
    const __nv_bfloat162* d = reinterpret_cast<const __nv_bfloat162*>(&v11);
#pragma unroll
    for (int k = 0; k < 4; ++k) rh[k] = __hmax2(__hmax2(a[k], bq[k]), __hmax2(c[k], d[k]));
    *reinterpret_cast<uint4*>(y + ((b * Ho + ho) * Wo + wo) * static_cast<long long>(y_pitch) + c8 * 8) = r;
  }
}

// one thread per 2x2 cell of the INPUT grid (ceil sizes) x 8 channels
__global__ void maxpool2_bwd_kernel(const __nv_bfloat16* __restrict__ dy, int dy_pitch, const __nv_bfloat16* __restrict__ x,
                                    int x_pitch, __nv_bfloat16* __restrict__ dx, int dx_pitch, int B, int H, int W,
                                    int C) {
  const int Ho = H / 2, Wo = W / 2, Hc = (H + 1) / 2, Wc = (W + 1) / 2, cv = C / 8;
  const long long total = static_cast<long long>(B) * Hc * Wc * cv;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c8 = static_cast<int>(i % cv);
    long long t = i / cv;
    const int wc = static_cast<int>(t % Wc); t /= Wc;
    const int hc = static_cast<int>(t % Hc);
    const long long b = t / Hc;
    const uint4 zero = make_uint4(0u, 0u, 0u, 0u);
    if (hc >= Ho || wc >= Wo) {   // cells no pooling window covers (odd H / W)
      for (int dyy = 0; dyy < 2; ++dyy)
        for (int dxx = 0; dxx < 2; ++dxx) {
          const int h = 2 * hc + dyy, w = 2 * wc + dxx;
          if (h < H && w < W)
            *reinterpret_cast<uint4*>(dx + ((b * H + h) * W + w) * static_cast<long long>(dx_pitch) + c8 * 8) = zero;
        }
      continue;
    }
    const __nv_bfloat16* px = x + ((b * H + 2 * hc) * W + 2 * wc) * static_cast<long long>(x_pitch) + c8 * 8;
    float v[4][8], g[8];
    unpack8(__ldg(reinterpret_cast<const uint4*>(px)), v[0]);
    unpack8(__ldg(reinterpret_cast<const uint4*>(px + x_pitch)), v[1]);
    unpack8(__ldg(reinterpret_cast<const uint4*>(px + static_cast<long long>(W) * x_pitch)), v[2]);
    unpack8(__ldg(reinterpret_cast<const uint4*>(px + static_cast<long long>(W + 1) * x_pitch)), v[3]);
    unpack8(__ldg(reinterpret_cast<const uint4*>(dy + ((b * Ho + hc) * Wo + wc) * static_cast<long long>(dy_pitch) + c8 * 8)), g);
    float o[4][8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      int arg = 0;
      float m = v[0][k];
#pragma unroll
      for (int q = 1; q < 4; ++q)
        if (v[q][k] > m) { m = v[q][k]; arg = q; }
#pragma unroll
      for (int q = 0; q < 4; ++q) o[q][k] = (q == arg) ? g[k] : 0.f;
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      uint4 r;
      __nv_bfloat162* rh = reinterpret_cast<__nv_bfloat162*>(&r);
#pragma unroll
      for (int k = 0; k < 4; ++k) rh[k] = __floats2bfloat162_rn(o[q][2 * k], o[q][2 * k + 1]);
      const int h = 2 * hc + (q >> 1), w = 2 * wc + (q & 1);
      *reinterpret_cast<uint4*>(dx + ((b * H + h) * W + w) * static_cast<long long>(dx_pitch) + c8 * 8) = r;
    }
  }
}

// Zero-insertion in front of the data-gradient conv of a stride-2 layer (transposed convolution):
// out[b][2i][2j][:] = dy[b][i][j][:], everything else 0, in one write-only pass (was a fill plus a strided copy).
// One CTA per output row (b, h): 32-bit index arithmetic only, odd rows are pure zero fills.
__global__ void __launch_bounds__(256)
zero_insert2_kernel(const __nv_bfloat16* __restrict__ dy, int dy_pitch, __nv_bfloat16* __restrict__ out, int Ho, int Wo,
                    int Hout, int Wout, int C8) {
  const int h = blockIdx.x % Hout;
  const long long b = blockIdx.x / Hout;
  uint4* __restrict__ row = reinterpret_cast<uint4*>(out) + (b * Hout + h) * static_cast<long long>(Wout) * C8;
  const int n = Wout * C8;
  const bool live = !(h & 1) && (h >> 1) < Ho;
  const __nv_bfloat16* __restrict__ src = dy + (b * Ho + (h >> 1)) * static_cast<long long>(Wo) * dy_pitch;
  for (int i = threadIdx.x; i < n; i += 256) {
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (live) {
      const int w = i / C8, c8 = i - w * C8;
      if (!(w & 1) && (w >> 1) < Wo)
        v = __ldg(reinterpret_cast<const uint4*>(src + static_cast<long long>(w >> 1) * dy_pitch + c8 * 8));
    }
    row[i] = v;
  }
}

// fp32 NCHW [B][c][H][W] (c <= 8: images, image pairs) -> bf16 NHWC [B][H][W][64], channels >= c zero: the operand
// the tensor-core convs take for a 3- or 6-channel input.  A CTA takes 256 pixels: every thread reads the c planes of
// ONE pixel (coalesced along the plane) and parks the packed 8 channels in shared memory; then the CTA writes its
// 256 x 128 B of output as consecutive 16-byte stores (7 of every 8 are zeros).  HBM-bound on the 128 B per pixel it
// writes.  `t` (optional, same layout as x): the tanh output whose backward this is -> x * (1 - t^2) is stored.
__global__ void __launch_bounds__(256)
nchw_to_nhwc64_kernel(const float* __restrict__ x, const float* __restrict__ t, __nv_bfloat16* __restrict__ y,
                      long long total_pix, int c, int HW) {
  __shared__ uint4 head[256];
  const long long pix0 = blockIdx.x * 256ll;
  const long long pix = pix0 + threadIdx.x;
  if (pix < total_pix) {
    const long long n = pix / HW;
    const int p = static_cast<int>(pix - n * HW);
    float f[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      f[j] = j < c ? __ldg(x + (n * c + j) * HW + p) : 0.f;
      if (t != nullptr && j < c) {
        const float tv = __ldg(t + (n * c + j) * HW + p);
        f[j] *= 1.f - tv * tv;
      }
    }
    head[threadIdx.x] = make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]), pack_bf16x2(f[6], f[7]));
  }
  __syncthreads();
  const long long left = total_pix - pix0;
  const int items = static_cast<int>(left < 256 ? left : 256) * 8;
  uint4* __restrict__ dst = reinterpret_cast<uint4*>(y + pix0 * 64);
  for (int i = threadIdx.x; i < items; i += 256) dst[i] = (i & 7) ? make_uint4(0u, 0u, 0u, 0u) : head[i >> 3];
}

// its backward: dx[n][j][h][w] = float(dy[n][h][w][j]) for j < c (the padded channels carry no gradient)
__global__ void __launch_bounds__(256)
nhwc64_to_nchw_kernel(const __nv_bfloat16* __restrict__ dy, int dy_pitch, float* __restrict__ dx, long long total_pix,
                      int c, int HW, int apply_tanh) {
  const long long pix = blockIdx.x * 256ll + threadIdx.x;
  if (pix >= total_pix) return;
  const uint4 u = __ldg(reinterpret_cast<const uint4*>(dy + pix * dy_pitch));
  float f[8];
  unpack8f(u, f);
  const long long n = pix / HW;
  const int p = static_cast<int>(pix - n * HW);
#pragma unroll
  for (int j = 0; j < 8; ++j)
    if (j < c) dx[(n * c + j) * HW + p] = apply_tanh ? tanhf(f[j]) : f[j];
}

}  // namespace s2p

extern "C" int s2p_maxpool2_fwd(const void* x, int x_pitch, void* y, int y_pitch, int B, int H, int W, int C,
                                void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(x && y && B > 0 && H >= 2 && W >= 2 && C > 0 && C % 8 == 0, "maxpool2_fwd: bad args");
  S2P_CHECK_ARG(x_pitch >= C && x_pitch % 8 == 0 && y_pitch >= C && y_pitch % 8 == 0, "maxpool2_fwd: pitch");
  const long long total = static_cast<long long>(B) * (H / 2) * (W / 2) * (C / 8);
  const int blocks = static_cast<int>((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
  maxpool2_fwd_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(x), x_pitch, static_cast<__nv_bfloat16*>(y), y_pitch, B, H, W, C);
  count_launch();
  return check_launch("maxpool2_fwd_kernel");
}

extern "C" int s2p_maxpool2_bwd(const void* dy, int dy_pitch, const void* x, int x_pitch, void* dx, int dx_pitch, int B,
                                int H, int W, int C, void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(dy && x && dx && B > 0 && H >= 2 && W >= 2 && C > 0 && C % 8 == 0, "maxpool2_bwd: bad args");
  S2P_CHECK_ARG(x_pitch >= C && x_pitch % 8 == 0 && dy_pitch >= C && dy_pitch % 8 == 0 && dx_pitch >= C && dx_pitch % 8 == 0,
                "maxpool2_bwd: pitch");
  const long long total = static_cast<long long>(B) * ((H + 1) / 2) * ((W + 1) / 2) * (C / 8);
  const int blocks = static_cast<int>((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
  maxpool2_bwd_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(dy), dy_pitch, static_cast<const __nv_bfloat16*>(x), x_pitch,
      static_cast<__nv_bfloat16*>(dx), dx_pitch, B, H, W, C);
  count_launch();
  return check_launch("maxpool2_bwd_kernel");
}

extern "C" int s2p_adam_step_multi(float* const* param, const float* const* grad, float* const* exp_avg,
                                   float* const* exp_avg_sq, const long long* n, int count, float lr, float beta1,
                                   float beta2, float eps, const int* step_dev, void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(param && grad && exp_avg && exp_avg_sq && n && step_dev, "adam_step_multi: null pointer");
  S2P_CHECK_ARG(count >= 1 && count <= kAdamMulti, "adam_step_multi: count=%d (1..%d)", count, kAdamMulti);
  AdamMultiArgs a;
  int blocks = 0;
  for (int i = 0; i < count; ++i) {
    S2P_CHECK_ARG(param[i] && grad[i] && exp_avg[i] && exp_avg_sq[i] && n[i] > 0, "adam_step_multi: tensor %d", i);
    a.p[i] = param[i]; a.g[i] = grad[i]; a.m[i] = exp_avg[i]; a.v[i] = exp_avg_sq[i]; a.n[i] = n[i];
    a.first_block[i] = blocks;
    const long long nb = (n[i] + kAdamChunk - 1) / kAdamChunk;
    S2P_CHECK_ARG(blocks + nb < (1ll << 30), "adam_step_multi: too many elements");
    blocks += static_cast<int>(nb);
  }
  for (int i = count; i < kAdamMulti; ++i) { a.p[i] = nullptr; a.g[i] = nullptr; a.m[i] = nullptr; a.v[i] = nullptr; a.n[i] = 0; a.first_block[i] = blocks; }
  a.first_block[kAdamMulti] = blocks;
  a.count = count;
  adam_step_multi_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(a, lr, beta1, beta2, eps, step_dev);
  count_launch();
  return check_launch("adam_step_multi_kernel");
}

// ---------------------------------------------------------------------------------------------------
// Multi-tensor scale (training glue): the MAT blend factors a = sigmoid(blend) scale whole parameter tensors
// (gamma/beta conv weights by 1-a, state-head rows by a; oracle MAT.forward).  Done tensor by tensor in torch
// this was ~480 tiny kernels per step (mul / sum of the forward and of MulBackward); here the forward is one
// launch per 64 tensors writing straight into the concatenated operand buffers, and the backward one launch
// producing the parameter gradients and fixed-order per-block partial dot products for the scale gradients.
namespace s2p {

constexpr int kScaleMulti = 64;
constexpr int kScaleBlocks = 32;
struct MultiScaleArgs {
  const float* src[kScaleMulti];   // parameter tensor
  float* dst[kScaleMulti];         // fwd: scaled copy;  bwd: gradient w.r.t. the parameter
  const float* dout[kScaleMulti];  // bwd: gradient w.r.t. the scaled copy
  long long n[kScaleMulti];
  int sidx[kScaleMulti];           // index into scales[]
};

__global__ void __launch_bounds__(256) multi_scale_fwd_kernel(const __grid_constant__ MultiScaleArgs a,
                                                              const float* __restrict__ scales) {
  const int t = blockIdx.y;
  const float s = scales[a.sidx[t]];
  const float* __restrict__ src = a.src[t];
  float* __restrict__ dst = a.dst[t];
  const long long n = a.n[t];
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * 256) dst[i] = src[i] * s;
}

// dst = dout * s;  partial[t][block] = sum over the block's elements of dout * src   (fixed order)
__global__ void __launch_bounds__(256) multi_scale_bwd_kernel(const __grid_constant__ MultiScaleArgs a,
                                                              const float* __restrict__ scales, float* __restrict__ partial) {
  __shared__ float red[8];
  const int t = blockIdx.y;
  const float s = scales[a.sidx[t]];
  const float* __restrict__ src = a.src[t];
  const float* __restrict__ dout = a.dout[t];
  float* __restrict__ dst = a.dst[t];
  const long long n = a.n[t];
  float acc = 0.f;
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * 256) {
    const float g = dout[i];
    dst[i] = g * s;
    acc = fmaf(g, src[i], acc);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float v = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) v += red[w];
    partial[t * kScaleBlocks + blockIdx.x] = v;
  }
}

}  // namespace s2p

static int multi_scale_fill(s2p::MultiScaleArgs& a, const float* const* src, float* const* dst, const float* const* dout,
                            const long long* n, const int* sidx, int count) {
  using namespace s2p;
  S2P_CHECK_ARG(src && dst && n && sidx && count >= 1 && count <= kScaleMulti, "multi_scale: bad args (count=%d)", count);
  for (int i = 0; i < kScaleMulti; ++i) {
    const bool live = i < count;
    if (live) S2P_CHECK_ARG(src[i] && dst[i] && n[i] > 0 && sidx[i] >= 0 && (!dout || dout[i]), "multi_scale: tensor %d", i);
    a.src[i] = live ? src[i] : nullptr;
    a.dst[i] = live ? dst[i] : nullptr;
    a.dout[i] = (live && dout) ? dout[i] : nullptr;
    a.n[i] = live ? n[i] : 0;
    a.sidx[i] = live ? sidx[i] : 0;
  }
  return S2P_OK;
}

extern "C" int s2p_multi_scale_blocks(void) { return s2p::kScaleBlocks; }

extern "C" int s2p_multi_scale_fwd(const float* const* src, float* const* dst, const long long* n, const int* scale_idx,
                                   int count, const float* scales, void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(scales != nullptr, "multi_scale_fwd: scales");
  MultiScaleArgs a;
  const int rc = multi_scale_fill(a, src, dst, nullptr, n, scale_idx, count);
  if (rc != S2P_OK) return rc;
  multi_scale_fwd_kernel<<<dim3(kScaleBlocks, count), 256, 0, static_cast<cudaStream_t>(stream)>>>(a, scales);
  count_launch();
  return check_launch("multi_scale_fwd_kernel");
}

extern "C" int s2p_multi_scale_bwd(const float* const* dout, const float* const* src, float* const* dsrc, const long long* n,
                                   const int* scale_idx, int count, const float* scales, float* partial, void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(scales && partial && dout, "multi_scale_bwd: null pointer");
  MultiScaleArgs a;
  const int rc = multi_scale_fill(a, src, dsrc, dout, n, scale_idx, count);
  if (rc != S2P_OK) return rc;
  multi_scale_bwd_kernel<<<dim3(kScaleBlocks, count), 256, 0, static_cast<cudaStream_t>(stream)>>>(a, scales, partial);
  count_launch();
  return check_launch("multi_scale_bwd_kernel");
}

// ---------------------------------------------------------------------------------------------------
// mean |a - b| over NHWC bf16 feature maps and its backward (feature-matching and VGG perceptual terms of the
// generator loss, oracle generator_losses / vgg_loss: F.l1_loss on fp32 copies).  One read of a and b, fp32 math,
// fixed-order two-stage reduction (bit-stable); the backward writes sign(a - b) * g / N in bf16 with the upstream
// gradient g read from device memory (no host sync, graph-capturable).
namespace s2p {

constexpr int kL1MaxChunks = 1184;

__global__ void __launch_bounds__(256)
l1_sum_kernel(const __nv_bfloat16* __restrict__ a, int a_pitch, const __nv_bfloat16* __restrict__ b, int b_pitch,
              long long rows, int groups, long long vec_per_cta, float* __restrict__ partial) {
  __shared__ float red[8];
  const long long total = rows * groups;                 // 8-channel vectors
  const long long v0 = blockIdx.x * vec_per_cta;
  const long long v1 = (v0 + vec_per_cta < total) ? v0 + vec_per_cta : total;
  float acc = 0.f;
  for (long long v = v0 + threadIdx.x; v < v1; v += 256) {
    const long long r = v / groups;
    const int c0 = static_cast<int>(v - r * groups) * 8;
    float x[8], y[8];
    unpack8f(__ldg(reinterpret_cast<const uint4*>(a + r * a_pitch + c0)), x);
    unpack8f(__ldg(reinterpret_cast<const uint4*>(b + r * b_pitch + c0)), y);
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) s += fabsf(x[j] - y[j]);
    acc += s;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += red[w];
    partial[blockIdx.x] = t;
  }
}

__global__ void l1_finalize_kernel(const float* __restrict__ partial, int chunks, double inv_n, float* __restrict__ out) {
  __shared__ double red[32];
  double s = 0.0;
  for (int k = threadIdx.x; k < chunks; k += 32) s += static_cast<double>(partial[k]);
  red[threadIdx.x] = s;
  __syncwarp();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int k = 0; k < 32; ++k) t += red[k];
    out[0] = static_cast<float>(t * inv_n);
  }
}

__global__ void __launch_bounds__(256)
l1_bwd_kernel(const __nv_bfloat16* __restrict__ a, int a_pitch, const __nv_bfloat16* __restrict__ b, int b_pitch,
              const float* __restrict__ gout, float inv_n, long long rows, int groups, __nv_bfloat16* __restrict__ da,
              int da_pitch) {
  const float g = gout[0] * inv_n;
  const long long total = rows * groups;
  for (long long v = blockIdx.x * 256ll + threadIdx.x; v < total; v += static_cast<long long>(gridDim.x) * 256) {
    const long long r = v / groups;
    const int c0 = static_cast<int>(v - r * groups) * 8;
    float x[8], y[8], o[8];
    unpack8f(__ldg(reinterpret_cast<const uint4*>(a + r * a_pitch + c0)), x);
    unpack8f(__ldg(reinterpret_cast<const uint4*>(b + r * b_pitch + c0)), y);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float d = x[j] - y[j];
      o[j] = d > 0.f ? g : (d < 0.f ? -g : 0.f);
    }
    *reinterpret_cast<uint4*>(da + r * da_pitch + c0) =
        make_uint4(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]), pack_bf16x2(o[4], o[5]), pack_bf16x2(o[6], o[7]));
  }
}

}  // namespace s2p

extern "C" int s2p_l1_mean_chunks(void) { return s2p::kL1MaxChunks; }

extern "C" int s2p_l1_mean_fwd(const void* a, int a_pitch, const void* b, int b_pitch, long long rows, int C,
                               float* partial, float* out, void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(a && b && partial && out && rows > 0 && C > 0 && C % 8 == 0, "l1_mean_fwd: bad args");
  S2P_CHECK_ARG(a_pitch >= C && a_pitch % 8 == 0 && b_pitch >= C && b_pitch % 8 == 0, "l1_mean_fwd: pitch");
  const int groups = C / 8;
  const long long total = rows * groups;
  long long chunks = (total + 2047) / 2048;
  if (chunks > kL1MaxChunks) chunks = kL1MaxChunks;
  const long long vpc = (total + chunks - 1) / chunks;
  chunks = (total + vpc - 1) / vpc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  l1_sum_kernel<<<static_cast<unsigned>(chunks), 256, 0, st>>>(static_cast<const __nv_bfloat16*>(a), a_pitch,
                                                              static_cast<const __nv_bfloat16*>(b), b_pitch, rows, groups, vpc, partial);
  count_launch();
  int rc = check_launch("l1_sum_kernel");
  if (rc != S2P_OK) return rc;
  l1_finalize_kernel<<<1, 32, 0, st>>>(partial, static_cast<int>(chunks), 1.0 / (static_cast<double>(rows) * C), out);
  count_launch();
  return check_launch("l1_finalize_kernel");
}

extern "C" int s2p_l1_mean_bwd(const void* a, int a_pitch, const void* b, int b_pitch, const float* gout, long long rows,
                               int C, void* da, int da_pitch, void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(a && b && gout && da && rows > 0 && C > 0 && C % 8 == 0, "l1_mean_bwd: bad args");
  S2P_CHECK_ARG(a_pitch >= C && a_pitch % 8 == 0 && b_pitch >= C && b_pitch % 8 == 0 && da_pitch >= C && da_pitch % 8 == 0,
                "l1_mean_bwd: pitch");
  const int groups = C / 8;
  const long long total = rows * groups;
  const int blocks = static_cast<int>((total + 255) / 256 < 148 * 8 ? (total + 255) / 256 : 148 * 8);
  l1_bwd_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(a), a_pitch, static_cast<const __nv_bfloat16*>(b), b_pitch, gout,
      static_cast<float>(1.0 / (static_cast<double>(rows) * C)), rows, groups, static_cast<__nv_bfloat16*>(da), da_pitch);
  count_launch();
  return check_launch("l1_bwd_kernel");
}

static int nchw_to_nhwc64_impl(const float* x_nchw, const float* tanh_out, void* y_nhwc64, int B, int C, int H, int W,
                               void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(x_nchw && y_nhwc64 && B > 0 && H > 0 && W > 0, "nchw_to_nhwc64: bad args");
  S2P_CHECK_ARG(C >= 1 && C <= 8, "nchw_to_nhwc64: C=%d (1..8 channels are padded to 64)", C);
  S2P_CHECK_ARG((reinterpret_cast<uintptr_t>(y_nhwc64) & 15) == 0, "nchw_to_nhwc64: 16-byte alignment");
  const long long total = static_cast<long long>(B) * H * W;
  const long long blocks = (total + 255) / 256;
  S2P_CHECK_ARG(blocks < (1ll << 31), "nchw_to_nhwc64: too many pixels");
  nchw_to_nhwc64_kernel<<<static_cast<unsigned>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      x_nchw, tanh_out, static_cast<__nv_bfloat16*>(y_nhwc64), total, C, H * W);
  count_launch();
  return check_launch("nchw_to_nhwc64_kernel");
}

extern "C" int s2p_nchw_to_nhwc64(const float* x_nchw, void* y_nhwc64, int B, int C, int H, int W, void* stream) {
  return nchw_to_nhwc64_impl(x_nchw, nullptr, y_nhwc64, B, C, H, W, stream);
}

extern "C" int s2p_tanh_bwd_nchw_to_nhwc64(const float* dy_nchw, const float* y_nchw, void* dx_nhwc64, int B, int C, int H,
                                           int W, void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(y_nchw != nullptr, "tanh_bwd_nchw_to_nhwc64: null tanh output");
  return nchw_to_nhwc64_impl(dy_nchw, y_nchw, dx_nhwc64, B, C, H, W, stream);
}

static int nhwc64_to_nchw_impl(const void* dy_nhwc, int dy_pitch, float* dx_nchw, int B, int C, int H, int W,
                               int apply_tanh, void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(dy_nhwc && dx_nchw && B > 0 && H > 0 && W > 0, "nhwc64_to_nchw: bad args");
  S2P_CHECK_ARG(C >= 1 && C <= 8 && dy_pitch >= 8 && dy_pitch % 8 == 0, "nhwc64_to_nchw: C=%d pitch=%d", C, dy_pitch);
  S2P_CHECK_ARG((reinterpret_cast<uintptr_t>(dy_nhwc) & 15) == 0, "nhwc64_to_nchw: 16-byte alignment");
  const long long total = static_cast<long long>(B) * H * W;
  const long long blocks = (total + 255) / 256;
  S2P_CHECK_ARG(blocks < (1ll << 31), "nhwc64_to_nchw: too many pixels");
  nhwc64_to_nchw_kernel<<<static_cast<unsigned>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(dy_nhwc), dy_pitch, dx_nchw, total, C, H * W, apply_tanh);
  count_launch();
  return check_launch("nhwc64_to_nchw_kernel");
}

extern "C" int s2p_nhwc64_to_nchw(const void* dy_nhwc, int dy_pitch, float* dx_nchw, int B, int C, int H, int W,
                                  void* stream) {
  return nhwc64_to_nchw_impl(dy_nhwc, dy_pitch, dx_nchw, B, C, H, W, 0, stream);
}

extern "C" int s2p_tanh_nhwc_to_nchw(const void* x_nhwc, int x_pitch, float* y_nchw, int B, int C, int H, int W,
                                     void* stream) {
  return nhwc64_to_nchw_impl(x_nhwc, x_pitch, y_nchw, B, C, H, W, 1, stream);
}

extern "C" int s2p_zero_insert2(const void* dy, int dy_pitch, void* out, int B, int Ho, int Wo, int Hout, int Wout, int C,
                                void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(dy && out && B > 0 && Ho > 0 && Wo > 0 && C > 0 && C % 8 == 0, "zero_insert2: bad args");
  S2P_CHECK_ARG(Hout >= 2 * Ho - 1 && Wout >= 2 * Wo - 1, "zero_insert2: output %dx%d too small for %dx%d", Hout, Wout, Ho, Wo);
  S2P_CHECK_ARG(dy_pitch >= C && dy_pitch % 8 == 0, "zero_insert2: pitch");
  S2P_CHECK_ARG(((reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(out)) & 15) == 0, "zero_insert2: 16-byte alignment");
  const long long rows = static_cast<long long>(B) * Hout;
  S2P_CHECK_ARG(rows < (1ll << 31) && static_cast<long long>(Wout) * (C / 8) < (1ll << 31), "zero_insert2: too large");
  zero_insert2_kernel<<<static_cast<unsigned>(rows), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __nv_bfloat16*>(dy), dy_pitch, static_cast<__nv_bfloat16*>(out), Ho, Wo, Hout, Wout, C / 8);
  count_launch();
  return check_launch("zero_insert2_kernel");
}

// ---------------------------------------------------------------------------------------------------
// nn.AvgPool2d(3, stride 2, padding 1, count_include_pad=False) on fp32 planes [P][H][W] and its backward: the
// down-sampling between the scales of the multiscale discriminator (oracle MultiscaleDiscriminator.forward).  One
// thread per output (forward) / input (backward) element; torch's backward kernel took 323 us for the 15 MB of a
// batch-128 step's [256, 6, 100, 100] input.
namespace s2p {

__device__ __forceinline__ int pool_count(int o, int n) {     // valid taps of the window centred at 2 * o along an axis of n
  const int lo = max(2 * o - 1, 0), hi = min(2 * o + 1, n - 1);
  return hi - lo + 1;
}

__global__ void __launch_bounds__(256)
avgpool3s2_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, long long planes, int H, int W, int Ho, int Wo) {
  const long long total = planes * Ho * Wo;
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < total; i += gridDim.x * 256ll) {
    const int wo = static_cast<int>(i % Wo);
    const long long t = i / Wo;
    const int ho = static_cast<int>(t % Ho);
    const float* xp = x + (t / Ho) * H * W;
    float s = 0.f;
    for (int dh = -1; dh <= 1; ++dh) {
      const int h = 2 * ho + dh;
      if (h < 0 || h >= H) continue;
      for (int dw = -1; dw <= 1; ++dw) {
        const int w = 2 * wo + dw;
        if (w >= 0 && w < W) s += __ldg(xp + static_cast<long long>(h) * W + w);
      }
    }
    y[i] = s / static_cast<float>(pool_count(ho, H) * pool_count(wo, W));
  }
}

__global__ void __launch_bounds__(256)
avgpool3s2_bwd_kernel(const float* __restrict__ dy, float* __restrict__ dx, long long planes, int H, int W, int Ho, int Wo) {
  const long long total = planes * H * W;
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < total; i += gridDim.x * 256ll) {
    const int w = static_cast<int>(i % W);
    const long long t = i / W;
    const int h = static_cast<int>(t % H);
    const float* dp = dy + (t / H) * Ho * Wo;
    // outputs whose window [2o - 1, 2o + 1] contains this coordinate: o = c / 2 for even c, (c - 1) / 2 and (c + 1) / 2 for odd c
    const int h0 = h >> 1, h1 = (h & 1) ? h0 + 1 : h0;
    const int w0 = w >> 1, w1 = (w & 1) ? w0 + 1 : w0;
    float s = 0.f;
    for (int ho = h0; ho <= h1; ++ho) {
      if (ho >= Ho) continue;
      for (int wo = w0; wo <= w1; ++wo)
        if (wo < Wo) s += __ldg(dp + static_cast<long long>(ho) * Wo + wo) / static_cast<float>(pool_count(ho, H) * pool_count(wo, W));
    }
    dx[i] = s;
  }
}

}  // namespace s2p

extern "C" int s2p_avgpool3s2_fwd(const float* x, float* y, long long planes, int H, int W, void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(x && y && planes > 0 && H > 0 && W > 0, "avgpool3s2_fwd: bad args");
  const int Ho = (H - 1) / 2 + 1, Wo = (W - 1) / 2 + 1;
  const long long total = planes * Ho * Wo;
  const int blocks = static_cast<int>((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
  avgpool3s2_fwd_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(x, y, planes, H, W, Ho, Wo);
  count_launch();
  return check_launch("avgpool3s2_fwd_kernel");
}

extern "C" int s2p_avgpool3s2_bwd(const float* dy, float* dx, long long planes, int H, int W, void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(dy && dx && planes > 0 && H > 0 && W > 0, "avgpool3s2_bwd: bad args");
  const int Ho = (H - 1) / 2 + 1, Wo = (W - 1) / 2 + 1;
  const long long total = planes * H * W;
  const int blocks = static_cast<int>((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
  avgpool3s2_bwd_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(dy, dx, planes, H, W, Ho, Wo);
  count_launch();
  return check_launch("avgpool3s2_bwd_kernel");
}

// out[c] = sum_r x[r][c] for a row-major fp32 [rows][cols] matrix (bias gradient of a Linear layer: the batch sum of dy).
// blockDim = (32 columns, 8 row slices), fp64 slices combined in a fixed order; torch's generic reduction took 24 us per
// [128, 512] matrix (27 launches per training step).
namespace s2p {
__global__ void __launch_bounds__(256)
colsum_rows_f32_kernel(const float* __restrict__ x, int pitch, int rows, int cols, float* __restrict__ out) {
  __shared__ double red[8][32];
  const int c = blockIdx.x * 32 + threadIdx.x;
  double s = 0.0;
  if (c < cols)
    for (int r = threadIdx.y; r < rows; r += 8) s += static_cast<double>(__ldg(x + static_cast<size_t>(r) * pitch + c));
  red[threadIdx.y][threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.y == 0 && c < cols) {
    double t = 0.0;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += red[k][threadIdx.x];
    out[c] = static_cast<float>(t);
  }
}
}  // namespace s2p

extern "C" int s2p_colsum_rows_f32(const float* x, int pitch, int rows, int cols, float* out, void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(x && out && rows > 0 && cols > 0 && pitch >= cols, "colsum_rows_f32: bad args");
  colsum_rows_f32_kernel<<<(cols + 31) / 32, dim3(32, 8), 0, static_cast<cudaStream_t>(stream)>>>(x, pitch, rows, cols, out);
  count_launch();
  return check_launch("colsum_rows_f32_kernel");
}

