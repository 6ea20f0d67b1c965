// SLAC window assembly (SURVEY.md section 8 f, rank 2): the consumer side of the generated frames.
// The reference keeps every 9-frame window as a Python list of CHW uint8 arrays on the host and, per
// sample, copies them into a staging array, uploads it and divides by 255
// (rlkit/torch/slac/buffer.py:127-147); the CHW transpose of the whole dataset happens at load time
// (rlkit/torch/slac/algo.py:174-175).  Here the dataset's frames stay in HBM exactly as stored on disk
// (uint8 HWC, 30 kB per frame: 1M rows x 2 arrays = 60 GB of the 180 GB) and ONE kernel does
// window gather + HWC->CHW + uint8 -> float/255 for a sampled batch.  Byte work, HBM-bound:
// algorithmic bytes per frame = H*W*C read + 4*H*W*C written.
#include <cuda_bf16.h>

#include "s2p_internal.h"

namespace s2p {
namespace sw {

constexpr int kPix = 1024;  // pixels per CTA
constexpr int kThreads = 256;

// codes[f] = 2 * row + which (0: frames, 1: frames_tp1);  out[f][c][p] = src[row][p][c] / 255
template <int C>
__global__ void __launch_bounds__(kThreads)
window_gather_kernel(const uint8_t* __restrict__ frames, const uint8_t* __restrict__ frames_tp1,
                     const long long* __restrict__ codes, float* __restrict__ out, int HW, long long n_rows) {
  __shared__ __align__(16) uint8_t tile[kPix * C];
  const long long f = blockIdx.y;
  long long code = codes[f];
  if (code < 0 || (code >> 1) >= n_rows) code = 0;   // never read outside the dataset (row 0 instead; host validates)
  const uint8_t* src = ((code & 1) ? frames_tp1 : frames) + (code >> 1) * static_cast<long long>(HW) * C;
  const int p0 = blockIdx.x * kPix;
  const int npix = min(kPix, HW - p0);
  const int nbytes = npix * C;
  const uint8_t* s = src + static_cast<long long>(p0) * C;
  if ((reinterpret_cast<uintptr_t>(s) & 15) == 0) {
    const int nvec = nbytes >> 4;
    for (int i = threadIdx.x; i < nvec; i += kThreads)
      reinterpret_cast<uint4*>(tile)[i] = __ldg(reinterpret_cast<const uint4*>(s) + i);
    for (int i = (nvec << 4) + threadIdx.x; i < nbytes; i += kThreads) tile[i] = __ldg(s + i);
  } else {
    for (int i = threadIdx.x; i < nbytes; i += kThreads) tile[i] = __ldg(s + i);
  }
  __syncthreads();
  float* o = out + f * static_cast<long long>(HW) * C + p0;
  const bool vec_ok = (HW % 4 == 0) && (p0 % 4 == 0);
#pragma unroll
  for (int c = 0; c < C; ++c) {
    float* oc = o + static_cast<long long>(c) * HW;
    if (vec_ok) {
      for (int q = threadIdx.x * 4; q < npix; q += kThreads * 4) {
        if (q + 4 <= npix) {
          float4 v;
          v.x = __fdiv_rn(static_cast<float>(tile[(q + 0) * C + c]), 255.f);
          v.y = __fdiv_rn(static_cast<float>(tile[(q + 1) * C + c]), 255.f);
          v.z = __fdiv_rn(static_cast<float>(tile[(q + 2) * C + c]), 255.f);
          v.w = __fdiv_rn(static_cast<float>(tile[(q + 3) * C + c]), 255.f);
          __stcs(reinterpret_cast<float4*>(oc + q), v);
        } else {
          for (int t = q; t < npix; ++t) oc[t] = __fdiv_rn(static_cast<float>(tile[t * C + c]), 255.f);
        }
      }
    } else {
      for (int q = threadIdx.x; q < npix; q += kThreads) oc[q] = __fdiv_rn(static_cast<float>(tile[q * C + c]), 255.f);
    }
  }
}

// Fast path (C == 3, H*W % 4 == 0, 4-byte aligned frames): no shared memory, no barrier.  A thread owns 4
// consecutive pixels: three 32-bit loads (12 contiguous bytes; a warp reads 384 contiguous bytes) and one
// 128-bit streaming store per colour plane.
__global__ void __launch_bounds__(kThreads)
window_gather_quad_kernel(const uint8_t* __restrict__ frames, const uint8_t* __restrict__ frames_tp1,
                          const long long* __restrict__ codes, float* __restrict__ out, int HW, long long n_rows) {
  const int q = blockIdx.x * kThreads + threadIdx.x;  // quad of pixels
  if (q * 4 >= HW) return;
  const long long f = blockIdx.y;
  long long code = __ldg(codes + f);
  if (code < 0 || (code >> 1) >= n_rows) code = 0;
  const uint8_t* src = ((code & 1) ? frames_tp1 : frames) + (code >> 1) * static_cast<long long>(HW) * 3;
  const uint32_t* s = reinterpret_cast<const uint32_t*>(src) + q * 3;
  const uint32_t w0 = __ldg(s), w1 = __ldg(s + 1), w2 = __ldg(s + 2);  // r0 g0 b0 r1 | g1 b1 r2 g2 | b2 r3 g3 b3
  float4 r, g, b;
  r.x = __fdiv_rn(static_cast<float>(w0 & 0xff), 255.f);
  g.x = __fdiv_rn(static_cast<float>((w0 >> 8) & 0xff), 255.f);
  b.x = __fdiv_rn(static_cast<float>((w0 >> 16) & 0xff), 255.f);
  r.y = __fdiv_rn(static_cast<float>(w0 >> 24), 255.f);
  g.y = __fdiv_rn(static_cast<float>(w1 & 0xff), 255.f);
  b.y = __fdiv_rn(static_cast<float>((w1 >> 8) & 0xff), 255.f);
  r.z = __fdiv_rn(static_cast<float>((w1 >> 16) & 0xff), 255.f);
  g.z = __fdiv_rn(static_cast<float>(w1 >> 24), 255.f);
  b.z = __fdiv_rn(static_cast<float>(w2 & 0xff), 255.f);
  r.w = __fdiv_rn(static_cast<float>((w2 >> 8) & 0xff), 255.f);
  g.w = __fdiv_rn(static_cast<float>((w2 >> 16) & 0xff), 255.f);
  b.w = __fdiv_rn(static_cast<float>(w2 >> 24), 255.f);
  float* o = out + f * static_cast<long long>(HW) * 3 + q * 4;
  __stcs(reinterpret_cast<float4*>(o), r);
  __stcs(reinterpret_cast<float4*>(o + HW), g);
  __stcs(reinterpret_cast<float4*>(o + 2 * HW), b);
}

// ---------------------------------------------------------------------------------------------------
// First SLAC encoder layer straight from the resident uint8 dataset: window gather + /255 + Conv2d(3, 32, 5, 2, 2)
// + LeakyReLU(0.2) (rlkit/torch/slac/network/latent.py:143-145) -> NHWC bf16 with a 64-channel pixel pitch
// (channels 32..63 written as zeros) so the tcgen05 implicit-GEMM kernel takes it as the next layer's operand.
// K = 75 is far too thin for UMMA: CUDA cores, one output pixel x 32 channels per thread, the
// (2*rows+3) x (W+4) x 3 input patch and the 75 x 32 weights in shared memory.  The float windows
// [B,9,3,100,100] of the reference (120 kB per frame written and re-read) never exist on this path.
constexpr int kC1Out = 32, kC1K = 5, kC1Taps = kC1K * kC1K * 3;

struct Conv1Params {
  const uint8_t* frames; const uint8_t* frames_tp1; const long long* codes;   // uint8 HWC source (codes != NULL)
  long long n_rows;                                                            // rows of frames / frames_tp1
  const float* x_nchw;                                                         // or float [n][3][H][W] in [0,1]
  const float* w;      // [32][3][5][5] (nn.Conv2d layout)
  const float* bias;   // [32]
  __nv_bfloat16* out; int out_pitch;
  int H, W, Ho, Wo, rows_per_cta;
};

__global__ void __launch_bounds__(256) slac_conv1_kernel(const Conv1Params p) {
  extern __shared__ __align__(16) float c1_smem[];
  float* wsm = c1_smem;                         // [75][32]: tap-major, channel-minor
  float* patch = c1_smem + kC1Taps * kC1Out;    // [2*rows+3][W+4][3]
  const long long f = blockIdx.y;
  const int oy0 = blockIdx.x * p.rows_per_cta;
  const int prow = 2 * p.rows_per_cta + 3, pcol = p.W + 4;
  for (int i = threadIdx.x; i < kC1Taps * kC1Out; i += 256) {
    const int co = i % kC1Out, t = i / kC1Out;          // t = (r*5 + s)*3 + c
    const int c = t % 3, rs = t / 3;
    wsm[i] = p.w[(co * 3 + c) * 25 + rs];
  }
  const int iy0 = 2 * oy0 - 2;
  const long long HW = static_cast<long long>(p.H) * p.W;
  const uint8_t* src8 = nullptr;
  if (p.codes) {
    long long code = p.codes[f];
    if (code < 0 || (code >> 1) >= p.n_rows) code = 0;
    src8 = ((code & 1) ? p.frames_tp1 : p.frames) + (code >> 1) * HW * 3;
  }
  for (int i = threadIdx.x; i < prow * pcol * 3; i += 256) {
    const int c = i % 3, col = (i / 3) % pcol, row = i / (3 * pcol);
    const int iy = iy0 + row, ix = col - 2;
    float v = 0.f;
    if (iy >= 0 && iy < p.H && ix >= 0 && ix < p.W) {
      if (src8) v = __fdiv_rn(static_cast<float>(src8[(static_cast<long long>(iy) * p.W + ix) * 3 + c]), 255.f);
      else v = p.x_nchw[(f * 3 + c) * HW + static_cast<long long>(iy) * p.W + ix];
    }
    patch[i] = v;
  }
  __syncthreads();
  const int px = threadIdx.x;
  const int oy = px / p.Wo, ox = px % p.Wo;
  if (oy >= p.rows_per_cta || oy0 + oy >= p.Ho) return;
  float acc[kC1Out];
#pragma unroll
  for (int j = 0; j < kC1Out; ++j) acc[j] = p.bias[j];
  for (int r = 0; r < kC1K; ++r) {
    const float* prow_ptr = patch + ((2 * oy + r) * pcol + 2 * ox) * 3;
#pragma unroll
    for (int sc = 0; sc < kC1K * 3; ++sc) {
      const float v = prow_ptr[sc];
      const float4* wv = reinterpret_cast<const float4*>(wsm + (r * kC1K * 3 + sc) * kC1Out);
#pragma unroll
      for (int j = 0; j < kC1Out / 4; ++j) {
        const float4 w4 = wv[j];
        acc[4 * j + 0] = fmaf(v, w4.x, acc[4 * j + 0]);
        acc[4 * j + 1] = fmaf(v, w4.y, acc[4 * j + 1]);
        acc[4 * j + 2] = fmaf(v, w4.z, acc[4 * j + 2]);
        acc[4 * j + 3] = fmaf(v, w4.w, acc[4 * j + 3]);
      }
    }
  }
  __nv_bfloat16* o = p.out + ((f * p.Ho + oy0 + oy) * p.Wo + ox) * static_cast<long long>(p.out_pitch);
#pragma unroll
  for (int j = 0; j < kC1Out / 8; ++j) {
    uint32_t pk[4];
#pragma unroll
    for (int h = 0; h < 4; ++h) {
      float a = acc[8 * j + 2 * h], b = acc[8 * j + 2 * h + 1];
      a = a > 0.f ? a : 0.2f * a;
      b = b > 0.f ? b : 0.2f * b;
      const __nv_bfloat162 t = __floats2bfloat162_rn(a, b);
      pk[h] = *reinterpret_cast<const uint32_t*>(&t);
    }
    reinterpret_cast<uint4*>(o)[j] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
  }
  for (int j = kC1Out / 8; j < 8; ++j) reinterpret_cast<uint4*>(o)[j] = make_uint4(0u, 0u, 0u, 0u);
}

}  // namespace sw
}  // namespace s2p

extern "C" int s2p_slac_window_gather(const uint8_t* frames, const uint8_t* frames_tp1, long long n_rows,
                                      const long long* codes, float* out, long long n_frames, int H, int W, int C,
                                      void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(frames && frames_tp1 && codes && out, "slac_window_gather: null pointer");
  S2P_CHECK_ARG(n_frames > 0 && n_frames <= 65535ll * 1024 && H > 0 && W > 0 && n_rows > 0, "slac_window_gather: bad dims");
  S2P_CHECK_ARG((reinterpret_cast<uintptr_t>(out) & 15) == 0, "slac_window_gather: out not 16-B aligned");
  const int HW = H * W;
  const cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int chunks = (HW + sw::kPix - 1) / sw::kPix;
  // gridDim.y is limited to 65535: split long frame lists into several launches
  for (long long f0 = 0; f0 < n_frames; f0 += 65535) {
    const long long nf = n_frames - f0 < 65535 ? n_frames - f0 : 65535;
    const dim3 grid(chunks, static_cast<unsigned>(nf));
    float* o = out + f0 * static_cast<long long>(HW) * C;
    const bool quad = C == 3 && HW % 4 == 0 && ((reinterpret_cast<uintptr_t>(frames) | reinterpret_cast<uintptr_t>(frames_tp1)) & 3) == 0;
    if (quad) {
      const dim3 qgrid((HW / 4 + sw::kThreads - 1) / sw::kThreads, static_cast<unsigned>(nf));
      sw::window_gather_quad_kernel<<<qgrid, sw::kThreads, 0, st>>>(frames, frames_tp1, codes + f0, o, HW, n_rows);
    } else if (C == 3) sw::window_gather_kernel<3><<<grid, sw::kThreads, 0, st>>>(frames, frames_tp1, codes + f0, o, HW, n_rows);
    else if (C == 1) sw::window_gather_kernel<1><<<grid, sw::kThreads, 0, st>>>(frames, frames_tp1, codes + f0, o, HW, n_rows);
    else return set_error(S2P_ERR_UNSUPPORTED, "slac_window_gather: C=%d (1 or 3)", C);
    count_launch();
    const int rc = check_launch("window_gather_kernel");
    if (rc != S2P_OK) return rc;
  }
  return S2P_OK;
}

extern "C" int s2p_slac_conv1_fwd(const uint8_t* frames, const uint8_t* frames_tp1, long long n_rows, const long long* codes,
                                  const float* x_nchw, const float* w_oihw, const float* bias, void* y, int y_pitch,
                                  long long n_frames, int H, int W, void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG((codes && frames && frames_tp1 && n_rows > 0 && !x_nchw) || (!codes && x_nchw), "slac_conv1: pass either "
                "(frames, frames_tp1, codes) or x_nchw");
  S2P_CHECK_ARG(w_oihw && bias && y, "slac_conv1: null pointer");
  S2P_CHECK_ARG(n_frames > 0 && n_frames <= 65535 && H >= 5 && W >= 5, "slac_conv1: bad dims");
  S2P_CHECK_ARG(y_pitch >= 64 && y_pitch % 8 == 0 && (reinterpret_cast<uintptr_t>(y) & 15) == 0, "slac_conv1: y_pitch / alignment");
  sw::Conv1Params p;
  p.frames = frames; p.frames_tp1 = frames_tp1; p.codes = codes; p.n_rows = n_rows; p.x_nchw = x_nchw; p.w = w_oihw; p.bias = bias;
  p.out = static_cast<__nv_bfloat16*>(y); p.out_pitch = y_pitch; p.H = H; p.W = W;
  p.Ho = (H + 4 - 5) / 2 + 1; p.Wo = (W + 4 - 5) / 2 + 1;
  S2P_CHECK_ARG(p.Wo <= 256, "slac_conv1: W=%d too wide", W);
  p.rows_per_cta = 256 / p.Wo;
  const size_t smem = (sw::kC1Taps * sw::kC1Out + static_cast<size_t>(2 * p.rows_per_cta + 3) * (W + 4) * 3) * 4;
  if (smem > 200 * 1024) return set_error(S2P_ERR_UNSUPPORTED, "slac_conv1: %zu B of shared memory needed", smem);
  static size_t configured[kMaxDevices] = {};   // per device (0 = the 48 KB default)
  const int dev = current_device();
  if (smem > 48 * 1024 && smem > configured[dev]) {
    S2P_CHECK_CUDA(cudaFuncSetAttribute(sw::slac_conv1_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured[dev] = smem;
  }
  const dim3 grid((p.Ho + p.rows_per_cta - 1) / p.rows_per_cta, static_cast<unsigned>(n_frames));
  sw::slac_conv1_kernel<<<grid, 256, smem, static_cast<cudaStream_t>(stream)>>>(p);
  count_launch();
  return check_launch("slac_conv1_kernel");
}
