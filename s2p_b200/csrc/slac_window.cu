// SLAC window assembly (SURVEY.md section 8 f, rank 2): the consumer side of the generated frames.
// The reference keeps every 9-frame window as a Python list of CHW uint8 arrays on the host and, per
// sample, copies them into a staging array, uploads it and divides by 255
// (rlkit/torch/slac/buffer.py:127-147); the CHW transpose of the whole dataset happens at load time
// (rlkit/torch/slac/algo.py:174-175).  Here the dataset's frames stay in HBM exactly as stored on disk
// (uint8 HWC, 30 kB per frame: 1M rows x 2 arrays = 60 GB of the 180 GB) and ONE kernel does
// window gather + HWC->CHW + uint8 -> float/255 for a sampled batch.  Byte work, HBM-bound:
// algorithmic bytes per frame = H*W*C read + 4*H*W*C written.
#include "s2p_internal.h"

namespace s2p {
namespace sw {

constexpr int kPix = 1024;  // pixels per CTA
constexpr int kThreads = 256;

// codes[f] = 2 * row + which (0: frames, 1: frames_tp1);  out[f][c][p] = src[row][p][c] / 255
template <int C>
__global__ void __launch_bounds__(kThreads)
window_gather_kernel(const uint8_t* __restrict__ frames, const uint8_t* __restrict__ frames_tp1,
                     const long long* __restrict__ codes, float* __restrict__ out, int HW) {
  __shared__ __align__(16) uint8_t tile[kPix * C];
  const long long f = blockIdx.y;
  const long long code = codes[f];
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
                          const long long* __restrict__ codes, float* __restrict__ out, int HW) {
  const int q = blockIdx.x * kThreads + threadIdx.x;  // quad of pixels
  if (q * 4 >= HW) return;
  const long long f = blockIdx.y;
  const long long code = __ldg(codes + f);
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

}  // namespace sw
}  // namespace s2p

extern "C" int s2p_slac_window_gather(const uint8_t* frames, const uint8_t* frames_tp1, const long long* codes,
                                      float* out, long long n_frames, int H, int W, int C, void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(frames && frames_tp1 && codes && out, "slac_window_gather: null pointer");
  S2P_CHECK_ARG(n_frames > 0 && n_frames <= 65535ll * 1024 && H > 0 && W > 0, "slac_window_gather: bad dims");
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
      sw::window_gather_quad_kernel<<<qgrid, sw::kThreads, 0, st>>>(frames, frames_tp1, codes + f0, o, HW);
    } else if (C == 3) sw::window_gather_kernel<3><<<grid, sw::kThreads, 0, st>>>(frames, frames_tp1, codes + f0, o, HW);
    else if (C == 1) sw::window_gather_kernel<1><<<grid, sw::kThreads, 0, st>>>(frames, frames_tp1, codes + f0, o, HW);
    else return set_error(S2P_ERR_UNSUPPORTED, "slac_window_gather: C=%d (1 or 3)", C);
    count_launch();
    const int rc = check_launch("window_gather_kernel");
    if (rc != S2P_OK) return rc;
  }
  return S2P_OK;
}
