// Probe for the next round's halo-tile convolution ("operand ingest"): can a K-major SWIZZLE_128B
// tcgen05 A-operand descriptor start at ANY 128-byte row of a tile that was written with the absolute-address 128B swizzle
// (the pattern TMA produces), i.e. not on a 1024-byte atom boundary?  If it can, the nine taps of a 3x3 convolution can
// read one halo tile of shared memory through nine row-shifted descriptors instead of nine TMA boxes.
//
// One CTA.  A: 192 rows x 64 bf16 (rows 128 B, chunk16 ^= row & 7 with a 1024-aligned base), B: 64 rows x 64.
// For each row shift j and each descriptor mode, D[m][n] = sum_k A[m + j][k] * B[n][k] (M=128, N=64, K=64 = 4 MMAs)
// is compared with the integer-exact host result.  mode 0: base_offset field 0; mode 1: base_offset = (start >> 7) & 7.
//
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 -I s2p_b200/csrc -o tools/_build/umma_rowshift_probe
//        tools/umma_rowshift_probe.cu          run: tools/_build/umma_rowshift_probe
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "ptx.cuh"

using namespace s2p;

constexpr int kRowsA = 192, kN = 64, kK = 64, kM = 128;

__global__ void __launch_bounds__(128, 1)
probe_kernel(const __nv_bfloat16* __restrict__ a, const __nv_bfloat16* __restrict__ b, float* __restrict__ d, int shift,
             int mode) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sa = smem;                         // 192 x 128 B
  uint8_t* sb = smem + kRowsA * 128;          // 64 x 128 B (24 KB offset: 1024-aligned)
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_base;
  const int tid = threadIdx.x, warp = tid / 32;

  for (int i = tid; i < (kRowsA + kN) * 8; i += 128) {      // 16-byte chunks, absolute-address swizzle
    const int row = i / 8, ch = i % 8;
    const bool is_a = row < kRowsA;
    const int r = is_a ? row : row - kRowsA;
    const uint4 v = *reinterpret_cast<const uint4*>((is_a ? a : b) + r * kK + ch * 8);
    uint8_t* base = is_a ? sa : sb;
    *reinterpret_cast<uint4*>(base + r * 128 + ((ch ^ (r & 7)) * 16)) = v;
  }
  if (tid == 0) {
    mbar_init(&bar, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc<64>(&tmem_base);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = tmem_base;

  if (warp == 0) {
    if (elect_one()) {
      const uint32_t a_addr = smem_u32(sa) + shift * 128;
      const uint32_t b_addr = smem_u32(sb);
      const uint32_t idesc = umma_idesc_bf16(kM, kN);
      for (int k = 0; k < kK / 16; ++k) {
        uint64_t da = umma_desc_sw128(a_addr + k * 32);
        if (mode == 1) da |= static_cast<uint64_t>((a_addr >> 7) & 7) << 49;
        const uint64_t db = umma_desc_sw128(b_addr + k * 32);
        umma_bf16(tm, da, db, idesc, k > 0 ? 1u : 0u);
      }
      umma_commit(&bar);
    }
    __syncwarp();
  }
  mbar_wait(&bar, 0);
  tc_fence_after();
  uint32_t v[32];
  for (int c = 0; c < kN; c += 32) {
    tmem_ld32(tm + (static_cast<uint32_t>(warp * 32) << 16) + c, v);
    tmem_ld_wait();
    for (int i = 0; i < 32; ++i) d[(warp * 32 + (tid % 32)) * kN + c + i] = __uint_as_float(v[i]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<64>(tm);
}

#define CK(x)                                                                      \
  do {                                                                             \
    cudaError_t e_ = (x);                                                          \
    if (e_ != cudaSuccess) {                                                       \
      printf("CUDA error %s at line %d\n", cudaGetErrorString(e_), __LINE__);      \
      return 2;                                                                    \
    }                                                                              \
  } while (0)

int main() {
  std::vector<__nv_bfloat16> ha(kRowsA * kK), hb(kN * kK);
  std::vector<float> fa(kRowsA * kK), fb(kN * kK);
  uint32_t s = 12345;
  auto rnd = [&]() { s = s * 1664525u + 1013904223u; return static_cast<float>(static_cast<int>((s >> 16) % 7) - 3); };
  for (size_t i = 0; i < ha.size(); ++i) { fa[i] = rnd(); ha[i] = __float2bfloat16(fa[i]); }
  for (size_t i = 0; i < hb.size(); ++i) { fb[i] = rnd(); hb[i] = __float2bfloat16(fb[i]); }
  __nv_bfloat16 *da, *db;
  float* dd;
  CK(cudaMalloc(&da, ha.size() * 2));
  CK(cudaMalloc(&db, hb.size() * 2));
  CK(cudaMalloc(&dd, kM * kN * 4));
  CK(cudaMemcpy(da, ha.data(), ha.size() * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(db, hb.data(), hb.size() * 2, cudaMemcpyHostToDevice));
  const int smem_bytes = (kRowsA + kN) * 128 + 1024;
  CK(cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
  const int shifts[] = {0, 8, 16, 1, 2, 3, 4, 5, 7, 9, 27, 53, 63};
  std::vector<float> hd(kM * kN);
  for (int mode = 0; mode < 2; ++mode) {
    for (int shift : shifts) {
      CK(cudaMemset(dd, 0xff, kM * kN * 4));
      probe_kernel<<<1, 128, smem_bytes>>>(da, db, dd, shift, mode);
      CK(cudaGetLastError());
      CK(cudaDeviceSynchronize());
      CK(cudaMemcpy(hd.data(), dd, kM * kN * 4, cudaMemcpyDeviceToHost));
      int bad = 0, first_bad_row = -1;
      for (int m = 0; m < kM; ++m)
        for (int n = 0; n < kN; ++n) {
          float ref = 0.f;
          for (int k = 0; k < kK; ++k) ref += fa[(m + shift) * kK + k] * fb[n * kK + k];
          if (hd[m * kN + n] != ref) {
            if (!bad) first_bad_row = m;
            ++bad;
          }
        }
      printf("mode=%d (base_offset %s) row shift %2d: %s (%d of %d wrong, first bad row %d)\n", mode,
             mode ? "= (start>>7)&7" : "= 0", shift, bad ? "MISMATCH" : "exact", bad, kM * kN, first_bad_row);
    }
  }
  return 0;
}
