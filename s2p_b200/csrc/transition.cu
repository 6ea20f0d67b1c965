// State-transition producer (SURVEY.md section 8 f, rank 1): the step immediately before the
// generator.  One fused kernel runs the reference's 7-member Gaussian ensemble
// (gaussian_ensemble.py:59-96) AND the per-row post-processing of its rollout script
// (state_transition_rollout.py:149,181-204) for a tile of dataset rows:
//
//   normalise obs -> cat(obs, action) -> E x [Linear+Swish x hidden_layers, Linear] ->
//   soft-clamped log-std, local-mode mean -> selected member de-normalised,
//   disagreement = max_e ||mu_e - mean_e mu||, aleatoric = max_e ||std_e||.
//
// fp32 on the CUDA cores, by choice: the whole producer is 2.2 MFLOP per row against the
// generator's 113 GFLOP per frame (2e-5 of the pipeline), and fp32 keeps it exact against the
// reference (1e-6) instead of TF32-close.  What matters is that nothing round-trips through HBM
// or the host: activations live in shared memory (rows are WARP-PRIVATE: warp w owns rows
// 8w..8w+7 of the 64-row tile through every layer, so there is no block-wide barrier after
// setup), the 4.4 MB of ensemble weights stream from L2 through a 4-stage TMA bulk-copy ring
// (cp.async.bulk + mbarrier, one producer warp), and the reference's per-trajectory Python loop
// (~2000 launches of 250 rows) becomes one launch over the whole dataset.
#include "ptx.cuh"
#include "s2p_internal.h"

namespace s2p {

namespace tr {
constexpr int kHidden = 256;    // hidden_features of the reference's call site (state_transition_rollout.py:46)
constexpr int kTileRows = 64;   // dataset rows per CTA
constexpr int kRowPitch = kTileRows + 4;  // conflict-free 128-bit column stores
constexpr int kInPad = 32;      // obs_dim + action_dim padded (zero weight rows)
constexpr int kOutPad = 64;     // 2 * (obs_dim + 1) padded (zero weight columns)
constexpr int kChunkK = 16;     // weight rows per ring stage
constexpr int kStages = 4;
constexpr int kConsumerWarps = 8;
constexpr int kThreads = (kConsumerWarps + 1) * 32;
constexpr int kStageFloats = kChunkK * kHidden;

struct Params {
  const float* obs; int obs_pitch;
  const float* action; int act_pitch;
  const float* obs_mean; const float* obs_std;
  const float* w; long long member_floats; int n_members;
  const float* min_logstd; const float* max_logstd;
  float* mean_out; float* std_out;
  const int* member_idx;
  const float* next_obs_mean; const float* next_obs_std; float reward_mean, reward_std;
  float* next_obs; float* reward; float* disagreement; float* aleatoric;
  long long T; int obs_dim, action_dim, hidden_layers;
};

__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(
                   smem_u32(smem_dst)),
               "l"(reinterpret_cast<uint64_t>(gsrc)), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// F.softplus(beta=1, threshold=20)
__device__ __forceinline__ float softplus(float x) { return x > 20.f ? x : log1pf(expf(x)); }

// dynamic shared memory carve-up (floats)
struct Smem {
  float* act;   // [kHidden][kRowPitch]   activations, k-major, updated in place (rows are warp-private)
  float* xin;   // [kInPad][kRowPitch]    normalised obs | action | zeros
  float* wst;   // [kStages][kChunkK][kHidden]
  float* mus;   // [E][kTileRows][od]     per-member means (kept for the ensemble statistics)
  float* sds;   // [kTileRows][od]        current member's std
  uint64_t* full;   // [kStages]
  uint64_t* empty;  // [kStages]
};

__device__ __forceinline__ Smem carve(float* base, int E, int od) {
  Smem s;
  s.act = base;
  s.xin = s.act + kHidden * kRowPitch;
  s.wst = s.xin + kInPad * kRowPitch;
  s.mus = s.wst + kStages * kStageFloats;
  s.sds = s.mus + E * kTileRows * od;
  float* end = s.sds + kTileRows * od;
  s.full = reinterpret_cast<uint64_t*>(end + ((2 - ((end - base) & 1)) & 1));
  s.empty = s.full + kStages;
  return s;
}

inline size_t smem_bytes(int E, int od) {
  const size_t floats = static_cast<size_t>(kHidden) * kRowPitch + kInPad * kRowPitch + kStages * kStageFloats +
                        static_cast<size_t>(E) * kTileRows * od + kTileRows * od + 2;
  return floats * 4 + 2 * kStages * sizeof(uint64_t);
}

__global__ void __launch_bounds__(kThreads, 1) ensemble_transition_kernel(const Params p) {
  extern __shared__ __align__(128) float smem_f[];
  const int od = p.obs_dim + 1;
  const int E = p.n_members;
  const Smem s = carve(smem_f, E, od);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const long long row0 = static_cast<long long>(blockIdx.x) * kTileRows;
  const int kin = p.obs_dim + p.action_dim;

  if (tid == 0) {
    for (int i = 0; i < kStages; ++i) {
      mbar_init(&s.full[i], 1);
      mbar_init(&s.empty[i], kConsumerWarps);
    }
    fence_barrier_init();
  }
  // input tile: x[k][row] = (obs - mean) / std | action | 0   (state_transition_rollout.py:149,180)
  for (int i = tid; i < kInPad * kTileRows; i += kThreads) {
    const int r = i / kInPad, k = i % kInPad;
    const long long row = row0 + r;
    float v = 0.f;
    if (row < p.T) {
      if (k < p.obs_dim) {
        v = p.obs[row * p.obs_pitch + k];
        if (p.obs_mean) v = (v - p.obs_mean[k]) / p.obs_std[k];
      } else if (k < kin) {
        v = p.action[row * p.act_pitch + (k - p.obs_dim)];
      }
    }
    s.xin[k * kRowPitch + r] = v;
  }
  __syncthreads();

  // chunk stream per member: layer 0 (kInPad rows) | hidden layers 1.. (kHidden rows each) | output layer
  const int chunks0 = kInPad / kChunkK, chunksH = kHidden / kChunkK;
  const int L = p.hidden_layers;

  if (warp == kConsumerWarps) {
    // ---------------- producer warp: TMA bulk copies of weight chunks into the ring
    if (lane == 0) {
      uint32_t g = 0;
      for (int e = 0; e < E; ++e) {
        const float* wm = p.w + e * p.member_floats;
        for (int layer = 0; layer <= L; ++layer) {
          const int ncols = layer == L ? kOutPad : kHidden;
          const int nchunks = layer == 0 ? chunks0 : chunksH;
          const uint32_t bytes = kChunkK * ncols * 4;
          for (int c = 0; c < nchunks; ++c, ++g) {
            const int st = g % kStages;
            if (g >= kStages) mbar_wait(&s.empty[st], ((g / kStages) - 1) & 1);
            mbar_arrive_expect_tx(&s.full[st], bytes);
            bulk_load(s.wst + st * kStageFloats, wm + static_cast<size_t>(c) * kChunkK * ncols, bytes, &s.full[st]);
          }
          wm += static_cast<size_t>(nchunks) * kChunkK * ncols + ncols;  // weights then bias
        }
      }
    }
    return;
  }

  // ---------------- consumer warps: rows rbase..rbase+7 of the tile, all columns
  const int rbase = warp * 8;
  uint32_t g = 0;
  float amax = 0.f;  // lanes 0..7: running max_e ||std_e|| of row rbase+lane
  for (int e = 0; e < E; ++e) {
    const float* wm = p.w + e * p.member_floats;
    for (int layer = 0; layer < L; ++layer) {
      const int nchunks = layer == 0 ? chunks0 : chunksH;
      const float* in = layer == 0 ? s.xin : s.act;
      float acc[8][8];  // [col j: lane + 32 j][row i]
#pragma unroll
      for (int j = 0; j < 8; ++j)
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[j][i] = 0.f;
      for (int c = 0; c < nchunks; ++c, ++g) {
        const int st = g % kStages;
        mbar_wait(&s.full[st], (g / kStages) & 1);
        const float* wt = s.wst + st * kStageFloats;
        const float* a_ptr = in + (c * kChunkK) * kRowPitch + rbase;
#pragma unroll 4
        for (int kk = 0; kk < kChunkK; ++kk) {
          const float4 a0 = *reinterpret_cast<const float4*>(a_ptr + kk * kRowPitch);
          const float4 a1 = *reinterpret_cast<const float4*>(a_ptr + kk * kRowPitch + 4);
          const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float b = wt[kk * kHidden + lane + 32 * j];
#pragma unroll
            for (int i = 0; i < 8; ++i) acc[j][i] = fmaf(a[i], b, acc[j][i]);
          }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&s.empty[st]);
      }
      // bias + Swish (gaussian_ensemble.py:9-11,86), in place: this warp's rows only
      const float* bias = wm + static_cast<size_t>(nchunks) * kChunkK * kHidden;
      __syncwarp();
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int col = lane + 32 * j;
        const float bj = __ldg(bias + col);
        float v[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float t = acc[j][i] + bj;
          v[i] = t * (1.f / (1.f + expf(-t)));
        }
        float* dst = s.act + col * kRowPitch + rbase;
        *reinterpret_cast<float4*>(dst) = make_float4(v[0], v[1], v[2], v[3]);
        *reinterpret_cast<float4*>(dst + 4) = make_float4(v[4], v[5], v[6], v[7]);
      }
      __syncwarp();
      wm = bias + kHidden;
    }
    // ---- output layer: kOutPad columns (2*od live), lane owns columns lane and lane+32
    {
      float acc[2][8];
#pragma unroll
      for (int j = 0; j < 2; ++j)
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[j][i] = 0.f;
      for (int c = 0; c < chunksH; ++c, ++g) {
        const int st = g % kStages;
        mbar_wait(&s.full[st], (g / kStages) & 1);
        const float* wt = s.wst + st * kStageFloats;
        const float* a_ptr = s.act + (c * kChunkK) * kRowPitch + rbase;
#pragma unroll 4
        for (int kk = 0; kk < kChunkK; ++kk) {
          const float4 a0 = *reinterpret_cast<const float4*>(a_ptr + kk * kRowPitch);
          const float4 a1 = *reinterpret_cast<const float4*>(a_ptr + kk * kRowPitch + 4);
          const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            const float b = wt[kk * kOutPad + lane + 32 * j];
#pragma unroll
            for (int i = 0; i < 8; ++i) acc[j][i] = fmaf(a[i], b, acc[j][i]);
          }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&s.empty[st]);
      }
      const float* bias = wm + static_cast<size_t>(chunksH) * kChunkK * kOutPad;
      float* mus_e = s.mus + static_cast<size_t>(e) * kTileRows * od;
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int col = lane + 32 * j;
        if (col >= 2 * od) continue;
        const float bj = __ldg(bias + col);
        if (col < od) {
          // mean; local mode adds the normalised observation to the predicted delta (:90-93)
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            float m = acc[j][i] + bj;
            if (col < p.obs_dim) m += s.xin[col * kRowPitch + rbase + i];
            mus_e[(rbase + i) * od + col] = m;
            const long long row = row0 + rbase + i;
            if (p.mean_out && row < p.T) p.mean_out[(static_cast<long long>(e) * p.T + row) * od + col] = m;
          }
        } else {
          const int d = col - od;
          const float lo = __ldg(p.min_logstd + d), hi = __ldg(p.max_logstd + d);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            float ls = acc[j][i] + bj;
            ls = hi - softplus(hi - ls);  // soft_clamp, gaussian_ensemble.py:13-19
            ls = lo + softplus(ls - lo);
            const float sd = expf(ls);
            s.sds[(rbase + i) * od + d] = sd;
            const long long row = row0 + rbase + i;
            if (p.std_out && row < p.T) p.std_out[(static_cast<long long>(e) * p.T + row) * od + d] = sd;
          }
        }
      }
      __syncwarp();
      if (lane < 8) {  // aleatoric: ||std_e|| over all od dims (state_transition_rollout.py:204)
        float ss = 0.f;
        for (int d = 0; d < od; ++d) {
          const float v = s.sds[(rbase + lane) * od + d];
          ss = fmaf(v, v, ss);
        }
        amax = fmaxf(amax, sqrtf(ss));
      }
      __syncwarp();
    }
  }

  // ---------------- per-row selection and ensemble statistics (state_transition_rollout.py:192-204)
  if (p.member_idx == nullptr) return;
  // ensemble average of the means into the (now free) std scratch: [row][d]
  for (int idx = lane; idx < 8 * p.obs_dim; idx += 32) {
    const int r = rbase + idx / p.obs_dim, d = idx % p.obs_dim;
    float avg = 0.f;
    for (int e = 0; e < E; ++e) avg += s.mus[(static_cast<size_t>(e) * kTileRows + r) * od + d];
    s.sds[r * od + d] = avg / static_cast<float>(E);
  }
  __syncwarp();
  const int r = rbase + (lane & 7);
  float dmax = 0.f;  // disagreement: max_e ||mu_e - avg|| over the obs dims (:199-203)
  for (int e = lane >> 3; e < E; e += 4) {
    const float* me = s.mus + (static_cast<size_t>(e) * kTileRows + r) * od;
    float ss = 0.f;
    for (int d = 0; d < p.obs_dim; ++d) {
      const float df = me[d] - s.sds[r * od + d];
      ss = fmaf(df, df, ss);
    }
    dmax = fmaxf(dmax, sqrtf(ss));
  }
  dmax = fmaxf(dmax, __shfl_xor_sync(0xffffffffu, dmax, 8));
  dmax = fmaxf(dmax, __shfl_xor_sync(0xffffffffu, dmax, 16));
  const long long row = row0 + r;
  if (lane < 8 && row < p.T) {
    const int sel = min(max(p.member_idx[row], 0), E - 1);   // a bad index must not read outside the tile's shared memory
    const float* msel = s.mus + (static_cast<size_t>(sel) * kTileRows + r) * od;
    for (int d = 0; d < p.obs_dim; ++d)
      p.next_obs[row * p.obs_dim + d] = msel[d] * p.next_obs_std[d] + p.next_obs_mean[d];  // :194
    p.reward[row] = msel[p.obs_dim] * p.reward_std + p.reward_mean;                        // :195
    p.disagreement[row] = dmax;
    p.aleatoric[row] = amax;
  }
}

// slac_observation_indices [N][num_seq+1] / slac_action_indices [N][num_seq]
// (generate_obs_act_indices, state_transition_rollout.py:105-132)
__global__ void slac_indices_kernel(const long long* __restrict__ traj_start, long long* __restrict__ obs_idx,
                                    long long* __restrict__ act_idx, long long N, int num_seq, long long sentinel) {
  const long long total = N * (num_seq + 1);
  for (long long t = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; t < total;
       t += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long i = t / (num_seq + 1);
    const int j = static_cast<int>(t % (num_seq + 1));
    const long long pos = i - traj_start[i];
    const long long v = pos < num_seq ? sentinel : i - num_seq + j;
    obs_idx[t] = v;
    if (j < num_seq) act_idx[i * num_seq + j] = v;
  }
}

}  // namespace tr
}  // namespace s2p

extern "C" long long s2p_transition_member_floats(int obs_dim, int action_dim, int hidden_layers) {
  using namespace s2p::tr;
  if (obs_dim <= 0 || action_dim < 0 || obs_dim + action_dim > kInPad || 2 * (obs_dim + 1) > kOutPad ||
      hidden_layers < 1)
    return -1;
  return static_cast<long long>(kInPad) * kHidden + kHidden +
         static_cast<long long>(hidden_layers - 1) * (kHidden * kHidden + kHidden) + kHidden * kOutPad + kOutPad;
}

extern "C" int s2p_ensemble_transition_fwd(const float* obs, int obs_pitch, const float* action, int act_pitch,
                                           const float* obs_mean, const float* obs_std, const float* w_packed,
                                           int n_members, const float* min_logstd, const float* max_logstd,
                                           float* mean_out, float* std_out, const int* member_idx,
                                           const float* next_obs_mean, const float* next_obs_std, float reward_mean,
                                           float reward_std, float* next_obs, float* reward, float* disagreement,
                                           float* aleatoric, long long T, int obs_dim, int action_dim,
                                           int hidden_layers, void* stream) {
  using namespace s2p;
  using namespace s2p::tr;
  S2P_CHECK_ARG(obs && action && w_packed && min_logstd && max_logstd, "ensemble_transition: null pointer");
  S2P_CHECK_ARG(T > 0 && obs_pitch >= obs_dim && act_pitch >= action_dim, "ensemble_transition: bad T / pitch");
  const long long mf = s2p_transition_member_floats(obs_dim, action_dim, hidden_layers);
  if (mf < 0)
    return set_error(S2P_ERR_UNSUPPORTED,
                     "ensemble_transition: obs_dim+action_dim <= %d, 2*(obs_dim+1) <= %d, hidden_layers >= 1 required",
                     kInPad, kOutPad);
  S2P_CHECK_ARG(n_members >= 1 && n_members <= 16, "ensemble_transition: n_members=%d", n_members);
  S2P_CHECK_ARG((obs_mean == nullptr) == (obs_std == nullptr), "ensemble_transition: obs_mean/obs_std");
  S2P_CHECK_ARG((mean_out == nullptr) == (std_out == nullptr), "ensemble_transition: mean_out/std_out");
  if (member_idx)
    S2P_CHECK_ARG(next_obs_mean && next_obs_std && next_obs && reward && disagreement && aleatoric,
                  "ensemble_transition: member_idx given but a selection output is NULL");
  S2P_CHECK_ARG(member_idx || mean_out, "ensemble_transition: nothing to compute");
  S2P_CHECK_ARG((reinterpret_cast<uintptr_t>(w_packed) & 15) == 0, "ensemble_transition: w_packed not 16-B aligned");
  const size_t smem = smem_bytes(n_members, obs_dim + 1);
  if (smem > 227 * 1024)
    return set_error(S2P_ERR_UNSUPPORTED, "ensemble_transition: %zu B of shared memory needed", smem);
  static bool attr_set[kMaxDevices] = {};   // per device
  const int dev = current_device();
  if (!attr_set[dev]) {
    S2P_CHECK_CUDA(cudaFuncSetAttribute(ensemble_transition_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        227 * 1024));
    attr_set[dev] = true;
  }
  Params p;
  p.obs = obs; p.obs_pitch = obs_pitch; p.action = action; p.act_pitch = act_pitch;
  p.obs_mean = obs_mean; p.obs_std = obs_std; p.w = w_packed; p.member_floats = mf; p.n_members = n_members;
  p.min_logstd = min_logstd; p.max_logstd = max_logstd; p.mean_out = mean_out; p.std_out = std_out;
  p.member_idx = member_idx; p.next_obs_mean = next_obs_mean; p.next_obs_std = next_obs_std;
  p.reward_mean = reward_mean; p.reward_std = reward_std; p.next_obs = next_obs; p.reward = reward;
  p.disagreement = disagreement; p.aleatoric = aleatoric; p.T = T; p.obs_dim = obs_dim; p.action_dim = action_dim;
  p.hidden_layers = hidden_layers;
  const long long tiles = (T + kTileRows - 1) / kTileRows;
  S2P_CHECK_ARG(tiles < (1ll << 31), "ensemble_transition: T too large for one launch");
  ensemble_transition_kernel<<<static_cast<unsigned>(tiles), kThreads, smem, static_cast<cudaStream_t>(stream)>>>(p);
  count_launch();
  return check_launch("ensemble_transition_kernel");
}

extern "C" int s2p_slac_indices(const long long* traj_start, long long* obs_idx, long long* act_idx, long long N,
                                int num_seq, long long sentinel, void* stream) {
  using namespace s2p;
  S2P_CHECK_ARG(traj_start && obs_idx && act_idx && N > 0 && num_seq >= 1, "slac_indices: bad args");
  const long long total = N * (num_seq + 1);
  const int blocks = static_cast<int>((total + 255) / 256 < 148 * 8 ? (total + 255) / 256 : 148 * 8);
  tr::slac_indices_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(traj_start, obs_idx, act_idx, N, num_seq,
                                                                              sentinel);
  count_launch();
  return check_launch("slac_indices_kernel");
}
