/* libs2p_sm100.so -- C ABI of the B200-native S2P generator hot path.
 *
 * The reference ships no native code and no FFI for this path: its generator
 * (`--netG=s2p`, /root/reference/README.md:33,59) is plain torch.nn modules whose
 * source is absent from the snapshot.  Each entry point below
 * therefore cites the torch.nn call it replaces inside the frozen restatement
 * (oracle/s2p_oracle.py, ORACLE_SPEC v1) and the SURVEY.md section-8 row it covers.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer owned by the caller (PyTorch allocates);
 *     kernels never allocate, free or synchronise; all work is enqueued on `stream`
 *     (a cudaStream_t passed as void*), so every call is CUDA-graph capturable;
 *   - activations: NHWC bf16, `pitch` = elements between consecutive pixels
 *     (>= C, multiple of 8) so channel slices of a wider tensor can be passed;
 *   - WIDE activations: a tensor that an instance norm or a residual add reads (the residual stream,
 *     a block's inner activation, the encoder's pre-norm maps) may be stored as a bf16 PAIR
 *     hi + lo (lo = bf16(v - hi): ~16 mantissa bits); the `*_lo` pointers below take the low halves
 *     (same pitch as the high halves; NULL = plain bf16).  Tensor-core operands are always bf16;
 *   - statistics / modulation vectors / state path: fp32;
 *   - conv weights: pre-packed [Cout][R][S][Cin] bf16 (s2p_pack_conv_weight);
 *   - return 0 on success, a negative S2P_ERR_* otherwise, text via
 *     s2p_last_error_string(); there is no CPU fallback anywhere.
 */
#ifndef S2P_B200_H_
#define S2P_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define S2P_OK 0
#define S2P_ERR_BAD_ARG (-1)
#define S2P_ERR_CUDA (-2)
#define S2P_ERR_UNSUPPORTED (-3)

#define S2P_ACT_NONE 0
#define S2P_ACT_RELU 1
#define S2P_ACT_LRELU 2 /* LeakyReLU(0.2) */

#define S2P_IMG_F32_NCHW 0 /* [-1,1] float, oracle/driver-facing */
#define S2P_IMG_U8_NHWC 1  /* uint8 HWC frames, rlkit/torch/slac/algo.py:174-175 */

int s2p_version(void);
const char* s2p_last_error_string(void);
/* number of kernel launches this library has issued in this process (bench.py's gpu_launches). */
long long s2p_launch_count(void);

/* ---- (a4) implicit-GEMM convolution on tcgen05/TMEM fed by TMA -----------------
 * replaces nn.Conv2d 3x3 s1/s2 and 1x1 (oracle: MATResBlk.conv_0/conv_1/conv_s,
 * MAT.mlp_shared/mlp_gamma/mlp_beta, ImageEncoder.conv2/conv3, S2PGenerator.head).
 * y = act(conv(x, w) + bias + residual); optionally emits per-image per-channel
 * (sum, sumsq) partials of the STORED bf16 values for the following instance norm:
 * stats_partial is [B][s2p_conv_stats_rows(Hout,Wout,Cout,ksize,stride)][Cout][2] fp32 (the row count follows the
 * tile geometry the convolution will use: 3x3 stride-1 convs with Cout <= 64 run in a horizontal-halo layout).
 * Requires Cin % 64 == 0, Cout % 64 == 0, ksize in 1..5 (3x3/1x1: generator; 4x4: discriminator; 5x5: SLAC),
 * stride in {1,2}, explicit symmetric zero padding `pad` (nn.Conv2d padding). */
int s2p_conv_stats_rows(int Hout, int Wout, int Cout, int ksize, int stride);
int s2p_conv_igemm_fwd(const void* x, int x_pitch, const void* w_packed, const float* bias,
                       const void* residual, int res_pitch, void* y, int y_pitch,
                       float* stats_partial, int B, int Hin, int Win, int Cin, int Cout, int ksize,
                       int stride, int pad, int act, void* stream);
/* the same convolution with a SECOND operand accumulated into it: y = act(conv(x, W[:, :k*k*Cin]) +
 * x2 . W[:, k*k*Cin:] + bias + residual), x2 NHWC bf16 [B][Hout][Wout][Cin2] (a 1x1 convolution over a tensor of the
 * output's size; stride 1, 'same' padding).  w_packed is [Cout][k*k*Cin + Cin2]: s2p_pack_conv_weight of the two
 * weights concatenated along the row.  Replaces `conv_1(h) + conv_s(x_s)` of the oracle's res-block (SPADEResnetBlock
 * lineage: learned shortcut) in one launch, without the bf16 round trip of the shortcut through HBM.  x2 == NULL: the
 * plain call.  residual_lo / y_lo: low halves of a wide residual / output (see Conventions); with y_lo the statistics
 * partials are those of the fp32 values.
 * cin_wrap > 0: SPLIT-PRECISION convolution.  x holds cin_wrap = 2*C channels [x_hi | x_lo] (a wide activation as one
 * tensor) and is walked as Cin = 3*C logical channels [x_hi | x_lo | x_hi] (channels >= cin_wrap re-read the tensor
 * from channel 0) against weights packed [W_hi | W_hi | W_lo] along Cin: one tensor-core pass computes the three
 * significant products of (W_hi + W_lo)(x_hi + x_lo), i.e. the convolution to ~2^-16 instead of 2^-8.  Used for the
 * image encoder's main-line convs when their weights do not fit fp16 (see operands_f16).
 * operands_f16 != 0: x (and x2) and w_packed hold fp16, not bf16 (s2p_pack_conv_weight f16 = 1): the same tcgen05
 * kind::f16 rate with an 11-bit significand.  For operands whose range is bounded by construction -- the
 * instance-normalised feature maps of the image encoder (|v| <= sqrt(H*W)), which feed conv2 / conv3 / head and
 * every MAT's mlp_shared: 13 % of the flops, upstream of every modulation. */
int s2p_conv_igemm_fwd2(const void* x, int x_pitch, const void* x2, int x2_pitch, int Cin2, const void* w_packed,
                        const float* bias, const void* residual, const void* residual_lo, int res_pitch, void* y,
                        void* y_lo, int y_pitch, float* stats_partial, int B, int Hin, int Win, int Cin, int Cout,
                        int ksize, int stride, int pad, int act, int cin_wrap, int operands_f16, void* stream);

/* ---- (a8) weight gradient of the convolution (training) -----------------------------
 * dW[co,ci,r,s] (+)= sum_{n,ho,wo} dy[n,ho,wo,co] * x[n, ho*stride-pad+r, wo*stride-pad+s, ci]
 * replaces the autograd backward of nn.Conv2d w.r.t. its weight (oracle: every conv of
 * S2PGenerator / MultiscaleDiscriminator).  x, dy: NHWC bf16; dW: fp32 OIHW (nn.Conv2d layout).
 * tcgen05 with MN-major operands straight from the NHWC tensors (no transpose pass); split-K
 * partials in `workspace` (s2p_conv_wgrad_workspace_bytes) combined in a fixed order.
 * Cin % 64 == 0, Cout % 8 == 0, ksize in 1..4, stride in {1,2}. */
long long s2p_conv_wgrad_workspace_bytes(int B, int Hin, int Win, int Cin, int Cout, int ksize, int stride,
                                         int pad);
int s2p_conv_wgrad(const void* x, int x_pitch, const void* dy, int dy_pitch, float* dw_oihw, void* workspace,
                   long long workspace_bytes, int B, int Hin, int Win, int Cin, int Cout, int ksize,
                   int stride, int pad, int accumulate, void* stream);

/* OIHW fp32 -> [Cout][R][S][Cin] bf16 (f16 = 0) or fp16 (f16 = 1) rows, each weight multiplied by `scale`
 * (used to fold the MAT blend factors at load time). */
int s2p_pack_conv_weight(const float* w_oihw, void* dst, int Cout, int Cin, int R, int S,
                         float scale, int f16, void* stream);

/* ---- (a3) param-free instance norm + MAT modulation + LeakyReLU, one fused pass ---
 * replaces F.instance_norm + the gamma/beta arithmetic of oracle MAT.forward + the
 * F.leaky_relu that follows it (and F.interpolate(nearest x2) in front when up==2).
 *   y[n,h,w,c] = act( (x[n,h/up,w/up,c]-mean)*rstd * (1 + gs[n,c] + gI[n,h,w,c]) + bs[n,c] + bI[n,h,w,c] )
 * gb_spatial holds gI in channels [0,C) and bI in [C,2C) of each pixel; gs/bs/gb_spatial
 * may be NULL (plain IN + act, oracle ImageEncoder).  C % 8 == 0.
 * x_lo: low halves of a wide x.  y_lo: the output is stored wide (with y and y_lo the two channel halves of one
 * tensor it is the [hi | lo] operand of the split-precision convolution, s2p_conv_igemm_fwd2 cin_wrap).
 * y_f16 != 0: y is stored as fp16 (the operand of an operands_f16 convolution; no y_lo then). */
int s2p_in_stats_finalize(const float* stats_partial, int rows, float* mean_rstd, int B, int C,
                          int HW, float eps, void* stream);
int s2p_mat_norm_fwd(const void* x, const void* x_lo, int x_pitch, int up, const float* mean_rstd, const float* gs,
                     const float* bs, int gsbs_pitch, const void* gb_spatial, int gb_pitch, void* y, void* y_lo,
                     int y_pitch, int B, int H, int W, int C, int act, int y_f16, void* stream);

/* ---- (a3 fused into a4) gamma/beta conv with the MAT norm applied in its epilogue ----
 * replaces, in ONE kernel, oracle MAT.mlp_gamma + MAT.mlp_beta (nn.Conv2d 3x3) and the
 * whole of F.instance_norm + modulation + F.leaky_relu (+ F.interpolate(x2) when up==2):
 * the spatial gamma/beta maps never leave TMEM.
 *   acc[pix, :] = conv(h, w_packed_gb)                         (2C columns per pixel)
 *   y[n,h,w,c]  = act( (x[n,h/up,w/up,c]*A + Bc) * (G1 + gamma_acc) + (B0 + beta_acc) )
 * with (A, Bc, G1, B0) = mod_table[n][c] from s2p_mat_prepare (A = rstd, Bc = -mean*rstd,
 * G1 = 1 + gs + bias_gamma, B0 = bs + bias_beta).
 * Weight rows are tile-interleaved: with ct = s2p_mat_fused_ct(C), rows
 * [2*ct*j, 2*ct*j+ct) are gamma channels [ct*j, ct*(j+1)) and the next ct rows the beta
 * channels of the same range (so one accumulator tile holds both for its channels).
 * bias_gb in s2p_mat_prepare is in plain [gamma(C) | beta(C)] order.
 * x reaches the epilogue through TMA (the tile's pixels as 64-channel boxes in slots of the operand ring), not
 * through per-thread global loads.  x_lo: low halves of a wide x (NULL = plain bf16); a wide x must be ONE tensor
 * [hi | lo], i.e. x_lo == x + C channels, x_pitch >= 2*C.
 * Requires Cin % 64 == 0, C == 64 or C % 128 == 0. */
int s2p_mat_fused_ct(int C);
int s2p_mat_prepare(const float* stats_partial, int rows, const float* gs, const float* bs, int gsbs_pitch,
                    const float* bias_gb, float* mod_table, int B, int C, int HW, float eps, void* stream);
int s2p_conv_mat_fused_fwd(const void* h, int h_pitch, const void* w_packed_gb, const void* x, const void* x_lo,
                           int x_pitch, int up, const float* mod_table, void* y, int y_pitch, int B, int H, int W,
                           int Cin, int C, int ksize, int act, void* stream);

/* backward of s2p_mat_norm_fwd (training; replaces autograd through F.instance_norm, the MAT
 * modulation, F.leaky_relu and F.interpolate(nearest x2) of oracle MAT / MATResBlk):
 *   d_gb  [B][H][W][2C] bf16 : gradient of the spatial (gI | bI) map          (NULL iff gb_spatial NULL)
 *   sums  [B][C][4] fp32     : (d gs, d bs, S1, S2) per (n, c)
 *   dx    [B][H/up][W/up][C] : gradient w.r.t. the pre-norm activation, plus `add` when given
 *   dxhat [B][H][W][C] bf16 and partial [B][s2p_mat_norm_bwd_chunks][C][4] fp32 are scratch. */
int s2p_mat_norm_bwd_chunks(int B, int H, int W, int C);
int s2p_mat_norm_bwd(const void* dy, int dy_pitch, const void* x, int x_pitch, int up, const float* mean_rstd,
                     const float* gs, const float* bs, int gsbs_pitch, const void* gb_spatial, int gb_pitch,
                     int act, void* d_gb, int dgb_pitch, void* dxhat, int dxh_pitch, float* partial,
                     float* sums, const void* add, int add_pitch, void* dx, int dx_pitch, int B, int H, int W,
                     int C, void* stream);

/* ---- (a5 head/tail) Cin=3 stem conv and Cout=3 ToRGB + tanh --------------------
 * stem: oracle ImageEncoder.conv1 (3->Cout 3x3 s1 p1) straight from the boundary
 * image (fp32 NCHW in [-1,1] or uint8 NHWC mapped by u/127.5-1) to NHWC bf16, with
 * statistics partials [B][s2p_stem_stats_rows(H,W)][Cout][2].
 * tail: tanh(conv3x3(Cin->3)(lrelu(x)) + bias) -> fp32 NCHW and/or uint8 NHWC
 * (floor((y+1)*127.5+0.5) clamped), oracle S2PGenerator.conv_img + to_uint8_hwc. */
int s2p_stem_stats_rows(int H, int W);
int s2p_stem_conv_fwd(const void* image, int image_fmt, const float* w_oihw, const float* bias,
                      void* y, void* y_lo, int y_pitch, float* stats_partial, int B, int H, int W, int Cout,
                      void* stream);
int s2p_torgb_tanh_fwd(const void* x, int x_pitch, const float* w_oihw, const float* bias,
                       float* out_f32_nchw, uint8_t* out_u8_nhwc, int B, int H, int W, int Cin,
                       void* stream);

/* ToRGB on the tensor cores: the same implicit-GEMM pipeline in its halo mode, with the three taps of a filter
 * row folded into the accumulator's columns: ONE 48-column UMMA group per (filter row, 64-channel chunk) over the
 * un-shifted activation box, and the epilogue adds the 16-column groups of rows m, m+1, m+2 (12 UMMAs per tile
 * instead of 36; a thin UMMA costs its A-operand read whatever N is).  x must already be activated (the producing conv
 * applies the LeakyReLU of oracle S2PGenerator.forward in its epilogue).  w_packed16 (name kept from the 16-column
 * version) is [48][3*Cin] bf16: row 16*s + c = filter column s of output channel c (c < 3, other rows zero),
 * column r*Cin + ci = filter row r of input channel ci.
 * out = tanh(conv3x3(x) + bias) -> fp32 NCHW and/or uint8 NHWC.  Cin % 64 == 0. */
int s2p_torgb_igemm_fwd(const void* x, int x_pitch, const void* w_packed16, const float* bias,
                        float* out_f32_nchw, uint8_t* out_u8_nhwc, int B, int H, int W, int Cin,
                        void* stream);

/* ---- (a2) state encoder ---------------------------------------------------------
 * pos_encode: oracle positional_encoding (nerf-pytorch embedding, include input).
 * linear: y = act(x W^T + b), fp32 (oracle StateEncoder.layers, MAT.fc_gamma/fc_beta). */
int s2p_pos_encode_fwd(const float* state, float* pe, int pe_pitch, int B, int S, int n_freqs,
                       void* stream);
/* K >= 256 runs split-K (4 ways; K/512 <= 16 ways from K = 1024, the backward of the batched state heads): the
 * layers are skinny (batch x 512) and latency-bound on a serial K walk.  Partial sums go to `workspace`
 * (s2p_linear_workspace_floats floats, 0 when no split) and are combined in a fixed order; the split is a
 * function of K only, so results never depend on the batch size.  workspace may be NULL when 0 floats are needed. */
long long s2p_linear_workspace_floats(int B, int K, int N);
int s2p_linear_fwd(const float* x, int x_pitch, const float* w, const float* bias, float* y,
                   int y_pitch, int B, int K, int N, int act, float* workspace, long long workspace_floats,
                   void* stream);

/* The WHOLE state path in one launch: gsbs = heads(MLP(pos_encode(state))), i.e. oracle StateEncoder.forward
 * (positional_encoding + (Linear + LeakyReLU(0.2)) x n_layers, width `width`) followed by every MAT's
 * fc_gamma / fc_beta batched into one Linear [heads_n, width] (no activation).  A persistent cooperative kernel
 * (grid <= SM count) whose phases (= layers) meet at a global-memory barrier; for small batches (the batch-16
 * latency case of BASELINE configs[0]/[1]) it replaces 1 + 2*(n_layers+1) launches.  Results are BIT-IDENTICAL to
 * s2p_pos_encode_fwd + s2p_linear_fwd chained (same split-K slices, same summation order), so which of the two a
 * batch size selects is invisible in the output.  layer_w / layer_b: HOST arrays of n_layers device pointers
 * (weights [width, K_l] row-major fp32, K_0 = S*(1+2*n_freqs) <= 512, width <= 512; bias may be NULL);
 * workspace: s2p_state_encoder_workspace_bytes(B, width) bytes, 256-B aligned, ZERO-INITIALISED once by the caller
 * (the kernel leaves its barrier counters zeroed) and not shared between concurrent launches. */
long long s2p_state_encoder_workspace_bytes(int B, int width);
int s2p_state_encoder_fwd(const float* state, int B, int S, int n_freqs, const float* const* layer_w,
                          const float* const* layer_b, int n_layers, int width, const float* heads_w,
                          const float* heads_b, int heads_n, float* gsbs, int gsbs_pitch, void* workspace,
                          long long workspace_bytes, void* stream);

/* s2p_state_encoder_fwd for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  No counterpart in the reference.  Same kernel,
 * same arguments minus the workspace, BIT-IDENTICAL results; the workspace is a file-scope device array of
 * S2P_STATE_ENCODER_STATIC_BYTES that every GPU loading the library holds until exit.  Limit: B * width <=
 * s2p_state_encoder_static_max_floats() = 32768 (B <= 64 at the generator's width 512); larger sizes return
 * S2P_ERR_UNSUPPORTED and launch nothing.  The call allocates nothing (no cudaMalloc / cudaFree, no library
 * allocation): one cudaMemsetAsync of the 256 B of barrier counters and one cooperative launch on `stream`.  ALL calls
 * on a device share the one array: calls that overlap on the GPU (two streams, two host threads) race -- serialise
 * them on one stream or use s2p_state_encoder_fwd with separate workspaces. */
#define S2P_STATE_ENCODER_STATIC_BYTES (256 + 2 * 32768 * 4)
long long s2p_state_encoder_static_max_floats(void);
int s2p_state_encoder_fwd_static(const float* state, int B, int S, int n_freqs, const float* const* layer_w,
                                 const float* const* layer_b, int n_layers, int width, const float* heads_w,
                                 const float* heads_b, int heads_n, float* gsbs, int gsbs_pitch, void* stream);

/* packed weight of the data-gradient convolution: dst[ci][R-1-r][S-1-s][co] = w[co][ci][r][s]
 * (bf16 rows of R*S*Cout), so dX = s2p_conv_igemm_fwd(dY (zero-inserted when stride 2), dst,
 * pad = k-1-pad) -- replaces autograd's conv backward w.r.t. the input. */
int s2p_pack_conv_weight_dgrad(const float* w_oihw, void* dst, int Cout, int Cin, int R, int S, void* stream);
/* both packed forms (s2p_pack_conv_weight with scale 1, s2p_pack_conv_weight_dgrad) of `count` (<= 48) weight
 * tensors in ONE launch; every argument is a HOST array of `count` entries, dst_dgrad[i] may be NULL when the layer
 * needs no data gradient.  Replaces the per-layer re-pack after each optimiser step of train.py (README.md:59). */
int s2p_pack_conv_weights_multi(const float* const* w_oihw, void* const* dst_fwd, void* const* dst_dgrad,
                                const int* Cout, const int* Cin, const int* R, const int* S, int count, void* stream);
/* dy_out = dy * act'(y) (act from the ACTIVATED output y; skipped when act == 0) and
 * bias_grad[c] = sum over `rows` pixels of dy_out[:, c]; partial: fp32 scratch of
 * s2p_act_bwd_bias_grad_chunks() * C floats; bias_grad/partial may both be NULL. */
int s2p_act_bwd_bias_grad_chunks(void);
int s2p_act_bwd_bias_grad(const void* dy, int dy_pitch, const void* y, int y_pitch, int act, void* dy_out,
                          int out_pitch, float* bias_grad, float* partial, long long rows, int C, void* stream);

/* ---- (a8) optimiser: torch.optim.Adam's update on fp32 tensors, one fused pass ------------ */
int s2p_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, long long n, float lr,
                  float beta1, float beta2, float eps, int step, void* stream);
/* the same update for `count` (<= 64) tensors in ONE launch: param/grad/exp_avg/exp_avg_sq/n are HOST arrays of
 * `count` device pointers / element counts (copied into the kernel's parameter space, so the call is
 * CUDA-graph capturable); step count read from device memory. */
int s2p_adam_step_multi(float* const* param, const float* const* grad, float* const* exp_avg, float* const* exp_avg_sq,
                        const long long* n, int count, float lr, float beta1, float beta2, float eps,
                        const int* step_dev, void* stream);
/* same update with the step count read from device memory (CUDA-graph replayable) */
int s2p_adam_step_dev(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, long long n, float lr,
                      float beta1, float beta2, float eps, const int* step_dev, void* stream);

/* ---- (f1) state-transition producer: the reference's Gaussian ensemble + rollout post-processing ----
 * replaces, in ONE launch over the whole dataset, EnsembleTransition.forward
 * (/root/reference/gaussian_ensemble.py:83-96; EnsembleLinear :37-48, Swish :9-11, soft_clamp :13-19;
 * mode='local', with_reward=True, hidden_features=256 as constructed at
 * state_transition_rollout.py:44-56) and the per-row arithmetic of the rollout loop
 * (state_transition_rollout.py:149 normalisation, :192-195 member selection + de-normalisation,
 * :199-204 disagreement / aleatoric uncertainty).  fp32 throughout.
 *   obs [T][obs_dim] (raw when obs_mean/obs_std given, already normalised when both NULL), action [T][action_dim]
 *   w_packed: n_members x s2p_transition_member_floats() floats, per member
 *       W0 [32][256] (rows >= obs_dim+action_dim zero) | b0 [256] | (W_l [256][256] | b_l [256]) x (hidden_layers-1)
 *       | W_out [256][64] (columns >= 2*(obs_dim+1) zero) | b_out [64]          ([in][out] = EnsembleLinear.weight[e])
 *   mean_out / std_out [n_members][T][obs_dim+1]  (Normal.mean / .stddev of forward()), both or neither
 *   member_idx [T] in [0, n_members): when given, also writes next_obs [T][obs_dim], reward [T],
 *       disagreement [T], aleatoric [T]. */
long long s2p_transition_member_floats(int obs_dim, int action_dim, int hidden_layers);
int s2p_ensemble_transition_fwd(const float* obs, int obs_pitch, const float* action, int act_pitch,
                                const float* obs_mean, const float* obs_std, const float* w_packed, int n_members,
                                const float* min_logstd, const float* max_logstd, float* mean_out, float* std_out,
                                const int* member_idx, const float* next_obs_mean, const float* next_obs_std,
                                float reward_mean, float reward_std, float* next_obs, float* reward,
                                float* disagreement, float* aleatoric, long long T, int obs_dim, int action_dim,
                                int hidden_layers, void* stream);
/* slac_observation_indices [N][num_seq+1] and slac_action_indices [N][num_seq] (int64) of
 * generate_obs_act_indices, state_transition_rollout.py:105-132: row i of a trajectory starting at
 * traj_start[i] gets i-num_seq+j, or `sentinel` (1e9) for the first num_seq rows of the trajectory. */
int s2p_slac_indices(const long long* traj_start, long long* obs_idx, long long* act_idx, long long N, int num_seq,
                     long long sentinel, void* stream);

/* ---- (f2) SLAC window assembly: gather + HWC->CHW + uint8 -> float/255 in one pass --------------
 * replaces the state half of ReplayBuffer.sample_latent / sample_sac / random_batch
 * (/root/reference/rlkit/torch/slac/buffer.py:127-147: per-sample host copies of 9 LazyFrames, upload,
 * .float().div_(255)) and the load-time np.transpose of rlkit/torch/slac/algo.py:174-175.
 * frames / frames_tp1: [n_rows][H][W][C] uint8 as stored on disk (image_observations, image_observations_tp1; a code
 * whose row is outside [0, n_rows) reads row 0 instead of faulting -- the host mirror validates its tables);
 * codes [n_frames] int64 = 2*row + which (0: frames, 1: frames_tp1); out [n_frames][C][H][W] fp32
 * (= [batch][num_seq+1][C][H][W] with n_frames = batch*(num_seq+1)).  C in {1, 3}. */
int s2p_slac_window_gather(const uint8_t* frames, const uint8_t* frames_tp1, long long n_rows, const long long* codes,
                           float* out, long long n_frames, int H, int W, int C, void* stream);

/* ---- (f3) SLAC encoder, first layer from the resident uint8 dataset --------------------------------
 * replaces nn.Conv2d(3, 32, 5, 2, 2) + nn.LeakyReLU(0.2) of Encoder.net
 * (/root/reference/rlkit/torch/slac/network/latent.py:143-145 / :126-128) fused with the window gather and
 * the /255 of ReplayBuffer.sample_latent (buffer.py:131-135): pass (frames, frames_tp1, codes) as for
 * s2p_slac_window_gather, or x_nchw = float [n_frames][3][H][W] (the reference-facing forward(x)).
 * y: NHWC bf16 [n_frames][Ho][Wo][y_pitch >= 64], channels 32..63 zero, ready for s2p_conv_igemm_fwd with
 * Cin = 64 (the remaining encoder layers, latent.py:146-160, run on that kernel; ksize up to 5 there). */
int s2p_slac_conv1_fwd(const uint8_t* frames, const uint8_t* frames_tp1, long long n_rows, const long long* codes, const float* x_nchw,
                       const float* w_oihw, const float* bias, void* y, int y_pitch, long long n_frames, int H, int W,
                       void* stream);

/* ---- (a8) 2x2 stride-2 max pool (floor mode) on NHWC bf16 and its backward -------------------------
 * replaces nn.MaxPool2d(2, 2) of the VGG19 feature extractor behind the perceptual loss of the training step
 * (/root/reference/rebuttal.md:71,131-135 "perceptual (VGG)"; oracle VGG19Features).  y: [B][H/2][W/2][C].
 * backward: dx gets dy at the first maximum of each window (nn.MaxPool2d tie rule) and zeros elsewhere,
 * including the row / column an odd H / W leaves uncovered.  C % 8 == 0. */
int s2p_maxpool2_fwd(const void* x, int x_pitch, void* y, int y_pitch, int B, int H, int W, int C, void* stream);
/* out[c] = sum over rows of a row-major fp32 [rows][cols] matrix with row pitch `pitch` (elements): the bias gradient
 * of the state encoder's Linear layers (autograd's dy.sum(0) in oracle StateEncoder / the MAT state heads). */
int s2p_colsum_rows_f32(const float* x, int pitch, int rows, int cols, float* out, void* stream);
/* nn.AvgPool2d(3, stride=2, padding=1, count_include_pad=False) on `planes` contiguous fp32 [H][W] planes
 * (x: [planes][H][W] -> y: [planes][(H-1)/2+1][(W-1)/2+1]) and its backward: the down-sampling between the scales of
 * the multiscale discriminator (oracle MultiscaleDiscriminator.forward; pix2pixHD's `downsample`). */
int s2p_avgpool3s2_fwd(const float* x, float* y, long long planes, int H, int W, void* stream);
int s2p_avgpool3s2_bwd(const float* dy, float* dx, long long planes, int H, int W, void* stream);
int s2p_maxpool2_bwd(const void* dy, int dy_pitch, const void* x, int x_pitch, void* dx, int dx_pitch, int B, int H,
                     int W, int C, void* stream);

/* ---- (a8) image layout for the tensor-core convs of the training step -------------------------------------------
 * fp32 NCHW [B][C][H][W] with C <= 8 (an image, or the discriminator's 6-channel image pair) -> bf16 NHWC
 * [B][H][W][64] with channels >= C zero, and the backward (the first C channels of a bf16 NHWC gradient -> fp32
 * NCHW).  Replaces the `permute + zero-pad` (torch.cat along the 3+61 channel axis) in front of the oracle's first
 * Conv2d of generator, discriminator and VGG19 extractor and autograd's slice / permute backward of it. */
int s2p_nchw_to_nhwc64(const float* x_nchw, void* y_nhwc64, int B, int C, int H, int W, void* stream);
int s2p_nhwc64_to_nchw(const void* dy_nhwc, int dy_pitch, float* dx_nchw, int B, int C, int H, int W, void* stream);
/* the generator's output head in training: y_nchw[b][c][h][w] = tanh(x_nhwc[b][h][w][c]) for the first C (<= 8)
 * channels of the padded ToRGB conv output, and its backward dx_nhwc64 = pad64(dy * (1 - y^2)) in bf16.  Replaces
 * slice + float + tanh + permute/contiguous (and autograd's backward of each) of `tanh(conv_img(x))`. */
int s2p_tanh_nhwc_to_nchw(const void* x_nhwc, int x_pitch, float* y_nchw, int B, int C, int H, int W, void* stream);
int s2p_tanh_bwd_nchw_to_nhwc64(const float* dy_nchw, const float* y_nchw, void* dx_nhwc64, int B, int C, int H, int W,
                                void* stream);

/* zero insertion in front of the data-gradient conv of a stride-2 layer: out [B][Hout][Wout][C] (contiguous) gets
 * dy[b][i][j] at (2i, 2j) and zeros elsewhere; Hout >= 2*Ho-1, Wout >= 2*Wo-1, C % 8 == 0. */
int s2p_zero_insert2(const void* dy, int dy_pitch, void* out, int B, int Ho, int Wo, int Hout, int Wout, int C,
                     void* stream);

/* ---- (a8) multi-tensor scale: the MAT blend factors applied to parameter tensors, and its backward ----
 * replaces the per-tensor `weight * sigmoid(blend)` of oracle MAT.forward and autograd's MulBackward for it.
 * fwd: dst[i][:] = src[i][:] * scales[scale_idx[i]] for `count` (<= 64) fp32 tensors in one launch (dst may point
 * into a concatenated operand buffer).  bwd: dsrc[i] = dout[i] * scale and partial[i][b] (b < s2p_multi_scale_blocks())
 * = per-block sums of dout[i] * src[i] in a fixed order (sum over b = gradient contribution to scales[scale_idx[i]]).
 * src/dst/dout/n/scale_idx are HOST arrays (copied into the kernel's parameter space: graph-capturable). */
int s2p_multi_scale_blocks(void);
int s2p_multi_scale_fwd(const float* const* src, float* const* dst, const long long* n, const int* scale_idx, int count,
                        const float* scales, void* stream);
int s2p_multi_scale_bwd(const float* const* dout, const float* const* src, float* const* dsrc, const long long* n,
                        const int* scale_idx, int count, const float* scales, float* partial, void* stream);

/* ---- (a8) mean |a - b| over NHWC bf16 feature maps and its backward ----------------------------------
 * replaces F.l1_loss(a.float(), b.float()) and its autograd backward in the feature-matching and VGG perceptual
 * terms (oracle generator_losses / vgg_loss).  a, b: [rows][pitch] bf16 with C live channels; out[0] = mean;
 * partial: fp32 scratch of s2p_l1_mean_chunks() floats (fixed-order reduction).  backward: da = sign(a - b) * gout[0]
 * / (rows * C) in bf16, gout read from device memory. */
int s2p_l1_mean_chunks(void);
int s2p_l1_mean_fwd(const void* a, int a_pitch, const void* b, int b_pitch, long long rows, int C, float* partial,
                    float* out, void* stream);
int s2p_l1_mean_bwd(const void* a, int a_pitch, const void* b, int b_pitch, const float* gout, long long rows, int C,
                    void* da, int da_pitch, void* stream);

/* ---- (f4) RAD image augmentations, one pass per call --------------------------------------------------
 * replaces /root/reference/examples/iql/data_augs.py (imported as `rad` by
 * rlkit/launchers/experiments/awac/finetune_rl.py:34): random_crop :8-23, random_translate :244-254,
 * random_cutout :71-89, random_cutout_color :91-113 (per-image host loops over numpy arrays) and random_flip
 * :117-146, random_rotation :150-187 (flip / rot90 copies of the whole batch blended with 0/1 masks).
 * in: [N][C][H][W], out: [N][C][Ho][Wo], elements uint8 (elem_bytes 1) or float32 (4); the per-image draws
 * are int32 DEVICE arrays of N entries, made on the host with np.random in the reference's order:
 *   S2P_RAD_CROP       p0 = h1, p1 = w1: out[y][x] = in[y+h1][x+w1]              (Ho <= H, Wo <= W)
 *   S2P_RAD_TRANSLATE  p0 = h1, p1 = w1: out[h1+y][w1+x] = in[y][x], else zero   (Ho >= H, Wo >= W)
 *   S2P_RAD_CUTOUT     p0 = h1, p1 = w1: rows [h1,2*h1) x columns [w1,2*w1) (the reference's slice, clipped to the
 *                      image) = fill[n][c] (same element type, DEVICE) or zero when fill == NULL; the rest copied
 *   S2P_RAD_FLIP       p0 = flag: flag ? in[y][W-1-x] : in[y][x]; p1 unused
 *   S2P_RAD_ROT90      p0 = quarter turns 0..3 as torch.rot90(k, [2,3]); square images; p1 unused
 * out must not alias in. */
#define S2P_RAD_CROP 0
#define S2P_RAD_TRANSLATE 1
#define S2P_RAD_CUTOUT 2
#define S2P_RAD_FLIP 3
#define S2P_RAD_ROT90 4
int s2p_rad_augment(const void* in, void* out, int elem_bytes, long long N, int C, int H, int W, int Ho, int Wo, int op,
                    const int* p0, const int* p1, const void* fill, void* stream);
/* random_grayscale (data_augs.py:26-66) with the reference's roundings: u8 = trunc(x*255); per frame (3 channels)
 * gray = trunc(u8_r*0.2989f + u8_g*0.587f + u8_b*0.114f); out = (mask[n] ? gray : u8) / 255.
 * in/out: float32 [N][frames*3][H][W]; mask: int32 DEVICE [N]. */
int s2p_rad_grayscale(const float* in, float* out, long long N, int frames, int H, int W, const int* mask, void* stream);
/* grayscale (data_augs.py:26-38), the public helper random_grayscale is built on: values taken as they are (uint8,
 * elem_bytes 1, or float32, 4), per frame gray = float(uint8(r*0.2989f + g*0.587f + b*0.114f)) (separately rounded fp32
 * products and sums, truncation), tiled over a new channel axis: out float32 [N][frames][3][H][W]. */
int s2p_rad_grayscale_plain(const void* in, int elem_bytes, float* out, long long N, int frames, int H, int W, void* stream);
/* random_convolution (data_augs.py:193-223): image n gets its own 3x3 convolution 3 -> 3 channels (padding 1, no bias;
 * weights [N][3][3][3][3] float32 = nn.Conv2d.weight per image, drawn by the host mirror with torch's RNG in the
 * reference's order), applied to each of its `frames` frames: in / out float32 [N][frames*3][H][W].  Replaces N
 * nn.Conv2d launches and N torch.cat copies of the growing result by one launch.  out must not alias in. */
int s2p_rad_random_convolution(const float* in, float* out, const float* weights, long long N, int frames, int H, int W,
                               void* stream);

#ifdef __cplusplus
}
#endif
#endif /* S2P_B200_H_ */
