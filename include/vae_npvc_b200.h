/*
 * vae_npvc_b200 - C ABI of the B200-native VQ-VAE hot path (sm_100a).
 *
 * The reference (Sinica-SLAM/vae_npvc) is pure Python/PyTorch and has NO native boundary;
 * this header is the thin layer the Python mirror classes (vae_npvc_b200/model/*.py,
 * vae_npvc_b200/trainer/basic.py) bind with ctypes.  Each entry point names the reference
 * torch-op sequence it replaces (paths relative to the reference repo root).
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless named host_*; nothing is allocated or kept
 *     by the library; every kernel is launched on `stream` (a cudaStream_t passed as void*)
 *   - return value: 0 on success, negative VNB_E_* otherwise; vnb_last_error() gives the
 *     thread-local message.  There is no CPU fallback anywhere.
 *   - activations are float32 CHANNELS-LAST: logical (B, C, T) [the reference's shape], element
 *     (b, c, t) stored at (b*T + t)*C + c, i.e. a contiguous (B, T, C) array.  The Python layer
 *     hands these out as (B, C, T) tensors with permuted strides, so the reference-facing shapes
 *     are unchanged; vnb_transpose converts a reference-layout (B, C, T)-contiguous tensor.
 *     Reason: TMA needs 16-byte-aligned innermost coordinates, so the +-1 frame shift of a k=3
 *     conv must be a ROW coordinate - frames are rows, channels are the contiguous dimension.
 *   - GEMM operands are bf16 "planes": (P, B, T, Cp) for activations, Cp = C rounded up to 8
 *     (padding never read as data);  P = 1: bf16 operands (fast path);  P = 3: hi/mid/lo split of
 *     the fp32 value, contracted with 6 plane-pair products = fp32-faithful on tensor cores
 *   - conv weights as GEMM operands: (P, taps, Cout, Cin_p) bf16, Cin_p = Cin rounded up to 8
 *   - indices are int64, like torch.argmin
 */
#ifndef VAE_NPVC_B200_H
#define VAE_NPVC_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VNB_OK 0
#define VNB_E_ARG (-1)      /* bad argument / unsupported shape */
#define VNB_E_CUDA (-2)     /* a CUDA runtime/driver call or launch failed */
#define VNB_E_NODEV (-3)    /* no sm_100 device / driver entry point missing */

#define VNB_BACKEND_AUTO 0
#define VNB_BACKEND_SIMT 1  /* fp32 CUDA-core GEMM: debug / cross-check backend */
#define VNB_BACKEND_TC 2    /* tcgen05 + TMEM + TMA GEMM */

#define VNB_ACT_NONE 0
#define VNB_ACT_LRELU 1     /* LeakyReLU(0.2) */
#define VNB_ACT_RELU 2

const char* vnb_last_error(void);
int vnb_version(void);
/* number of kernels launched by this library since the last reset (bench "gpu_launches") */
int64_t vnb_launch_count(void);
void vnb_launch_count_reset(void);
/* 0 = auto (TC where the shape allows), 1 = SIMT everywhere, 2 = TC (error if impossible) */
void vnb_set_backend(int backend);
int vnb_get_backend(void);
/* runtime options; initial value = environment variable of the same name (VNB_EW_V2, VNB_FAST_GATE, VNB_WN_V2), else the
 * compiled default.
 *   VNB_OPT_EW_V2     second-generation streaming kernels for make_planes / dec_gate / gn_bwd_apply /
 *                     gate_bwd_apply (csrc/ew_v2.cuh); same results as the first versions
 *   VNB_OPT_FAST_GATE with EW_V2 and P == 1 (bf16 operand planes) only: tanh.approx / __expf in the gate, whose
 *                     2^-11 error is below the bf16 rounding of the result; P == 3 always uses tanhf / expf
 *   VNB_OPT_WN_V2     warp-per-row weight-norm backward over the flat buffers (csrc/weightnorm.cu) */
#define VNB_OPT_EW_V2 0
#define VNB_OPT_FAST_GATE 1
#define VNB_OPT_WN_V2 2
/*   VNB_OPT_GEMM_V3   third-generation conv / weight-gradient GEMMs (CTA pairs, all operand planes of a chunk per
 *                     pipeline stage; csrc/tc_conv3.cuh): 0 off, 1 where the cost model prefers them (default), 2 always
 *   VNB_OPT_TC3_BN / _PAIR / _STAGES   tuning overrides of those kernels (0 / -1 / 0 = automatic)
 *   VNB_OPT_TC_DBG    9: the tcgen05 conv kernels record the clock64 timeline of CTA (0,0,0) (vnb_debug_tc_timeline) */
#define VNB_OPT_GEMM_V3 3
#define VNB_OPT_TC3_BN 4
#define VNB_OPT_TC3_PAIR 5
#define VNB_OPT_TC3_STAGES 6
#define VNB_OPT_TC_DBG 7
/*   VNB_OPT_TC3_SMEM_KB  cap on the dynamic shared memory of the third-generation GEMMs (0 = none): <= 110 lets two CTAs
 *                     share an SM, so that the next kernel's prologue overlaps this kernel's epilogue */
#define VNB_OPT_TC3_SMEM_KB 8
/*   VNB_OPT_FAST_EPI  conv GEMM epilogue of full tiles without residual / accumulate: bias and GroupNorm statistics in
 *                     registers on the way out of TMEM, rows stored by the bulk-copy engine.  Default 0: measured no faster
 *                     (4.2k vs 3.9k cycles for a 128 x 128 fp32 tile, profiles/r02f_fast_epilogue.txt) - the epilogue is
 *                     bound by writing 8 MB of fp32 output into L2 from 128 CTAs at once, not by how the threads do it */
#define VNB_OPT_FAST_EPI 9
/*   VNB_OPT_VQ_FUSED  fused VQ lookup (frame preparation + three-plane-product tensor-core shortlist with a rigorous error window + exact fp32 re-rank +
 *                     gather + loss sums + histogram in one kernel, csrc/tc_vq_fused.cuh; default 1) */
#define VNB_OPT_VQ_FUSED 10
/*   VNB_OPT_WG_FUSE   tap-fused weight-gradient tiles for k = 3 convs (one N-192 accumulator per 64 columns x 3 taps,
 *                     csrc/tc_wgrad.cuh; default 0: faster per launch, slower under data parallelism - DESIGN.md 5.4) */
#define VNB_OPT_WG_FUSE 11
#define VNB_OPT_COUNT_ 12
/* diagnostics: buf[slot] = %globaltimer (ns) when `stream` reaches this launch; usable inside captured graphs */
int vnb_debug_stamp(int64_t* buf, int slot, void* stream);
void vnb_set_option(int which, int value);
int vnb_get_option(int which);
/* debug: how many bounded mbarrier waits of the tcgen05 kernels timed out since load (must stay 0);
 * synchronises the device */
int vnb_debug_tc_timeouts(void);

/* ---- weight-norm materialisation: nn.utils.weight_norm pre-hook, vqvae.py:203-208,329-334
 * v: (d0,d1,k) fp32, g: (d0) fp32 or NULL (plain weight).  transposed=0: Conv1d (co=d0,ci=d1);
 * transposed=1: ConvTranspose1d stride 1 (ci=d0, co=d1, conv tap = k-1-j).
 * Optional outputs (NULL to skip): wf (P,k,co,ci_p) forward operand; wb (P,k,ci,co_p) data-grad
 * operand (taps flipped); w32 (d0,d1,k) fp32 effective weight. */
int vnb_wn_prepare(const float* v, const float* g, int d0, int d1, int k, int transposed, int P,
                   void* wf, void* wb, float* w32, void* stream);
/* backward of the same: dw (d0,d1,k) -> dv (d0,d1,k), dg (d0).   */
int vnb_wn_backward(const float* dw, const float* v, const float* g, int d0, int d1, int k,
                    float* dv, float* dg, void* stream);

/* Batched forms over the trainer's flat fp32 parameter / gradient buffers: all conv layers of the model (55 in
 * the VCC2020 recipe) in two launches (row scales g/||v||, then 64 x 64 x k tiles transposed through shared
 * memory).  desc: int64 [n_layers][12] = {v_off, g_off (-1 = plain weight), d0, d1, k, transposed, wf_off,
 * wb_off, w32_off (-1 = none), row0, tile0, 0} with element offsets into flat_p / wbuf (bf16) / w32buf;
 * row2layer: int32 [total_rows] (total_rows = sum of d0); tile2layer: int32 [total_tiles] (a layer has
 * ceil(d0/64)*ceil(d1/64) tiles, numbered row-major from tile0); max_k = largest k; row_scale: scratch of
 * total_rows floats.
 * backward: on entry the v slot of flat_g holds dw (written there by vnb_conv_wgrad), on exit dv; the
 * g slot receives dg. */
int vnb_wn_prepare_batch(const float* flat_p, const int64_t* desc, const int* row2layer, int total_rows,
                         const int* tile2layer, int total_tiles, int max_k, int P, float* row_scale,
                         void* wbuf, float* w32buf, void* stream);
int vnb_wn_backward_batch(const float* flat_p, float* flat_g, const int64_t* desc, const int* row2layer,
                          int total_rows, void* stream);

/* ---- stride-1 'same' conv as GEMM: nn.Conv1d / nn.ConvTranspose1d call sites layers.py:152,165,
 * 229,231,244; vqvae.py:155,174,256,283,285.  Also the data-gradient (pass the wb operand).
 * out[b,co,t] = beta*out + bias[b*bias_bstride+co] + residual[b,co,t]
 *             + sum_{tap,ci} W[tap][co][ci] * in[b,ci,t+(tap-(taps-1)/2)*dil]   (zero padded) */
int vnb_conv_fwd(const void* in_planes, int P, int B, int Cin, int T,
                 const void* w_planes, int taps, int dil, int Cout,
                 const float* bias, int bias_bstride, const float* residual,
                 float* out, int accumulate, double* gn_sums, int gn_groups, void* stream);
/* gn_sums (optional, (B, gn_groups, 2) doubles ZEROED by the caller): the epilogue also accumulates
 * sum and sum-of-squares of the final output values per (sample, channel group) - the GroupNorm
 * statistics of layers.py:154,201 without a second pass over the activation.
 * vnb_gn_finalize turns them into mean / rstd (biased variance, eps), n = values per (sample, group) */
/* vnb_conv_fwd with an activation backward folded into the epilogue (the data gradient of a conv whose input went
 * through LeakyReLU / ReLU: layers.py:152,165 stack, vqvae.py:281-286 final layers):
 *   out = mask_scale * conv(in) * act'(mask * mask_scale) [+ residual]      mask, residual, out: (B, T, Cout) fp32
 * = vnb_conv_fwd followed by vnb_act_bwd_add, without the extra pass over the result. */
int vnb_conv_fwd_mask(const void* in_planes, int P, int B, int Cin, int T, const void* w_planes, int taps,
                      int dil, int Cout, const float* residual, float* out, const float* mask, int mask_act,
                      float mask_scale, void* stream);
int vnb_gn_finalize(const double* gn_sums, int BG, double n, float eps, float* mean, float* rstd, void* stream);
/* weight gradient in the reference parameter layout (d0,d1,k):
 * conv tap j: dW[co][ci][j] = sum_{b,t} dy[b,co,t] * u[b,ci,t+(j-(taps-1)/2)*dil];
 * transposed=0 -> dw[co][ci][j];  transposed=1 -> dw[ci][co][taps-1-j] */
int vnb_conv_wgrad(const void* dy_planes, const void* u_planes, int P, int B, int Cout, int Cin,
                   int T, int taps, int dil, int transposed, float* dw, void* stream);

/* ---- layout conversion: (B, R, S) contiguous -> (B, S, R) contiguous (reference NCT <-> channels-last) */
int vnb_transpose(const float* in, int B, int R, int S, float* out, void* stream);
/* ---- operand planes from fp32 activations: out_id <- x*scale, out_act <- act(x*scale).
 * csum (optional, C floats ZEROED by the caller): += per-channel sums over all frames of x*scale (the bias
 * gradient of a conv whose output gradient is x, without a second pass) */
int vnb_make_planes(const float* x, int B, int C, int T, int P, float scale, int act,
                    void* out_id, void* out_act, float* csum, void* stream);
/* channel-concat [xa (B,Ca,T); xb (B,Cb,T)] -> planes (P,B,T,(Ca+Cb)p); csum (optional, Ca+Cb floats, zeroed by
 * the caller): += per-channel sums.  xa may be null (= zeros). */
int vnb_concat_planes(const float* xa, int Ca, const float* xb, int Cb, int B, int T, int P,
                      void* out, float* csum, void* stream);

/* ---- nn.GroupNorm statistics (layers.py:154,201): mean,rstd (B,groups), biased var, eps.
 * ws: scratch of vnb_gn_stats_ws_floats(B, groups) floats (chunk-wise mean/M2, merged exactly) */
int vnb_gn_stats_ws_floats(int B, int groups);
int vnb_gn_stats(const float* h, int B, int C, int T, int groups, float eps, float* ws, float* mean,
                 float* rstd, void* stream);
/* vnb_gn_stats WITHOUT a caller-provided workspace.  No counterpart in the reference: added
 * for callers that cannot allocate device memory (code under CUDA-graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  The chunk-wise workspace is a file-scope static __device__ array of
 * vnb_gn_stats_static_max_floats() = 65536 floats (256 KB, held by every GPU that loads the library until exit);
 * shapes with B * groups > 512 are rejected with VNB_E_ARG (use vnb_gn_stats).  The call allocates nothing and
 * returns what vnb_gn_stats returns (to fp32 rounding: the chunk sums use float atomics).  ALL calls share the one array (it keeps the last call's
 * contents; nothing in it needs clearing): calls that overlap on the GPU - different streams without an ordering
 * between them - race. */
int vnb_gn_stats_static_max_floats(void);
int vnb_gn_stats_static(const float* h, int B, int C, int T, int groups, float eps, float* mean, float* rstd,
                        void* stream);
/* encoder residual block epilogue (layers.py:176): y = GN1(h)*gamma+beta (+ s); optional outputs:
 * y fp32, planes of y, planes of LeakyReLU_0.2(y) */
int vnb_enc_out(const float* h, const float* s, const float* mean, const float* rstd,
                const float* gamma, const float* beta, int B, int C, int T, int P,
                float* y, void* y_planes, void* u_planes, void* stream);
/* decoder gate (layers.py:236-242): hn = GN2(a); g = tanh(hn[:C])*sigmoid(hn[C:]) -> planes of g */
int vnb_dec_gate(const float* a, const float* mean, const float* rstd, const float* gamma,
                 const float* beta, int B, int C, int T, int P, void* g_planes, void* stream);
/* vnb_enc_out / vnb_dec_gate that FINALISE the conv's fused GroupNorm statistics themselves (gn_sums as accumulated by
 * vnb_conv_fwd: (B, groups, 2) fp64 sum / sum of squares over C*T resp. C*T per half) and store (mean, rstd) for the
 * backward pass - one launch less per normalisation; same arithmetic as vnb_gn_finalize (bit-identical values) */
int vnb_enc_out_s(const float* h, const float* s, const double* gn_sums, float eps, const float* gamma,
                  const float* beta, int B, int C, int T, int P, float* y, void* y_planes, void* u_planes,
                  float* mean_out, float* rstd_out, void* stream);
int vnb_dec_gate_s(const float* a, const double* gn_sums, float eps, const float* gamma, const float* beta, int B,
                   int C, int T, int P, void* g_planes, float* mean_out, float* rstd_out, void* stream);
/* decoder residual/skip split (layers.py:246-247, vqvae.py:313): r (B,C+S,T);
 * x_out = r[:C]+x (fp32 + planes);  skip_sum (+)= r[C:]  (first=1 assigns) */
int vnb_dec_out(const float* r, const float* x, int B, int C, int S, int T, int P,
                float* x_out, void* x_planes, float* skip_sum, int first, void* stream);

/* ---- backward reductions.  rows: per (b,c) sums over t */
/* R[b,c,0..2] = sum_t dn, sum_t dn*xhat, sum_t xhat   with xhat=(h-mean)*rstd (GroupNorm bwd) */
int vnb_gn_bwd_rows(const float* dn, const float* h, const float* mean, const float* rstd,
                    int B, int C, int T, int groups, float* R, void* stream);
/* same for the gated decoder block: dn is recomputed from dg (B,C,T) through tanh*sigmoid;
 * a (B,2C,T) pre-norm */
int vnb_gate_bwd_rows(const float* dg, const float* a, const float* mean, const float* rstd,
                      const float* gamma, const float* beta, int B, int C, int T, float* R, void* stream);
/* same, told the operand mode of the step: with P == 1 (bf16 planes) and VNB_OPT_FAST_GATE the gate is re-evaluated
 * with the approximate functions the forward and the apply kernel use in that mode; P == 3 is vnb_gate_bwd_rows */
int vnb_gate_bwd_rows_p(const float* dg, const float* a, const float* mean, const float* rstd,
                        const float* gamma, const float* beta, int B, int C, int T, int P, float* R, void* stream);
/* row sums as vnb_gn_bwd_rows / vnb_gate_bwd_rows_p, ALSO producing the per-(sample, group) sums
 * S (B, groups, 2) = sum_{c in group} gamma_c * R[b,c,0..1] that vnb_gn_bwd_apply / vnb_gate_bwd_apply read, so that
 * vnb_gn_bwd_finalize (parameter gradients only) leaves the critical path.  S64: B * groups * 2 + 1 ZEROED doubles
 * (fp64 accumulators + a ticket slot); 32 must divide the channels per group. */
int vnb_gn_bwd_rows_s(const float* dn, const float* h, const float* mean, const float* rstd, const float* gamma,
                      int B, int C, int T, int groups, float* R, double* S64, float* S, void* stream);
int vnb_gate_bwd_rows_s(const float* dg, const float* a, const float* mean, const float* rstd,
                        const float* gamma, const float* beta, int B, int C, int T, int P, float* R, double* S64,
                        float* S, void* stream);
/* from R: dgamma[c], dbeta[c], S[b,group,0..1] = sum_c gamma*R0, sum_c gamma*R1, and (optional)
 * dpre_bsum[b,c] = sum_t dpre, dpre_csum[c] = sum_{b,t} dpre  (dpre = grad wrt the GN input);
 * dbeta2 / dpre_csum2: optional second copies (two biases that receive the same gradient, written
 * straight into their gradient slots) */
int vnb_gn_bwd_finalize(const float* R, const float* gamma, const float* rstd, int B, int C, int T,
                        int groups, float* dgamma, float* dbeta, float* S, float* dpre_bsum,
                        float* dpre_csum, float* dbeta2, float* dpre_csum2, void* stream);
/* dpre = rstd*(gamma*dn - S0/n - xhat*S1/n) -> planes (+ optional planes of dn itself) */
int vnb_gn_bwd_apply(const float* dn, const float* h, const float* mean, const float* rstd,
                     const float* gamma, const float* S, int B, int C, int T, int groups, int P,
                     void* dpre_planes, void* dn_planes, void* stream);
int vnb_gate_bwd_apply(const float* dg, const float* a, const float* mean, const float* rstd,
                       const float* gamma, const float* beta, const float* S, int B, int C, int T,
                       int P, void* da_planes, void* stream);
/* out = add + scale * act'(c*scale) * du   (e.g. encoder block input gradient: dskip_path +
 * lrelu'(c)*du) ; act as VNB_ACT_* ; add may be NULL */
int vnb_act_bwd_add(const float* du, const float* c, const float* add, int64_t n, int act, float scale,
                    float* out, void* stream);
/* out[c] = sum_{b,t} x[b,c,t]  (bias gradients) */
int vnb_channel_sum(const float* x, int B, int C, int T, float* out, void* stream);

/* ---- speaker conditioning (layers.py:42-54 Conditions; layers.py:231 conv_cond on a
 * time-invariant input, vqvae.py:309-312) */
int vnb_embed_fwd(const float* table, const int64_t* idx, int n, int dim, float* out, void* stream);
int vnb_embed_bwd(const float* dout, const int64_t* idx, int n, int dim, float* dtable, void* stream);
/* p[b,o] = bias[o] + bias2[o] + sum_i W[o,i]*y[b,i]   (either bias may be NULL) */
int vnb_linear_fwd(const float* y, const float* W, const float* bias, const float* bias2, int B, int I,
                   int O, float* p, void* stream);
/* dW[o,i] = sum_b dp[b,o]*y[b,i];  dy[b,i] (+)= sum_o W[o,i]*dp[b,o] */
int vnb_linear_bwd(const float* dp, const float* y, const float* W, int B, int I, int O, float* dW,
                   float* dy, int accumulate_dy, void* stream);
/* the conditioning 1x1 convs of ALL gated decoder blocks on the same time-invariant speaker embedding in one launch
 * (vqvae.py:309-312, layers.py:231).  ptrs: L x 4 int64 DEVICE addresses {W_l (O x I fp32), bias_l | 0, bias2_l | 0,
 * dW_l | 0}.  fwd: p (L, B, O) = y W_l^T + bias_l + bias2_l.  bwd: dW_l = dp[l]^T y written to ptrs[.][3] (skipped when 0);
 * dy (B, I) += sum_l dp[l] W_l, accumulated into a ZEROED buffer (NULL: skipped) */
int vnb_linear_fwd_batch(const float* y, const int64_t* ptrs, int L, int B, int I, int O, float* p, void* stream);
int vnb_linear_bwd_batch(const float* dp, const float* y, const int64_t* ptrs, int L, int B, int I, int O, float* dy,
                         void* stream);

/* ---- vector quantiser (layers_vq.py:79-150 VectorQuantizer.forward, :268-323 EMA) */
/* codebook: normalize!=0 -> E <- E/||E|| IN PLACE (embed_norm, :28-33) then e_hat = E/||E||;
 * else e_hat = E.  e_planes (3,K,Dp) hi/mid/lo, e_sq[k] = sum e_hat^2 */
int vnb_vq_prep_codebook(float* E, int K, int D, int normalize, int inplace_renorm, float* e_hat,
                         void* e_planes, float* e_sq, void* stream);
/* frames: zn = z/||z|| (or z); z_planes (3,B,T,Dp); z_sq[n], z_norm[n] with n = b*T+t */
int vnb_vq_prep_z(const float* z, int B, int D, int T, int normalize, float* zn,
                  void* z_planes, float* z_sq, float* z_norm, void* stream);
/* idx[n] = argmin_k ((z_sq[n] + e_sq[k]) - 2 zn[n].e_hat[k]), ties -> lowest k (:103-108) */
int vnb_vq_lookup(const float* zn, const void* z_planes, const float* z_sq, const float* e_hat,
                  const void* e_planes, const float* e_sq, int B, int D, int T, int K,
                  int64_t* idx, void* stream);
/* layers_vq.py:85-150 in one launch: z (B*T, D) channels-last frames -> [normalised frames zn_out, norms z_norm_out,]
 * idx (int64), gathered rows q_out, sums[0] += sum (q - zn)^2, sums[1] += sum (zn - z)^2 (normalize only), counts[idx]++.
 * e_planes: the bf16 planes of vnb_vq_prep_codebook (plane 0 is used).  Indices = the fp32 argmin of
 * (|zn|^2 + |e|^2) - 2 zn.e wherever its top-2 gap exceeds fp32 rounding (bf16 tensor-core shortlist + exact fp32
 * re-rank).  Returns 1 WITHOUT launching when the shape is not supported (caller falls back to prep_z + lookup + gather). */
int vnb_vq_fused(const float* z, int B, int D, int T, int normalize, const float* e_hat, const void* e_planes,
                 const float* e_sq, int K, float* zn_out, float* z_norm_out, int64_t* idx, float* q_out,
                 float* sums, int* counts, void* stream);
/* q = e_hat[idx] -> q_out (B,D,T); sums[0] += sum (q-zn)^2, sums[1] += sum (zn-z)^2;
 * counts[k] += #{idx==k}  (sums, counts must be zeroed by the caller).  zn may be null for a pure gather
 * (sums and z null): Model.decode, layers_vq.py:62-77 */
int vnb_vq_gather(const int64_t* idx, const float* e_hat, const float* zn, const float* z, int B, int D,
                  int T, int K, float* q_out, float* sums, int* counts, void* stream);
/* scal[0]=l_q, scal[1]=l_e, scal[2]=perplexity(eps) from sums/counts; scale = 1/(B*T) etc. */
int vnb_vq_finalize(const float* sums, const int* counts, int K, int64_t N, float scale, int normalize,
                    float eps, float* scal, void* stream);
/* backward.  g_out (B,D,T) or NULL; g_lq,g_le device scalars or NULL;
 * scale as in finalize.  dz (B,D,T); dE (K,D) must be zeroed by the caller (atomic adds) */
int vnb_vq_bwd(const float* g_out, const float* g_lq, const float* g_le, const int64_t* idx,
               const float* z, const float* zn, const float* z_norm, const float* e_hat, const float* E,
               int B, int D, int T, int K, float scale, int normalize, float* dz, float* dE_hat,
               void* stream);
/* dE = (I - e e^T) dE_hat / ||E||  (normalised codebook) ; else copy */
int vnb_vq_bwd_codebook(const float* dE_hat, const float* E, int K, int D, int normalize, float* dE,
                        void* stream);
/* EMA update (layers_vq.py:203-233) in two steps so that a data-parallel caller can all-reduce
 * zsum/counts in between.  scatter: zsum[k,:] += z frames with idx==k, counts[k] += 1 (both
 * zeroed by the caller). */
int vnb_vq_ema_scatter(const int64_t* idx, const float* z, int B, int D, int T, int K, float* zsum,
                       int* counts, void* stream);
/* update: emb_sum/emb_elem EMA, codebook <- emb_sum/emb_elem where emb_elem >= threshold else
 * rand_rows (K,D); n_frames = total frames behind counts;
 * stats[0..3] = entropy, used_curr, usage, diff_emb */
int vnb_vq_ema_update(const float* zsum, const int* counts, int K, int D, int64_t n_frames, float mu,
                      float threshold, const float* rand_rows, float* emb_sum, float* emb_elem,
                      float* embeddings, float* stats, void* stream);
/* gather rows of the frame matrix: out[k,:] = z[:, :, .] at flat frame perm[k] % N (+ optional
 * noise) - EMA init / restart candidates (:183-201,212-213) */
int vnb_vq_frame_rows(const float* z, const int64_t* perm, int B, int D, int T, int K, float* out,
                      void* stream);
/* Jitter (layers_vq.py:353-379): out[:,:,t] = in[:,:,src[t]] */
int vnb_jitter_fwd(const float* in, const int64_t* src, int B, int D, int T, float* out, void* stream);
/* grad: zero where src[t] != t */
int vnb_jitter_bwd(const float* dout, const int64_t* src, int B, int D, int T, float* din, void* stream);
/* the per-frame draws of Jitter.forward (vae_npvc/model/layers_vq.py:360-378) on the device: src[t] in {t-1, t, t+1}
 * with the reference's distribution; state = {seed, call counter} (int64[2], counter incremented by the kernel), so a
 * captured CUDA graph draws a fresh table on every replay */
int vnb_jitter_sources(int T, float p, int64_t* state, int64_t* src, void* stream);

/* ---- log-loss (layers.py:283-296): acc[0] += scale * sum 0.5*(log2pi + (xhat-x)^2) (zeroed by caller) */
int vnb_logloss_fwd(const float* xhat, const float* x, int64_t n, float scale, float* acc, void* stream);
/* dxhat = g * (xhat-x) * scale ; g device scalar (NULL = 1) */
int vnb_logloss_bwd(const float* xhat, const float* x, const float* g, int64_t n, float scale,
                    float* dxhat, void* stream);

/* ---- optimiser (trainer/basic.py:62-72): global L2 norm -> clip coef -> Adam, on flat buffers */
/* acc[0] += sum g^2 (zeroed by caller) */
int vnb_sqnorm(const float* g, int64_t n, float* acc, void* stream);
/* state[0] = Adam step count, state[1] = StepLR step count (floats, both incremented on device);
 * lr = lr*gamma^floor(state[1]/step_size) (step_size<=0: constant);
 * clip: coef = min(1, max_norm/(sqrt(sqn[0])+1e-6)) when max_norm > 0 */
int vnb_adam_step(float* p, const float* g, float* m, float* v, int64_t n, const float* sqn,
                  float max_norm, float lr, double beta1, double beta2, float eps, float* state,
                  int step_size, float gamma, void* stream);

/* ---- packed variable-length conversion / token extraction.  The reference converts one
 * utterance per forward call (decoder/basic.py:58-75, bin/extract_bnf.py:47-67); here utterances of different
 * lengths share one zero-padded (B, T, C) batch, lengths[b] (int32, device) valid frames each.
 * mask_rows: zero the frames t >= lengths[b] of an fp32 activation x (B,T,C) and of up to three operand-plane
 * buffers (P,B,T,Cp) (any of the four may be NULL) - a k=3 conv then reads zero padding past each utterance's end,
 * exactly as it does for a lone utterance (layers.py:152,229). */
int vnb_mask_rows(float* x, void* planes_a, void* planes_b, void* planes_c, int B, int C, int T, int P,
                  const int* lengths, void* stream);
/* GroupNorm over the valid frames only (layers.py:154,201 on a lone utterance): take the share of the frames
 * t >= lengths[b] of h (B,C,T) out of the (B, groups, 2) fp64 sums the conv epilogue accumulated (vnb_conv_fwd
 * gn_sums), then finalise with n = channels_per_group * lengths[b] (lengths[b] == 0: mean = rstd = 0) */
int vnb_gn_sums_unpad(const float* h, int B, int C, int T, int groups, const int* lengths, double* gn_sums,
                      void* stream);
int vnb_gn_finalize_masked(const double* gn_sums, int B, int groups, int channels_per_group, int T,
                           const int* lengths, float eps, float* mean, float* rstd, void* stream);
/* 'csid' tokens of bin/extract_bnf.py:59-60 (Tensor.unique_consecutive) for every utterance of a batch:
 * idx (B,T) int64 -> out[b, 0..out_len[b]) = ids of utterance b's first lengths[b] frames (NULL: all T) with
 * consecutive repeats collapsed, out[b, out_len[b]..T) = -1 */
int vnb_csid(const int64_t* idx, int B, int T, const int* lengths, int64_t* out, int* out_len, void* stream);

/* ---- data-parallel gradient mean over NVSwitch multicast memory (collective 1).  The reference
 * has no multi-GPU path (train.py drives one device); this replaces the library all-reduce a DDP wrapper would
 * issue.  mc_base = multicast address of the symmetric fp32 gradient buffer, [offset, offset+count) elements
 * (both multiples of 4) are summed over all ranks by the switch, scaled and broadcast back in place.  flags =
 * device array of `world` pointers to each rank's zero-initialised flag words (>= nblocks*world uint32).  Every
 * rank launches the same sequence of calls. */
int vnb_allreduce_multimem(void* mc_base, int64_t offset, int64_t count, const void* flags, int rank, int world,
                           float scale, int nblocks, void* stream);

/* ---- hierarchical VQ-VAE2 + global style tokens (vqvae2.py, layers_gst.py).  Channels-last rows.
 * The strided down-sampling convs (vqvae2.py:204-212) need no entry of their own: a stride-s, kernel-2s conv on
 * channels-last frames is vnb_conv_fwd with 3 taps over rows of s frames (functional.strided_conv1d). */
/* out[b, t, off + c] = src[b, min(t / (T / Ts), Ts - 1), c]: Model.upsample (vqvae2.py:130-143: repeat each frame
 * T // Ts times, cut or pad with the last frame) written into the channel slice [off, off + C) of the (B, T, Ct)
 * concatenation (vqvae2.py:104,110,114).  C, Ct, off multiples of 4. */
int vnb_upsample_cat_fwd(const float* src, int B, int C, int Ts, float* out, int Ct, int off, int T, void* stream);
/* dsrc[b, ts, c] = sum of dout[b, t, off + c] over the frames t that read (b, ts, c) */
int vnb_upsample_cat_bwd(const float* dout, int B, int Ct, int off, int T, float* dsrc, int C, int Ts, void* stream);
/* out[b, c] = mean_t x[b, t, c] (vqvae2.py:96) and its backward dx[b, t, c] = dy[b, c] / T */
int vnb_time_mean_fwd(const float* x, int B, int C, int T, float* out, void* stream);
int vnb_time_mean_bwd(const float* dy, int B, int C, int T, float* dx, void* stream);
/* style-token attention core (layers_gst.py:128-147, one query per utterance, no mask, dropout 0):
 * q (B, H*dk), K, V (N, H*dk) -> attn (B, H, N) = softmax_n(q_h . K_nh / sqrt(dk)), ctx (B, H*dk) = attn V.  N <= 32. */
int vnb_gst_attn_fwd(const float* q, const float* K, const float* V, int B, int N, int H, int dk, float* attn,
                     float* ctx, void* stream);
/* backward: ds (B, H, N) scratch; dq (B, H*dk); dK, dV (N, H*dk) summed over the batch (deterministic order) */
int vnb_gst_attn_bwd(const float* dctx, const float* q, const float* K, const float* V, const float* attn, int B,
                     int N, int H, int dk, float* ds, float* dq, float* dK, float* dV, void* stream);

#ifdef __cplusplus
}
#endif
#endif
