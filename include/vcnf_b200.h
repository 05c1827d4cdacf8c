/* libvcnf_b200.so -- C ABI of the B200 (sm_100a) bijector kernels.
 *
 * The reference (telegraphroad/VCNF, a `normflow` fork) is pure Python on PyTorch: there is no
 * FFI to bind to, the "operator interface" of the hot path is the set of ATen op chains listed
 * in SURVEY.md 2.3 (K1..K11).  Each entry point below replaces one such chain and cites it
 * (paths relative to the reference root).  Conventions:
 *   - every pointer is a DEVICE pointer valid on the current device unless named *_host;
 *   - `dtype`: 0 = float32, 1 = float64 (all floating-point buffers of one call share it);
 *   - `stream` is a cudaStream_t (NULL = legacy default stream); nothing allocates, nothing
 *     synchronises, kernels are enqueued on `stream` and the call returns immediately;
 *   - leading dimensions (`ld*`) are in ELEMENTS; rows are the batch (sample) axis;
 *   - return value: 0 on success, a positive cudaError_t from the launch, or a negative
 *     VCNF_E_* for invalid arguments.  Numerical failure is reported like the reference does,
 *     by NaN propagation, plus the optional device-side `err_flag`.
 */
#ifndef VCNF_B200_H_
#define VCNF_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VCNF_F32 0
#define VCNF_F64 1

#define VCNF_E_BADARG (-1)      /* null pointer / negative size / unknown enum */
#define VCNF_E_UNSUPPORTED (-2) /* e.g. more bins than the kernels are instantiated for */

/* Column kinds for vcnf_rqs_layer_*: col_map[c] = (kind << 16) | index_within_segment */
#define VCNF_COL_PASS 0   /* y[:, c] = x[:, c], log-det 0 */
#define VCNF_COL_COND 1   /* spline with per-sample parameters cond_params[b, index*Pc ...] */
#define VCNF_COL_SHARED 2 /* spline with batch-shared parameters shared_params[index*Ps ...] */

/* Per-feature derivative layout (tails), see vcnf_b200/csrc/rqs_math.cuh DerivMode */
#define VCNF_TAIL_LINEAR 0
#define VCNF_TAIL_CIRCULAR 1
#define VCNF_TAIL_LIST_LINEAR 2
#define VCNF_TAIL_LIST_CIRCULAR 3
#define VCNF_TAIL_NONE 4          /* tails=None: bounded spline on [0, 2 bound] (bound 0.5 = the reference's [0, 1]), K+1 free derivatives */

int vcnf_version(void);
/* Number of SMs the library sizes persistent grids for (148 on B200); 0 if no device. */
int vcnf_sm_count(void);

/* ---- K4/K5: rational-quadratic spline layer ------------------------------------------------
 * Replaces normflow/utils/splines.py:20-193 as called from
 * flows/neural_spline/coupling.py:70-125,147-159,211-240,309-343 (coupling: conditional
 * transform on `transform_features` + unconditional CDF on `identity_features`) and
 * flows/neural_spline/autoregressive.py:92-131 (all features conditional), INCLUDING the
 * index_select split, the scatter merge and utils.sum_except_batch: one launch transforms
 * whole rows and reduces the per-sample log|det|.
 *   x, y          [B, D] (ldx, ldy)             y may not alias x
 *   logdet        [B]; overwritten, or added to when accumulate != 0
 *   col_map       [D] int32, see VCNF_COL_*
 *   cond_params   [B, Tc*Pc] (ld_cond), layout feature-major: [uw(K) | uh(K) | ud(Pc-2K)]
 *                 (coupling.py:155,310-312); uw, uh are multiplied by wh_scale (=1/sqrt(hidden)
 *                 on the coupling path, coupling.py:314-316; 1 on the autoregressive path)
 *   shared_params [Ts, Ps] same layout, not scaled (coupling.py:165-246)
 *   *_modes       uint8 per segment feature (VCNF_TAIL_*), *_bounds tail bound per feature
 *   inverse       0: spline forward (splines.py:178-193), 1: inverse (splines.py:152-177)
 *   outside_zero  1: list tails -> out-of-range inputs give 0 (reference quirk, splines.py:50-57)
 *   bin_out       optional int8 [B, D]: bin index, -1 out of range, -2 pass-through column
 *   err_flag      optional device int32, bit 0 set on a negative discriminant (splines.py:164)
 */
int vcnf_rqs_layer_fwd(const void* x, int64_t ldx, void* y, int64_t ldy, void* logdet, int accumulate,
                       int64_t B, int D, int K, const int32_t* col_map,
                       const void* cond_params, int64_t ld_cond, int Tc, int Pc,
                       const uint8_t* cond_modes, const void* cond_bounds, double wh_scale,
                       const void* shared_params, int Ts, int Ps,
                       const uint8_t* shared_modes, const void* shared_bounds,
                       int inverse, int outside_zero, int8_t* bin_out, int32_t* err_flag,
                       int dtype, void* stream);

/* Bytes of scratch vcnf_rqs_layer_bwd needs for the deterministic two-stage reduction of the
 * shared-parameter gradients (0 when Ts == 0). */
int64_t vcnf_rqs_layer_bwd_workspace(int Ts, int Ps, int dtype);

/* Backward of vcnf_rqs_layer_fwd (hand-written adjoint): given gy [B, D] and
 * glogdet [B] writes gx [B, D], g_cond [B, Tc*Pc] (already multiplied by wh_scale where the
 * forward scaled) and g_shared [Ts, Ps] (summed over the batch; overwritten). */
int vcnf_rqs_layer_bwd(const void* x, int64_t ldx, const void* gy, int64_t ldgy, const void* glogdet,
                       void* gx, int64_t ldgx, int64_t B, int D, int K, const int32_t* col_map,
                       const void* cond_params, int64_t ld_cond, int Tc, int Pc,
                       const uint8_t* cond_modes, const void* cond_bounds, double wh_scale,
                       void* g_cond, int64_t ld_gcond,
                       const void* shared_params, int Ts, int Ps,
                       const uint8_t* shared_modes, const void* shared_bounds,
                       void* g_shared, void* workspace,
                       int inverse, int outside_zero, int dtype, void* stream);

/* dst[b, j] = src[b, cols[j]]  (index_select split, coupling.py:78-79; Permute shuffle,
 * flows/mixing.py:32-33,44-45).  n <= 1024. */
int vcnf_gather_cols(const void* src, int64_t lds, void* dst, int64_t ldd, int64_t B, int n,
                     const int32_t* cols, int dtype, void* stream);
/* dst[b, cols[j]] (+)= src[b, j]; columns not listed are zero-filled when accumulate == 0 and
 * D_dst > 0 (adjoint of vcnf_gather_cols). */
int vcnf_scatter_cols(const void* src, int64_t lds, void* dst, int64_t ldd, int64_t B, int n,
                      int D_dst, const int32_t* cols, int accumulate, int dtype, void* stream);

/* ---- K1: affine coupling block ---------------------------------------------------------------
 * Replaces Split -> AffineCoupling -> Merge of normflow/flows/affine/coupling.py:113-168,225-259
 * (+ flows/reshape.py:25-29,50-55) after the conditioner: copies the conditioning half, applies
 * the affine map to the other half and reduces the per-sample log|det| in one pass.
 *   z, out [B, C, S] contiguous (S = H*W, or 1 for tabular data); the transformed channels are
 *   [c2_start, c2_start + C2); param [B, np*C2, S] with shift = param[:, 0::2], raw scale =
 *   param[:, 1::2] (coupling.py:122-123), np = 1 when scale_map == 3.
 *   scale_map: 0 exp, 1 sigmoid, 2 sigmoid_inv, 3 none (shift only), 4 MAF sigmoid(u+2)+1e-3
 *   (flows/affine/autoregressive.py:76-90, use swap_ts = 1: params are (scale, shift) pairs).
 *   inverse: 0 = the layer's forward(), 1 = its inverse().  logdet [B] is overwritten. */
int vcnf_affine_coupling_fwd(const void* z, const void* param, void* out, void* logdet, int64_t B, int C, int S,
                             int c2_start, int C2, int scale_map, int inverse, int swap_ts, int dtype,
                             void* stream);
int vcnf_affine_coupling_bwd(const void* z, const void* param, const void* gout, const void* glogdet, void* gz,
                             void* gparam, int64_t B, int C, int S, int c2_start, int C2, int scale_map,
                             int inverse, int swap_ts, int dtype, void* stream);

/* ---- K2: masked affine flow (normflow/flows/affine/coupling.py:202-222) -----------------------
 * out = b*z + (1-b)*(z*exp(s)+t)  /  b*z + (1-b)*(z-t)*exp(-s); non-finite s, t become NaN
 * (coupling.py:205-208); logdet[b] = +-sum (1-b)*s.  z, s, t, out [B, n]; mask [n]; s or t may be
 * NULL (the reference's zero lambdas, coupling.py:192-198). */
int vcnf_masked_affine_fwd(const void* z, const void* mask, const void* s, const void* t, void* out, void* logdet,
                           int64_t B, int n, int inverse, int dtype, void* stream);
int vcnf_masked_affine_bwd(const void* z, const void* mask, const void* s, const void* t, const void* gout,
                           const void* glogdet, void* gz, void* gs, void* gt, int64_t B, int n, int inverse,
                           int dtype, void* stream);

/* ---- K3: constant affine / ActNorm (coupling.py:37-53, flows/normalization.py:20-38) ----------
 * out = z*exp(s)+t / (z-t)*exp(-s) with s, t [Pn] broadcast over z [B, Pn, S].  The scalar
 * log-det (+-S*sum(s)) is formed on the host side from the parameters.  The backward also
 * reduces the parameter gradients over batch and spatial positions (two-stage, deterministic). */
int vcnf_const_affine_fwd(const void* z, const void* s, const void* t, void* out, int64_t B, int Pn, int S,
                          int inverse, int dtype, void* stream);
int64_t vcnf_const_affine_bwd_workspace(int Pn, int dtype);
int vcnf_const_affine_bwd(const void* z, const void* s, const void* t, const void* gout, void* gz, void* gs,
                          void* gt, void* workspace, int64_t B, int Pn, int S, int inverse, int dtype,
                          void* stream);

/* Same result as vcnf_const_affine_bwd without a caller workspace and without any allocation (no cudaMalloc /
 * cudaFree / allocating library call), for callers that may not allocate: code under CUDA-graph capture, tight
 * serving loops, harnesses that forbid it.  NO COUNTERPART IN THE REFERENCE.
 * The per-chunk partial sums go to a file-scope `static __device__` array of
 * 592 * 2 * VCNF_CONST_AFFINE_STATIC_MAX_PN * 8 bytes (2.3 MiB) that every GPU loading the library holds until exit;
 * its device address is taken once per device (cudaGetSymbolAddress with that device current, on the first call
 * there).  Pn > VCNF_CONST_AFFINE_STATIC_MAX_PN returns VCNF_E_UNSUPPORTED.  All calls share the array and it keeps the
 * last call's contents; each call writes every slot it later reads, so nothing has to be cleared -- but two calls
 * that overlap on the GPU (different streams, or graphs replayed concurrently) RACE on it: keep them on one stream. */
#define VCNF_CONST_AFFINE_STATIC_MAX_PN 256
int vcnf_const_affine_bwd_static(const void* z, const void* s, const void* t, const void* gout, void* gz, void* gs,
                                 void* gt, int64_t B, int Pn, int S, int inverse, int dtype, void* stream);

/* dst[b, j, s] = src[b, perm[j], s]  (Permute, normflow/flows/mixing.py:32-54). */
int vcnf_permute_channels(const void* src, void* dst, int64_t B, int C, int S, const int32_t* perm, int dtype,
                          void* stream);
/* Squeeze (normflow/flows/reshape.py:93-117).  inverse == 0: src [B, 4c, h, w] -> dst [B, c, 2h, 2w];
 * inverse == 1: src [B, c, 2h, 2w] -> dst [B, 4c, h, w]. */
int vcnf_squeeze(const void* src, void* dst, int64_t B, int c_small, int h_small, int w_small, int inverse,
                 int dtype, void* stream);

/* ---- K8/K10: small dense per-row linear map ---------------------------------------------------
 * y[r, out_idx[o]] = sum_i W[o, i] * x[r, in_idx[i]] + bias[o]   (sub_bias_first: W (x - bias))
 * rows are samples (S == 1, x [rows, D]) or pixels of an NCHW tensor (S = H*W, rows = B*S).
 * LULinearPermute (normflow/flows/mixing.py:398-429,484-492) and Invertible1x1Conv
 * (mixing.py:101-128); D <= 64; W [D, D] row-major; in_idx/out_idx/bias may be NULL. */
int vcnf_small_linear(const void* x, void* y, const void* W, const void* bias, const int32_t* in_idx,
                      const int32_t* out_idx, int64_t rows, int D, int S, int sub_bias_first, int dtype,
                      void* stream);

/* Backward of vcnf_small_linear for row-major samples (S == 1, y = W x[in_idx] + bias scattered to out_idx), one
 * pass over (x, gy): gx [rows, D] (may be NULL), gW [D, D], gb [D] (each may be NULL) -- autograd of
 * normflow/flows/mixing.py:398-410 behind the permutation (mixing.py:484-492).  Per-CTA partial sums of the
 * parameter gradients are summed in a fixed order (deterministic).  `workspace` must hold
 * vcnf_small_linear_bwd_workspace(D, dtype) bytes. */
int64_t vcnf_small_linear_bwd_workspace(int D, int dtype);
int vcnf_small_linear_bwd(const void* x, const void* gy, const void* W, const int32_t* in_idx, const int32_t* out_idx,
                          void* gx, void* gW, void* gb, void* workspace, int64_t rows, int D, int dtype, void* stream);

/* Parameter gradients of the same map on image tensors x, gy [B, D, S] (autograd of Invertible1x1Conv's conv2d,
 * mixing.py:101-128): gW[o, i] = sum_{b,s} gy[b, out_idx[o], s] x[b, in_idx[i], s], gb[o] = sum gy[b, out_idx[o], s]
 * (either output may be NULL).  Operands are read in place; fixed-order reduction.  D <= 64. */
int64_t vcnf_small_linear_wgrad_img_workspace(int D, int dtype);
int vcnf_small_linear_wgrad_img(const void* x, const void* gy, const int32_t* in_idx, const int32_t* out_idx, void* gW,
                                void* gb, void* workspace, int64_t B, int D, int S, int dtype, void* stream);

/* ---- LU-parametrised weight (normflow/flows/mixing.py:352-396, _create_lower_upper / weight / logabsdet) --
 * W[D, D] = L U and *logabsdet = sum_i log diag(U)_i from the parameter vectors, one launch:
 *   lower_entries [D(D-1)/2]  strictly-lower part of L in np.tril_indices(D, -1) order (L has a unit diagonal)
 *   upper_entries [D(D-1)/2]  strictly-upper part of U in np.triu_indices(D, 1) order
 *   diag_raw      [D]         diag(U) = softplus(diag_raw) + eps   (mixing.py:382-384)
 * D <= 64.  Replaces the four index_put calls, softplus, the D x D matmul and the log-sum. */
int vcnf_lu_compose_fwd(const void* lower_entries, const void* upper_entries, const void* diag_raw, double eps, int D,
                        void* W, void* logabsdet, int dtype, void* stream);
/* Adjoint: gW [D, D] (may be NULL = zeros) and glogabsdet [1] (may be NULL) -> gradients of the three
 * parameter vectors (overwritten). */
int vcnf_lu_compose_bwd(const void* lower_entries, const void* upper_entries, const void* diag_raw, double eps, int D,
                        const void* gW, const void* glogabsdet, void* g_lower, void* g_upper, void* g_diag, int dtype,
                        void* stream);

/* The same for n <= 32 layers of equal D in one launch each way (one CTA per layer; arrays of n device pointers, read
 * on the host at call time): the LU weights of a model do not depend on its activations, so a training step forms all
 * of them before the first layer and back-propagates all of them after the last (mixing.py:352-396 x n). */
int vcnf_lu_compose_fwd_batched(const void* const* lower_entries, const void* const* upper_entries,
                                const void* const* diag_raw, double eps, int D, int n, void* const* W,
                                void* const* logabsdet, int dtype, void* stream);
/* (L U)^-1 and log|det L U| of n <= 32 layers of equal D in one launch (grid D x n): a sampling pass of a model inverts
 * every LU layer once, and none of them depends on the samples (mixing.py:412-429 x n; vcnf_lu_inverse per layer). */
int vcnf_lu_inverse_batched(const void* const* lower_entries, const void* const* upper_entries,
                            const void* const* diag_raw, double eps, int D, int n, void* const* Winv,
                            void* const* logabsdet, int dtype, void* stream);
int vcnf_lu_compose_bwd_batched(const void* const* lower_entries, const void* const* upper_entries,
                                const void* const* diag_raw, double eps, int D, int n, const void* const* gW,
                                const void* const* glogabsdet, void* const* g_lower, void* const* g_upper,
                                void* const* g_diag, int dtype, void* stream);

/* Sampling direction of the same layer (mixing.py:412-429, two triangular solves per call): Winv[D, D] =
 * U^-1 L^-1 and *logabsdet = log|det W| from the parameter vectors, one launch; substitutions in fp64 for
 * either dtype.  No gradient (callers that differentiate through the sampling direction keep the solves). */
int vcnf_lu_inverse(const void* lower_entries, const void* upper_entries, const void* diag_raw, double eps, int D,
                    void* Winv, void* logabsdet, int dtype, void* stream);

/* ---- K6/K7: conditioner dense layer on tcgen05 tensor cores, fp32-faithful (3xTF32) -------------
 * Replaces F.linear (+ ReLU / residual add / ReLU-backward mask) of the conditioner nets:
 * normflow/nets/resnet.py:38-58,76-106 (ResidualBlock / ResidualNet), nets/mlp.py:7-58,
 * nets/made.py:79-80 (after pre-masking the weight), forward and input-gradient direction.
 *   C[m, n] = [mask[m, n] > 0] * ( sum_k act(A[m, k]) * W[n*sw_n + k*sw_k] + bias[n] ) + residual[m, n]
 * A [M, K] (lda), C [M, N] (ldc), float32 only; act = ReLU when relu_in != 0 (pre-activation blocks,
 * resnet.py:49-55); bias / residual / mask may be NULL.  Forward: W = weight [N, K], sw_n = K, sw_k = 1.
 * Input gradient: A = grad-output [M, out], W = weight [out, in] read as [n = in, k = out]:
 * sw_n = 1, sw_k = in; mask = the saved pre-activation input.  K <= 128.  C must not alias A.
 * relu_in bit 1 (value 2): W / bias were written by the preceding kernel on this stream (MADE's W * mask) -- the
 * kernel then orders its W fetch after the programmatic-dependent-launch wait instead of overlapping it.
 * relu_in bit 2 (value 4): C = [mask > 0] (act(A) W^T + bias + residual) -- the mask applies after the residual (last
 * slice of an input gradient whose contraction runs over more than 128 output features: residual = C of the earlier
 * slices). */
int vcnf_linear_tf32x3(const void* A, int64_t lda, const void* W, int64_t sw_n, int64_t sw_k,
                       const void* bias, const void* residual, int64_t ldr, const void* mask, int64_t ldm,
                       void* C, int64_t ldc, int64_t M, int N, int K, int relu_in, void* stream);

/* Weight / bias gradient of the same dense layer (autograd of F.linear, the "nt" sgemm + column-sum):
 *   gW[o, i] = sum_b gY[b, o] * act(X[b, i])      gb[o] = sum_b gY[b, o]   (gb may be NULL)
 * gY [B, N] (ldg), X [B, K] (ldx), gW [N, K] contiguous, float32, K <= 128.  Split over the batch
 * across the SMs, partials summed in a fixed order (deterministic).  `workspace` must hold
 * vcnf_linear_wgrad_workspace(B, N, K) bytes. */
int64_t vcnf_linear_wgrad_workspace(int64_t B, int N, int K);
int vcnf_linear_wgrad_tf32x3(const void* gY, int64_t ldg, const void* X, int64_t ldx, void* gW, void* gb,
                             void* workspace, int64_t B, int N, int K, int relu_in, void* stream);

/* ---- K9: convolutional conditioner of the Glow coupling (normflow/nets/cnn.py:7-47: conv3x3 -> LeakyReLU ->
 * conv1x1 -> LeakyReLU -> conv3x3, called from flows/affine/glow.py:41-56; the reference runs nn.Conv2d = cuDNN) ----
 *
 * Patch matrix:  col[p, t*C + c] = act(src[b, c, h + s*dy(t), w + s*dx(t)]), zero outside the image and in the pad
 * columns [ksize^2 * C, Kp);  p = (b*H + h)*W + w, t = ky*3 + kx, (dy, dx) = (ky-1, kx-1), s = +1, or -1 when flip != 0
 * (patch matrix of the transposed convolution, i.e. of F.conv2d's input gradient).  src is fp32 with element strides
 * (sb, sc, sh, sw) -- NCHW or NHWC; act != 0 applies LeakyReLU(slope) while reading.  col is fp32 (out_bf16 = 0: the
 * fp32-faithful route feeds it to vcnf_linear_tf32x3) or bf16 (operand of vcnf_conv_gemm_bf16).  ksize is 1 or 3
 * (stride 1, padding ksize/2: what cnn.py:29-44 builds). */
int vcnf_im2col(const void* src, int64_t sb, int64_t sc, int64_t sh, int64_t sw, void* col, int64_t ldcol, int64_t B,
                int C, int H, int W, int ksize, int Kp, int flip, int act, double slope, int out_bf16, void* stream);
/* Adjoint of vcnf_im2col (fp32): gsrc[b, c, h, w] = act'(src[b, c, h, w]) * sum_t gcol[p - s*off(t), t*C + c]; every
 * element is written once (gather form, deterministic).  src is only read when act != 0. */
int vcnf_col2im(const void* gcol, int64_t ldcol, const void* src, void* gsrc, int64_t sb, int64_t sc, int64_t sh,
                int64_t sw, int64_t B, int C, int H, int W, int ksize, int flip, int act, double slope, void* stream);
/* out[c] (fp32) = sum over rows of X[., c]: bias gradients of the convolutions (autograd of nn.Conv2d's bias).
 * layout 0: X bf16 [rows, ld_or_S] row-major; layout 1: X fp32 NCHW [rows = B, C, ld_or_S = H*W].  Two stages, fixed
 * order.  `workspace`: vcnf_colsum_workspace(C) bytes. */
int64_t vcnf_colsum_workspace(int C);
int vcnf_colsum(const void* X, int layout, int64_t rows, int C, int64_t ld_or_S, void* out, void* workspace,
                void* stream);
/* All bf16 weight layouts of one ConvNet2d (nets/cnn.py:29-47: conv3x3 C0->Hd, conv1x1 Hd->Hd, conv3x3 Hd->C3; fp32
 * weights w1 [Hd,C0,3,3], w2 [Hd,Hd,1,1], w3 [C3,Hd,3,3]) in one launch.  Segments of out_bf16, in order, with
 * K1p = ceil64(9 C0), Np3 = ceil16(C3), K3p = ceil64(9 C3), Np1 = ceil16(C0), t = 3 ky + kx, zero padding:
 *   w1p [Hd][K1p] (w1p[o][t C0 + c] = w1[o][c][ky][kx]) | w2p [Hd][Hd] | w3p [9][Np3][Hd] (w3p[t][o][i] = w3[o][i][ky][kx])
 *   and, when with_bwd != 0, the operands of the input-gradient convolutions:
 *   wd3 [Hd][K3p] (wd3[i][t C3 + o] = w3[o][i][ky][kx]) | wd2 [Hd][Hd] = w2^T | wd1 [9][Np1][Hd] (wd1[t][c][o] =
 *   w1[o][c][2-ky][2-kx]).  b3p [Np3] fp32 = b3 (may be NULL) zero-padded.  vcnf_convnet3_pack_elems = length of out_bf16. */
int64_t vcnf_convnet3_pack_elems(int C0, int Hd, int C3, int with_bwd);
int vcnf_convnet3_pack(const void* w1, const void* w2, const void* w3, const void* b3, int C0, int Hd, int C3, int with_bwd,
                       void* out_bf16, void* b3p, void* stream);

/* Implicit-GEMM convolution on the tcgen05 tensor cores, bf16 operands, fp32 accumulation (replaces F.conv2d and,
 * with transposed / flipped weights, its input gradient):
 *   out[p, n] = epi( sum_{t < taps} sum_{c < Cp} act[p (+) t, c] * Wp[t][n][c] + bias[n] )
 * act [B, H, W, Cp] NHWC bf16, Cp % 64 == 0; taps = 1 (1x1, or a 3x3 whose patches were laid out by vcnf_im2col) or 9
 * (3x3, padding 1: nine shifted TMA loads of the same tensor, rows / columns outside the image read as zeros);
 * Wp [taps][Np][Cp] bf16 (Np % 16 == 0); bias fp32 [Np] or NULL.  epi: 0 none, 1 LeakyReLU(slope), 2 multiply by the
 * LeakyReLU derivative taken from `mask` = the layer's saved bf16 output [P, ldm] (> 0 ? 1 : slope).
 * out_mode 0: NHWC bf16 [P, ldo] (all Np channels); 1: NCHW fp32 [B, Cout, H, W] (channels n < Cout).
 * W must divide 128 and H*W be a multiple or a divisor of 128 / a power of two (tiles of 128 pixels). */
int vcnf_conv_gemm_bf16(const void* act, int64_t B, int H, int W, int Cp, int taps, const void* Wp, int Np,
                        const void* bias, const void* mask, int64_t ldm, double slope, int epi, void* out, int out_mode,
                        int Cout, int64_t ldo, void* stream);
/* Weight gradient of the same convolutions: G[m, n] (fp32, [Mx, Ny]) = sum_p X[p, m] * Y[p, n], X [P, ldx] and
 * Y [P, ldy] bf16 (pixel-major activations / gradients / patch matrices), Mx % 64 == 0, Ny % 64 == 0.  Split over the
 * pixels across the SMs, partial sums added in a fixed order (deterministic).  `workspace`:
 * vcnf_conv_wgrad_workspace(P, Mx, Ny) bytes. */
int64_t vcnf_conv_wgrad_workspace(int64_t P, int Mx, int Ny);
int vcnf_conv_wgrad_bf16(const void* X, int64_t ldx, int Mx, const void* Y, int64_t ldy, int Ny, int64_t P, void* G,
                         void* workspace, void* stream);

/* ---- K6 (small nets): MLP conditioner evaluated in registers, one launch per direction --------------
 * normflow/nets/mlp.py:7-58: Linear + LeakyReLU(slope) ... Linear, nl <= 4 linear layers, every width <= 32
 * (the [1,32,32,2] net of the README Real NVP, the [15,30,15] / [1,4,1] nets of the dequantisation flows).
 * W / b are HOST arrays of nl device pointers (W[l] is [dims[l+1], dims[l]] row-major, b[l] may be NULL),
 * dims a HOST array of nl+1 widths; slope >= 0.  x [B, dims[0]] (ldx), y [B, dims[nl]] (ldy). */
int vcnf_mlp_fwd(const void* x, int64_t ldx, void* y, int64_t ldy, const void* const* W, const void* const* b,
                 const int32_t* dims, int nl, double slope, int64_t B, int dtype, void* stream);
/* Backward (autograd of the nn.Sequential above): gx [B, dims[0]] (may be NULL) and the parameter gradients
 * gparams = for l = 0..nl-1: [gW_l (dims[l+1] x dims[l]) | gb_l (dims[l+1])], summed over the batch in a fixed
 * order (deterministic).  `workspace` must hold vcnf_mlp_bwd_workspace(dims, nl, dtype) bytes.  Returns
 * VCNF_E_UNSUPPORTED when the per-CTA accumulators do not fit in shared memory (wide fp64 nets). */
int64_t vcnf_mlp_bwd_workspace(const int32_t* dims, int nl, int dtype);
int vcnf_mlp_bwd(const void* x, int64_t ldx, const void* gy, int64_t ldgy, const void* const* W, const void* const* b,
                 const int32_t* dims, int nl, double slope, void* gx, int64_t ldgx, void* gparams, void* workspace,
                 int64_t B, int dtype, void* stream);

/* ---- K11: variational dequantisation of the categorical columns, all columns in one launch ---------------------
 * Replaces the per-column loop of normflow/core.py:42-54 around VariationalDequantization.forward
 * (normflow/nets/varquant.py:22-25,33-43,64-83) when every layer of the dequantisers' flow stacks is an elementwise
 * affine map on the [B, 1] column: ActNorm / AffineConstFlow (flows/affine/coupling.py:37-53), MaskedAffineFlow with a
 * one-element mask (coupling.py:202-222).  tab [n, L, 2] fp64 = (log-scale a_l, shift c_l) of layer l of column k's
 * stack, evaluated by the host from the parameters.
 *   x [B, ldx] (first n columns = the categorical values, dtype), noise [B, n] fp64 uniform(0,1) (varquant.py:70),
 *   quants [n] fp64, alpha (varquant.py:55);  zc [B, n] (dtype) = dequantised logits (core.py:48),
 *   logq_tot [B] fp32 = what core.py:52 accumulates in log_qt: column k's log-det counted (n - k) times (the
 *   accumulator is passed by reference and mutated in place), logq_run [B] fp32 (optional) = the plain sum. */
int vcnf_vdeq_fwd(const void* x, int64_t ldx, const void* noise, const void* tab, const void* quants, double alpha,
                  int64_t B, int n, int L, void* zc, void* logq_tot, void* logq_run, int dtype, void* stream);
/* Gradient of the table: g_tab [n, L, 2] fp64 from g_zc [B, n] (dtype, may be NULL) and g_logq [B] fp32 (gradient of
 * logq_tot, may be NULL); fixed-order reduction over the batch.  workspace: vcnf_vdeq_bwd_workspace(n, L) bytes. */
int64_t vcnf_vdeq_bwd_workspace(int n, int L);
int vcnf_vdeq_bwd(const void* x, int64_t ldx, const void* noise, const void* tab, const void* quants, double alpha,
                  int64_t B, int n, int L, const void* g_zc, const void* g_logq, void* g_tab, void* workspace,
                  int dtype, void* stream);
/* Sampling direction (varquant.py:26-30 driven by core.py:157-165) for all columns: out[b, k] =
 * clamp(floor(sigmoid(z[b, k]) * quants[k]), 0, quants[k] - 1); logq [B] (dtype) is updated in place the way the
 * reference does (log-dets added, then doubled after every column because the returned log-det aliases it). */
int vcnf_vdeq_decode(const void* z, int64_t ldz, const void* quants, int64_t B, int n, void* out, void* logq, int dtype,
                     void* stream);

/* ---- f1: base-distribution ends of the flow -----------------------------------------------------------------
 * Diagonal Gaussian, optionally class-conditional (normflow/distributions/base.py:609-652 DiagGaussian.forward /
 * log_prob; :715-775 ClassCondDiagGaussian), one launch per call instead of ~10 ATen launches:
 *   mode 0: logp[b] = (accum ? accum[b] : 0) - n/2 log(2 pi) - sum_j (l_j + ((zin[b,j] - loc_j) exp(-l_j))^2 / 2)
 *   mode 1: zout[b,j] = loc_j + exp(l_j) zin[b,j]  (zin = standard-normal draws), logp from the same pass
 * with loc_j = loc[j*j_stride + y[b]*c_stride], l_j = log_scale[same] + log_temp (temperature annealing,
 * base.py:636-640).  DiagGaussian: j_stride 1, c_stride 0, y NULL; ClassCondDiagGaussian (parameters
 * [*shape, classes]): j_stride = classes, c_stride 1, y int64 [B].  n = number of event elements. */
int vcnf_gauss_fwd(const void* zin, const void* loc, const void* log_scale, int64_t j_stride, int64_t c_stride,
                   const int64_t* y, double log_temp, const void* accum, void* zout, void* logp, int64_t B, int n,
                   int mode, int dtype, void* stream);
/* Backward: mode 0: gz [B, n] from g_logp [B]; mode 1: from g_zout [B, n] and g_logp.  g_loc / g_log_scale (both or
 * neither) are DENSE [classes, n] sums over the batch (fixed order); workspace: vcnf_gauss_bwd_workspace bytes. */
int64_t vcnf_gauss_bwd_workspace(int n, int classes, int dtype);
int vcnf_gauss_bwd(const void* zin, const void* loc, const void* log_scale, int64_t j_stride, int64_t c_stride,
                   const int64_t* y, double log_temp, const void* g_logp, const void* g_zout, void* gz, void* g_loc,
                   void* g_log_scale, void* workspace, int64_t B, int n, int classes, int mode, int dtype, void* stream);
/* Mixture of K diagonal Gaussians (base.py:1023-1034): logp[b] = logsumexp_k( log softmax(weight_scores)_k +
 * log N(z[b]; mu_k, exp(log_scale_k)) ) (+ accum[b]).  mu, log_scale [K, d]; weight_scores [K]. */
int vcnf_gmm_logprob(const void* z, const void* mu, const void* log_scale, const void* weight_scores,
                     const void* accum, void* logp, int64_t B, int K, int d, int dtype, void* stream);
/* Backward: gz [B, d] (may be NULL) and g_params = [g_mu (K*d) | g_log_scale (K*d) | g_weight_scores (K)], fixed-order
 * reduction over the batch; workspace: vcnf_gmm_bwd_workspace(K, d, dtype) bytes. */
int64_t vcnf_gmm_bwd_workspace(int K, int d, int dtype);
int vcnf_gmm_logprob_bwd(const void* z, const void* mu, const void* log_scale, const void* weight_scores,
                         const void* g_logp, void* gz, void* g_params, void* workspace, int64_t B, int K, int d,
                         int dtype, void* stream);

/* ---- f4 / f2: preprocessing either side of the path, checkerboard split -----------------------------------------
 * Logit transform (normflow/transforms.py:7-39, the optional MultiscaleFlow.transform; utils/preprocessing.py:5-22):
 *   to_logit = 1: u = a0 + beta z, out = log u - log(1 - u), logdet[b] = n log beta - sum_j (log u + log(1 - u))
 *   to_logit = 0: out = (sigmoid(z) - a0) / beta, logdet[b] = -n log beta + sum_j (logsigmoid(z) + logsigmoid(-z))
 * z, out [B, n]; logdet [B] or NULL.  transforms.Logit(alpha): a0 = alpha, beta = 1 - 2 alpha. */
int vcnf_logit_fwd(const void* z, void* out, void* logdet, int64_t B, int n, double a0, double beta, int to_logit,
                   int dtype, void* stream);
int vcnf_logit_bwd(const void* z, const void* gy, const void* gld, void* gz, int64_t B, int n, double a0, double beta,
                   int to_logit, int dtype, void* stream);
/* out = a x + b + c noise (noise may be NULL; out may alias x): Scale, Jitter (utils/preprocessing.py:25-56), Shift
 * (transforms.py:42-62, in place like the reference). */
int vcnf_axpb(const void* x, const void* noise, void* out, int64_t total, double a, double b, double c, int dtype,
              void* stream);
/* Split / Merge in the 'checkerboard' modes (normflow/flows/reshape.py:30-44,56-72): z [B, M, W] (M = product of the
 * dims between batch and last, W even) <-> z1, z2 [B, M, W/2]; z1 takes the positions whose non-batch indices sum to an
 * odd number (even when inv != 0).  mid_parity [M] uint8 = parity of the middle indices.  merge 0: z -> z1, z2;
 * merge 1: z1, z2 -> z. */
int vcnf_checkerboard(void* z, void* z1, void* z2, const uint8_t* mid_parity, int64_t B, int M, int W, int inv, int merge,
                      int dtype, void* stream);

/* ---- Glow's LU-parametrised 1x1 convolution weight (normflow/flows/mixing.py:83-102 `_assemble_W`, :117-118) ----
 * W[C, C] = P (tril(L, -1) + I) (triu(U, 1) + diag(sign_S exp(log_S))) and *logdet = ld_scale * sum(log_S) in one
 * launch (ld_scale = +-H W of the call site, mixing.py:123).  P, L, U are the module's full C x C tensors, sign_S /
 * log_S [C].  C <= 64.  Replaces tril, triu, diag, exp, two matmuls, sum and the scalar multiply.
 * With an_s != NULL (the ActNorm that precedes the convolution in a GlowBlock's density direction,
 * flows/affine/glow.py:59-75 + flows/normalization.py:30-38: z -> (z - t) e^-s) the two per-pixel maps are folded:
 * W[o][i] is multiplied by e^{-an_s[i]} (apply it as vcnf_small_linear(..., bias = t, sub_bias_first = 1)) and
 * *logdet = ld_scale (sum(log_S) - sum(an_s)).  Wt (may be NULL) receives the transpose (input-gradient operand). */
int vcnf_glow_w_fwd(const void* P, const void* L, const void* U, const void* sign_S, const void* log_S, const void* an_s,
                    int C, double ld_scale, void* W, void* Wt, void* logdet, int dtype, void* stream);
/* Adjoint: gW [C, C] (may be NULL) and g_logdet [1] (may be NULL) -> gL, gU [C, C] (zero outside the strict triangles,
 * like autograd through tril / triu) and g_log_S [C].  Folded form (an_s != NULL): gW is the raw product
 * sum_pixels gy[o] z[i] and gb[o] = sum_pixels gy[o] of vcnf_small_linear_wgrad_img, W2 the forward's folded weight,
 * an_t the ActNorm shift; g_an_s / g_an_t [C] receive the ActNorm parameter gradients. */
int vcnf_glow_w_bwd(const void* P, const void* L, const void* U, const void* sign_S, const void* log_S, const void* an_s,
                    const void* an_t, const void* W2, const void* gb, int C, double ld_scale, const void* gW,
                    const void* g_logdet, void* gL, void* gU, void* g_log_S, void* g_an_s, void* g_an_t, int dtype,
                    void* stream);

/* ---- K7: autoregressive layers in the sampling direction, all D passes in one launch ----------------------------
 * Replaces the D-pass loop of normflow/flows/affine/autoregressive.py:29-36 (each pass: MADE, nets/made.py:288-299, then
 * the elementwise inverse, flows/affine/autoregressive.py:92-103 or flows/neural_spline/autoregressive.py:92-131) for a
 * MADE with residual blocks, ReLU, sequential (possibly permuted) degrees: pass t evaluates only the hidden units of
 * degree t (they are final from then on) and the P output rows of the feature of degree t + 1 -- one MADE evaluation
 * in total instead of D.  Host-packed operands (fp32, device, 16 B aligned), one contiguous block per pass t so that a
 * single bulk copy brings it on chip (n_layers = 1 + 2 blocks <= 5: initial layer, then first / second of each block):
 *   band [D][(D + (n_layers-1) H) 8 + n_layers 8]: [Wb_0[t] (D x 8) | Wb_1[t] (H x 8) | ... | bb_0[t] (8) | ...] =
 *        masked weights (transposed: contraction index major) and biases of layer l's hidden units of degree t,
 *        8 slots per degree; band_idx [D][8] = their unit indices (-1 = padding); band_units = the largest number
 *        of units any degree has (slots >= band_units are padding everywhere: the kernel skips their products)
 *   fin  [D][(H + 1) Pp]: [Wf[t] (H x Pp) | bf[t] (Pp)] = the final masked layer's rows of feature order[t],
 *        transposed, Pp = P rounded up to a multiple of 4
 *   order [D]: feature fixed in pass t (ascending input degree)
 * map 0: MAF map, P = 2 (unconstrained scale, shift); map 1: RQ spline, P = 3K - 1 / 3K / 3K + 1 with per-feature tail
 * `modes` (uint8, VCNF_TAIL_*) and `bounds` (fp32).  y, x [B, D] fp32; logdet [B] = the inverse's log|det|.
 * D and H must be multiples of 16. */
int vcnf_made_inverse(const void* y, void* x, void* logdet, int64_t B, int D, int H, int P, int n_layers,
                      const void* band, const void* fin, const int32_t* band_idx, const int32_t* order, int band_units,
                      const void* bounds, const uint8_t* modes, int K, int map, int outside_zero, int32_t* err_flag,
                      void* stream);

/* ---- K10: optimiser step on flat buffers -------------------------------------------------------
 * Replaces torch.optim.Adam(...).step() of the reference's training loops (runcred.py:192,257; run.py:139:
 * lr, weight_decay, default betas / eps, amsgrad off) when parameters, gradients and both moments each live in
 * ONE contiguous buffer of n elements (16 B aligned): one launch, 28 B per fp32 parameter.
 *   g' = g + weight_decay p;  m = beta1 m + (1 - beta1) g';  v = beta2 v + (1 - beta2) g'^2;
 *   p -= lr / (1 - beta1^t) * m / (sqrt(v) / sqrt(1 - beta2^t) + eps),   t = *step + 1.
 * `step` (device fp32 [1], steps taken so far) is advanced by the kernel itself after every CTA has read it
 * (`ticket`: device uint32 [1], zero before the first call), so a captured launch replays correctly. */
int vcnf_adam_flat(void* p, const void* g, void* m, void* v, void* step, void* ticket, int64_t n, double lr,
                   double beta1, double beta2, double eps, double weight_decay, int dtype, void* stream);

/* The same step for training loops replayed from a CUDA graph / data-parallel loops: the learning rate is read
 * from a device fp64 scalar (edits take effect on the next replay), gradients are multiplied by grad_scale first
 * (1 / world_size after a summing all-reduce), and when `guard` (device scalar of guard_dtype, the step's loss) is
 * NaN or Inf the whole update is skipped -- parameters, moments and the step counter keep their values -- which is
 * what the reference's loops do on the host (`if ~(isnan(loss) | isinf(loss))`, runcred.py:249-250). */
int vcnf_adam_flat_guarded(void* p, const void* g, void* m, void* v, void* step, void* ticket, int64_t n,
                           const void* lr_dev, double beta1, double beta2, double eps, double weight_decay,
                           double grad_scale, const void* guard, int guard_dtype, int dtype, void* stream);

/* Adamax on the same flat buffers (torch.optim.Adamax of example/glow.ipynb:88): u = max(beta2 u, |g'| + eps),
 * p -= lr / (1 - beta1^t) * m / u; arguments as vcnf_adam_flat_guarded (`u` in place of `v`). */
int vcnf_adamax_flat_guarded(void* p, const void* g, void* m, void* u, void* step, void* ticket, int64_t n,
                             const void* lr_dev, double beta1, double beta2, double eps, double weight_decay,
                             double grad_scale, const void* guard, int guard_dtype, int dtype, void* stream);

/* ---- fused conditioner + spline (no-grad directions of a spline coupling layer) -----------------------------------
 * Replaces, for `sample` / `log_prob` / evaluation, the whole chain normflow/nets/resnet.py:76-106 (ResidualNet.forward:
 * initial layer, num_blocks pre-activation residual blocks, final layer) -> flows/neural_spline/coupling.py:70-96,309-343
 * (parameter reshape, 1/sqrt(hidden) scaling, utils/splines.py:88-193) with ONE launch per layer: a 128-row tile goes
 * from x to (y, log|det|) without the hidden activations or the [B, Tc * Pc] parameters touching HBM.  3xTF32 on
 * tcgen05 (fp32-faithful, see vcnf_linear_tf32x3).  Supported: fp32, hidden = 128, ReLU, no context / dropout /
 * batch norm, num_blocks <= 4, 8 bins, D <= 32, n_in <= 32, Tc <= 32 with ceil(Tc / (128 / Pc')) <= 4 (Pc' = 24 or 32);
 * anything else returns VCNF_E_UNSUPPORTED (callers keep the layer-by-layer kernels).
 *
 * vcnf_resnet_spline_pack re-tiles the net's parameters for the kernel (TF32 hi / lo planes in the tensor-core operand
 * layout, 16 KB stages): W / b are HOST arrays of 2 + 2 num_blocks DEVICE pointers in forward order
 * (initial [128, n_in]; per block the two [128, 128] layers; final [Tc * Pc, 128]), b entries may be null.
 * `packed` holds vcnf_resnet_spline_pack_bytes(...) bytes (16 B aligned) and must be re-packed when parameters change. */
int64_t vcnf_resnet_spline_pack_bytes(int n_in, int hidden, int num_blocks, int Tc, int Pc);
int vcnf_resnet_spline_pack(const void* const* W, const void* const* b, int n_in, int hidden, int num_blocks, int Tc,
                            int Pc, void* packed, void* stream);
/*   x, y [B, D] fp32 (ldx, ldy), y may not alias x; logdet [B] overwritten or added to (accumulate)
 *   id_cols [n_in], tr_cols [Tc]: HOST arrays, the columns feeding the net (in the net's input order) and the
 *     transformed columns (feature f of the parameters acts on column tr_cols[f]); every other column and the id
 *     columns are copied; modes (VCNF_TAIL_*) / bounds: HOST arrays per transformed feature
 *   wh_scale, inverse, outside_zero, err_flag: as vcnf_rqs_layer_fwd
 *   shared_params [Ts, Ps] (device; Ts = 0 or n_in): the unconditional spline of id column j (coupling.py:165-246,
 *     `apply_unconditional_transform`), shared_modes / shared_bounds HOST arrays; density direction: the net sees the
 *     raw id columns and the output carries their CDF; sampling direction: the id columns are inverted first and the
 *     net sees the result (coupling.py:98-125)
 *   params (optional, [B, Tc * Pc] with row pitch ldp) receives the spline parameters, acts (optional,
 *     [1 + 2 num_blocks][B, 128] contiguous) the pre-activation output of every linear layer, xid (optional,
 *     [B, n_in] with row pitch ldxid) the gathered net inputs x[:, id_cols] (density direction: the raw columns). */
int vcnf_resnet_spline_fwd(const void* x, int64_t ldx, void* y, int64_t ldy, void* logdet, int accumulate,
                           int64_t B, int D, int K, const int32_t* id_cols, int n_in, const int32_t* tr_cols,
                           int Tc, int Pc, int hidden, int num_blocks, const void* packed,
                           const uint8_t* modes, const float* bounds, double wh_scale, int inverse,
                           int outside_zero, int32_t* err_flag, const void* shared_params, int Ts, int Ps,
                           const uint8_t* shared_modes, const float* shared_bounds, void* params, int64_t ldp,
                           void* acts, void* xid, int64_t ldxid, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* VCNF_B200_H_ */
