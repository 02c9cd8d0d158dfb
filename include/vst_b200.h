/*
 * vst_b200.h — C ABI of libvst_b200.so: the sm_100a hot path of tanmud/video_style_transfer
 * (AnimateDiff-XL motion module + UnZipLoRA content/style attention).
 *
 * Conventions:
 *   - every pointer is a raw DEVICE pointer unless the name ends in _host;
 *   - `stream` is a cudaStream_t passed as void*; no entry point synchronises it or allocates;
 *   - return 0 on success, negative on error (VST_E_*); vst_last_error() gives the text;
 *   - dtype tags: VST_F32 / VST_BF16.  bf16 storage, fp32 accumulation everywhere.
 * Each entry point names the reference interface (file:line under the reference tree) whose
 * arithmetic it replaces.
 */
#ifndef VST_B200_H
#define VST_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VST_F32 0
#define VST_BF16 1

#define VST_ACT_NONE 0
#define VST_ACT_GELU 1
#define VST_ACT_GEGLU 2

#define VST_OK 0
#define VST_E_BADARG (-1)
#define VST_E_UNSUPPORTED (-2)
#define VST_E_CUDA (-3)

/* library version (major*10000 + minor*100 + patch) and last error text (thread-local). */
int vst_version(void);
const char* vst_last_error(void);
/* Workspace size of any op (SURVEY.md §8b `vst_workspace_bytes(op, shape...)`): dims are the
 * arguments of the op's own *_workspace_bytes function, in order; ops without a workspace return 0,
 * an unknown op or too few dims a negative status. */
#define VST_OP_REARRANGE 0
#define VST_OP_GROUPNORM5D 1
#define VST_OP_LAYERNORM 2
#define VST_OP_LORA_GEMM 3
#define VST_OP_ATTN_FWD 4
#define VST_OP_TEMPORAL_ATTN_FWD 5
#define VST_OP_GEGLU 6
#define VST_OP_GEMM 7               /* dims: M, N, K1, K2 */
#define VST_OP_COLSUM 8             /* dims: M, N */
#define VST_OP_ATTN_BWD 9           /* dims: Bt, heads, Sq, Sk, kv_group */
#define VST_OP_TEMPORAL_ATTN_BWD 10
#define VST_OP_LAYERNORM_BWD 11     /* dims: rows, C */
#define VST_OP_GROUPNORM5D_BWD 12   /* dims: B, C, F, groups */
#define VST_OP_LORA_GEMM_BWD_DX 13
#define VST_OP_LORA_GEMM_BWD_PARAMS 14   /* dims: M, N, K, Klow, R */
int64_t vst_workspace_bytes(int op, const int64_t* dims, int ndims);
/* SM count / compute capability of the current device. */
int vst_device_info(int* sm_count, int* cc_major, int* cc_minor);

/* ---- a9: layout rearrange (+ sinusoidal position add) -------------------------------------
 * animatediff/temporal_transformer.py:129-133: (B,C,F,H,W) -> (B*H*W, F, C), then x + pe[:, :F].
 * x is addressed as x[b*sB + c*sC + f*sF + hw] (element strides), so both the contiguous
 * (B,C,F,H,W) tensor of config 1 and the live (B*F,C,H,W) layout are accepted.
 * pe: optional fp32 table, row stride C, at least F rows (PositionalEncoding.pe, :11-18).
 * The add is done in fp32 and rounded once to out_dtype (bit-exact vs torch). */
int vst_rearrange_fwd(const void* x, void* y, const float* pe, int B, int C, int F, int HW,
                      int64_t sB, int64_t sC, int64_t sF, int in_dtype, int out_dtype,
                      void* stream);
/* animatediff/temporal_transformer.py:143-144: (B*H*W, F, C) -> out[b*sB + c*sC + f*sF + hw]
 * (+ residual at the same addressing: the motion module's closing `+ residual`). */
int vst_rearrange_bwd(const void* y, void* out, const void* residual, int B, int C, int F,
                      int HW, int64_t sB, int64_t sC, int64_t sF, int in_dtype, int out_dtype,
                      void* stream);

/* ---- a10: GroupNorm over the (B,C,F,H,W) view, fused with the forward rearrange ------------
 * diffusers TransformerTemporalModel.norm = GroupNorm(32, C, eps 1e-6) (un-vendored); in-tree
 * analogue call animatediff/attention_processor.py:50-51.
 * stats: per (b, group) partial (count, mean, M2) triples -> workspace `partials`
 *        (fp32, vst_groupnorm5d_partials(...) triples per (b,group));
 * finalize: Chan-combines the partials into mean / rstd ([B*groups] fp32 each);
 * apply_rearrange: y[(b,hw), f, c] = (x - mean) * rstd * gamma[c] + beta[c]. */
int vst_groupnorm5d_partials(int C, int F, int HW, int groups);
int vst_groupnorm5d_stats(const void* x, float* partials, int B, int C, int F, int HW,
                          int64_t sB, int64_t sC, int64_t sF, int groups, int dtype,
                          void* stream);
int vst_groupnorm5d_finalize(const float* partials, int nparts, int B, int groups, float eps,
                             float* mean, float* rstd, void* stream);
int vst_groupnorm5d_apply_rearrange(const void* x, void* y, const float* mean,
                                    const float* rstd, const void* gamma, const void* beta,
                                    int B, int C, int F, int HW, int64_t sB, int64_t sC,
                                    int64_t sF, int groups, int in_dtype, int out_dtype,
                                    int param_dtype, void* stream);

/* ---- K9: LayerNorm over the last dim, optional fused `+ pe[row % F]` -----------------------
 * animatediff/temporal_transformer.py:44,52,69,74,111,140 (LN) and :27 (position add, which
 * diffusers' BasicTransformerBlock applies right after norm1/norm2). */
int vst_layernorm(const void* x, void* y, const void* gamma, const void* beta, const float* pe,
                  int64_t rows, int C, int F, float eps, int in_dtype, int out_dtype,
                  int param_dtype, void* stream);

/* ---- a8: frame-axis temporal self-attention ------------------------------------------------
 * animatediff/temporal_transformer.py:45-50,69-71 / diffusers motion attn1, attn2.
 * q,k,v,o: (N, F, heads*d) bf16 with row strides ldq/ldk/ldv/ldo (elements), F <= 32,
 * softmax(q k^T / sqrt(d)) v per (sequence, head). */
int vst_temporal_attn_fwd(const void* q, const void* k, const void* v, void* o, int64_t N,
                          int F, int heads, int d, int64_t ldq, int64_t ldk, int64_t ldv,
                          int64_t ldo, float scale, void* stream);

/* ---- a3/a4/a5/a7: fused LoRA GEMM ----------------------------------------------------------
 * D[M,N] = (act(A1[M,K1] * B1[N,K1]^T + A2[M,K2] * B2[N,K2]^T + bias[N]) + residual[M,N]) * out_scale
 * act: VST_ACT_NONE, VST_ACT_GELU (exact erf GELU: temporal_transformer.py:53-55), or
 * VST_ACT_GEGLU (diffusers FeedForward("geglu"), un-vendored: B1 is (2N, K1) = [W_hidden ; W_gate],
 * bias has 2N entries, D[:, j] = (x W_hidden^T + b)[j] * gelu((x W_gate^T + b)[N + j]); the
 * 2N-wide pre-activation stays in TMEM.  N % 128 == 0, bf16 output, no low-rank pair / residual).
 * bf16 operands, fp32 accumulation in TMEM (tcgen05), bf16 or fp32 output.
 * It is the single GEMM behind
 *   unziplora_unet/lora_linear.py:74-81 + unziplora_linear_layer.py:298-346 (A1=x, B1=W,
 *   A2=[u_c|u_s], B2=scale*[m_c.B_c | m_s.B_s]), lora_unzip.py:66-75 (u from other streams),
 *   animatediff/temporal_lora.py:29-32 (A2=u, B2=(alpha/r) B), and the u = x A^T pre-pass
 *   (K2 = 0).  K1, K2 multiples of 8; N multiple of 8; lda/ldb/ldd in elements.
 * k2_group_n > 0 groups the low-rank term for projections fused along N (q|k|v in one GEMM):
 * output columns [g*k2_group_n, (g+1)*k2_group_n) read A2 columns [g*K2, (g+1)*K2) and their own
 * rows of B2[N,K2]; A2 is then [M, G*K2].  k2_group_n must be a multiple of 64 that tiles N.
 * bias is bias_dtype (VST_F32 / VST_BF16); residual has the output dtype.
 * residual/out_scale fuse attention_processor.py:91-94. */
int vst_lora_gemm(const void* A1, int64_t lda1, const void* B1, int64_t ldb1, int K1,
                  const void* A2, int64_t lda2, const void* B2, int64_t ldb2, int K2,
                  int k2_group_n, const void* bias, int bias_dtype, const void* residual, int64_t ldr,
                  float out_scale, int act, void* D, int64_t ldd, int64_t M, int N,
                  int out_dtype, void* stream);
/* Which kernel instantiation vst_lora_gemm runs a shape on: 1000 * pair + BN (pair = 1: the CTA-pair
 * kernel with 256 x BN tiles, 0: 128 x BN tiles), or VST_E_UNSUPPORTED.  Host-only, launches nothing;
 * bench.py uses it to attribute event-timed launches to the dominant kernel. */
int vst_lora_gemm_plan(int64_t M, int N, int K2, int k2_group_n, int act);
/* ---- the single-launch fused LoRA projection (north_star: "ONE fused GEMM") ---------------------------
 * D[M,N] = act(X[M,K1] W[N,K1]^T + (X A_down^T) Bp^T + bias) * out_scale with u = X A_down^T computed INSIDE the
 * kernel, once per (256-row tile, projection group): a leading tcgen05 MMA into a spare accumulator stage, bf16
 * into shared memory as the A operand of the K2 extra k-blocks.  No second launch, no u round trip through HBM.
 * unziplora_linear_layer.py:221-263, 298-346 + lora_linear.py:74-81 (A_down = [A_c ; A_s] (K2, K1), Bp = packed
 * scale * [m_c.mask_c.B_c | m_s.mask_s.B_s] (N, K2)); temporal_lora.py:29-32; grouped along N (k2_group_n) for
 * q|k|v: A_down is then (G*K2, K1) and output columns of group g use rows [g*K2, (g+1)*K2).
 * U_out (optional, bf16 [M, G*K2], row stride ldu): u is also written out for the backward.
 * Shapes: vst_lora_gemm_fused_supported(...) != 0 (M >= 2048, K2 in {64, 128}, group width a multiple of the tile
 * width it returns); otherwise VST_E_UNSUPPORTED and the caller uses vst_lora_gemm twice.  bf16 output. */
int vst_lora_gemm_fused_supported(int64_t M, int N, int K1, int K2, int k2_group_n);
int vst_lora_gemm_fused(const void* X, int64_t ldx, const void* W, int64_t ldw, int K1, const void* A_down,
                        int64_t lda, const void* Bp, int64_t ldbp, int K2, int k2_group_n, const void* bias,
                        int bias_dtype, float out_scale, int act, void* U_out, int64_t ldu, void* D, int64_t ldd,
                        int64_t M, int N, void* stream);
/* A'[br * rp + j, :] = A_br[j, :] as bf16 (rows j in [r, rp) zero): the down-projections of up to two branches
 * stacked for vst_lora_gemm_fused / the u = x A^T pre-pass.  unziplora_linear_layer.py:311,319 ({key}_down). */
int vst_lora_pack_down(const void* Ac, const void* As, void* out, int r, int rp, int K, int param_dtype,
                       void* stream);
/* B2'[n, j] = scale * m[n] * mask[n] * B[n, j] for up to two branches packed side by side
 * (columns [0,r) content, [r,2r) style); m/mask may be NULL (=1).  unziplora_linear_layer.py
 * :313,:321 (merger), :236-237 (hard mask), lora_linear.py:81 (scale). */
int vst_lora_pack_up(const void* Bc, const void* mc, const uint8_t* maskc, const void* Bs,
                     const void* ms, const uint8_t* masks, float scale, void* out, int N, int r,
                     int param_dtype, void* stream);

/* ---- a3/a4/a5/a7 backward: GEMM with transposed operands, deterministic split-K ----------------
 * D[M,N] (+)= alpha * ( opA(A1) opB(B1)^T + opA(A2) opB(B2)^T )
 *   opA(A) is [M,K]: transA = 0 -> A stored [M,K] (row stride lda), 1 -> stored [K,M];
 *   opB(B) is [N,K]: transB = 0 -> B stored [N,K] (row stride ldb), 1 -> stored [K,N].
 * bf16 operands, fp32 accumulation (tcgen05), bf16 or fp32 output; accumulate != 0 adds into D.
 * It is the arithmetic autograd derives for unziplora_linear_layer.py:221-263,
 * temporal_lora.py:29-32 and lora_linear.py:74-81:
 *   dx = dy W + du [A_c; A_s] (transB=1, two pairs), du = dy B' (transB=1),
 *   dB' = dy^T u, dA = du^T x, dW = dy^T x (transA = transB = 1, contraction over tokens).
 * Small outputs with a long contraction are split over k; the fp32 partials go to `workspace`
 * (vst_gemm_workspace_bytes; 0 = no split needed) and are summed in a fixed order.  With a NULL or
 * short workspace the call still succeeds, unsplit. */
int64_t vst_gemm_workspace_bytes(int64_t M, int N, int64_t K1, int64_t K2);
int vst_gemm(const void* A1, int64_t lda1, const void* B1, int64_t ldb1, int64_t K1,
             const void* A2, int64_t lda2, const void* B2, int64_t ldb2, int64_t K2, int transA,
             int transB, void* D, int64_t ldd, int64_t M, int N, float alpha, int accumulate,
             int out_dtype, void* workspace, int64_t workspace_bytes, void* stream);

/* Adjoint of vst_lora_pack_up.  dBp: fp32 gradient of the packed operand, row stride ld, branch br
 * occupying columns [br*rp, br*rp + r) (rp = rank padded to 8).  Outputs (fp32, any may be NULL):
 *   dB_br[n,j] = scale m_br[n] mask_br[n] dBp[n, br*rp+j],  dm_br[n] = scale mask_br[n] sum_j dBp B_br. */
int vst_lora_pack_up_bwd(const float* dBp, int64_t ld, int rp, const void* Bc, const void* mc,
                         const uint8_t* maskc, const void* Bs, const void* ms,
                         const uint8_t* masks, float scale, float* dBc, float* dmc, float* dBs,
                         float* dms, int N, int r, int param_dtype, void* stream);
/* ---- one LoRA linear at MODULE level (raw parameters in): forward / dX / parameter gradients ----
 * vst_lora_gemm_fwd:  y = (x W^T + bias + u B'^T + residual) * out_scale,  u = [x_c A_c^T | x_s A_s^T],
 *   B' = scale * [m_c.mask_c.B_c | m_s.mask_s.B_s]  (lora_linear.py:74-81, lora_unzip.py:66-75,
 *   unziplora_linear_layer.py:221-263, 298-346).  x_c / x_s NULL = x (single stream); a block-masked branch
 *   (:308,:316) is passed as A = NULL; mask_* are the hard column masks of the training layer (:236-237) or
 *   NULL; mode VST_LORA_CONTENT / _STYLE run one branch without its merger (:324-345).  A_* (rank, Klow),
 *   B_* (N, rank), m_* (N) in param_dtype; rank a multiple of 8.  u is computed INSIDE the main kernel
 *   (vst_lora_gemm_fused) when the streams coincide and the shape is a fused one, by a pre-pass otherwise, and
 *   written to u_out (bf16 [M, branches*rank]) when given -- vst_lora_gemm_bwd_params needs it.
 *   workspace: vst_lora_gemm_fwd_workspace_bytes(M, N, Klow, rank, branches, u_out != NULL), 256-byte aligned
 *   (packed bf16 A', B' and, without u_out, u).
 * The two backward calls are the arithmetic torch autograd derives for the same modules (the reference
 * trains through autograd, train_animatediff.py:314), expressed on the fused operands of the forward:
 *   y = x W^T + u B'^T,  u = x_low A^T,  A = [A_c ; A_s] (R, Klow),  B' = vst_lora_pack_up (N, R).
 * vst_lora_gemm_bwd_dx:     dU[M,R] = dY B' (bf16, kept for bwd_params) and
 *                           dX[M,K] (+)= alpha (dY W + dU A)  -- A is added only when the adapter read
 *                           the same stream as the base projection (single stream, Klow = K); for the
 *                           three-stream form pass R = 0 here and get the streams' own gradient
 *                           dX_low = dU A from vst_gemm.  R = 0: plain dX = alpha dY W.
 * vst_lora_gemm_bwd_params: dW[N,K] = dY^T X, dA[R,Klow] = dU^T Xlow, dB'[N,R] = dY^T U (fp32, dense
 *                           row-major, any may be NULL; contraction over the M tokens, deterministic
 *                           split-K through `workspace`).  dB_c / dB_s / d-merger follow from dB' by
 *                           vst_lora_pack_up_bwd.  Xlow NULL = X. */
#define VST_LORA_BOTH 0
#define VST_LORA_CONTENT 1
#define VST_LORA_STYLE 2
int64_t vst_lora_gemm_fwd_workspace_bytes(int64_t M, int N, int Klow, int rank, int branches, int have_u_out);
int vst_lora_gemm_fwd(const void* x, int64_t ldx, const void* x_c, int64_t ldxc, const void* x_s, int64_t ldxs,
                      const void* W, int64_t ldw, const void* bias, int bias_dtype, const void* A_c,
                      const void* B_c, const void* m_c, const uint8_t* mask_c, const void* A_s, const void* B_s,
                      const void* m_s, const uint8_t* mask_s, int rank, int param_dtype, float scale, int mode,
                      const void* residual, int64_t ldr, float out_scale, void* u_out, int64_t ldu, void* D,
                      int64_t ldd, int64_t M, int N, int K, int Klow, int out_dtype, void* workspace,
                      int64_t workspace_bytes, void* stream);
int vst_lora_gemm_bwd_dx(const void* dY, int64_t lddy, const void* W, int64_t ldw, const void* Bp,
                         int64_t ldbp, const void* A, int64_t lda, int R, void* dU, int64_t lddu,
                         void* dX, int64_t lddx, int64_t M, int N, int K, float alpha, int accumulate,
                         int out_dtype, void* stream);
int64_t vst_lora_gemm_bwd_params_workspace_bytes(int64_t M, int N, int K, int Klow, int R);
int vst_lora_gemm_bwd_params(const void* dY, int64_t lddy, const void* X, int64_t ldx,
                             const void* Xlow, int64_t ldxl, const void* U, int64_t ldu,
                             const void* dU, int64_t lddu, float* dW, float* dA, float* dBp,
                             int64_t M, int N, int K, int Klow, int R, int accumulate,
                             void* workspace, int64_t workspace_bytes, void* stream);
/* out[n] = alpha * sum_m x[m, n] (bf16 in, fp32 out): the bias gradient.  Deterministic two-stage
 * sum through a caller workspace of vst_colsum_workspace_bytes(M, N) bytes. */
int64_t vst_colsum_workspace_bytes(int64_t M, int N);
int vst_colsum(const void* x, int64_t ld, int64_t M, int N, float alpha, float* out,
               void* workspace, int64_t workspace_bytes, void* stream);
/* vst_colsum without a caller workspace and without any allocation (no cudaMalloc / cudaFree / allocating
 * library call).  It has no counterpart in the reference and was added
 * for callers that cannot allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc.  The fp32 partials live in a static device array of VST_COLSUM_NOALLOC_SCRATCH_BYTES (every GPU
 * that loads the library holds it until exit); a shape with vst_colsum_workspace_bytes(M, N) beyond that is
 * refused with VST_E_UNSUPPORTED and a message.  All calls share the array: nothing needs clearing between
 * calls (every slot read is written first), but calls that overlap on the GPU (two streams or threads) race. */
#define VST_COLSUM_NOALLOC_SCRATCH_BYTES (1 << 20)
int vst_colsum_noalloc(const void* x, int64_t ld, int64_t M, int N, float alpha, float* out, void* stream);

/* ---- §8f-1 (next): GEGLU gate of the motion feed-forward ------------------------------------
 * out[r, j] = hg[r, j] * gelu(hg[r, inner + j]); diffusers FeedForward("geglu") (un-vendored),
 * exact GELU as animatediff/temporal_transformer.py:55.  bf16. */
int vst_geglu(const void* hg, void* out, int64_t rows, int inner, int64_t ld_in, int64_t ld_out,
              void* stream);

/* ---- §8f-4 (next): CFG combine + Euler update of the sampler -------------------------------------
 * inference_animatediff.py:122,131 (EulerDiscreteScheduler.step, epsilon prediction):
 *   eps = eps_u + g (eps_c - eps_u);  latents_out = latents + (sigma_next - sigma) eps;
 *   scaled_next (optional) = latents_out / sqrt(sigma_next^2 + 1)  (next scale_model_input).
 * eps_* / scaled_next in eps_dtype (VST_BF16 | VST_F32), latents in latent_dtype. */
int vst_cfg_euler_step(const void* eps_uncond, const void* eps_cond, const void* latents,
                       void* latents_out, void* scaled_next, int64_t n, float guidance, float sigma,
                       float sigma_next, int eps_dtype, int latent_dtype, void* stream);

/* ---- a2: spatial self / text cross attention (flash, tcgen05) -------------------------------
 * animatediff/attention_processor.py:74-83 (head split, SDPA, head merge) in one kernel.
 * q,o: (Bt, Sq, heads*64) bf16; k,v: (Bk, Sk, heads*64) bf16 where batch b reads k/v batch
 * b / kv_group (kv_group = Bt/Bk implements the text repeat_interleave of :62-66 without
 * materialising it).  lse (optional, fp32 (Bt,heads,Sq)) is saved for the backward.
 * mask (optional, NULL on the AnimateDiff-XL path): the additive attention mask of :46-48 /
 * F.scaled_dot_product_attention(attn_mask=...), fp32, element (b, h, i, j) at
 * mask[b*mask_sb + h*mask_sh + i*mask_sq + j]; a stride of 0 broadcasts that axis (a key-padding
 * mask has mask_sh = mask_sq = 0).  A row whose keys are all -inf yields NaN, as torch does. */
int vst_attn_fwd(const void* q, const void* k, const void* v, void* o, float* lse, int Bt,
                 int heads, int Sq, int Sk, int d, int64_t ldq, int64_t ldk, int64_t ldv,
                 int64_t ldo, int kv_group, float scale, const float* mask, int64_t mask_sb,
                 int64_t mask_sh, int64_t mask_sq, void* stream);

/* ---- a8 backward: gradients of vst_temporal_attn_fwd (probabilities recomputed; no saved state) --
 * q,k,v,d_o: (N,F,heads*d) bf16, 16-byte aligned rows; dq,dk,dv: same shape, bf16. */
int vst_temporal_attn_bwd(const void* q, const void* k, const void* v, const void* d_o, void* dq,
                          void* dk, void* dv, int64_t N, int F, int heads, int d, int64_t ldq,
                          int64_t ldk, int64_t ldv, int64_t lddo, int64_t lddq, int64_t lddk,
                          int64_t lddv, float scale, void* stream);

/* ---- K9 / §8f-1 / a10 backward (config 4: the motion module trains its norms and projections,
 * animatediff/utils.py:66-96) ------------------------------------------------------------------
 * layernorm_bwd: x, dy, dx bf16 contiguous (rows, C), C <= 1280; dgamma / dbeta fp32 (C), may be
 *   NULL; statistics are recomputed from x.  Workspace: vst_layernorm_bwd_workspace_bytes.
 * geglu_bwd: dhg[:, :inner] = dout * gelu(b), dhg[:, inner:] = dout * a * gelu'(b).
 * groupnorm5d_bwd: x, dy, dx (and the optional `add`, summed into dx: the motion module's residual
 *   branch) are frames-major contiguous (B*F, C, HW) bf16; mean / rstd from vst_groupnorm5d_finalize. */
int64_t vst_layernorm_bwd_workspace_bytes(int64_t rows, int C);
int vst_layernorm_bwd(const void* x, const void* dy, const void* gamma, void* dx, float* dgamma,
                      float* dbeta, int64_t rows, int C, float eps, int param_dtype,
                      void* workspace, int64_t workspace_bytes, void* stream);
int vst_geglu_bwd(const void* hg, const void* dout, void* dhg, int64_t rows, int inner,
                  int64_t ld_in, int64_t ld_do, int64_t ld_out, void* stream);
/* dz = alpha * dy * gelu'(z): the VST_ACT_GELU epilogue of vst_lora_gemm, differentiated. */
int vst_gelu_bwd(const void* z, const void* dy, void* dz, int64_t rows, int N, int64_t ld_z,
                 int64_t ld_dy, int64_t ld_dz, float alpha, void* stream);
int64_t vst_groupnorm5d_bwd_workspace_bytes(int B, int C, int F, int groups);
int vst_groupnorm5d_bwd(const void* x, const void* dy, const void* add, void* dx, const float* mean,
                        const float* rstd, const void* gamma, float* dgamma, float* dbeta, int B,
                        int C, int F, int HW, int groups, int param_dtype, void* workspace,
                        int64_t workspace_bytes, void* stream);

/* ---- a2 backward ------------------------------------------------------------------------------
 * Gradients of vst_attn_fwd (autograd through F.scaled_dot_product_attention,
 * animatediff/attention_processor.py:78-80).  o / lse are the forward's outputs, d_o the incoming
 * gradient; dq (Bt,Sq,C), dk / dv (Bt/kv_group,Sk,C) are written in bf16 (the text K/V gradients
 * are summed over the kv_group frames that shared them, :62-66).  `workspace`: 256-byte aligned,
 * vst_attn_bwd_workspace_bytes(...) bytes (D = rowsum(dO.O), the fp32 dQ accumulator and, for
 * kv_group > 1, fp32 dK / dV accumulators).  mask / mask_s*: the forward's additive mask (no
 * gradient is produced for it). */
int64_t vst_attn_bwd_workspace_bytes(int Bt, int heads, int Sq, int Sk, int kv_group);
int vst_attn_bwd(const void* q, const void* k, const void* v, const void* o, const void* d_o,
                 const float* lse, void* dq, void* dk, void* dv, int Bt, int heads, int Sq, int Sk,
                 int d, int64_t ldq, int64_t ldk, int64_t ldv, int64_t ldo, int64_t lddo,
                 int64_t lddq, int64_t lddk, int64_t lddv, int kv_group, float scale,
                 const float* mask, int64_t mask_sb, int64_t mask_sh, int64_t mask_sq,
                 void* workspace, int64_t workspace_bytes, void* stream);

/* ---- §8e: one video, sequence-parallel over the GPUs of an NVSwitch box -------------------------
 * Spatial work is sharded over frames, temporal work over positions (the reference
 * itself has no collective).  Each rank owns a "symmetric" device buffer (data + `world` 32-bit flag
 * words, zeroed) that every peer maps through CUDA IPC: vst_comm_ipc_get / _open / _close move the
 * 64-byte handles, which the host exchanges with whatever it has (torch.distributed here).
 * vst_comm_init takes, for every rank r, the address AS MAPPED ON THIS GPU of r's data buffer and
 * flag words, and returns the communicator.
 * The two all-to-alls of a motion module are the motion module's own layout kernels addressing
 * peer memory (no separate exchange pass):
 *   vst_a2a_frames_to_hw: this rank's frames x[(b, f_local), c, hw] (strides sB, sC, sF) ->
 *       GroupNorm-apply (when mean != NULL) -> token rows [(b, hw_local), f, c] STORED into the
 *       buffer of the rank that owns position hw, at byte offset dst_offset_bytes, laid out
 *       (B, HW/world, F_total, C) bf16;
 *   vst_a2a_hw_to_frames: token rows LOADED from the owners' buffers (src_offset_bytes) ->
 *       out[(b, f_local), c, hw] (+ residual, local).
 * Neither call synchronises ranks: bracket them with vst_comm_barrier (a flag barrier over NVLink,
 * st.release.sys / ld.acquire.sys; it traps after 20 s instead of hanging if a peer is gone).
 * bf16 only; HW must be a multiple of world x 64; F_total = world x F_local. */
int vst_comm_alloc(int64_t bytes, void** out);   /* cudaMalloc + zero fill: the symmetric buffer */
int vst_comm_free(void* ptr);
int vst_comm_ipc_get(const void* dev_ptr, void* handle64);
int vst_comm_ipc_open(const void* handle64, void** out);
int vst_comm_ipc_close(void* ptr);
int vst_comm_init(int rank, int world, void* const* bufs, void* const* flags, void** comm_out);
int vst_comm_fini(void* comm);
int vst_comm_barrier(void* comm, void* stream);
/* rank p's GroupNorm partials (rows, nparts, 3) fp32 -> columns [p*nparts, (p+1)*nparts) of a
 * (rows, world*nparts, 3) array at dst_offset_bytes of EVERY rank's symmetric buffer (peer stores
 * over NVLink; bracket with vst_comm_barrier).  The flag block of a rank must be >= 256 bytes:
 * one word per peer plus the rank's barrier count (kept on the device so that a captured CUDA
 * graph of a sequence-parallel step replays correctly). */
int vst_comm_allgather_partials(void* comm, const float* partials, int rows, int nparts,
                                int64_t dst_offset_bytes, void* stream);
int vst_a2a_frames_to_hw(void* comm, const void* x, int64_t dst_offset_bytes, const float* mean,
                         const float* rstd, const void* gamma, const void* beta, int B, int C,
                         int F_local, int F_total, int HW, int64_t sB, int64_t sC, int64_t sF,
                         int groups, int param_dtype, void* stream);
int vst_a2a_hw_to_frames(void* comm, int64_t src_offset_bytes, void* out, const void* residual, int B,
                         int C, int F_local, int F_total, int HW, int64_t sB, int64_t sC, int64_t sF,
                         void* stream);
/* GroupNorm affine applied IN PLACE on a token tensor (B, rows_per_batch, C) bf16: the same fp32
 * coefficients and single fma + rounding as vst_groupnorm5d_apply_rearrange.  Used by the sequence-
 * parallel path, whose push sends raw activations and partial statistics under ONE barrier. */
int vst_groupnorm_tokens_apply(void* tok, const float* mean, const float* rstd, const void* gamma,
                               const void* beta, int B, int64_t rows_per_batch, int C, int groups,
                               int param_dtype, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* VST_B200_H */
