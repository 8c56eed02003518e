/*
 * taco_b200_train.h -- C ABI of the device primitives of the teacher-forced TRAINING step
 * (BASELINE.json config 4).
 *
 * The reference builds the training graph with TensorFlow 1.9 (Tacotron.initialize(is_training=True),
 * add_loss, add_optimizer: TacotronModel/modules/Tacotron.py:46-349, 351-446, 448-540) and lets TF's
 * autodiff derive the backward pass.  Here the step is a host-side composition
 * (tacotron_prosody_multi_gpu_b200/train/step.py) of the stateless primitives below; each cites the
 * reference graph segment whose forward or gradient it computes.
 *
 * Conventions: every pointer is DEVICE memory (fp32 / int32, innermost dimension contiguous; leading
 * strides are passed where views are allowed); return 0 on success, -1 bad argument, -3 CUDA error;
 * asynchronous on `stream` (cudaStream_t as void*); no state, no allocation, no CPU fallback.
 * Random keep bits: Philox4x32-10, counter (idx >> 2, stream_id, idx >> 34, "TRN1"), key = seed,
 * word idx & 3 compared with keep_thresh = floor(p * 2^32) (restated in oracle/philox.py).
 */
#ifndef TACO_B200_TRAIN_H
#define TACO_B200_TRAIN_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { TACO_ACT_NONE = 0, TACO_ACT_RELU = 1, TACO_ACT_TANH = 2 };
enum { TACO_EW_COPY = 0, TACO_EW_ADD = 1, TACO_EW_MUL = 2, TACO_EW_SCALE = 3, TACO_EW_TANH = 4, TACO_EW_TANH_BWD = 5 };

/* Process-wide operand precision of the THROUGHPUT-bound contractions (conv1d fwd / dgrad / wgrad, the GEMMs over all
 * S*B rows): TACO_TRAIN_FP32 = fp32 SIMT FMA (parity mode, default); TACO_TRAIN_BF16 = operands rounded to bf16,
 * mma.sync tensor cores, fp32 accumulation.  The per-step recurrent contractions always run in fp32. */
enum { TACO_TRAIN_FP32 = 0, TACO_TRAIN_BF16 = 1 };
int taco_op_set_precision(int precision);

/* C[M,N] = op(A) op(B) + beta*C + bias[N]; op(A) = A[M,K] (lda) or A[K,M]^T when ta; same for B.
 * tf.layers.dense / LSTMCell / GRUCell matmuls and all their gradients. */
int taco_op_gemm(const float* A, long long lda, const float* B, long long ldb, float* C, long long ldc, int M, int N, int K,
                 int ta, int tb, float beta, const float* bias, void* stream);
/* Fused pairs of the per-step (M = batch) contractions, always accumulating into their outputs (split-K, atomics):
 *   C += [A1 | A2] . B          -- x- and h-part of an LSTM gate contraction in one launch (B = kernel rows [K1+K2, N])
 *   [C1 | C2] += A . B^T        -- its gradient w.r.t. both inputs in one launch (B = the same kernel rows [N1+N2, K]) */
int taco_op_gemm_pair_acc(const float* A1, long long lda1, int K1, const float* A2, long long lda2, int K2, const float* B, long long ldb,
                          float* C, long long ldc, int M, int N, void* stream);
int taco_op_gemm_nt_split_acc(const float* A, long long lda, const float* B, long long ldb, float* C1, long long ldc1, int N1, float* C2,
                              long long ldc2, int N2, int M, int K, void* stream);
/* C += A . B (tb = 0: B [K,N]) or C += A . B^T (tb = 1: B [N,K]), M = batch rows: the recurrent contraction of the LSTM loops */
int taco_op_gemm_acc(const float* A, long long lda, const float* B, long long ldb, float* C, long long ldc, int M, int N, int K, int tb,
                     void* stream);
/* tf.layers.conv1d(padding='same') NWC, kernel [k,Ci,Co] (Encoder.py:159-164) and its two gradients. */
int taco_op_conv1d_fwd(const float* x, const float* w, const float* bias, float* y, int B, int T, int Ci, int Co, int k, void* stream);
int taco_op_conv1d_dgrad(const float* dy, const float* w, float* dx, int B, int T, int Ci, int Co, int k, void* stream);
int taco_op_conv1d_wgrad(const float* x, const float* dy, float* dw, int B, int T, int Ci, int Co, int k, void* stream);
/* tf.layers.conv2d(3x3, strides 2, 'same') NHWC, kernel [3,3,Ci,Co] (Encoder.py:173-178) and gradients. */
int taco_op_conv2d_fwd(const float* x, const float* w, const float* bias, float* y, int B, int H, int W, int Ci, int Co, void* stream);
int taco_op_conv2d_dgrad(const float* dy, const float* w, float* dx, int B, int H, int W, int Ci, int Co, void* stream);
int taco_op_conv2d_wgrad(const float* x, const float* dy, float* dw, int B, int H, int W, int Ci, int Co, void* stream);
/* out[C] = sum over M rows of x[M,C] (row stride ld) + beta*out: bias gradients. */
int taco_op_colsum(const float* x, long long ld, int M, int C, float* out, float beta, void* stream);

/* tf.layers.batch_normalization(training=True) axis -1 eps 1e-3 -> activation -> dropout(keep)
 * (Encoder.py:165-167); x,y [M,C]; writes the batch mean and 1/sqrt(var+eps). */
int taco_op_bn_fwd(const float* x, const float* gamma, const float* beta, int act, uint32_t keep_thresh, float keep,
                   uint64_t seed, uint32_t stream_id, float* y, float* mean, float* rstd, int M, int C, void* stream);
int taco_op_bn_bwd(const float* dy, const float* x, const float* y, const float* mean, const float* rstd, const float* gamma,
                   int act, uint32_t keep_thresh, float keep, uint64_t seed, uint32_t stream_id, float* dx, float* dgamma,
                   float* dbeta, int M, int C, void* stream);
/* UPDATE_OPS of batch normalisation (Tacotron.py:513): moving = momentum*moving + (1-momentum)*batch. */
int taco_op_bn_update_moving(const float* mean, const float* rstd, float* mov_mean, float* mov_var, float momentum, int C, void* stream);

/* ZoneoutLSTMCell (Encoder.py:118-142) pointwise part.  gates [B,4H] (+pre [B,4H]) (+bias [4H]) are the
 * pre-activations i,j,f,o; on return gates holds sigmoid(i), tanh(j), sigmoid(f+1), sigmoid(o).
 * lens != NULL: rows with s >= lens[b] copy the state through and output 0 (dynamic_rnn sequence_length).
 * train (zo_keep_thresh = floor((1-zoneout)*2^32) passed explicitly): Bernoulli keep masks from streams (stream_c,
 * stream_h), element (row0 + b)*H + u; train == 0: deterministic (1-zoneout)/zoneout blend. */
int taco_op_lstm_fwd(float* gates, const float* pre, const float* bias, const float* c_prev, const float* h_prev,
                     const int32_t* lens, int s, float zoneout, int train, uint32_t zo_keep_thresh, uint64_t seed, uint32_t stream_c,
                     uint32_t stream_h, int row0, float* cnew, float* cz, float* hraw, float* hz, int B, int H, void* stream);
int taco_op_lstm_bwd(const float* acts, const float* c_prev, const float* h_prev, const float* cnew, const int32_t* lens, int s,
                     float zoneout, int train, uint32_t zo_keep_thresh, uint64_t seed, uint32_t stream_c, uint32_t stream_h, int row0,
                     const float* d_hraw, const float* d_cz, const float* d_hz, float* dgates, float* d_cprev, float* d_hprev,
                     int B, int H, void* stream);
/* bidirectional_dynamic_rnn position bookkeeping: pos_b = s (forward) or len_b-1-s (reverse), valid iff s < len_b.
 * gather: out[b,:] = src[b,pos_b,:] (0 for invalid rows); scatter: dst[b,pos_b,:] = src[b,:] for valid rows. */
/* The same two kernels for the encoder time loops with the sequence gather / scatter fused in (Encoder.py:118-142 over the valid
 * length): forward adds row pos of xw [B,T,4H] (strides xw_sb, xw_st; pos = s, or len - 1 - s when reverse) to the gates and
 * writes the raw output to out[b,pos,:H]; backward reads d(raw output) from dout[b,pos,:H] and also writes d(gates) to
 * dxw[b,pos,:4H].  Rows with s >= len: state carried through, nothing scattered. */
int taco_op_lstm_fwd_seq(float* gates, const float* xw, long long xw_sb, long long xw_st, const float* c_prev, const float* h_prev,
                         const int32_t* lens, int s, int reverse, float zoneout, int train, uint32_t zo_keep_thresh, uint64_t seed,
                         uint32_t stream_c, uint32_t stream_h, int row0, float* cnew, float* cz, float* hraw, float* hz, float* out,
                         long long out_sb, long long out_st, int B, int H, void* stream);
int taco_op_lstm_bwd_seq(const float* acts, const float* c_prev, const float* h_prev, const float* cnew, const int32_t* lens, int s,
                         int reverse, float zoneout, int train, uint32_t zo_keep_thresh, uint64_t seed, uint32_t stream_c, uint32_t stream_h,
                         int row0, const float* dout, long long dout_sb, long long dout_st, const float* d_cz, const float* d_hz,
                         float* dgates, float* d_cprev, float* d_hprev, float* dxw, long long dxw_sb, long long dxw_st, int B, int H,
                         void* stream);
int taco_op_gather_rows(const float* src, long long sb, long long st, const int32_t* lens, int s, int reverse, float* out,
                        int B, int N, void* stream);
int taco_op_scatter_rows(const float* src, const int32_t* lens, int s, int reverse, float* dst, long long db, long long dt,
                         int B, int N, void* stream);

/* One LocationSensitiveAttention step (Attention.py:169-210,345-401), fp32: energy / softmax / context kernels tiled over
 * (positions x rows) and (value columns x rows). */
int taco_op_attn_fwd(const float* q, const float* cum_prev, const float* keys, const float* values, const int32_t* lens,
                     const float* conv_w, const float* conv_b, const float* Wl, const float* va, const float* ba,
                     float* align, float* cum, float* ctx, int B, int T, int A, int V, int NF, int KW, void* stream);
/* Its gradient given d(ctx_t), d(cum_t): dq, dcum_prev overwritten; dkeys and the per-row weight partials
 * p_* [B, ...] accumulated (d(values) is a post-loop contraction done with taco_op_gemm). */
int taco_op_attn_bwd(const float* dctx, const float* dcum, const float* q, const float* cum_prev, const float* align,
                     const float* keys, const float* values, const int32_t* lens, const float* conv_w, const float* conv_b,
                     const float* Wl, const float* va, const float* ba, float* dq, float* dcum_prev, float* dkeys,
                     float* p_conv_w, float* p_conv_b, float* p_Wl, float* p_va, float* p_ba, float* scratch /* >= B*T*NF floats */,
                     int B, int T, int A, int V, int NF, int KW, void* stream);

/* taco_op_attn_bwd for callers that cannot allocate (CUDA graph capture, tight serving / training loops, harnesses that forbid
 * cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same arguments and results as
 * taco_op_attn_bwd minus the scratch pointer: the d(location features) scratch [B,T,NF] lives in a file-scope
 * `static __device__ __align__(256)` array of TACO_ATTN_BWD_NOALLOC_MAX_FLOATS floats (4 MiB, held by every GPU that loads the
 * library until exit), whose device address is resolved with cudaGetSymbolAddress once per device (the device current at the
 * call).  Calls no cudaMalloc / cudaFree / allocating library routine.  B*T*NF beyond the limit: returns
 * TACO_OP_ERR_SCRATCH_LIMIT (and says so on stderr) without launching anything.  All calls on a device share the array and it
 * keeps the last call's contents (every entry a call reads is written by that call first, so nothing has to be cleared): calls
 * that overlap on the GPU (different streams, or graph branches without an edge between them) RACE -- serialise them. */
#define TACO_ATTN_BWD_NOALLOC_MAX_FLOATS (1 << 20)
#define TACO_OP_ERR_SCRATCH_LIMIT (-4)
int taco_op_attn_bwd_noalloc(const float* dctx, const float* dcum, const float* q, const float* cum_prev, const float* align,
                             const float* keys, const float* values, const int32_t* lens, const float* conv_w, const float* conv_b,
                             const float* Wl, const float* va, const float* ba, float* dq, float* dcum_prev, float* dkeys,
                             float* p_conv_w, float* p_conv_b, float* p_Wl, float* p_va, float* p_ba,
                             int B, int T, int A, int V, int NF, int KW, void* stream);

/* The same attention step for the time loops, fused (csrc/train_attn_fused.cu): the location convolution and the location dense
 * layer folded into one filter bank U [KW,A] (+ c0 [A]) once per training step; forward = ONE launch per decoder step (query
 * projection q = h2raw . Wq, energies, masked softmax, cumulative update, context) that also saves tanh(pre) [B,T,A]; backward
 * = ONE launch per step that produces only what the recurrence needs (d(q) and d(h2raw) += d(q) Wq^T, d(cum_{t-1}), the total
 * d(ctx_t) = dctx + carry written back to dctx, and the saved d(energy) [B,T]); d(keys) and the weight gradients of all S steps
 * come from ONE pass after the loop (taco_op_attn_bwd_hoisted; results are ADDED to g_* and dkeys; work >= B*(KW+2)*A floats).
 * taco_op_attn_fused_supported: 1 when these kernels hold the shape (A = 128, NF = 32, KW = 31, T <= 256-ish), else 0. */
int taco_op_attn_fused_supported(int T, int D, int A, int V, int NF, int KW);
int taco_op_attn_fold(const float* conv_w, const float* conv_b, const float* Wl, float* U, float* c0, void* stream);
int taco_op_attn_fwd_fused(const float* h2raw, const float* Wq, const float* U, const float* c0, const float* ba, const float* va,
                           const float* cum_prev, const float* keys, const float* values, const int32_t* lens, float* q, float* th,
                           float* align, float* cum, float* ctx, int B, int T, int D, int V, void* stream);
int taco_op_attn_bwd_fused(float* dctx, const float* carry, const float* dcum, const float* align, const float* th, const float* values,
                           const int32_t* lens, const float* U, const float* va, const float* Wq, float* de, float* dq, float* dh2,
                           float* dcum_prev, float* scratch /* >= B*T floats */, int B, int T, int D, int V, void* stream);
/* d(values)[b,t,:] = sum_s align[s,b,t] d(ctx)[s,b,:] for all rows in one launch (align [S,B,T], dctx [S,B,V], dvalues [B,T,V]
 * overwritten; V % 64 == 0, 16-byte aligned dctx / dvalues) -- the gradient of the context sums of all steps (Attention.py:208-210). */
int taco_op_attn_dvalues(const float* align, const float* dctx, float* dvalues, int S, int B, int T, int V, void* stream);
int taco_op_attn_bwd_hoisted(const float* de, const float* th, const float* cum, const int32_t* lens, const float* conv_w,
                             const float* conv_b, const float* Wl, const float* va, float* dkeys, float* g_conv_w, float* g_conv_b,
                             float* g_Wl, float* g_va, float* g_ba, float* work, int S, int B, int T, void* stream);

/* GRUCell (Tacotron.py:201; reset before the candidate matmul) pointwise parts and gradients. */
int taco_op_gru_pre(float* g, const float* h, float* rh, int B, int R, void* stream);
int taco_op_gru_out(float* cpre, const float* ru, const float* h, float* hnew, int B, int R, void* stream);
int taco_op_gru_out_bwd(const float* dhnew, const float* ru, const float* h, const float* c, float* dcpre, float* dg, float* dh,
                        int B, int R, void* stream);
int taco_op_gru_pre_bwd(const float* drh, const float* ru, const float* h, float* dg, float* dh, int B, int R, void* stream);

/* MultiheadAttention mlp_attention with weight-normalised v (Attention.py:243-341). q [B,U], k [N,U],
 * v,b [U/nh], g [1], tt = tanh(tokens) [N,D]; wts [B,nh,N], style [B,nh*D]. */
int taco_op_gst_fwd(const float* q, const float* k, const float* v, const float* g, const float* b, const float* tt,
                    float* wts, float* style, int B, int N, int U, int D, int nh, void* stream);
int taco_op_gst_bwd(const float* dstyle, const float* q, const float* k, const float* v, const float* g, const float* b,
                    const float* tt, const float* wts, float* dq, float* dk, float* dv, float* dg, float* db, float* dtt,
                    float* scratch /* >= U/nh floats */, int B, int N, int U, int D, int nh, void* stream);

int taco_op_ew(int op, float* out, const float* a, const float* b, float alpha, long long n, void* stream);
int taco_op_clip_fwd(const float* x, float lo, float hi, float* y, long long n, void* stream);
int taco_op_clip_bwd(const float* dy, const float* x, float lo, float hi, float* dx, long long n, void* stream);
/* tf.losses.mean_squared_error / mean sigmoid_cross_entropy_with_logits (Tacotron.py:390-396) + gradients. */
int taco_op_mse(const float* pred, const float* tgt, float* loss, float* dpred, long long n, void* stream);
int taco_op_sigmoid_ce(const float* logits, const float* labels, float* loss, float* dlogits, long long n, void* stream);
int taco_op_embedding_fwd(const int32_t* ids, const float* table, float* out, int n_tok, int dim, int n_sym, void* stream);
int taco_op_embedding_bwd(const int32_t* ids, const float* dout, float* dtable, int n_tok, int dim, int n_sym, void* stream);
/* Prenet activation (Decoder.py:342-352): relu + dropout with the decoder's Philox stream (seed, step, row, layer, unit). */
int taco_op_prenet_act_fwd(const float* z, int layer, uint64_t seed, int step0, float keep, float* out, int S, int B, int U, void* stream);
int taco_op_prenet_act_bwd(const float* dout, const float* out, int layer, uint64_t seed, int step0, float keep, float* dz,
                           int S, int B, int U, void* stream);
int taco_op_mask_rows(float* x, const int32_t* lens, int B, int T, int C, void* stream);
int taco_op_tile_style(const float* style, float* memory, int B, int T, int SD, int MD, int col0, void* stream);
int taco_op_sum_style(const float* dmemory, float* dstyle, int B, int T, int SD, int MD, int col0, void* stream);

/* add_loss regularisation (Tacotron.py:419-422) and add_optimizer (:448-515). */
int taco_op_reg_loss(const float* p, const float* regw, float* out, long long n, void* stream);
int taco_op_grad_finalize(float* g, const float* p, const float* regw, float inv_world, float* sumsq, long long n, void* stream);
int taco_op_adam_step(float* p, const float* g, float* m, float* v, const float* sumsq, float clip_norm, float lr_t, float b1,
                      float b2, float eps, long long n, void* stream);

/* ---- whole time loops, launched from C (no host-language overhead per step) ------------------------------------
 * The same per-step primitives as above, sequenced for all S decoder steps / all T encoder positions. */
typedef struct TacoDecLoop {
    int32_t B, T, S, D, MD, A, U2, NF, KW;   /* batch, tokens, steps, LSTM units 256, memory width 768, attention dim 128,
                                                prenet width 256, location filters 32, location kernel 31 */
    float zoneout;
    uint32_t zo_keep_thresh;
    uint64_t seed;
    uint32_t stream1, stream2;               /* Philox streams of the cell masks of LSTM 1 / 2 (hidden masks: +1) */
    const float *W1, *W2, *b2, *Wq;          /* [U2+MD+D,4D], [2D,4D], [4D], [D,A] */
    const float *conv_w, *conv_b, *Wl, *va, *ba;
    const float *keys, *values;              /* [B,T,A], [B,T,MD] */
    const int32_t* lens;
    const float* G1x;                        /* [S,B,4D] prenet term of the LSTM-1 gates (+bias), hoisted */
    float *acts1, *acts2;                    /* [S,B,4D] saved gate activations */
    float *c1new, *c2new, *h1raw, *h2raw;    /* [S,B,D] */
    float *c1z, *h1z, *c2z, *h2z;            /* [S+1,B,D] zoned state, slot 0 = zero state */
    float *ctx, *cum;                        /* [S+1,B,MD], [S+1,B,T] */
    float *align, *q;                        /* [S,B,T], [S,B,A] */
    /* backward only */
    float *dH2, *dCtx;                       /* [S,B,D], [S,B,MD]: in = gradients from the projections, out = totals */
    float *dG1, *dG2, *dQ;                   /* [S,B,4D], [S,B,4D], [S,B,A] out */
    float* dkeys;                            /* [B,T,A] accumulated (caller zeroes) */
    float *p_conv_w, *p_conv_b, *p_Wl, *p_va, *p_ba;   /* per-row partials of the attention weight gradients (caller zeroes) */
    float* work;                             /* scratch, >= B*(8*D + 2*T + T*NF) + S*B*D + (S+1)*B*MD floats */
    /* fused attention path (taco_op_attn_*_fused; used when all three are set and taco_op_attn_fused_supported): */
    float* th;                               /* [S,B,T,A] tanh(pre) saved by the forward loop */
    float* de;                               /* [S,B,T] d(energy) saved by the backward loop */
    float* attn_work;                        /* >= (KW+1)*A + B*(KW+2)*A floats: folded filter bank | hoisted accumulators */
} TacoDecLoop;
int taco_op_decoder_loop_fwd(const TacoDecLoop* p, void* stream);
int taco_op_decoder_loop_bwd(const TacoDecLoop* p, void* stream);

typedef struct TacoLstmLoop {
    int32_t B, T, H, C, reverse;             /* one direction of the encoder BiLSTM; C = input width */
    float zoneout;
    uint32_t zo_keep_thresh;
    uint64_t seed;
    uint32_t stream_c, stream_h;
    const float* K;                          /* [C+H, 4H] */
    const int32_t* lens;
    const float* xw;                         /* [B,T,4H] hoisted input projection (+bias) */
    float *acts, *cnew, *cz, *hz;            /* [T,B,4H], [T,B,H], [T+1,B,H], [T+1,B,H] */
    float* out;                              /* fwd: raw h scattered to out[b,pos,:H]; bwd: d(out) gathered from it */
    long long out_sb, out_st;                /* batch / time strides of `out` (a column block of [B,T,2H] or [B,T,MD]) */
    float *dxw, *dg;                         /* bwd: [B,T,4H] (caller zeroes), [T,B,4H] */
    float* work;                             /* scratch, >= B*9*H floats */
} TacoLstmLoop;
int taco_op_lstm_loop_fwd(const TacoLstmLoop* p, void* stream);
int taco_op_lstm_loop_bwd(const TacoLstmLoop* p, void* stream);

/* The GRU of the reference encoder (Tacotron.py:201, GRUCell over the Tp frames of the convolution stack) sequenced in C:
 * forward per step  GX[s] += h . Wg[In:], (r|u) = sigmoid, rh = r * h, CX[s] += rh . Wc[In:], h' = u * h + (1 - u) * tanh(CX[s]);
 * backward the mirrored chain.  GX / CX hold the hoisted input projections (+ bias) on entry and the activations on exit. */
typedef struct TacoGruLoop {
    int32_t B, Tp, In, R;                    /* batch, steps, input width, units */
    const float *Wg, *Wc;                    /* [In+R, 2R], [In+R, R] */
    float *GX, *CX, *hg, *rh;                /* [Tp,B,2R], [Tp,B,R], [Tp+1,B,R] (slot 0 = initial state), [Tp,B,R] */
    /* backward only */
    float *dGX, *dCX;                        /* [Tp,B,2R], [Tp,B,R] out */
    float *dh;                               /* [B,R]: in = d(final state), out = d(initial state) */
    float* work;                             /* >= 2*B*R floats */
} TacoGruLoop;
int taco_op_gru_loop_fwd(const TacoGruLoop* p, void* stream);
int taco_op_gru_loop_bwd(const TacoGruLoop* p, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* TACO_B200_TRAIN_H */
