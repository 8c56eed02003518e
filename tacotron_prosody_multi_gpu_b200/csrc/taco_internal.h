// Internal declarations shared by the .cu translation units of libtaco_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include "../../include/taco_b200.h"

// ---------------------------------------------------------------------------------------------
// decoder.cu
// ---------------------------------------------------------------------------------------------
struct DecParams {
    int B, T, RBP, GB, GN, U, NS;
    int S_max, S_cap, t0, mode;
    int stop_at_any, cumulative, smoothing, stop_sigmoid;
    float zoneout, keep_scale;
    // resident weights
    const float* blobs;  // [GN][blob_floats]
    int blob_floats;
    const float *Wl, *conv_w, *conv_b, *v_a, *b_a;
    // encoder memory
    const float *values, *keys;
    const int* lens;
    // exchange buffers (XT layout [GB][K][RBP] unless noted)
    float *xT, *p1T, *p2T, *ctxT, *h1zT, *h1rawT, *h2zT, *h2rawT;
    float *q;    // [B][128]
    float *pre;  // [B][T][128]
    float *cum;  // [B][T]
    float *c1, *c2;  // [B][256]
    // inputs
    const float* targets;
    int F_tgt;
    const uint8_t* masks;
    uint64_t seed;
    // outputs
    float *frames, *stop, *hist;
    int* steps_out;
    int* stop_count;  // [S_max], zeroed
    unsigned* barrier;
};

int taco_decoder_blob_floats(int U);
size_t taco_decoder_smem_bytes(int U);
cudaError_t taco_launch_decoder(const DecParams& p, cudaStream_t stream);
cudaError_t taco_launch_transpose_alignments(const float* hist, float* out, int S, int B, int T, cudaStream_t stream);

// ---------------------------------------------------------------------------------------------
// decoder3.cu : cluster persistent decoder (16-CTA clusters, DSMEM exchanges, bf16 mma.sync contractions)
// ---------------------------------------------------------------------------------------------
#define TACO_DEC3_CLUSTER 16
#define TACO_DEC3_MAX_CLUSTERS 8
#define TACO_DEC3_TMAX 512
#define TACO_DEC3_RES_WORDS (221 * 128 + 176)
#define TACO_DEC3_STREAM_TILES 320
#define TACO_DEC3_ACAP 1392   // shared-memory words for the alignment rows of one cluster: rows per cluster * round_up(T, 4) <= this

struct Dec3Params {
    int B, T, NCL, R;       // NCL clusters; cluster i owns batch rows [i*R, min(B, i*R + R))
    int TP;                 // round_up(T, 4): row stride of the per-cluster alignment rows (R * TP <= TACO_DEC3_ACAP)
    int S_max, S_cap, t0, mode;
    int stop_at_any, cumulative, smoothing, stop_sigmoid;
    int pre_reg;            // location pre-activations live in registers (T <= 128, or any T split into parts) -- else in pre_scratch
    int parts;              // attention work items per row: 1, or ceil(T / 128) position slices of 128 owned by different CTAs
    int dbg;                // experiment bitmask (TACO_DEC3_DBG): 1 = never consume the stop flags
    int cont;               // continuation launch (reset_state == 0): pre-activations are loaded from pre_scratch
    float zoneout, keep_scale;
    const uint32_t* res_blobs;   // [16 ranks][TACO_DEC3_RES_WORDS] A-fragment tiles resident in shared memory + fp32 biases
    const uint4* stream_blobs;   // [16 ranks][320 tiles][32 lanes] A-fragment tiles streamed from L2: W1c [4][48], W1h [4][16], W2z [4][16]
    const uint4* stop_tiles;     // [64 k-steps][32 lanes] A-fragment tiles of the stop projection (rows 0..1 of the m16 tile)
    float stop_bias[2];
    const uint32_t *g_hi, *g_lo; // G = conv31 . Wl, B-fragment-major [2 ks][16 nt][32 lanes][2], hi / lo bf16 split
    const float *va, *bconst;    // [128] attention_variable; attention_bias + conv_bias . Wl
    const uint16_t* values;      // bf16 [B][T][768]
    const float* keys;           // fp32 [B][T][128]
    const int* lens;
    // recurrent state, fp32 row-major (TacotronDecoderCellState, Decoder.py:419-438)
    float *c1, *h1z, *c2, *h2z;  // [B][256]
    float* ctx_state;            // [B][768]
    float* cum;                  // [B][T]
    float* h2raw;                // [B][256] raw LSTM-2 output of the last step (input of the folded prenet dense_1)
    float* pre_scratch;          // [B][T][128] location pre-activations: working copy (!pre_reg) / state saved between launches
    const float* targets;
    int F_tgt;
    const uint8_t* masks;
    uint64_t seed;
    uint32_t* act_hist;          // [S_max][B][512] bf16 pairs: projection input rows [h2raw | context] of every step
    float *stop, *hist;
    int* steps_out;
    int* flags;       // [S_max][NCL]: 0 not ready, 1 continue, 2 stop (free-running)
    int* abort_flag;
    long long* timing;  // optional [grid][16] per-CTA stage nanoseconds (thread 0), or null
    long long* trace;   // optional [grid][64][2] (clock64, globaltimer) stamps of step trace_step (TACO_DEC3_TRACE), or null
    int trace_step;
};
size_t taco_decoder3_smem_bytes();
// frames[b][2s + f][m] = act_hist[s][b][:] . Wo[:, 80 f + m] + bo  for s < *steps (FrameProjection, Decoder.py:371-380)
cudaError_t taco_launch_frame_projection(const uint32_t* act_hist, const uint32_t* wo_t, const float* bo, const int* steps, int S_max, int B,
                                         int S_cap, float* frames, cudaStream_t stream);
cudaError_t taco_decoder3_max_clusters(int* out);
cudaError_t taco_launch_f32_to_bf16(const float* in, void* out, size_t n, cudaStream_t stream);   // frame_proj.cu; n % 4 == 0
cudaError_t taco_launch_decoder3(const Dec3Params& p, cudaStream_t stream);

// ---------------------------------------------------------------------------------------------
// conv.cu : conv1d('same') / dense as implicit GEMM, fused scale/shift + activation epilogue
// ---------------------------------------------------------------------------------------------
enum { ACT_NONE = 0, ACT_RELU = 1, ACT_TANH = 2 };

struct ConvParams {
    const float* x;       // [B][T][Cin]; row stride x_stride (floats) between t, batch stride x_bstride
    long long x_stride, x_bstride;
    const float* w;       // [k][Cin][Cout]
    const float* scale;   // [Cout] or null (=1)
    const float* shift;   // [Cout] or null (=0)
    float* y;             // [B][T][Cout]; row stride y_stride, batch stride y_bstride
    long long y_stride, y_bstride;
    const float* residual;  // optional [B][T][Cout] dense, added before clip
    int B, T, Cin, Cout, k, act;
    int clip;             // clamp to [clip_lo, clip_hi] after residual
    float clip_lo, clip_hi;
    float in_clip_lo, in_clip_hi;
    int in_clip;          // clamp the INPUT while loading (decoder_output clip, Tacotron.py:282-284)
};
cudaError_t taco_launch_conv1d_f32(const ConvParams& p, cudaStream_t stream);

// ---------------------------------------------------------------------------------------------
// conv_tc.cu : tcgen05 / TMEM / TMA implicit-GEMM conv1d('same') / dense, bf16 operands
// ---------------------------------------------------------------------------------------------
struct ConvTcParams {
    const void* x;        // bf16 [B][F][Cin], Cin multiple of 64 (channel-padded)
    const void* wt;       // bf16 [Cout][taps*Cin]  (K-major, pre-transposed at load time)
    int B, F, Cin, taps, Cout;
    const float* scale;   // [Cout] or null
    const float* shift;   // [Cout] or null
    int act;
    void* y_bf16;         // bf16 [B][F][y_stride] (ignored when out_f32_only)
    int y_stride;
    float* y_f32;         // optional fp32 copy [B][F][Cout]; the only output when out_f32_only
    int out_f32_only;     // dense + residual + clip epilogue, Cout in {80, 128}
    const float* residual;
    int clip;
    float clip_lo, clip_hi;
};
cudaError_t taco_launch_conv_tc(const ConvTcParams& p, int n_sm, cudaStream_t stream);
cudaError_t taco_launch_transpose_bf16(const float* in, void* out, int B, int T, int C, int Cpad, int Tp, int margin, int nshift, int shift0,
                                       cudaStream_t st);
cudaError_t taco_launch_conv_wgrad_tc(const void* xT, const void* dyT, float* dw, int K, int Ci, int Ci_pad, int Co, int taps, int n_sm,
                                      cudaStream_t st);
cudaError_t taco_launch_prep_bf16(const float* in, float* out32, void* out16, int B, int F, long long in_bstride, int C,
                                  int Cpad, float lo, float hi, int clip, cudaStream_t stream);

// ---------------------------------------------------------------------------------------------
// encoder.cu
// ---------------------------------------------------------------------------------------------
cudaError_t taco_launch_embedding(const int32_t* ids, const float* table, float* out, int n_tok, int dim, int n_symbols,
                                  cudaStream_t stream);
struct BiLstmParams {
    int B, T, RBP, GB, GN, U;  // GN unit tiles per direction; grid = 2 * GB * GN
    const float* blobs;        // [2][GN][256*4U + 4U]  recurrent kernel shard [256][4U] + zero pad
    int blob_floats;
    const float* xw;           // [2][B*T][1024] input projection (+bias), gate order i,j,f,o
    const int* lens;
    float* hT;                 // [2 dir][2 parity][GB][256][RBP]
    float* out;                // [B][T][out_stride]; fw at +0, bw at +256
    int out_stride;
    float zoneout;
    unsigned* barrier;         // [2] one counter per direction
};
int taco_bilstm_blob_floats(int U);
cudaError_t taco_launch_bilstm(const BiLstmParams& p, cudaStream_t stream);
cudaError_t taco_launch_style_fill(float* memory, const float* style, int B, int T, int style_dim, int mem_dim,
                                   int style_bstride, cudaStream_t stream);

// ---------------------------------------------------------------------------------------------
// refenc.cu : reference encoder (conv2d s2 + BN + ReLU stack, GRU, dense tanh) and GST attention
// ---------------------------------------------------------------------------------------------
cudaError_t taco_launch_conv2d_s2(const float* x, const float* w, const float* scale, const float* shift, float* y, int B,
                                  int H, int W, int Cin, int Cout, cudaStream_t stream);
cudaError_t taco_launch_ref_gru(const float* x, int B, int Tp, int In, const float* Wg, const float* bg, const float* Wc,
                                const float* bc, const float* Wd, const float* bd, float* ref_out, cudaStream_t stream);
cudaError_t taco_launch_gst_attention(const float* ref, const float* tokens, const float* Wq, const float* bq,
                                      const float* Wk, const float* bk, const float* v, const float* g, const float* b,
                                      float* style, int B, int n_tok, int tok_dim, int heads, int units,
                                      cudaStream_t stream);
cudaError_t taco_launch_style_random(const float* gst_weights, const float* tokens, float* style, int heads, int n_tok,
                                     int tok_dim, cudaStream_t stream);
