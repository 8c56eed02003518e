/*
 * taco_b200.h -- C ABI of the B200-native Tacotron-2 mel-generation hot path.
 *
 * The reference (Thien223/Tacotron_Prosody_multi_GPU) has no FFI: its operator boundary is the
 * Python class surface of TacotronModel/modules/*.py driven by Utils/Tacotron_synthesizer.py.
 * Each entry point below replaces the TF-1.9 graph segment cited beside it; the Python classes
 * of tacotron_prosody_multi_gpu_b200/ (same names as the reference's) bind these symbols with
 * ctypes and pass raw device pointers (torch tensors are only the memory carrier).
 *
 * Conventions: return 0 on success, negative TACO_ERR_* otherwise (never throws); the caller
 * owns every I/O buffer (contiguous fp32 / int32 DEVICE memory unless stated); the handle owns
 * weight copies and workspaces; one handle per GPU; calls are asynchronous on `stream`
 * (a cudaStream_t passed as void*); a handle is not thread-safe.  There is no CPU fallback.
 */
#ifndef TACO_B200_H
#define TACO_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct TacoHandle TacoHandle;

enum {
    TACO_OK = 0,
    TACO_ERR_INVALID = -1,      /* bad argument / shape */
    TACO_ERR_UNSUPPORTED = -2,  /* hyper-parameter combination the kernels do not cover */
    TACO_ERR_CUDA = -3,         /* CUDA runtime error (see taco_last_error) */
    TACO_ERR_MISSING_WEIGHT = -4,
    TACO_ERR_STATE = -5         /* call order (e.g. decode before weights are finalised) */
};

enum { TACO_MODE_FREE = 0, TACO_MODE_TEACHER = 1 };
enum { TACO_DTYPE_F32 = 0 };
enum { TACO_PRECISION_FP32 = 0, TACO_PRECISION_BF16_CONV = 1, TACO_PRECISION_BF16 = 2 };

/* Subset of Utils/Hyperparams.py:4-239 that shapes the path (network dims :43-63, GST :7-14,
 * run params :157-172).  Field names follow the reference. */
typedef struct TacoHParams {
    int32_t n_symbols;          /* len(hangul_symbol), Tacotron.py:79-88,180 */
    int32_t embedding_dim;      /* 512 */
    int32_t conv_num_layers;    /* 3 */
    int32_t conv_kernel_size;   /* 5 */
    int32_t conv_channels;      /* 512 */
    int32_t enc_lstm_hidden_size; /* 256 */
    int32_t use_gst;            /* 1 */
    int32_t num_gst;            /* 10 */
    int32_t num_heads;          /* 4 */
    int32_t style_embed_depth;  /* 256 */
    int32_t style_att_dim;      /* 128 */
    int32_t reference_depth;    /* 128 */
    int32_t n_reference_filters;/* 6 */
    int32_t reference_filters[8];
    int32_t attention_dim;      /* 128 */
    int32_t attention_filters;  /* 32 */
    int32_t attention_kernel;   /* 31 */
    int32_t prenet_layers[2];   /* 256, 256 */
    int32_t decoder_lstm_units; /* 256 */
    int32_t num_mels;           /* 80 */
    int32_t outputs_per_step;   /* 2 */
    int32_t postnet_num_layers; /* 5 */
    int32_t postnet_kernel_size;/* 5 */
    int32_t postnet_channels;   /* 512 (+ style_embed_depth = 768 wide, Postnet.py:22) */
    int32_t mask_encoder;       /* 1 */
    int32_t smoothing;          /* 0 */
    int32_t cumulative_weights; /* 1 */
    int32_t stop_at_any;        /* 1 */
    int32_t clip_outputs;       /* 1 */
    int32_t symmetric_mels;     /* 1 */
    float zoneout_rate;         /* 0.1 */
    float dropout_rate;         /* 0.5 */
    float max_abs_value;        /* 4.0 */
    float lower_bound_decay;    /* 0.1 */
} TacoHParams;

/* Replaces graph construction in Tacotron.initialize (Tacotron.py:46-349). */
int taco_create(const TacoHParams* hp, int device, TacoHandle** out);
void taco_destroy(TacoHandle* h);
const char* taco_last_error(const TacoHandle* h);
/* ABI / build identification ("taco_b200 <version> sm_100a"). */
const char* taco_version(void);

/* Replaces tf.train.Saver.restore (Utils/Tacotron_synthesizer.py:49-50): one call per variable,
 * `name` = reference variable scope name below model/inference/ (see weights.py), TF layouts:
 * Dense [in,out]; conv1d [k,in,out]; conv2d [kh,kw,in,out]; LSTM kernel [in+h,4h] (i,j,f,o);
 * GRU gates [in+h,2h] / candidate [in+h,h].  `data` may be host or device memory. */
int taco_set_weight(TacoHandle* h, const char* name, const void* data, const int64_t* shape, int rank, int dtype);
/* Folds BN, packs the per-CTA resident decoder shards; errors if a variable is missing. */
int taco_finalize_weights(TacoHandle* h);
/* TACO_PRECISION_FP32: fp32 SIMT everywhere.  TACO_PRECISION_BF16_CONV: encoder-conv / postnet
 * contractions on tcgen05 tensor cores with bf16 operands, fp32 accumulation.
 * TACO_PRECISION_BF16: additionally the decoder step's contractions (prenet, LSTM gates, query,
 * frame/stop projection) run on tensor cores with bf16 operands and the attention values are
 * streamed as bf16 (dataflow persistent kernel); recurrent state, softmax, outputs stay fp32. */
int taco_set_precision(TacoHandle* h, int precision);

/* Embedding -> EncoderConvolution -> EncoderLSTM -> (reference encoder -> GST | random-weight
 * GST) -> concat (Tacotron.py:179-223; Encoder.py:40-46,68-77; Attention.py:243-341).
 *  ids [B,T] int32 (pad 0), lens [B] int32; ref_mel [B,Tm,80] or NULL; gst_weights [heads,tokens]
 *  (used when ref_mel == NULL: the injected stand-in for tf.random_uniform, Tacotron.py:217);
 *  memory_out [B,T,768] = `encoder_outputs` (Tacotron.py:224). */
int taco_encode(TacoHandle* h, const int32_t* ids, const int32_t* lens, int B, int T,
                const float* ref_mel, int Tm, const float* gst_weights, float* memory_out, void* stream);

/* dynamic_decode(CustomDecoder(TacotronDecoderCell, helper)) as ONE persistent cooperative kernel
 * (Tacotron.py:262-284,312; Decoder.py:117-146,246-292; Attention.py:153-210,345-401;
 * Utils/Helpers.py:36-59,115-128).
 *  mode TACO_MODE_FREE: TacoTestHelper, batch-global stop (Helpers.py:51-52), <= S_max steps.
 *  mode TACO_MODE_TEACHER: TacoTrainingHelper with ratio 1.0 / GTA; mel_targets [B,F_tgt,80].
 *  prenet_masks: uint8 [S_max,B,2,256] keep bits or NULL -> Philox4x32-10 keyed by philox_seed
 *   (prenet dropout is always on, Decoder.py:350-351).
 *  reset_state != 0: start from zero_state (Decoder.py:212-244); 0: continue from the state kept
 *   in the handle by the previous taco_decode call (single-step `cell(inputs, state)` use).
 *  Outputs (strides are for S_max): frames [B, 2*S_max, 80] raw cell outputs (unclipped),
 *  stop [B, 2*S_max] (sigmoid unless logits_only), align_hist [S_max, B, T] step-major alignments,
 *  steps_out: device int32, number of steps executed S (= first any-stop step + 1, or S_max). */
int taco_decode(TacoHandle* h, int mode, const float* memory, const int32_t* lens, int B, int T,
                const float* mel_targets, int F_tgt, int S_max,
                const uint8_t* prenet_masks, uint64_t philox_seed, int reset_state, int logits_only,
                float* frames, float* stop, float* align_hist, int32_t* steps_out, void* stream);

/* alignments = transpose(alignment_history.stack(), [1,2,0]) (Tacotron.py:312):
 * align_hist [S_stride>=S rows used: S, B, T] -> out [B, T, S]. */
int taco_transpose_alignments(TacoHandle* h, const float* align_hist, int S, int B, int T, float* out, void* stream);

/* reshape + clip + Postnet + postnet_projection + residual + clip (Tacotron.py:280-309;
 * Postnet.py:28-36).  frames [B, F_stride, 80] raw decoder frames, first F used;
 * decoder_output [B,F,80] (clipped) and mel_outputs [B,F,80]. */
int taco_postnet(TacoHandle* h, const float* frames, int B, int F, int F_stride,
                 float* decoder_output, float* mel_outputs, void* stream);

/* Number of kernels launched by this handle since creation (bench.py's gpu_launches). */
int64_t taco_launch_count(const TacoHandle* h);

/* Grid of the persistent BiLSTM kernel of taco_encode (EncoderLSTM, Encoder.py:68-77).  0 (default): the lowest-latency grid
 * (128 CTAs at B = 64, cooperative launch -- cannot start while a cluster decoder of another handle occupies 112 SMs);
 * max_ctas > 0: at most that many CTAs (32 at B = 64: 16 units per CTA, ~2x the latency, but it runs on the 36 SMs a running
 * decoder leaves free, so handles that keep several batches in flight overlap the whole encoder with the decoder). */
int taco_set_encoder_ctas(TacoHandle* h, int max_ctas);

/* Device time of the most recent run of a stage, from CUDA events recorded on the caller's stream
 * around it (blocks until the stage has finished).  TACO_STAGE_DECODER_KERNEL brackets only the
 * persistent decoder kernel launch -- the figure bench.py's roofline uses. */
enum { TACO_STAGE_ENCODE = 0, TACO_STAGE_DECODER_KERNEL = 1, TACO_STAGE_POSTNET = 2, TACO_N_STAGES = 3 };
int taco_stage_ms(TacoHandle* h, int stage, float* ms_out);

/* Diagnostics: copies a named internal workspace buffer to HOST memory (synchronises the device).
 * Used by the profiling scripts (e.g. "dec2_timing": per-CTA per-stage nanoseconds of the dataflow
 * decoder when the environment variable TACO_DEC2_TIMING is set). */
int taco_debug_read_workspace(TacoHandle* h, const char* name, void* host_dst, size_t bytes);

/* Decoder recurrent state kept in the handle (TacotronDecoderCellState, Decoder.py:419-438):
 * copies c1,h1,c2,h2 [B,256], attention [B,768], alignments (cumulative) [B,T] out to / in from
 * caller DEVICE buffers.  Any pointer may be NULL to skip that field. */
int taco_get_decoder_state(TacoHandle* h, int B, int T, float* c1, float* h1, float* c2, float* h2,
                           float* attention, float* alignments, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* TACO_B200_H */
