/* taco_b200_audio.h -- C ABI of the mel -> waveform step, sm_100a.
 *
 * Replaces, in the reference (Python, no FFI of its own):
 *   Utils/AudioProcessing/AudioPreprocess.py:152-186  mel_to_audio_serie(mel, hparams)
 *       :469-484 _denormalize, :447 _db_to_amp, :163-165 pinv(mel basis) . amp, :174 ** power,
 *       :175-185 random phases + griffin_lim_iters x (librosa.stft -> unit phases -> librosa.istft),
 *       :45-48   inv_preemphasis (scipy.signal.lfilter([1], [1, -k]))
 * called from Utils/Tacotron_synthesizer.py (inference mode) and TacotronModel/train.py:124.
 *
 * Plain pointers and sizes only.  All *_dev pointers are device memory on the plan's device; the *_host entry point
 * takes host memory and does its own copies.  Every function returns 0 on success, a negative code otherwise
 * (-1 bad argument, -2 out of memory, -3 CUDA error; taco_gl_last_error() has the text).
 *
 * Layouts (row-major, fp32):
 *   mel    [B, F, n_mels]      normalised mel frames as the synthesizer emits them (F = padded frame count)
 *   mag    [B, F, n_bins]      n_bins = n_fft/2 + 1; linear magnitudes already raised to `power`
 *   turns  [B, F, n_bins]      optional initial phases in turns (phase = 2*pi*turns); NULL = Philox (see below)
 *   audio  [B, audio_stride]   utterance b holds hop * (n_frames[b] - 1) samples, zero beyond
 *   n_frames [B] int32         valid frames per utterance (NULL = all F); the reference runs one utterance per call,
 *                              so every utterance is processed with its own length, never the batch's padded one
 * Initial phases when turns == NULL (the reference draws them from the unseeded np.random.rand, :175): element
 *   idx = k * n_frames[b] + t uses word idx & 3 of Philox4x32-10(counter = (idx >> 2, b + utterance0, idx >> 34, 0x474C494D),
 *   key = seed); turns = (word >> 8) * 2^-24.   oracle/griffin_lim.py restates it.
 */
#ifndef TACO_B200_AUDIO_H
#define TACO_B200_AUDIO_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct TacoGl TacoGl;

/* hparams of the de-normalisation (Utils/Hyperparams.py:77-89) */
typedef struct TacoMelNorm {
    int32_t signal_normalization;            /* 0: mel is already dB - ref_level */
    int32_t allow_clipping_in_normalization;
    int32_t symmetric_mels;
    float max_abs_value;                     /* 4.0 */
    float min_level_db;                      /* -100 */
    float ref_level_db;                      /* 20 */
    float power;                             /* 1.2 */
} TacoMelNorm;

/* n_fft must be 2048 (hparams.n_fft; the in-shared-memory FFT is built for it), 0 < hop <= win <= n_fft, win even.
 * inv_mel_basis: host, [n_bins, n_mels] = pinv(librosa.filters.mel(...)) (:163-164), copied. */
int taco_gl_create(int device, int n_fft, int hop, int win, int n_mels, const float* inv_mel_basis, TacoGl** out);
void taco_gl_destroy(TacoGl* g);
const char* taco_gl_last_error(const TacoGl* g);

/* D = max(1e-10, inv_basis . 10^((denorm(mel) + ref_level_db) / 20)) ^ power       (:157-165,174) */
int taco_gl_mel_to_mag(TacoGl* g, const float* mel_dev, const int32_t* n_frames_dev, int B, int F, const TacoMelNorm* norm,
                       float* mag_dev, void* stream);

/* istft(mag * e^{i phase0}); iters x { stft; phase <- angle; istft }; optional inverse pre-emphasis (k = 0: off).
 * Needs hop * (n_frames[b] - 1) > n_fft / 2 (librosa's reflect padding).  Workspace grows inside the plan. */
int taco_gl_griffin_lim(TacoGl* g, const float* mag_dev, const float* turns_dev, const int32_t* n_frames_dev, uint64_t seed,
                        int utterance0, int B, int F, int iters, float preemphasis_k, float* audio_dev, long long audio_stride,
                        void* stream);

/* Relative spectral distance || |stft(audio)| - mag || / || mag || per utterance (the convergence measure of the iteration);
 * out_dev [B]. */
int taco_gl_spectral_distance(TacoGl* g, const float* audio_dev, long long audio_stride, const float* mag_dev,
                              const int32_t* n_frames_dev, int B, int F, float* out_dev, void* stream);

/* Host-buffer form of the whole step: what mel_to_audio_serie binds.  n_frames_host may be NULL. */
int taco_mel_to_audio(TacoGl* g, const float* mel_host, const int32_t* n_frames_host, int B, int F, const TacoMelNorm* norm,
                      uint64_t seed, int utterance0, int iters, float preemphasis_k, float* audio_host, long long audio_stride);

#ifdef __cplusplus
}
#endif
#endif
