// train_loops.cu -- the time loops of the training step sequenced from C: all S decoder steps (forward and
// BPTT) and all T positions of one encoder-LSTM direction.  Same primitives and the same order as the host
// composition in train/step.py (which stays the readable restatement and is what the CPU tests check); moving the
// loops here removes ~10 000 host-language launches per step.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/taco_b200_train.h"

#define OK(x) do { int rc_ = (x); if (rc_) return rc_; } while (0)

extern "C" {

int taco_op_decoder_loop_fwd(const TacoDecLoop* p, void* st) {
    if (!p || !p->W1 || !p->W2 || !p->G1x || !p->acts1 || !p->acts2 || !p->ctx || !p->cum) return -1;
    const int B = p->B, T = p->T, S = p->S, D = p->D, MD = p->MD, A = p->A, U2 = p->U2, G = 4 * p->D;
    const float* W1c = p->W1 + (size_t)U2 * G;          // rows of the context
    const float* W1h = p->W1 + (size_t)(U2 + MD) * G;   // rows of h1
    const float* W2h = p->W2 + (size_t)D * G;
    // the per-step contractions are split-K with atomic accumulation: zero their outputs once for all steps instead of
    // one memset per step
    cudaStream_t cs = (cudaStream_t)st;
    if (cudaMemsetAsync(p->acts1, 0, sizeof(float) * (size_t)S * B * G, cs) != cudaSuccess ||
        cudaMemsetAsync(p->acts2, 0, sizeof(float) * (size_t)S * B * G, cs) != cudaSuccess ||
        cudaMemsetAsync(p->q, 0, sizeof(float) * (size_t)S * B * A, cs) != cudaSuccess)
        return -3;
    // fused attention (train_attn_fused.cu): query projection + attention step in one launch, tanh(pre) saved for the backward pass
    const bool fused = p->th && p->attn_work && taco_op_attn_fused_supported(T, D, A, MD, p->NF, p->KW);
    float* Uf = p->attn_work;
    float* c0 = Uf ? Uf + (size_t)p->KW * A : nullptr;
    if (fused) OK(taco_op_attn_fold(p->conv_w, p->conv_b, p->Wl, Uf, c0, st));
    for (int s = 0; s < S; ++s) {
        const size_t oD = (size_t)s * B * D, oG = (size_t)s * B * G, nD = oD + (size_t)B * D;
        float* a1 = p->acts1 + oG;
        float* a2 = p->acts2 + oG;
        OK(taco_op_gemm_pair_acc(p->ctx + (size_t)s * B * MD, MD, MD, p->h1z + oD, D, D, W1c, G, a1, G, B, G, st));   // W1c | W1h adjacent rows
        OK(taco_op_lstm_fwd(a1, p->G1x + oG, nullptr, p->c1z + oD, p->h1z + oD, nullptr, 0, p->zoneout, 1, p->zo_keep_thresh, p->seed,
                            p->stream1, p->stream1 + 1, s * B, p->c1new + oD, p->c1z + nD, p->h1raw + oD, p->h1z + nD, B, D, st));
        OK(taco_op_gemm_pair_acc(p->h1raw + oD, D, D, p->h2z + oD, D, D, p->W2, G, a2, G, B, G, st));
        OK(taco_op_lstm_fwd(a2, nullptr, p->b2, p->c2z + oD, p->h2z + oD, nullptr, 0, p->zoneout, 1, p->zo_keep_thresh, p->seed,
                            p->stream2, p->stream2 + 1, s * B, p->c2new + oD, p->c2z + nD, p->h2raw + oD, p->h2z + nD, B, D, st));
        float* q = p->q + (size_t)s * B * A;
        if (fused) {
            OK(taco_op_attn_fwd_fused(p->h2raw + oD, p->Wq, Uf, c0, p->ba, p->va, p->cum + (size_t)s * B * T, p->keys, p->values, p->lens,
                                      q, p->th + (size_t)s * B * T * A, p->align + (size_t)s * B * T, p->cum + (size_t)(s + 1) * B * T,
                                      p->ctx + (size_t)(s + 1) * B * MD, B, T, D, MD, st));
            continue;
        }
        OK(taco_op_gemm(p->h2raw + oD, D, p->Wq, A, q, A, B, A, D, 0, 0, 1.f, nullptr, st));
        OK(taco_op_attn_fwd(q, p->cum + (size_t)s * B * T, p->keys, p->values, p->lens, p->conv_w, p->conv_b, p->Wl, p->va, p->ba,
                            p->align + (size_t)s * B * T, p->cum + (size_t)(s + 1) * B * T, p->ctx + (size_t)(s + 1) * B * MD, B, T, A, MD,
                            p->NF, p->KW, st));
    }
    return 0;
}

int taco_op_decoder_loop_bwd(const TacoDecLoop* p, void* st) {
    if (!p || !p->dH2 || !p->dCtx || !p->dG1 || !p->dG2 || !p->dQ || !p->dkeys || !p->work) return -1;
    const int B = p->B, T = p->T, S = p->S, D = p->D, MD = p->MD, A = p->A, U2 = p->U2, G = 4 * p->D;
    const float* W1c = p->W1 + (size_t)U2 * G;
    const float* W1h = p->W1 + (size_t)(U2 + MD) * G;
    const float* W2h = p->W2 + (size_t)D * G;
    const size_t BD = (size_t)B * D;
    float* w = p->work;
    // work layout: 8 state vectors [B,D] | dcum x2 [B,T] | attention scratch [B,T,NF] | per-step d(h1raw) [S,B,D] |
    // per-step d(ctx) carry [S+1,B,MD]  (per-step slots are zeroed once: their producers accumulate with split-K atomics)
    const size_t n_work = (size_t)B * (8 * D + 2 * T + (size_t)T * p->NF) + (size_t)S * B * D + (size_t)(S + 1) * B * MD;
    if (cudaMemsetAsync(w, 0, sizeof(float) * n_work, (cudaStream_t)st) != cudaSuccess) return -3;
    float *dc1 = w, *dh1z = w + BD, *dc2 = w + 2 * BD, *dh2z = w + 3 * BD;
    float *dc1n = w + 4 * BD, *dh1zn = w + 5 * BD, *dc2n = w + 6 * BD, *dh2zn = w + 7 * BD;
    float* dcum = w + 8 * BD;
    float* dcum_n = dcum + (size_t)B * T;
    float* attn_scratch = dcum_n + (size_t)B * T;      // [B][T][NF]
    float* dh1raw_all = attn_scratch + (size_t)B * T * p->NF;
    float* dcarry_all = dh1raw_all + (size_t)S * B * D;   // slot s+1 = carry INTO step s (slot S = 0)
    const bool fused = p->th && p->de && p->attn_work && taco_op_attn_fused_supported(T, D, A, MD, p->NF, p->KW);
    float* Uf = p->attn_work;               // folded by the forward loop of this step
    for (int s = S - 1; s >= 0; --s) {
        const size_t oD = (size_t)s * B * D, oG = (size_t)s * B * G;
        float* dctx = p->dCtx + (size_t)s * B * MD;
        float* dh1raw = dh1raw_all + (size_t)s * B * D;
        const float* dctx_carry = dcarry_all + (size_t)(s + 1) * B * MD;
        float* dctx_carry_out = dcarry_all + (size_t)s * B * MD;
        float* dq = p->dQ + (size_t)s * B * A;
        float* dh2 = p->dH2 + oD;
        if (fused) {
            OK(taco_op_attn_bwd_fused(dctx, dctx_carry, dcum, p->align + (size_t)s * B * T, p->th + (size_t)s * B * T * A, p->values, p->lens,
                                      Uf, p->va, p->Wq, p->de + (size_t)s * B * T, dq, dh2, dcum_n, attn_scratch, B, T, D, MD, st));
        } else {
            OK(taco_op_ew(TACO_EW_ADD, dctx, dctx, dctx_carry, 1.f, (long long)B * MD, st));
            OK(taco_op_attn_bwd(dctx, dcum, p->q + (size_t)s * B * A, p->cum + (size_t)s * B * T, p->align + (size_t)s * B * T, p->keys,
                                p->values, p->lens, p->conv_w, p->conv_b, p->Wl, p->va, p->ba, dq, dcum_n, p->dkeys, p->p_conv_w,
                                p->p_conv_b, p->p_Wl, p->p_va, p->p_ba, attn_scratch, B, T, A, MD, p->NF, p->KW, st));
            OK(taco_op_gemm(dq, A, p->Wq, A, dh2, D, B, D, A, 0, 1, 1.f, nullptr, st));
        }
        { float* t_ = dcum; dcum = dcum_n; dcum_n = t_; }
        OK(taco_op_lstm_bwd(p->acts2 + oG, p->c2z + oD, p->h2z + oD, p->c2new + oD, nullptr, 0, p->zoneout, 1, p->zo_keep_thresh, p->seed,
                            p->stream2, p->stream2 + 1, s * B, dh2, dc2, dh2z, p->dG2 + oG, dc2n, dh2zn, B, D, st));
        OK(taco_op_gemm_nt_split_acc(p->dG2 + oG, G, p->W2, G, dh1raw, D, D, dh2zn, D, D, B, G, st));
        OK(taco_op_lstm_bwd(p->acts1 + oG, p->c1z + oD, p->h1z + oD, p->c1new + oD, nullptr, 0, p->zoneout, 1, p->zo_keep_thresh, p->seed,
                            p->stream1, p->stream1 + 1, s * B, dh1raw, dc1, dh1z, p->dG1 + oG, dc1n, dh1zn, B, D, st));
        OK(taco_op_gemm_nt_split_acc(p->dG1 + oG, G, W1c, G, dctx_carry_out, MD, MD, dh1zn, D, D, B, G, st));
        float* t_;
        t_ = dc1; dc1 = dc1n; dc1n = t_;
        t_ = dh1z; dh1z = dh1zn; dh1zn = t_;
        t_ = dc2; dc2 = dc2n; dc2n = t_;
        t_ = dh2z; dh2z = dh2zn; dh2zn = t_;
    }
    // d(keys) and the attention weight gradients of all S steps in one pass (row 0 of the per-row partial buffers)
    if (fused)
        OK(taco_op_attn_bwd_hoisted(p->de, p->th, p->cum, p->lens, p->conv_w, p->conv_b, p->Wl, p->va, p->dkeys, p->p_conv_w, p->p_conv_b,
                                    p->p_Wl, p->p_va, p->p_ba, Uf + (size_t)(p->KW + 1) * A, S, B, T, st));
    return 0;
}

int taco_op_lstm_loop_fwd(const TacoLstmLoop* p, void* st) {
    if (!p || !p->K || !p->lens || !p->xw || !p->acts || !p->cnew || !p->cz || !p->hz || !p->out || !p->work) return -1;
    const int B = p->B, T = p->T, H = p->H, G = 4 * p->H;
    const float* Kh = p->K + (size_t)p->C * G;
    float* pre = p->work;
    float* hraw = p->work + (size_t)B * G;
    if (cudaMemsetAsync(p->acts, 0, sizeof(float) * (size_t)T * B * G, (cudaStream_t)st) != cudaSuccess) return -3;
    for (int s = 0; s < T; ++s) {
        const size_t oH = (size_t)s * B * H, nH = oH + (size_t)B * H;
        float* a = p->acts + (size_t)s * B * G;
        // 2 launches per position: recurrent contraction, then the pointwise step with the gather of the input-projection row and
        // the scatter of the raw output fused in
        OK(taco_op_gemm_acc(p->hz + oH, H, Kh, G, a, G, B, G, H, 0, st));
        OK(taco_op_lstm_fwd_seq(a, p->xw, (long long)T * G, G, p->cz + oH, p->hz + oH, p->lens, s, p->reverse, p->zoneout, 1, p->zo_keep_thresh,
                                p->seed, p->stream_c, p->stream_h, s * B, p->cnew + oH, p->cz + nH, hraw, p->hz + nH, p->out, p->out_sb,
                                p->out_st, B, H, st));
    }
    return 0;
}

int taco_op_lstm_loop_bwd(const TacoLstmLoop* p, void* st) {
    if (!p || !p->K || !p->lens || !p->acts || !p->cnew || !p->cz || !p->hz || !p->out || !p->dxw || !p->dg || !p->work) return -1;
    const int B = p->B, T = p->T, H = p->H, G = 4 * p->H;
    const float* Kh = p->K + (size_t)p->C * G;
    const size_t BH = (size_t)B * H;
    if (cudaMemsetAsync(p->work, 0, sizeof(float) * 5 * BH, (cudaStream_t)st) != cudaSuccess) return -3;
    float *dc = p->work, *dh = p->work + BH, *dcn = p->work + 2 * BH, *dhn = p->work + 3 * BH, *d_hraw = p->work + 4 * BH;
    for (int s = T - 1; s >= 0; --s) {
        const size_t oH = (size_t)s * B * H;
        float* dg = p->dg + (size_t)s * B * G;
        OK(taco_op_lstm_bwd_seq(p->acts + (size_t)s * B * G, p->cz + oH, p->hz + oH, p->cnew + oH, p->lens, s, p->reverse, p->zoneout, 1,
                                p->zo_keep_thresh, p->seed, p->stream_c, p->stream_h, s * B, p->out, p->out_sb, p->out_st, dc, dh, dg, dcn, dhn,
                                p->dxw, (long long)T * G, G, B, H, st));
        OK(taco_op_gemm_acc(dg, G, Kh, G, dhn, H, B, H, G, 1, st));
        float* t_;
        t_ = dc; dc = dcn; dcn = t_;
        t_ = dh; dh = dhn; dhn = t_;
    }
    return 0;
}

int taco_op_gru_loop_fwd(const TacoGruLoop* p, void* st) {
    if (!p || !p->Wg || !p->Wc || !p->GX || !p->CX || !p->hg || !p->rh || p->B < 1 || p->Tp < 1 || p->R < 1) return -1;
    const int B = p->B, R = p->R;
    const float* Wgh = p->Wg + (size_t)p->In * 2 * R;
    const float* Wch = p->Wc + (size_t)p->In * R;
    for (int s = 0; s < p->Tp; ++s) {
        float* gx = p->GX + (size_t)s * B * 2 * R;
        float* cx = p->CX + (size_t)s * B * R;
        const float* h = p->hg + (size_t)s * B * R;
        float* rh = p->rh + (size_t)s * B * R;
        OK(taco_op_gemm_acc(h, R, Wgh, 2 * R, gx, 2 * R, B, 2 * R, R, 0, st));
        OK(taco_op_gru_pre(gx, h, rh, B, R, st));
        OK(taco_op_gemm_acc(rh, R, Wch, R, cx, R, B, R, R, 0, st));
        OK(taco_op_gru_out(cx, gx, h, p->hg + (size_t)(s + 1) * B * R, B, R, st));
    }
    return 0;
}

int taco_op_gru_loop_bwd(const TacoGruLoop* p, void* st) {
    if (!p || !p->Wg || !p->Wc || !p->GX || !p->CX || !p->hg || !p->dGX || !p->dCX || !p->dh || !p->work) return -1;
    const int B = p->B, R = p->R;
    const float* Wgh = p->Wg + (size_t)p->In * 2 * R;
    const float* Wch = p->Wc + (size_t)p->In * R;
    float* dh = p->dh;
    float* dhd = p->work;
    float* drh = p->work + (size_t)B * R;
    for (int s = p->Tp - 1; s >= 0; --s) {
        const float* gx = p->GX + (size_t)s * B * 2 * R;
        const float* cx = p->CX + (size_t)s * B * R;
        const float* h = p->hg + (size_t)s * B * R;
        float* dgx = p->dGX + (size_t)s * B * 2 * R;
        float* dcx = p->dCX + (size_t)s * B * R;
        OK(taco_op_gru_out_bwd(dh, gx, h, cx, dcx, dgx, dhd, B, R, st));
        OK(taco_op_gemm(dcx, R, Wch, R, drh, R, B, R, R, 0, 1, 0.f, nullptr, st));
        OK(taco_op_gru_pre_bwd(drh, gx, h, dgx, dhd, B, R, st));
        OK(taco_op_gemm(dgx, 2 * R, Wgh, 2 * R, dhd, R, B, R, 2 * R, 0, 1, 1.f, nullptr, st));
        float* t_ = dh; dh = dhd; dhd = t_;
    }
    if (dh != p->dh && cudaMemcpyAsync(p->dh, dh, sizeof(float) * (size_t)B * R, cudaMemcpyDeviceToDevice, (cudaStream_t)st) != cudaSuccess) return -3;
    return 0;
}

}  // extern "C"
