// C ABI of libtaco_b200.so (include/taco_b200.h): handle, weight ingestion / packing, and the
// host-side orchestration of the sm_100a kernels.  No torch types, no CPU compute fallback.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "taco_internal.h"

namespace {

struct Weight {
    std::vector<float> host;
    std::vector<int64_t> shape;
    float* dev = nullptr;
};

struct Buf {
    void* p = nullptr;
    size_t bytes = 0;
};

const char* ENC_CONV = "encoder_convolutions/conv_layer_%d_encoder_convolutions";
const char* POST_CONV = "postnet_convolutions/conv_layer_%d_postnet_convolutions";
const char* ENC_FW = "encoder_LSTM/bidirectional_rnn/fw/encoder_fw_lstm";
const char* ENC_BW = "encoder_LSTM/bidirectional_rnn/bw/encoder_bw_lstm";
const char* REF_CONV = "ref_encoder/conv2d_%d";
const char* REF_GRU = "ref_encoder/rnn/gru_cell";
const char* REF_DENSE = "ref_encoder/dense";
const char* MHA = "Multihead-attention";
const char* PRENET = "decoder/decoder_prenet/dense_%d";
const char* DEC_LSTM1 = "decoder/decoder_LSTM/multi_rnn_cell/cell_0/decoder_LSTM_1";
const char* DEC_LSTM2 = "decoder/decoder_LSTM/multi_rnn_cell/cell_1/decoder_LSTM_2";
const char* LSA = "decoder/Location_Sensitive_Attention";
const char* MEMORY_LAYER = "decoder/memory_layer";
const char* FRAME_PROJ = "decoder/linear_transform_projection/projection_linear_transform_projection";
const char* STOP_PROJ = "decoder/stop_token_projection/projection_stop_token_projection";
const char* POST_PROJ = "postnet_projection/projection_postnet_projection";

std::string fmt(const char* f, int a) {
    char buf[256];
    snprintf(buf, sizeof buf, f, a);
    return buf;
}

}  // namespace

struct TacoHandle {
    TacoHParams hp;
    int device = 0;
    std::string err;
    std::map<std::string, Weight> w;
    std::map<std::string, Buf> ws;
    std::vector<void*> owned;
    bool finalized = false;
    int precision = TACO_PRECISION_FP32;
    int64_t launches = 0;
    int n_sm = 0;
    // folded BN scale/shift
    float *enc_scale[8] = {}, *enc_shift[8] = {};
    float *post_scale[8] = {}, *post_shift[8] = {};
    float *ref_scale[8] = {}, *ref_shift[8] = {};
    // packed resident shards
    // tensor-core operands: bf16 K-major [Cout][taps*Cin_pad] (conv_tc.cu)
    void* post_wt[8] = {};
    int post_cin_pad[8] = {};
    void* post_proj_wt = nullptr;
    void* enc_wt[8] = {};
    void* enc_xw_wt[2] = {};   // LSTM input projection kernel[0:C] of fw / bw, bf16 K-major
    void* keys_wt = nullptr;    // memory_layer kernel, bf16 K-major [128][768]
    float* dec_blob[2] = {};     // U = 2 (GN 128), U = 4 (GN 64)
    // fused location matrix of the cluster decoder (hi / lo bf16, B-fragment order)
    uint32_t *dec2_ghi = nullptr, *dec2_glo = nullptr;
    bool warned_dec_fallback = false;
    float* dec2_bconst = nullptr;
    // cluster decoder (decoder3.cu): per-rank resident A-fragment tiles + L2-streamed tiles
    uint32_t* dec3_res = nullptr;
    uint32_t* dec3_stream = nullptr;
    uint32_t* dec3_stop_tiles = nullptr;   // [64][128] A-fragment tiles of the stop projection
    uint32_t* dec3_wo_t = nullptr;         // frame projection kernel, bf16 pairs along k: [160][512]
    float* dec3_bo = nullptr;              // frame projection bias [160]
    float dec3_stop_bias[2] = {0.f, 0.f};
    int dec3_max_clusters = -1;  // cudaOccupancyMaxActiveClusters of decoder3_kernel (queried once)
    int dec_kind = 0;            // which decoder kernel owns the running state (1 = decoder.cu fp32, 3 = decoder3.cu bf16 cluster kernel)
    float* bilstm_blob[3] = {};  // U = 2, 4, 16 units per CTA, [2 dir][GN][256*4U]
    int encoder_max_ctas = 0;    // 0: lowest-latency BiLSTM grid; > 0: at most this many CTAs (taco_set_encoder_ctas)
    // decoder running state
    int dec_t = 0, dec_B = 0, dec_T = 0, dec_RBP = 0, dec_GB = 0;
    // CUDA events bracketing each stage on the caller's stream (taco_stage_ms)
    cudaEvent_t ev[TACO_N_STAGES][2] = {};
    bool ev_valid[TACO_N_STAGES] = {};
};

namespace {

#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) {                                                                   \
            h->err = std::string(#call) + ": " + cudaGetErrorString(e_);                           \
            return TACO_ERR_CUDA;                                                                  \
        }                                                                                          \
    } while (0)

#define FAIL(code, msg)  \
    do {                 \
        h->err = (msg);  \
        return (code);   \
    } while (0)

int get_ws(TacoHandle* h, const std::string& name, size_t bytes, void** out) {
    Buf& b = h->ws[name];
    if (b.bytes < bytes) {
        if (b.p) {
            CK(cudaDeviceSynchronize());
            CK(cudaFree(b.p));
            b.p = nullptr;
            b.bytes = 0;
        }
        size_t want = bytes + bytes / 8 + 256;
        CK(cudaMalloc(&b.p, want));
        b.bytes = want;
    }
    *out = b.p;
    return TACO_OK;
}

#define WS(ptr, type, name, count)                                              \
    type* ptr;                                                                  \
    {                                                                           \
        void* v_;                                                               \
        int rc_ = get_ws(h, name, (size_t)(count) * sizeof(type), &v_);         \
        if (rc_ != TACO_OK) return rc_;                                         \
        ptr = reinterpret_cast<type*>(v_);                                      \
    }

const Weight* find_w(TacoHandle* h, const std::string& name) {
    auto it = h->w.find(name);
    return it == h->w.end() ? nullptr : &it->second;
}

int upload(TacoHandle* h, const std::vector<float>& v, float** out) {
    void* d;
    CK(cudaMalloc(&d, v.size() * sizeof(float)));
    CK(cudaMemcpy(d, v.data(), v.size() * sizeof(float), cudaMemcpyHostToDevice));
    h->owned.push_back(d);
    *out = reinterpret_cast<float*>(d);
    return TACO_OK;
}

uint16_t f32_to_bf16_rne(float f) {
    uint32_t u;
    memcpy(&u, &f, 4);
    if ((u & 0x7FFFFFFFu) > 0x7F800000u) return (uint16_t)((u >> 16) | 0x40);  // NaN
    u += 0x7FFFu + ((u >> 16) & 1u);
    return (uint16_t)(u >> 16);
}

// TF conv1d kernel [k][Cin][Cout] fp32 -> K-major bf16 [Cout][k*Cin_pad] (zero-padded channels)
int pack_conv_bf16(TacoHandle* h, const Weight* w, int k, int cin, int cin_pad, int cout, void** out) {
    std::vector<uint16_t> t((size_t)cout * k * cin_pad, 0);
    for (int tap = 0; tap < k; ++tap)
        for (int ci = 0; ci < cin; ++ci) {
            const float* src = w->host.data() + ((size_t)tap * cin + ci) * cout;
            for (int co = 0; co < cout; ++co) t[(size_t)co * k * cin_pad + (size_t)tap * cin_pad + ci] = f32_to_bf16_rne(src[co]);
        }
    void* d;
    CK(cudaMalloc(&d, t.size() * 2));
    CK(cudaMemcpy(d, t.data(), t.size() * 2, cudaMemcpyHostToDevice));
    h->owned.push_back(d);
    *out = d;
    return TACO_OK;
}

// BN(inference) + conv bias folded:  y = conv*scale + shift
int fold_bn(TacoHandle* h, const std::string& scope, const char* conv, float eps, float** scale, float** shift) {
    const Weight* bias = find_w(h, scope + "/" + conv + "/bias");
    const Weight* g = find_w(h, scope + "/batch_normalization/gamma");
    const Weight* b = find_w(h, scope + "/batch_normalization/beta");
    const Weight* mu = find_w(h, scope + "/batch_normalization/moving_mean");
    const Weight* var = find_w(h, scope + "/batch_normalization/moving_variance");
    if (!bias || !g || !b || !mu || !var) FAIL(TACO_ERR_MISSING_WEIGHT, "missing BN/bias variables under " + scope);
    size_t c = g->host.size();
    std::vector<float> sc(c), sh(c);
    for (size_t i = 0; i < c; ++i) {
        float s = g->host[i] / sqrtf(var->host[i] + eps);
        sc[i] = s;
        sh[i] = b->host[i] + (bias->host[i] - mu->host[i]) * s;
    }
    int rc = upload(h, sc, scale);
    if (rc) return rc;
    return upload(h, sh, shift);
}

int need(TacoHandle* h, const std::string& name, std::initializer_list<int64_t> shape, const Weight** out) {
    const Weight* w = find_w(h, name);
    if (!w) FAIL(TACO_ERR_MISSING_WEIGHT, "missing variable " + name);
    if (w->shape.size() != shape.size()) FAIL(TACO_ERR_INVALID, "rank mismatch for " + name);
    size_t i = 0;
    for (int64_t s : shape) {
        if (w->shape[i++] != s) FAIL(TACO_ERR_INVALID, "shape mismatch for " + name);
    }
    *out = w;
    return TACO_OK;
}

int pack_decoder_blobs(TacoHandle* h) {
    const Weight *k1, *b1, *k2, *b2, *kp1, *bp1, *kp2, *bp2, *kq, *kf, *bf, *kst, *bst;
    int rc;
#define NEED(var, name, ...)                              \
    if ((rc = need(h, name, {__VA_ARGS__}, &var))) return rc;
    NEED(k1, std::string(DEC_LSTM1) + "/kernel", 1280, 1024)
    NEED(b1, std::string(DEC_LSTM1) + "/bias", 1024)
    NEED(k2, std::string(DEC_LSTM2) + "/kernel", 512, 1024)
    NEED(b2, std::string(DEC_LSTM2) + "/bias", 1024)
    NEED(kp1, fmt(PRENET, 1) + "/kernel", 80, 256)
    NEED(bp1, fmt(PRENET, 1) + "/bias", 256)
    NEED(kp2, fmt(PRENET, 2) + "/kernel", 256, 256)
    NEED(bp2, fmt(PRENET, 2) + "/bias", 256)
    NEED(kq, std::string(LSA) + "/query_layer/kernel", 256, 128)
    NEED(kf, std::string(FRAME_PROJ) + "/kernel", 1024, 160)
    NEED(bf, std::string(FRAME_PROJ) + "/bias", 160)
    NEED(kst, std::string(STOP_PROJ) + "/kernel", 1024, 2)
    NEED(bst, std::string(STOP_PROJ) + "/bias", 2)
    for (int cfg = 0; cfg < 2; ++cfg) {
        const int U = cfg == 0 ? 2 : 4, GN = 256 / U, NCL = 4 * U;
        const int PCN = 256 / GN, QCN = 128 / GN;
        const int bf_ = taco_decoder_blob_floats(U);
        std::vector<float> blob((size_t)GN * bf_, 0.f);
        for (int nt = 0; nt < GN; ++nt) {
            float* p = blob.data() + (size_t)nt * bf_;
            float* W1 = p;
            float* W2 = W1 + 1280 * NCL;
            float* Wp1 = W2 + 512 * NCL;
            float* Wp2 = Wp1 + 80 * 4;
            float* Wq = Wp2 + 256 * 4;
            float* Wo = Wq + 256 * 4;
            float* bias = Wo + 1024 * 4;
            for (int g = 0; g < 4; ++g)
                for (int ul = 0; ul < U; ++ul) {
                    int col = g * 256 + nt * U + ul, lc = g * U + ul;
                    for (int k = 0; k < 1280; ++k) W1[k * NCL + lc] = k1->host[(size_t)k * 1024 + col];
                    for (int k = 0; k < 512; ++k) W2[k * NCL + lc] = k2->host[(size_t)k * 1024 + col];
                    bias[lc] = b1->host[col];
                    bias[NCL + lc] = b2->host[col];
                }
            for (int j = 0; j < PCN; ++j) {
                int col = nt * PCN + j;
                for (int k = 0; k < 80; ++k) Wp1[k * 4 + j] = kp1->host[(size_t)k * 256 + col];
                for (int k = 0; k < 256; ++k) Wp2[k * 4 + j] = kp2->host[(size_t)k * 256 + col];
                bias[2 * NCL + j] = bp1->host[col];
                bias[2 * NCL + 4 + j] = bp2->host[col];
            }
            for (int j = 0; j < QCN; ++j) {
                int col = nt * QCN + j;
                for (int k = 0; k < 256; ++k) Wq[k * 4 + j] = kq->host[(size_t)k * 128 + col];
            }
            for (int j = 0; j < 4; ++j) {
                int col = nt * 4 + j;
                if (col < 160) {
                    for (int k = 0; k < 1024; ++k) Wo[k * 4 + j] = kf->host[(size_t)k * 160 + col];
                    bias[2 * NCL + 8 + j] = bf->host[col];
                } else if (col < 162) {
                    for (int k = 0; k < 1024; ++k) Wo[k * 4 + j] = kst->host[(size_t)k * 2 + (col - 160)];
                    bias[2 * NCL + 8 + j] = bst->host[col - 160];
                }
            }
        }
        if ((rc = upload(h, blob, &h->dec_blob[cfg]))) return rc;
    }
    // BiLSTM recurrent shards (rows 512..767 of the [768,1024] kernels)
    const Weight *kfw, *kbw, *bfw, *bbw;
    NEED(kfw, std::string(ENC_FW) + "/kernel", 768, 1024)
    NEED(kbw, std::string(ENC_BW) + "/kernel", 768, 1024)
    NEED(bfw, std::string(ENC_FW) + "/bias", 1024)
    NEED(bbw, std::string(ENC_BW) + "/bias", 1024)
    for (int cfg = 0; cfg < 3; ++cfg) {
        const int U = cfg == 0 ? 2 : (cfg == 1 ? 4 : 16), GN = 256 / U, NCL = 4 * U;
        const int bf_ = taco_bilstm_blob_floats(U);
        std::vector<float> blob((size_t)2 * GN * bf_, 0.f);
        for (int d = 0; d < 2; ++d) {
            const Weight* kk = d == 0 ? kfw : kbw;
            for (int nt = 0; nt < GN; ++nt) {
                float* p = blob.data() + ((size_t)d * GN + nt) * bf_;
                for (int g = 0; g < 4; ++g)
                    for (int ul = 0; ul < U; ++ul) {
                        int col = g * 256 + nt * U + ul, lc = g * U + ul;
                        for (int k = 0; k < 256; ++k) p[k * NCL + lc] = kk->host[(size_t)(512 + k) * 1024 + col];
                    }
            }
        }
        if ((rc = upload(h, blob, &h->bilstm_blob[cfg]))) return rc;
    }
#undef NEED
    return TACO_OK;
}

int upload_words(TacoHandle* h, const std::vector<uint32_t>& v, uint32_t** out) {
    void* d;
    CK(cudaMalloc(&d, v.size() * sizeof(uint32_t)));
    CK(cudaMemcpy(d, v.data(), v.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
    h->owned.push_back(d);
    *out = reinterpret_cast<uint32_t*>(d);
    return TACO_OK;
}

inline uint32_t pack2(float lo, float hi) { return (uint32_t)f32_to_bf16_rne(lo) | ((uint32_t)f32_to_bf16_rne(hi) << 16); }
inline float bf16_round(float f) {
    uint32_t u = (uint32_t)f32_to_bf16_rne(f) << 16;
    float r;
    memcpy(&r, &u, 4);
    return r;
}

// Fused location term of the cluster decoder (Attention.py:193-197): conv1d(k=31, 32 filters) followed by Dense(128, no bias) is
// one [31 x 128] matrix G, split into hi/lo bf16 parts in mma B-fragment order; bconst = attention_bias + conv_bias . Wl.
int pack_location(TacoHandle* h) {
    const Weight *cw, *cb, *wl, *ba;
    int rc;
#define NEED(var, name, ...) \
    if ((rc = need(h, name, {__VA_ARGS__}, &var))) return rc;
    NEED(cw, std::string(LSA) + "/location_features_convolution/kernel", 31, 1, 32)
    NEED(cb, std::string(LSA) + "/location_features_convolution/bias", 32)
    NEED(wl, std::string(LSA) + "/location_features_layer/kernel", 32, 128)
    NEED(ba, std::string(LSA) + "/attention_bias", 128)
#undef NEED
    // G[k][d] = sum_c conv[k][c] Wl[c][d] (k < 31; row 31 = 0), hi/lo bf16 split, B-fragment-major [ks][nt][lane][2]
    std::vector<double> G((size_t)32 * 128, 0.0);
    for (int k = 0; k < 31; ++k)
        for (int d = 0; d < 128; ++d) {
            double a = 0;
            for (int c = 0; c < 32; ++c) a += (double)cw->host[(size_t)k * 32 + c] * (double)wl->host[(size_t)c * 128 + d];
            G[(size_t)k * 128 + d] = a;
        }
    std::vector<uint32_t> ghi((size_t)2 * 16 * 64), glo((size_t)2 * 16 * 64);
    for (int ks = 0; ks < 2; ++ks)
        for (int nt = 0; nt < 16; ++nt)
            for (int lane = 0; lane < 32; ++lane) {
                const int g = lane >> 2, t = lane & 3;
                for (int j = 0; j < 2; ++j) {
                    const int k = ks * 16 + 2 * t + 8 * j, d = nt * 8 + g;
                    const float a0 = (float)G[(size_t)k * 128 + d], a1 = (float)G[(size_t)(k + 1) * 128 + d];
                    const float h0 = bf16_round(a0), h1 = bf16_round(a1);
                    const size_t idx = ((size_t)(ks * 16 + nt) * 32 + lane) * 2 + j;
                    ghi[idx] = pack2(h0, h1);
                    glo[idx] = pack2((float)(G[(size_t)k * 128 + d] - h0), (float)(G[(size_t)(k + 1) * 128 + d] - h1));
                }
            }
    if ((rc = upload_words(h, ghi, &h->dec2_ghi))) return rc;
    if ((rc = upload_words(h, glo, &h->dec2_glo))) return rc;
    std::vector<float> bconst(128);
    for (int d = 0; d < 128; ++d) {
        double a = ba->host[d];
        for (int c = 0; c < 32; ++c) a += (double)cb->host[c] * (double)wl->host[(size_t)c * 128 + d];
        bconst[d] = (float)a;
    }
    return upload(h, bconst, &h->dec2_bconst);
}


// Shards for decoder3.cu: per cluster rank c (units 16c..16c+15) the mma.sync m16n8k16 A-fragments (weights^T:
// rows = output columns, cols = k) of its tiles: word [lane = 4g+t][reg] = bf16 pair
// (W[k0 + 2t + 8(reg>>1)][m = g + 8(reg&1)], W[k0 + 2t + 8(reg>>1) + 1][m]).
int pack_decoder3(TacoHandle* h) {
    const Weight *k1, *b1, *k2, *b2, *kp1, *bp1, *kp2, *bp2, *kq, *kf, *bf, *kst, *bst;
    int rc;
#define NEED(var, name, ...) \
    if ((rc = need(h, name, {__VA_ARGS__}, &var))) return rc;
    NEED(k1, std::string(DEC_LSTM1) + "/kernel", 1280, 1024)
    NEED(b1, std::string(DEC_LSTM1) + "/bias", 1024)
    NEED(k2, std::string(DEC_LSTM2) + "/kernel", 512, 1024)
    NEED(b2, std::string(DEC_LSTM2) + "/bias", 1024)
    NEED(kp1, fmt(PRENET, 1) + "/kernel", 80, 256)
    NEED(bp1, fmt(PRENET, 1) + "/bias", 256)
    NEED(kp2, fmt(PRENET, 2) + "/kernel", 256, 256)
    NEED(bp2, fmt(PRENET, 2) + "/bias", 256)
    NEED(kq, std::string(LSA) + "/query_layer/kernel", 256, 128)
    NEED(kf, std::string(FRAME_PROJ) + "/kernel", 1024, 160)
    NEED(bf, std::string(FRAME_PROJ) + "/bias", 160)
    NEED(kst, std::string(STOP_PROJ) + "/kernel", 1024, 2)
    NEED(bst, std::string(STOP_PROJ) + "/bias", 2)
#undef NEED
    auto pack_tile = [](uint32_t* dst, auto get) {   // get(k in 0..15, m in 0..15)
        for (int lane = 0; lane < 32; ++lane) {
            const int g = lane >> 2, t = lane & 3;
            for (int reg = 0; reg < 4; ++reg) {
                const int m = g + 8 * (reg & 1), k = 2 * t + 8 * (reg >> 1);
                dst[lane * 4 + reg] = pack2(get(k, m), get(k + 1, m));
            }
        }
    };
    // Folded feedback: x(s+1) = [h2raw;ctx].Wo[:,80:160] + bo[80:160] (TacoTestHelper.next_inputs, Helpers.py:56) feeds prenet
    // dense_1 (Decoder.py:346) => A = Wo[:,80:160].Wp1 [1024 x 256], bfold = bo[80:160].Wp1 + bp1, composed in double.
    std::vector<float> fold((size_t)1024 * 256), bfold(256);
    for (int k = 0; k < 1024; ++k)
        for (int n = 0; n < 256; ++n) {
            double a = 0.0;
            for (int m = 0; m < 80; ++m) a += (double)kf->host[(size_t)k * 160 + 80 + m] * (double)kp1->host[(size_t)m * 256 + n];
            fold[(size_t)k * 256 + n] = (float)a;
        }
    for (int n = 0; n < 256; ++n) {
        double a = bp1->host[n];
        for (int m = 0; m < 80; ++m) a += (double)bf->host[80 + m] * (double)kp1->host[(size_t)m * 256 + n];
        bfold[n] = (float)a;
    }
    const int RW = TACO_DEC3_RES_WORDS, SW = TACO_DEC3_STREAM_TILES * 128;
    std::vector<uint32_t> res((size_t)16 * RW, 0u), str((size_t)16 * SW, 0u);
    for (int c = 0; c < 16; ++c) {
        uint32_t* p = res.data() + (size_t)c * RW;
        for (int ks = 0; ks < 64; ++ks)   // folded dense_1 of [LSTM output ; context]
            pack_tile(p + ks * 128, [&](int k, int m) { return fold[(size_t)(ks * 16 + k) * 256 + 16 * c + m]; });
        p += 64 * 128;
        for (int ks = 0; ks < 5; ++ks) pack_tile(p + ks * 128, [&](int k, int m) { return kp1->host[(size_t)(ks * 16 + k) * 256 + 16 * c + m]; });
        p += 5 * 128;
        for (int ks = 0; ks < 16; ++ks) pack_tile(p + ks * 128, [&](int k, int m) { return kp2->host[(size_t)(ks * 16 + k) * 256 + 16 * c + m]; });
        p += 16 * 128;
        for (int mt = 0; mt < 4; ++mt)   // resident: first 32 k-steps of the LSTM-1 context rows [256, 256 + 512)
            for (int ks = 0; ks < 32; ++ks)
                pack_tile(p + (mt * 32 + ks) * 128, [&](int k, int m) { return k1->host[(size_t)(256 + ks * 16 + k) * 1024 + mt * 256 + 16 * c + m]; });
        p += 128 * 128;
        for (int mt = 0; mt < 8; ++mt) pack_tile(p + mt * 128, [&](int k, int m) { return kq->host[(size_t)(16 * c + k) * 128 + mt * 16 + m]; });
        p += 8 * 128;
        float* bias = reinterpret_cast<float*>(p);
        for (int gt = 0; gt < 4; ++gt)
            for (int u = 0; u < 16; ++u) {
                bias[gt * 16 + u] = b1->host[gt * 256 + 16 * c + u];
                bias[64 + gt * 16 + u] = b2->host[gt * 256 + 16 * c + u];
            }
        for (int u = 0; u < 16; ++u) {
            bias[128 + u] = bp1->host[16 * c + u];
            bias[144 + u] = bp2->host[16 * c + u];
            bias[160 + u] = bfold[16 * c + u];
        }
        uint32_t* q = str.data() + (size_t)c * SW;
        // streamed sets [4 mt][16 ks]: 0 LSTM-1 prenet rows [0,256); 1 LSTM-2 h1 rows [0,256); 2 LSTM-1 h1 rows [1024,1280);
        // 3 LSTM-2 h2 rows [256,512); 4 LSTM-1 context rows [768,1024) (context k-steps 32..47)
        for (int mt = 0; mt < 4; ++mt)
            for (int ks = 0; ks < 16; ++ks) {
                auto col = [&](int m) { return mt * 256 + 16 * c + m; };
                pack_tile(q + (0 + mt * 16 + ks) * 128, [&](int k, int m) { return k1->host[(size_t)(ks * 16 + k) * 1024 + col(m)]; });
                pack_tile(q + (64 + mt * 16 + ks) * 128, [&](int k, int m) { return k2->host[(size_t)(ks * 16 + k) * 1024 + col(m)]; });
                pack_tile(q + (128 + mt * 16 + ks) * 128, [&](int k, int m) { return k1->host[(size_t)(1024 + ks * 16 + k) * 1024 + col(m)]; });
                pack_tile(q + (192 + mt * 16 + ks) * 128, [&](int k, int m) { return k2->host[(size_t)(256 + ks * 16 + k) * 1024 + col(m)]; });
                pack_tile(q + (256 + mt * 16 + ks) * 128, [&](int k, int m) { return k1->host[(size_t)(768 + ks * 16 + k) * 1024 + col(m)]; });
            }
    }
    // stop projection (Decoder.py:404): one m16 tile per k-step, rows 0..1 = the two stop units
    std::vector<uint32_t> stp((size_t)64 * 128, 0u);
    for (int ks = 0; ks < 64; ++ks)
        pack_tile(stp.data() + ks * 128, [&](int k, int m) { return m < 2 ? kst->host[(size_t)(ks * 16 + k) * 2 + m] : 0.f; });
    h->dec3_stop_bias[0] = bst->host[0];
    h->dec3_stop_bias[1] = bst->host[1];
    // frame projection for the post-launch contraction (frame_proj.cu): Wo^T as bf16 pairs along k
    std::vector<uint32_t> wot((size_t)160 * 512);
    for (int n = 0; n < 160; ++n)
        for (int kk = 0; kk < 512; ++kk) wot[(size_t)n * 512 + kk] = pack2(kf->host[(size_t)(2 * kk) * 160 + n], kf->host[(size_t)(2 * kk + 1) * 160 + n]);
    if ((rc = upload_words(h, res, &h->dec3_res))) return rc;
    if ((rc = upload_words(h, stp, &h->dec3_stop_tiles))) return rc;
    if ((rc = upload_words(h, wot, &h->dec3_wo_t))) return rc;
    if ((rc = upload(h, bf->host, &h->dec3_bo))) return rc;
    return upload_words(h, str, &h->dec3_stream);
}

const float* dev_w(TacoHandle* h, const std::string& name) {
    const Weight* w = find_w(h, name);
    return w ? w->dev : nullptr;
}

__global__ void clip_copy_kernel(const float* __restrict__ in, float* __restrict__ out, int B, int F, int F_stride, int C,
                                 float lo, float hi, int clip) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    size_t n = (size_t)B * F * C;
    if (i >= n) return;
    size_t c = i % C, f = (i / C) % F, b = i / ((size_t)C * F);
    float v = in[(b * F_stride + f) * C + c];
    out[i] = clip ? fminf(fmaxf(v, lo), hi) : v;
}

// XT layout [GB][K][RBP] -> row-major [B][K]
__global__ void untranspose_xt_kernel(const float* __restrict__ xt, float* __restrict__ out, int B, int K, int RBP) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B * K) return;
    int b = i / K, k = i % K;
    out[i] = xt[((size_t)(b / RBP) * K + k) * RBP + (b % RBP)];
}

// teacher-forced continuation (`cell(inputs, state)`): x(t0) = targets[:, r*t0 - 1] into the XT exchange block
__global__ void load_xt_from_targets_kernel(const float* __restrict__ targets, int F_tgt, int fidx, float* __restrict__ xt, int B, int RBP) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B * 80) return;
    const int b = i / 80, k = i % 80;
    xt[((size_t)(b / RBP) * 80 + k) * RBP + (b % RBP)] = fidx < F_tgt ? targets[((size_t)b * F_tgt + fidx) * 80 + k] : 0.f;
}

int conv(TacoHandle* h, const float* x, long long x_stride, long long x_bstride, const float* w, const float* scale,
         const float* shift, float* y, long long y_stride, long long y_bstride, const float* residual, int B, int T, int Cin,
         int Cout, int k, int act, bool clip, float lo, float hi, cudaStream_t st) {
    ConvParams p;
    memset(&p, 0, sizeof p);
    p.x = x; p.x_stride = x_stride; p.x_bstride = x_bstride;
    p.w = w; p.scale = scale; p.shift = shift;
    p.y = y; p.y_stride = y_stride; p.y_bstride = y_bstride;
    p.residual = residual;
    p.B = B; p.T = T; p.Cin = Cin; p.Cout = Cout; p.k = k; p.act = act;
    p.clip = clip ? 1 : 0; p.clip_lo = lo; p.clip_hi = hi;
    CK(taco_launch_conv1d_f32(p, st));
    h->launches++;
    return TACO_OK;
}

// keys = memory_layer(values) (Attention.py:153-158) for the bf16 decoders: tcgen05 dense, bf16 operands, fp32 out
int keys_tc(TacoHandle* h, const void* vals_bf16, float* keys, int B, int T, cudaStream_t st) {
    ConvTcParams c;
    memset(&c, 0, sizeof c);
    c.x = vals_bf16; c.wt = h->keys_wt;
    c.B = B; c.F = T; c.Cin = 768; c.taps = 1; c.Cout = 128;
    c.act = ACT_NONE;
    c.y_f32 = keys; c.out_f32_only = 1;
    CK(taco_launch_conv_tc(c, h->n_sm, st));
    h->launches++;
    return TACO_OK;
}

// taco_decode through the cluster kernel (decoder3.cu) -- TACO_PRECISION_BF16 when 16-CTA clusters can be scheduled
int decode_v3(TacoHandle* h, int mode, const float* memory, const int32_t* lens, int B, int T, const float* mel_targets,
              int F_tgt, int S_max, const uint8_t* prenet_masks, uint64_t philox_seed, int reset_state, int logits_only,
              float* frames, float* stop, float* align_hist, int32_t* steps_out, cudaStream_t st) {
    const TacoHParams& hp = h->hp;
    Dec3Params p;
    memset(&p, 0, sizeof p);
    p.B = B; p.T = T;
    // batch rows are split over clusters; <= 8 rows per cluster keeps one n-tile, more clusters shorten every exchange
    int ncl = (B + 7) / 8;
    if (getenv("TACO_DEC3_NCL")) ncl = atoi(getenv("TACO_DEC3_NCL"));
    if (ncl > h->dec3_max_clusters) ncl = h->dec3_max_clusters;
    if (ncl < (B + 15) / 16) FAIL(TACO_ERR_UNSUPPORTED, "taco_decode: batch needs more co-resident 16-CTA clusters than this device schedules");
    if (ncl < 1) ncl = 1;
    // the alignment rows of a cluster live in every one of its CTAs: rows per cluster * round_up(T, 4) <= TACO_DEC3_ACAP words
    p.TP = (T + 3) & ~3;
    // T > 128: a row is split into position slices of 128, one attention work item (CTA) each, so that the location
    // pre-activations stay in registers; rows per cluster * slices <= 16 CTAs.  (Not with cumulative_weights = False: the
    // non-incremental location term needs the whole row in one CTA -- that case keeps the global scratch path.)
    int parts = (T > 128 && hp.cumulative_weights && !getenv("TACO_DEC3_NO_PARTS")) ? (T + 127) / 128 : 1;
    auto rows_of = [&](int n) { return (B + n - 1) / n; };
    auto fits = [&](int n, int pp) { return rows_of(n) * p.TP <= TACO_DEC3_ACAP && rows_of(n) * pp <= TACO_DEC3_CLUSTER; };
    int n2 = ncl;
    while (n2 < h->dec3_max_clusters && !fits(n2, parts)) ++n2;
    if (!fits(n2, parts)) {
        parts = 1;
        n2 = ncl;
        while (n2 < h->dec3_max_clusters && !fits(n2, 1)) ++n2;
        if (!fits(n2, 1))
            FAIL(TACO_ERR_UNSUPPORTED, "taco_decode: B x T too large for the cluster decoder (rows per cluster x T exceeds its alignment buffer)");
    }
    ncl = n2;
    p.parts = parts;
    p.R = (B + ncl - 1) / ncl;
    p.NCL = (B + p.R - 1) / p.R;
    p.S_max = S_max; p.S_cap = S_max; p.mode = mode;
    p.stop_at_any = hp.stop_at_any; p.cumulative = hp.cumulative_weights; p.smoothing = hp.smoothing;
    p.stop_sigmoid = logits_only ? 0 : 1;
    p.pre_reg = (T <= 128 || p.parts > 1) ? 1 : 0;
    p.zoneout = hp.zoneout_rate;
    p.keep_scale = 1.0f / (1.0f - hp.dropout_rate);
    p.res_blobs = h->dec3_res;
    p.stream_blobs = reinterpret_cast<const uint4*>(h->dec3_stream);
    p.stop_tiles = reinterpret_cast<const uint4*>(h->dec3_stop_tiles);
    p.stop_bias[0] = h->dec3_stop_bias[0]; p.stop_bias[1] = h->dec3_stop_bias[1];
    p.g_hi = h->dec2_ghi; p.g_lo = h->dec2_glo;
    p.va = dev_w(h, std::string(LSA) + "/attention_variable");
    p.bconst = h->dec2_bconst;
    if (!reset_state && (h->dec_kind != 3 || h->dec_B != B || h->dec_T != T))
        FAIL(TACO_ERR_STATE, "taco_decode: continuing with a different batch shape or precision");
    WS(lens_eff, int32_t, "dec_lens", B)
    if (hp.mask_encoder) {
        CK(cudaMemcpyAsync(lens_eff, lens, B * sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
    } else {
        std::vector<int32_t> full(B, T);
        CK(cudaMemcpyAsync(lens_eff, full.data(), B * sizeof(int32_t), cudaMemcpyHostToDevice, st));
        CK(cudaStreamSynchronize(st));
    }
    p.lens = lens_eff;
    WS(keys, float, "dec_keys", (size_t)B * T * 128)
    WS(vals, uint16_t, "dec2_values", (size_t)B * T * 768)
    if (reset_state) {
        // values as bf16 (the per-step stream), keys = values . W_mem on tcgen05 from the same bf16 copy
        CK(taco_launch_f32_to_bf16(memory, vals, (size_t)B * T * 768, st));
        h->launches++;
        int rc = keys_tc(h, vals, keys, B, T, st);
        if (rc) return rc;
    }
    p.keys = keys;
    p.values = vals;
    // recurrent state (fp32 row-major): c1 h1z c2 h2z [B][256], ctx [B][768], (unused [B][80]), cum [B][T]; h2raw [B][256] after it
    const size_t n_state = (size_t)B * (4 * 256 + 768 + 80 + T + 256);
    WS(state, float, "dec2_state", n_state)
    if (reset_state) {
        CK(cudaMemsetAsync(state, 0, n_state * sizeof(float), st));
        h->dec_t = 0;
    }
    float* s = state;
    p.c1 = s; s += (size_t)B * 256;
    p.h1z = s; s += (size_t)B * 256;
    p.c2 = s; s += (size_t)B * 256;
    p.h2z = s; s += (size_t)B * 256;
    p.ctx_state = s; s += (size_t)B * 768;
    s += (size_t)B * 80;
    p.cum = s; s += (size_t)B * T;
    p.h2raw = s;
    WS(act, uint32_t, "dec3_act_hist", (size_t)S_max * B * 512)
    p.act_hist = act;
    WS(pre, float, "dec3_pre", (size_t)B * T * 128)
    p.pre_scratch = pre;
    p.cont = reset_state ? 0 : 1;
    if (getenv("TACO_DEC3_DBG")) p.dbg = atoi(getenv("TACO_DEC3_DBG"));
    p.t0 = h->dec_t;
    p.targets = mel_targets; p.F_tgt = F_tgt;
    p.masks = prenet_masks; p.seed = philox_seed;
    p.stop = stop; p.hist = align_hist; p.steps_out = steps_out;
    const size_t n_flags = (size_t)S_max * p.NCL + 8;
    WS(flags, int, "dec3_flags", n_flags)
    CK(cudaMemsetAsync(flags, 0, n_flags * sizeof(int), st));
    p.flags = flags;
    p.abort_flag = flags + (size_t)S_max * p.NCL + 4;
    if (getenv("TACO_DEC2_TIMING")) {
        WS(timing, long long, "dec2_timing", (size_t)TACO_DEC3_CLUSTER * TACO_DEC3_MAX_CLUSTERS * 24)
        CK(cudaMemsetAsync(timing, 0, (size_t)TACO_DEC3_CLUSTER * TACO_DEC3_MAX_CLUSTERS * 24 * sizeof(long long), st));
        p.timing = timing;
    }
    if (getenv("TACO_DEC3_TRACE")) {
        WS(trace, long long, "dec3_trace", (size_t)TACO_DEC3_CLUSTER * TACO_DEC3_MAX_CLUSTERS * 128)
        CK(cudaMemsetAsync(trace, 0, (size_t)TACO_DEC3_CLUSTER * TACO_DEC3_MAX_CLUSTERS * 128 * sizeof(long long), st));
        p.trace = trace;
        p.trace_step = atoi(getenv("TACO_DEC3_TRACE"));
    }
    CK(cudaEventRecord(h->ev[TACO_STAGE_DECODER_KERNEL][0], st));
    CK(taco_launch_decoder3(p, st));
    // frame projection of all executed steps (hoisted out of the recurrence: nothing in the loop reads the frames)
    CK(taco_launch_frame_projection(act, h->dec3_wo_t, h->dec3_bo, steps_out, S_max, B, S_max, frames, st));
    h->launches++;
    CK(cudaEventRecord(h->ev[TACO_STAGE_DECODER_KERNEL][1], st));
    h->ev_valid[TACO_STAGE_DECODER_KERNEL] = true;
    h->launches++;
    h->dec_t += S_max;
    h->dec_B = B; h->dec_T = T; h->dec_kind = 3;
    return TACO_OK;
}

bool hp_supported(const TacoHParams& hp, std::string* why) {
#define REQ(cond)                                  \
    if (!(cond)) {                                 \
        *why = "unsupported hparams: " #cond;      \
        return false;                              \
    }
    REQ(hp.embedding_dim == 512)
    REQ(hp.conv_channels == 512)
    REQ(hp.enc_lstm_hidden_size == 256)
    REQ(hp.style_embed_depth == 256)
    REQ(hp.use_gst == 1)
    REQ(hp.num_heads * (hp.style_embed_depth / hp.num_heads) == hp.style_embed_depth)
    REQ(hp.reference_depth == 128)
    REQ(hp.n_reference_filters >= 1 && hp.n_reference_filters <= 8)
    REQ(hp.attention_dim == 128)
    REQ(hp.attention_filters == 32)
    REQ(hp.attention_kernel == 31)
    REQ(hp.prenet_layers[0] == 256 && hp.prenet_layers[1] == 256)
    REQ(hp.decoder_lstm_units == 256)
    REQ(hp.num_mels == 80)
    REQ(hp.outputs_per_step == 2)
    REQ(hp.conv_num_layers >= 1 && hp.conv_num_layers <= 8)
    REQ(hp.postnet_num_layers >= 1 && hp.postnet_num_layers <= 8)
    REQ((hp.postnet_channels + hp.style_embed_depth) % 16 == 0)
    REQ(hp.conv_kernel_size % 2 == 1 && hp.postnet_kernel_size % 2 == 1)
#undef REQ
    return true;
}

}  // namespace

extern "C" {

const char* taco_version(void) { return "taco_b200 0.1 sm_100a"; }

int taco_create(const TacoHParams* hp, int device, TacoHandle** out) {
    if (!hp || !out) return TACO_ERR_INVALID;
    *out = nullptr;
    TacoHandle* h = new TacoHandle();
    h->hp = *hp;
    h->device = device;
    *out = h;  // returned even on failure so the caller can read taco_last_error
    std::string why;
    if (!hp_supported(*hp, &why)) FAIL(TACO_ERR_UNSUPPORTED, why);
    CK(cudaSetDevice(device));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, device));
    h->n_sm = prop.multiProcessorCount;
    if (prop.major < 10) FAIL(TACO_ERR_UNSUPPORTED, "sm_100a device required (no fallback path)");
    if (!prop.cooperativeLaunch) FAIL(TACO_ERR_UNSUPPORTED, "cooperative launch unsupported");
    if (h->n_sm < 128) FAIL(TACO_ERR_UNSUPPORTED, "persistent decoder needs >= 128 SMs");
    for (int i = 0; i < TACO_N_STAGES; ++i)
        for (int j = 0; j < 2; ++j) CK(cudaEventCreate(&h->ev[i][j]));
    return TACO_OK;
}

void taco_destroy(TacoHandle* h) {
    if (!h) return;
    cudaSetDevice(h->device);
    cudaDeviceSynchronize();
    for (auto& kv : h->w)
        if (kv.second.dev) cudaFree(kv.second.dev);
    for (auto& kv : h->ws)
        if (kv.second.p) cudaFree(kv.second.p);
    for (void* p : h->owned) cudaFree(p);
    for (int i = 0; i < TACO_N_STAGES; ++i)
        for (int j = 0; j < 2; ++j)
            if (h->ev[i][j]) cudaEventDestroy(h->ev[i][j]);
    delete h;
}

const char* taco_last_error(const TacoHandle* h) { return h ? h->err.c_str() : "null handle"; }

int64_t taco_launch_count(const TacoHandle* h) { return h ? h->launches : 0; }

int taco_set_precision(TacoHandle* h, int precision) {
    if (!h) return TACO_ERR_INVALID;
    if (precision != TACO_PRECISION_FP32 && precision != TACO_PRECISION_BF16_CONV && precision != TACO_PRECISION_BF16)
        FAIL(TACO_ERR_INVALID, "bad precision");
    if (precision != TACO_PRECISION_FP32) {
        const int PC = h->hp.postnet_channels + h->hp.style_embed_depth;
        if (PC % 256 != 0 || h->hp.num_mels != 80) FAIL(TACO_ERR_UNSUPPORTED, "tcgen05 postnet needs a channel width multiple of 256 and 80 mels");
    }
    h->precision = precision;
    return TACO_OK;
}

int taco_set_weight(TacoHandle* h, const char* name, const void* data, const int64_t* shape, int rank, int dtype) {
    if (!h || !name || !data || rank < 0 || rank > 4) return TACO_ERR_INVALID;
    if (dtype != TACO_DTYPE_F32) FAIL(TACO_ERR_UNSUPPORTED, "only fp32 weights are accepted");
    CK(cudaSetDevice(h->device));
    size_t n = 1;
    for (int i = 0; i < rank; ++i) n *= (size_t)shape[i];
    Weight& w = h->w[name];
    w.shape.assign(shape, shape + rank);
    w.host.resize(n);
    CK(cudaMemcpy(w.host.data(), data, n * sizeof(float), cudaMemcpyDefault));
    if (w.dev) {
        CK(cudaFree(w.dev));
        w.dev = nullptr;
    }
    CK(cudaMalloc(&w.dev, (n + 4) * sizeof(float)));
    CK(cudaMemcpy(w.dev, w.host.data(), n * sizeof(float), cudaMemcpyHostToDevice));
    h->finalized = false;
    return TACO_OK;
}

int taco_finalize_weights(TacoHandle* h) {
    if (!h) return TACO_ERR_INVALID;
    CK(cudaSetDevice(h->device));
    for (void* p : h->owned) cudaFree(p);
    h->owned.clear();
    const float eps = 1e-3f;  // tf.layers.batch_normalization default
    int rc;
    for (int i = 0; i < h->hp.conv_num_layers; ++i)
        if ((rc = fold_bn(h, fmt(ENC_CONV, i + 1), "conv1d", eps, &h->enc_scale[i], &h->enc_shift[i]))) return rc;
    for (int i = 0; i < h->hp.postnet_num_layers; ++i)
        if ((rc = fold_bn(h, fmt(POST_CONV, i + 1), "conv1d", eps, &h->post_scale[i], &h->post_shift[i]))) return rc;
    for (int i = 0; i < h->hp.n_reference_filters; ++i)
        if ((rc = fold_bn(h, fmt(REF_CONV, i), "conv2d", eps, &h->ref_scale[i], &h->ref_shift[i]))) return rc;
    if ((rc = pack_decoder_blobs(h))) return rc;
    if ((rc = pack_location(h))) return rc;
    if ((rc = pack_decoder3(h))) return rc;
    const char* required[] = {"inputs_embedding", "style_tokens"};
    for (const char* r : required)
        if (!find_w(h, r)) FAIL(TACO_ERR_MISSING_WEIGHT, std::string("missing variable ") + r);
    std::string req2[] = {std::string(MEMORY_LAYER) + "/kernel", std::string(LSA) + "/location_features_convolution/kernel",
                          std::string(LSA) + "/location_features_convolution/bias", std::string(LSA) + "/location_features_layer/kernel",
                          std::string(LSA) + "/attention_variable", std::string(LSA) + "/attention_bias",
                          std::string(POST_PROJ) + "/kernel", std::string(POST_PROJ) + "/bias",
                          std::string(REF_GRU) + "/gates/kernel", std::string(REF_GRU) + "/gates/bias",
                          std::string(REF_GRU) + "/candidate/kernel", std::string(REF_GRU) + "/candidate/bias",
                          std::string(REF_DENSE) + "/kernel", std::string(REF_DENSE) + "/bias",
                          std::string(MHA) + "/conv1d/kernel", std::string(MHA) + "/conv1d/bias",
                          std::string(MHA) + "/conv1d_1/kernel", std::string(MHA) + "/conv1d_1/bias",
                          std::string(MHA) + "/attention_v", std::string(MHA) + "/attention_g", std::string(MHA) + "/attention_b"};
    for (const std::string& r : req2)
        if (!find_w(h, r)) FAIL(TACO_ERR_MISSING_WEIGHT, "missing variable " + r);
    for (int i = 0; i < h->hp.conv_num_layers; ++i)
        if (!find_w(h, fmt(ENC_CONV, i + 1) + "/conv1d/kernel")) FAIL(TACO_ERR_MISSING_WEIGHT, "missing encoder conv kernel");
    for (int i = 0; i < h->hp.postnet_num_layers; ++i)
        if (!find_w(h, fmt(POST_CONV, i + 1) + "/conv1d/kernel")) FAIL(TACO_ERR_MISSING_WEIGHT, "missing postnet conv kernel");
    for (int i = 0; i < h->hp.n_reference_filters; ++i)
        if (!find_w(h, fmt(REF_CONV, i) + "/conv2d/kernel")) FAIL(TACO_ERR_MISSING_WEIGHT, "missing reference conv2d kernel");
    {
        const TacoHParams& hp = h->hp;
        const int PC = hp.postnet_channels + hp.style_embed_depth;
        int cin = hp.num_mels;
        for (int i = 0; i < hp.postnet_num_layers; ++i) {
            const Weight* w = find_w(h, fmt(POST_CONV, i + 1) + "/conv1d/kernel");
            const int cpad = (cin + 63) / 64 * 64;
            if ((int64_t)w->shape[0] != hp.postnet_kernel_size || (int64_t)w->shape[1] != cin || (int64_t)w->shape[2] != PC)
                FAIL(TACO_ERR_INVALID, "postnet conv kernel shape");
            h->post_cin_pad[i] = cpad;
            if ((rc = pack_conv_bf16(h, w, hp.postnet_kernel_size, cin, cpad, PC, &h->post_wt[i]))) return rc;
            cin = PC;
        }
        const Weight* wp = find_w(h, std::string(POST_PROJ) + "/kernel");
        if ((int64_t)wp->shape[0] != PC || (int64_t)wp->shape[1] != hp.num_mels) FAIL(TACO_ERR_INVALID, "postnet projection shape");
        if ((rc = pack_conv_bf16(h, wp, 1, PC, PC, hp.num_mels, &h->post_proj_wt))) return rc;
        cin = hp.embedding_dim;
        for (int i = 0; i < hp.conv_num_layers; ++i) {
            const Weight* w = find_w(h, fmt(ENC_CONV, i + 1) + "/conv1d/kernel");
            if ((rc = pack_conv_bf16(h, w, hp.conv_kernel_size, cin, cin, hp.conv_channels, &h->enc_wt[i]))) return rc;
            cin = hp.conv_channels;
        }
        for (int d = 0; d < 2; ++d) {
            const Weight* w = find_w(h, std::string(d == 0 ? ENC_FW : ENC_BW) + "/kernel");
            if ((rc = pack_conv_bf16(h, w, 1, hp.conv_channels, hp.conv_channels, 4 * hp.enc_lstm_hidden_size, &h->enc_xw_wt[d]))) return rc;
        }
        const Weight* wm = find_w(h, std::string(MEMORY_LAYER) + "/kernel");
        if ((rc = pack_conv_bf16(h, wm, 1, (int)wm->shape[0], (int)wm->shape[0], hp.attention_dim, &h->keys_wt))) return rc;
    }
    h->finalized = true;
    return TACO_OK;
}

int taco_encode(TacoHandle* h, const int32_t* ids, const int32_t* lens, int B, int T, const float* ref_mel, int Tm,
                const float* gst_weights, float* memory_out, void* stream) {
    if (!h) return TACO_ERR_INVALID;
    if (!h->finalized) FAIL(TACO_ERR_STATE, "taco_finalize_weights has not been called");
    if (!ids || !lens || !memory_out || B < 1 || T < 1) FAIL(TACO_ERR_INVALID, "taco_encode: bad arguments");
    if (!ref_mel && !gst_weights) FAIL(TACO_ERR_INVALID, "taco_encode: neither reference mel nor GST weights given");
    if (B > 64) FAIL(TACO_ERR_UNSUPPORTED, "taco_encode: B > 64 per call (shard the batch on the host)");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    CK(cudaSetDevice(h->device));
    const TacoHParams& hp = h->hp;
    const int E = hp.embedding_dim, C = hp.conv_channels, H = hp.enc_lstm_hidden_size, SD = hp.style_embed_depth;
    const int MD = 2 * H + SD;
    const size_t M = (size_t)B * T;
    CK(cudaEventRecord(h->ev[TACO_STAGE_ENCODE][0], st));
    WS(bufA, float, "enc_a", M * 512)
    WS(bufB, float, "enc_b", M * 512)
    WS(xw, float, "enc_xw", 2 * M * 1024)
    CK(taco_launch_embedding(ids, dev_w(h, "inputs_embedding"), bufA, (int)M, E, hp.n_symbols, st));
    h->launches++;
    float *cur = bufA, *nxt = bufB;
    int cin = E, rc;
    const bool enc_tc = h->precision != TACO_PRECISION_FP32 && E % 64 == 0 && C % 256 == 0 && (4 * H) % 256 == 0;
    if (enc_tc) {
        // bf16-operand tcgen05 path: conv stack (BN folded, ReLU) bf16 -> bf16, then both LSTM input projections
        // (+bias) with an fp32 copy for the recurrent kernel
        const size_t wide = (size_t)(4 * H > C ? 4 * H : C);
        WS(ta, uint16_t, "enc_tc_a", M * wide)
        WS(tb, uint16_t, "enc_tc_b", M * wide)
        CK(taco_launch_f32_to_bf16(bufA, ta, M * (size_t)E, st));
        h->launches++;
        uint16_t *c16 = ta, *n16 = tb;
        for (int i = 0; i < hp.conv_num_layers; ++i) {
            ConvTcParams c;
            memset(&c, 0, sizeof c);
            c.x = c16; c.wt = h->enc_wt[i];
            c.B = B; c.F = T; c.Cin = cin; c.taps = hp.conv_kernel_size; c.Cout = C;
            c.scale = h->enc_scale[i]; c.shift = h->enc_shift[i];
            c.act = ACT_RELU;
            c.y_bf16 = n16; c.y_stride = C;
            CK(taco_launch_conv_tc(c, h->n_sm, st));
            h->launches++;
            std::swap(c16, n16);
            cin = C;
        }
        for (int d = 0; d < 2; ++d) {
            ConvTcParams c;
            memset(&c, 0, sizeof c);
            c.x = c16; c.wt = h->enc_xw_wt[d];
            c.B = B; c.F = T; c.Cin = C; c.taps = 1; c.Cout = 4 * H;
            c.shift = dev_w(h, std::string(d == 0 ? ENC_FW : ENC_BW) + "/bias");
            c.act = ACT_NONE;
            c.y_bf16 = n16; c.y_stride = 4 * H;          // bf16 copy is scratch; the BiLSTM reads the fp32 copy
            c.y_f32 = xw + (size_t)d * M * 1024;
            CK(taco_launch_conv_tc(c, h->n_sm, st));
            h->launches++;
        }
    }
    for (int i = 0; i < hp.conv_num_layers && !enc_tc; ++i) {
        rc = conv(h, cur, cin, (long long)T * cin, dev_w(h, fmt(ENC_CONV, i + 1) + "/conv1d/kernel"), h->enc_scale[i],
                  h->enc_shift[i], nxt, C, (long long)T * C, nullptr, B, T, cin, C, hp.conv_kernel_size, ACT_RELU, false, 0, 0, st);
        if (rc) return rc;
        std::swap(cur, nxt);
        cin = C;
    }
    // hoisted input projection of both LSTM directions: xw[d] = conv_out . W[0:512] + bias
    for (int d = 0; d < 2 && !enc_tc; ++d) {
        const char* scope = d == 0 ? ENC_FW : ENC_BW;
        rc = conv(h, cur, C, (long long)T * C, dev_w(h, std::string(scope) + "/kernel"), nullptr,
                  dev_w(h, std::string(scope) + "/bias"), xw + (size_t)d * M * 1024, 1024, (long long)T * 1024, nullptr, B, T, C,
                  4 * H, 1, ACT_NONE, false, 0, 0, st);
        if (rc) return rc;
    }
    CK(cudaMemsetAsync(memory_out, 0, M * MD * sizeof(float), st));
    {
        BiLstmParams p;
        memset(&p, 0, sizeof p);
        p.B = B; p.T = T;
        p.RBP = B <= 4 ? 4 : (B <= 8 ? 8 : (B <= 16 ? 16 : 32));
        p.GB = (B + p.RBP - 1) / p.RBP;
        int cfg = p.GB == 1 ? 0 : 1;
        if (h->encoder_max_ctas > 0 && p.RBP == 32 && p.GB * (256 / (cfg == 0 ? 2 : 4)) > h->encoder_max_ctas) cfg = 2;   // 16 units per CTA
        p.U = cfg == 0 ? 2 : (cfg == 1 ? 4 : 16);
        p.GN = 256 / p.U;
        p.blobs = h->bilstm_blob[cfg];
        p.blob_floats = taco_bilstm_blob_floats(p.U);
        p.xw = xw;
        p.lens = lens;
        const size_t hT_floats = (size_t)2 * 2 * p.GB * 256 * p.RBP;
        WS(hT, float, "enc_hT", hT_floats)
        WS(bar, unsigned, "enc_bar", 4)
        CK(cudaMemsetAsync(hT, 0, hT_floats * sizeof(float), st));
        CK(cudaMemsetAsync(bar, 0, 4 * sizeof(unsigned), st));
        p.hT = hT;
        p.out = memory_out;
        p.out_stride = MD;
        p.zoneout = hp.zoneout_rate;
        p.barrier = bar;
        CK(taco_launch_bilstm(p, st));
        h->launches++;
    }
    // style embedding
    WS(style, float, "enc_style", (size_t)B * SD)
    const int heads = hp.num_heads, tokd = SD / heads;
    if (ref_mel) {
        if (Tm < 1) FAIL(TACO_ERR_INVALID, "taco_encode: reference mel length");
        int Hh = Tm, Ww = hp.num_mels, cin2 = 1;
        size_t maxel = 0;
        {
            int hh = Hh, ww = Ww;
            for (int i = 0; i < hp.n_reference_filters; ++i) {
                hh = (hh + 1) / 2; ww = (ww + 1) / 2;
                size_t el = (size_t)B * hh * ww * hp.reference_filters[i];
                if (el > maxel) maxel = el;
            }
        }
        WS(rA, float, "ref_a", maxel)
        WS(rB, float, "ref_b", maxel)
        const float* src = ref_mel;
        float *dst = rA, *other = rB;
        for (int i = 0; i < hp.n_reference_filters; ++i) {
            int cout = hp.reference_filters[i];
            CK(taco_launch_conv2d_s2(src, dev_w(h, fmt(REF_CONV, i) + "/conv2d/kernel"), h->ref_scale[i], h->ref_shift[i], dst, B,
                                     Hh, Ww, cin2, cout, st));
            h->launches++;
            Hh = (Hh + 1) / 2; Ww = (Ww + 1) / 2; cin2 = cout;
            src = dst;
            std::swap(dst, other);
        }
        WS(refo, float, "ref_out", (size_t)B * 128)
        const int In = Ww * cin2;
        const Weight* gk = find_w(h, std::string(REF_GRU) + "/gates/kernel");
        if ((int64_t)gk->shape[0] != In + 128) FAIL(TACO_ERR_INVALID, "reference GRU kernel does not match conv stack output width");
        CK(taco_launch_ref_gru(src, B, Hh, In, dev_w(h, std::string(REF_GRU) + "/gates/kernel"),
                               dev_w(h, std::string(REF_GRU) + "/gates/bias"), dev_w(h, std::string(REF_GRU) + "/candidate/kernel"),
                               dev_w(h, std::string(REF_GRU) + "/candidate/bias"), dev_w(h, std::string(REF_DENSE) + "/kernel"),
                               dev_w(h, std::string(REF_DENSE) + "/bias"), refo, st));
        h->launches++;
        CK(taco_launch_gst_attention(refo, dev_w(h, "style_tokens"), dev_w(h, std::string(MHA) + "/conv1d/kernel"),
                                     dev_w(h, std::string(MHA) + "/conv1d/bias"), dev_w(h, std::string(MHA) + "/conv1d_1/kernel"),
                                     dev_w(h, std::string(MHA) + "/conv1d_1/bias"), dev_w(h, std::string(MHA) + "/attention_v"),
                                     dev_w(h, std::string(MHA) + "/attention_g"), dev_w(h, std::string(MHA) + "/attention_b"), style, B,
                                     hp.num_gst, tokd, heads, hp.style_att_dim, st));
        h->launches++;
        CK(taco_launch_style_fill(memory_out, style, B, T, SD, MD, SD, st));
    } else {
        CK(taco_launch_style_random(gst_weights, dev_w(h, "style_tokens"), style, heads, hp.num_gst, tokd, st));
        h->launches++;
        CK(taco_launch_style_fill(memory_out, style, B, T, SD, MD, 0, st));
    }
    h->launches++;
    CK(cudaEventRecord(h->ev[TACO_STAGE_ENCODE][1], st));
    h->ev_valid[TACO_STAGE_ENCODE] = true;
    return TACO_OK;
}

int taco_decode(TacoHandle* h, int mode, const float* memory, const int32_t* lens, int B, int T, const float* mel_targets,
                int F_tgt, int S_max, const uint8_t* prenet_masks, uint64_t philox_seed, int reset_state, int logits_only,
                float* frames, float* stop, float* align_hist, int32_t* steps_out, void* stream) {
    if (!h) return TACO_ERR_INVALID;
    if (!h->finalized) FAIL(TACO_ERR_STATE, "taco_finalize_weights has not been called");
    if (!memory || !lens || !frames || !stop || !align_hist || !steps_out || B < 1 || T < 1 || S_max < 1)
        FAIL(TACO_ERR_INVALID, "taco_decode: bad arguments");
    if (mode != TACO_MODE_FREE && mode != TACO_MODE_TEACHER) FAIL(TACO_ERR_INVALID, "taco_decode: bad mode");
    if (mode == TACO_MODE_TEACHER && (!mel_targets || F_tgt < 1)) FAIL(TACO_ERR_INVALID, "taco_decode: teacher mode needs mel targets");
    if (B > 64) FAIL(TACO_ERR_UNSUPPORTED, "taco_decode: B > 64 per call (shard the batch on the host)");
    if (T > 1024) FAIL(TACO_ERR_UNSUPPORTED, "taco_decode: T > 1024");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    CK(cudaSetDevice(h->device));
    const TacoHParams& hp = h->hp;
    if (h->precision == TACO_PRECISION_BF16 && h->dec3_max_clusters < 0) {
        int n = 0;
        if (taco_decoder3_max_clusters(&n) != cudaSuccess) {
            cudaGetLastError();
            n = 0;
        }
        h->dec3_max_clusters = n > TACO_DEC3_MAX_CLUSTERS ? TACO_DEC3_MAX_CLUSTERS : n;
    }
    const bool dec3_fits = h->dec3_max_clusters > 0 && T <= TACO_DEC3_TMAX && h->dec3_max_clusters * 16 >= B &&
                           ((B + h->dec3_max_clusters - 1) / h->dec3_max_clusters) * ((T + 3) & ~3) <= TACO_DEC3_ACAP;
    if (h->precision == TACO_PRECISION_BF16 && dec3_fits)
        return decode_v3(h, mode, memory, lens, B, T, mel_targets, F_tgt, S_max, prenet_masks, philox_seed, reset_state, logits_only,
                         frames, stop, align_hist, steps_out, st);
    if (h->precision == TACO_PRECISION_BF16 && !h->warned_dec_fallback) {
        // never silent: the bf16 cluster decoder holds the alignment rows of a cluster in shared memory (rows x T <= TACO_DEC3_ACAP
        // words, T <= TACO_DEC3_TMAX, 16-CTA clusters schedulable); beyond that the exact fp32 kernel below decodes this batch
        fprintf(stderr, "[taco_b200] taco_decode: B=%d T=%d does not fit the bf16 cluster decoder (T <= %d, rows per cluster x T <= %d, "
                        "%d clusters): this batch runs on the fp32 persistent kernel (3x slower per step)\n",
                B, T, TACO_DEC3_TMAX, TACO_DEC3_ACAP, h->dec3_max_clusters);
        h->warned_dec_fallback = true;
    }

    DecParams p;
    memset(&p, 0, sizeof p);
    p.B = B; p.T = T;
    p.RBP = B <= 4 ? 4 : (B <= 8 ? 8 : (B <= 16 ? 16 : 32));
    p.GB = (B + p.RBP - 1) / p.RBP;
    const int cfg = p.GB == 1 ? 0 : 1;
    p.U = cfg == 0 ? 2 : 4;
    p.GN = 256 / p.U;
    const int nCta = p.GB * p.GN;
    static const int ns_opts[] = {96, 64, 48, 32, 24, 16, 12, 8, 6, 4, 3, 2, 1};
    p.NS = 1;
    for (int ns : ns_opts)
        if (B * ns <= nCta) { p.NS = ns; break; }
    p.S_max = S_max; p.S_cap = S_max; p.mode = mode;
    p.stop_at_any = hp.stop_at_any; p.cumulative = hp.cumulative_weights; p.smoothing = hp.smoothing;
    p.stop_sigmoid = logits_only ? 0 : 1;
    p.zoneout = hp.zoneout_rate;
    p.keep_scale = 1.0f / (1.0f - hp.dropout_rate);
    p.blobs = h->dec_blob[cfg];
    p.blob_floats = taco_decoder_blob_floats(p.U);
    p.Wl = dev_w(h, std::string(LSA) + "/location_features_layer/kernel");
    p.conv_w = dev_w(h, std::string(LSA) + "/location_features_convolution/kernel");
    p.conv_b = dev_w(h, std::string(LSA) + "/location_features_convolution/bias");
    p.v_a = dev_w(h, std::string(LSA) + "/attention_variable");
    p.b_a = dev_w(h, std::string(LSA) + "/attention_bias");

    if (!reset_state && (h->dec_kind != 1 || h->dec_B != B || h->dec_T != T))
        FAIL(TACO_ERR_STATE, "taco_decode: continuing with a different batch shape or precision");
    // lengths: mask_encoder False => every row attends over all T positions (Attention.py:150)
    WS(lens_eff, int32_t, "dec_lens", B)
    if (hp.mask_encoder) {
        CK(cudaMemcpyAsync(lens_eff, lens, B * sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
    } else {
        std::vector<int32_t> full(B, T);
        CK(cudaMemcpyAsync(lens_eff, full.data(), B * sizeof(int32_t), cudaMemcpyHostToDevice, st));
        CK(cudaStreamSynchronize(st));
    }
    p.lens = lens_eff;
    // keys = memory_layer(values), computed once per utterance batch (Attention.py:153-158)
    WS(keys, float, "dec_keys", (size_t)B * T * 128)
    if (reset_state) {
        int rc = conv(h, memory, 768, (long long)T * 768, dev_w(h, std::string(MEMORY_LAYER) + "/kernel"), nullptr, nullptr, keys, 128,
                      (long long)T * 128, nullptr, B, T, 768, 128, 1, ACT_NONE, false, 0, 0, st);
        if (rc) return rc;
    }
    p.values = memory;
    p.keys = keys;
    // exchange buffers: one allocation, zeroed on reset
    const size_t R = p.RBP, GBs = p.GB;
    const size_t n_x = GBs * 80 * R, n_256 = GBs * 256 * R, n_768 = GBs * 768 * R;
    const size_t n_q = (size_t)B * 128, n_cum = (size_t)B * T, n_c = (size_t)B * 256;
    const size_t total = n_x + 2 * n_256 /*p1,p2*/ + n_768 + 2 * n_256 /*h1z*/ + n_256 + 2 * n_256 /*h2z*/ + n_256 + n_q + n_cum + 2 * n_c;
    WS(ex, float, "dec_exchange", total + 64)
    WS(pre, float, "dec_pre", (size_t)B * T * 128)
    if (reset_state) {
        CK(cudaMemsetAsync(ex, 0, (total + 64) * sizeof(float), st));
        h->dec_t = 0;
    }
    float* c = ex;
    p.xT = c; c += n_x;
    p.p1T = c; c += n_256;
    p.p2T = c; c += n_256;
    p.ctxT = c; c += n_768;
    p.h1zT = c; c += 2 * n_256;
    p.h1rawT = c; c += n_256;
    p.h2zT = c; c += 2 * n_256;
    p.h2rawT = c; c += n_256;
    p.q = c; c += n_q;
    p.cum = c; c += n_cum;
    p.c1 = c; c += n_c;
    p.c2 = c; c += n_c;
    p.pre = pre;
    p.t0 = h->dec_t;
    p.targets = mel_targets;
    p.F_tgt = F_tgt;
    p.masks = prenet_masks;
    p.seed = philox_seed;
    p.frames = frames; p.stop = stop; p.hist = align_hist; p.steps_out = steps_out;
    if (mode == TACO_MODE_TEACHER && !reset_state && p.t0 > 0) {
        load_xt_from_targets_kernel<<<(B * 80 + 255) / 256, 256, 0, st>>>(mel_targets, F_tgt, p.t0 * hp.outputs_per_step - 1, p.xT, B, p.RBP);
        CK(cudaGetLastError());
        h->launches++;
    }
    WS(cnt, int, "dec_stop_count", (size_t)S_max + 8)
    CK(cudaMemsetAsync(cnt, 0, ((size_t)S_max + 8) * sizeof(int), st));
    p.stop_count = cnt;
    p.barrier = reinterpret_cast<unsigned*>(cnt + S_max + 4);
    CK(cudaEventRecord(h->ev[TACO_STAGE_DECODER_KERNEL][0], st));
    CK(taco_launch_decoder(p, st));
    CK(cudaEventRecord(h->ev[TACO_STAGE_DECODER_KERNEL][1], st));
    h->ev_valid[TACO_STAGE_DECODER_KERNEL] = true;
    h->launches++;
    h->dec_t += S_max;  // exact for S_max == 1 (cell stepping) and for runs to completion
    h->dec_B = B; h->dec_T = T; h->dec_RBP = p.RBP; h->dec_GB = p.GB; h->dec_kind = 1;
    return TACO_OK;
}

int taco_transpose_alignments(TacoHandle* h, const float* align_hist, int S, int B, int T, float* out, void* stream) {
    if (!h) return TACO_ERR_INVALID;
    if (!align_hist || !out || S < 1 || B < 1 || T < 1) FAIL(TACO_ERR_INVALID, "taco_transpose_alignments: bad arguments");
    CK(cudaSetDevice(h->device));
    CK(taco_launch_transpose_alignments(align_hist, out, S, B, T, reinterpret_cast<cudaStream_t>(stream)));
    h->launches++;
    return TACO_OK;
}

int taco_postnet(TacoHandle* h, const float* frames, int B, int F, int F_stride, float* decoder_output, float* mel_outputs,
                 void* stream) {
    if (!h) return TACO_ERR_INVALID;
    if (!h->finalized) FAIL(TACO_ERR_STATE, "taco_finalize_weights has not been called");
    if (!frames || !decoder_output || !mel_outputs || B < 1 || F < 1 || F_stride < F) FAIL(TACO_ERR_INVALID, "taco_postnet: bad arguments");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    CK(cudaSetDevice(h->device));
    const TacoHParams& hp = h->hp;
    const int M = hp.num_mels, PC = hp.postnet_channels + hp.style_embed_depth;
    const float hi = hp.max_abs_value;
    const float lo = (hp.symmetric_mels ? -hp.max_abs_value : 0.f) - hp.lower_bound_decay;
    CK(cudaEventRecord(h->ev[TACO_STAGE_POSTNET][0], st));
    if (h->precision != TACO_PRECISION_FP32) {
        // tcgen05 path: bf16 operands, fp32 accumulation; activations stay bf16 between layers
        const int C0 = h->post_cin_pad[0];
        const size_t wide = (size_t)(PC > C0 ? PC : C0);
        WS(ta, uint16_t, "post_tc_a", (size_t)B * F * wide)
        WS(tb, uint16_t, "post_tc_b", (size_t)B * F * wide)
        CK(taco_launch_prep_bf16(frames, decoder_output, ta, B, F, (long long)F_stride * M, M, C0, lo, hi, hp.clip_outputs, st));
        h->launches++;
        uint16_t *cur = ta, *nxt = tb;
        for (int i = 0; i < hp.postnet_num_layers; ++i) {
            ConvTcParams c;
            memset(&c, 0, sizeof c);
            c.x = cur; c.wt = h->post_wt[i];
            c.B = B; c.F = F; c.Cin = h->post_cin_pad[i]; c.taps = hp.postnet_kernel_size; c.Cout = PC;
            c.scale = h->post_scale[i]; c.shift = h->post_shift[i];
            c.act = i < hp.postnet_num_layers - 1 ? ACT_TANH : ACT_NONE;
            c.y_bf16 = nxt; c.y_stride = PC;
            CK(taco_launch_conv_tc(c, h->n_sm, st));
            h->launches++;
            std::swap(cur, nxt);
        }
        ConvTcParams c;
        memset(&c, 0, sizeof c);
        c.x = cur; c.wt = h->post_proj_wt;
        c.B = B; c.F = F; c.Cin = PC; c.taps = 1; c.Cout = M;
        c.shift = dev_w(h, std::string(POST_PROJ) + "/bias");
        c.act = ACT_NONE;
        c.y_f32 = mel_outputs; c.out_f32_only = 1; c.residual = decoder_output;
        c.clip = hp.clip_outputs != 0; c.clip_lo = lo; c.clip_hi = hi;
        CK(taco_launch_conv_tc(c, h->n_sm, st));
        h->launches++;
        CK(cudaEventRecord(h->ev[TACO_STAGE_POSTNET][1], st));
        h->ev_valid[TACO_STAGE_POSTNET] = true;
        return TACO_OK;
    }
    {
        size_t n = (size_t)B * F * M;
        clip_copy_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(frames, decoder_output, B, F, F_stride, M, lo, hi, hp.clip_outputs);
        CK(cudaGetLastError());
        h->launches++;
    }
    WS(pa, float, "post_a", (size_t)B * F * PC)
    WS(pb, float, "post_b", (size_t)B * F * PC)
    const float* cur = decoder_output;
    float *dst = pa, *other = pb;
    int cin = M, rc;
    for (int i = 0; i < hp.postnet_num_layers; ++i) {
        int act = i < hp.postnet_num_layers - 1 ? ACT_TANH : ACT_NONE;
        rc = conv(h, cur, cin, (long long)F * cin, dev_w(h, fmt(POST_CONV, i + 1) + "/conv1d/kernel"), h->post_scale[i], h->post_shift[i],
                  dst, PC, (long long)F * PC, nullptr, B, F, cin, PC, hp.postnet_kernel_size, act, false, 0, 0, st);
        if (rc) return rc;
        cur = dst;
        std::swap(dst, other);
        cin = PC;
    }
    rc = conv(h, cur, PC, (long long)F * PC, dev_w(h, std::string(POST_PROJ) + "/kernel"), nullptr,
              dev_w(h, std::string(POST_PROJ) + "/bias"), mel_outputs, M, (long long)F * M, decoder_output, B, F, PC, M, 1, ACT_NONE,
              hp.clip_outputs != 0, lo, hi, st);
    if (rc) return rc;
    CK(cudaEventRecord(h->ev[TACO_STAGE_POSTNET][1], st));
    h->ev_valid[TACO_STAGE_POSTNET] = true;
    return TACO_OK;
}

int taco_debug_read_workspace(TacoHandle* h, const char* name, void* host_dst, size_t bytes) {
    if (!h || !name || !host_dst) return TACO_ERR_INVALID;
    auto it = h->ws.find(name);
    if (it == h->ws.end() || it->second.bytes < bytes) FAIL(TACO_ERR_STATE, std::string("no such workspace: ") + name);
    CK(cudaSetDevice(h->device));
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(host_dst, it->second.p, bytes, cudaMemcpyDeviceToHost));
    return TACO_OK;
}

int taco_stage_ms(TacoHandle* h, int stage, float* ms_out) {
    if (!h || !ms_out || stage < 0 || stage >= TACO_N_STAGES) return TACO_ERR_INVALID;
    if (!h->ev_valid[stage]) FAIL(TACO_ERR_STATE, "taco_stage_ms: stage has not run");
    CK(cudaSetDevice(h->device));
    CK(cudaEventSynchronize(h->ev[stage][1]));
    CK(cudaEventElapsedTime(ms_out, h->ev[stage][0], h->ev[stage][1]));
    return TACO_OK;
}

int taco_set_encoder_ctas(TacoHandle* h, int max_ctas) {
    if (!h || max_ctas < 0) return TACO_ERR_INVALID;
    h->encoder_max_ctas = max_ctas;
    return TACO_OK;
}

int taco_get_decoder_state(TacoHandle* h, int B, int T, float* c1, float* h1, float* c2, float* h2, float* attention,
                           float* alignments, void* stream) {
    if (!h) return TACO_ERR_INVALID;
    if (h->dec_B != B || h->dec_T != T) FAIL(TACO_ERR_STATE, "taco_get_decoder_state: no decoder state of that shape");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    CK(cudaSetDevice(h->device));
    if (h->dec_kind == 3) {
        auto it2 = h->ws.find("dec2_state");
        if (it2 == h->ws.end()) FAIL(TACO_ERR_STATE, "no decoder state");
        const float* s = reinterpret_cast<const float*>(it2->second.p);
        const size_t nb = (size_t)B * 256 * sizeof(float);
        if (c1) CK(cudaMemcpyAsync(c1, s, nb, cudaMemcpyDeviceToDevice, st));
        if (h1) CK(cudaMemcpyAsync(h1, s + (size_t)B * 256, nb, cudaMemcpyDeviceToDevice, st));
        if (c2) CK(cudaMemcpyAsync(c2, s + (size_t)B * 512, nb, cudaMemcpyDeviceToDevice, st));
        if (h2) CK(cudaMemcpyAsync(h2, s + (size_t)B * 768, nb, cudaMemcpyDeviceToDevice, st));
        if (attention) CK(cudaMemcpyAsync(attention, s + (size_t)B * 1024, (size_t)B * 768 * sizeof(float), cudaMemcpyDeviceToDevice, st));
        if (alignments)
            CK(cudaMemcpyAsync(alignments, s + (size_t)B * (1024 + 768 + 80), (size_t)B * T * sizeof(float), cudaMemcpyDeviceToDevice, st));
        return TACO_OK;
    }
    auto it = h->ws.find("dec_exchange");
    if (it == h->ws.end()) FAIL(TACO_ERR_STATE, "no decoder state");
    float* ex = reinterpret_cast<float*>(it->second.p);
    const size_t R = h->dec_RBP, GBs = h->dec_GB;
    const size_t n_x = GBs * 80 * R, n_256 = GBs * 256 * R, n_768 = GBs * 768 * R;
    const size_t n_q = (size_t)B * 128, n_cum = (size_t)B * T, n_c = (size_t)B * 256;
    float* ctxT = ex + n_x + 2 * n_256;
    float* h1zT = ctxT + n_768;
    float* h2zT = h1zT + 3 * n_256;
    float* cum = h2zT + 3 * n_256 + n_q;
    float* c1s = cum + n_cum;
    float* c2s = c1s + n_c;
    const int par = h->dec_t & 1;  // buffer the NEXT step will read
    const int nb = (B * 768 + 255) / 256;
    if (c1) CK(cudaMemcpyAsync(c1, c1s, n_c * sizeof(float), cudaMemcpyDeviceToDevice, st));
    if (c2) CK(cudaMemcpyAsync(c2, c2s, n_c * sizeof(float), cudaMemcpyDeviceToDevice, st));
    if (alignments) CK(cudaMemcpyAsync(alignments, cum, n_cum * sizeof(float), cudaMemcpyDeviceToDevice, st));
    if (h1) untranspose_xt_kernel<<<nb, 256, 0, st>>>(h1zT + par * n_256, h1, B, 256, (int)R);
    if (h2) untranspose_xt_kernel<<<nb, 256, 0, st>>>(h2zT + par * n_256, h2, B, 256, (int)R);
    if (attention) untranspose_xt_kernel<<<nb, 256, 0, st>>>(ctxT, attention, B, 768, (int)R);
    CK(cudaGetLastError());
    return TACO_OK;
}

}  // extern "C"
