// Handle / model objects, weight packing, workspace layout, AVM forward driver and the CUDA-core
// validation path (gn_dbg_*).
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "gn_common.cuh"

// ================================================================================================
// handle
// ================================================================================================
struct gn_handle_tables {
    float2* tw1;       // [15][128] {wr, wi} of W_2048^(t k1), k1 = 1..15   (gn_fft2.cuh, pass 1)
    float2* tw2;       // [15][8]   {wr, wi} of W_128^(n3 k2), k2 = 1..15   (pass 2)
    float* hann;       // [2048] periodic hann
};
static gn_handle_tables* tables_of(gn_handle* h) { return reinterpret_cast<gn_handle_tables*>(h + 1); }

struct gn_mfcc_tables { const float2* tw1; const float2* tw2; const float* hann; };
gn_mfcc_tables gn_get_mfcc_tables(gn_handle* h) {
    gn_handle_tables* t = tables_of(h);
    return gn_mfcc_tables{t->tw1, t->tw2, t->hann};
}

extern "C" int gn_abi_version(void) { return GN_ABI_VERSION; }

extern "C" const char* gn_status_string(int s) {
    switch (s) {
        case GN_OK: return "ok";
        case GN_E_ARG: return "invalid argument";
        case GN_E_SHAPE: return "unsupported shape";
        case GN_E_WORKSPACE: return "workspace too small";
        case GN_E_NODEVICE: return "no usable CUDA device";
        case GN_E_RANGE: return "integer range exceeded";
        case GN_E_STATIC_LIMIT: return "problem exceeds the static scratch limit of the allocation-free entry point";
        default: break;
    }
    if (s > 0) return cudaGetErrorString((cudaError_t)s);
    return "unknown status";
}

extern "C" int gn_create(gn_handle** out, int device) {
    if (!out) return GN_E_ARG;
    *out = nullptr;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0) { cudaGetLastError(); return GN_E_NODEVICE; }
    if (device < 0 || device >= count) return GN_E_ARG;
    GN_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    GN_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) return GN_E_NODEVICE;   // sm_100a code only
    gn_handle* h = reinterpret_cast<gn_handle*>(calloc(1, sizeof(gn_handle) + sizeof(gn_handle_tables)));
    if (!h) return GN_E_ARG;
    h->device = device;
    h->num_sms = prop.multiProcessorCount;
    h->launches = 0;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess || !fn) {
        cudaGetLastError();
        free(h);
        return GN_E_NODEVICE;
    }
    h->encode_tiled = fn;
    // FFT tables, computed in double on the host (one allocation: tw1 | tw2 | hann)
    gn_handle_tables* t = tables_of(h);
    const int n1 = 15 * 128, n2 = 15 * 8;
    float* host = reinterpret_cast<float*>(malloc(sizeof(float) * (size_t)(2 * (n1 + n2) + 2048)));
    if (!host) { free(h); return GN_E_ARG; }
    float* p1 = host;
    float* p2 = host + 2 * n1;
    float* hn = host + 2 * (n1 + n2);
    for (int k1 = 1; k1 < 16; ++k1)
        for (int tt = 0; tt < 128; ++tt) {
            const double a = -2.0 * M_PI * (double)(tt * k1) / 2048.0;
            float* e = p1 + 2 * ((k1 - 1) * 128 + tt);
            e[0] = (float)cos(a);
            e[1] = (float)sin(a);
        }
    for (int k2 = 1; k2 < 16; ++k2)
        for (int n3 = 0; n3 < 8; ++n3) {
            const double a = -2.0 * M_PI * (double)(n3 * k2) / 128.0;
            float* e = p2 + 2 * ((k2 - 1) * 8 + n3);
            e[0] = (float)cos(a);
            e[1] = (float)sin(a);
        }
    for (int i = 0; i < 2048; ++i) hn[i] = (float)(0.5 - 0.5 * cos(2.0 * M_PI * (double)i / 2048.0));
    float* dev = nullptr;
    const size_t bytes = sizeof(float) * (size_t)(2 * (n1 + n2) + 2048);
    cudaError_t e = cudaMalloc(&dev, bytes);
    if (e == cudaSuccess) e = cudaMemcpy(dev, host, bytes, cudaMemcpyHostToDevice);
    free(host);
    if (e != cudaSuccess) { if (dev) cudaFree(dev); free(h); return (int)e; }
    t->tw1 = reinterpret_cast<float2*>(dev);
    t->tw2 = reinterpret_cast<float2*>(dev + 2 * n1);
    t->hann = dev + 2 * (n1 + n2);
    *out = h;
    return GN_OK;
}

extern "C" int gn_destroy(gn_handle* h) {
    if (!h) return GN_OK;
    gn_handle_tables* t = tables_of(h);
    cudaFree(t->tw1);          // tw1 | tw2 | hann are one allocation
    gn_preprocess_free_plans(h);
    free(h);
    return GN_OK;
}

extern "C" int64_t gn_launch_count(const gn_handle* h) { return h ? h->launches : -1; }

// ================================================================================================
// weight packing (device kernels; inputs are the reference state_dict tensors, fp32)
// ================================================================================================
namespace {

// conv weight [O][I][3][3] -> K-major bf16 [O][(ky*3+kx)*I + ci]
__global__ void pack_conv_kmajor(const float* __restrict__ w, int O, int I, __nv_bfloat16* __restrict__ out) {
    const int64_t total = (int64_t)O * I * 9;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int o = (int)(i / (9 * I));
        const int r = (int)(i - (int64_t)o * 9 * I);
        const int tap = r / I, ci = r - tap * I;
        out[i] = __float2bfloat16(w[((int64_t)o * I + ci) * 9 + tap]);
    }
}
// conv1 weight [64][3][3][3] -> fp32 [(ky*3+kx)*3 + ci][64]
__global__ void pack_conv1(const float* __restrict__ w, float* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < 27 * 64) {
        const int k = i / 64, co = i - k * 64;
        const int tap = k / 3, ci = k - tap * 3;
        out[i] = w[(co * 3 + ci) * 9 + tap];
    }
}
// linear5 weight [512][C*PP] with column c*PP + p -> bf16, k = p*C + c, K-blocked [K/64][512][64] (gn_w5_off)
__global__ void pack_linear5(const float* __restrict__ w, int C, int PP, __nv_bfloat16* __restrict__ out) {
    const int64_t K = (int64_t)C * PP, total = 512 * K;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t o = i / K, k = i - o * K;
        const int p = (int)(k / C), c = (int)(k - (int64_t)p * C);
        out[gn_w5_off(o, k)] = __float2bfloat16(w[o * K + (int64_t)c * PP + p]);
    }
}
// AudBl conv2 weight [128][64][3] -> TF32 [128][tap*64 + ci]
__global__ void pack_aconv2(const float* __restrict__ w, float* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < 128 * 192) {
        const int co = i / 192, k = i - co * 192, tap = k / 64, ci = k - tap * 64;
        out[i] = gn_round_tf32(w[(co * 64 + ci) * 3 + tap]);
    }
}
// AudBl linear3 weight [128][128*L2] with column c*L2 + l -> TF32 [128][l*128 + c]
__global__ void pack_alinear3(const float* __restrict__ w, int L2, float* __restrict__ out) {
    const int K = 128 * L2;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < 128 * K) {
        const int j = i / K, k = i - j * K, l = k / 128, c = k - l * 128;
        out[i] = gn_round_tf32(w[j * K + c * L2 + l]);
    }
}
__global__ void cast_tf32(const float* __restrict__ w, int64_t total, float* __restrict__ out) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x)
        out[i] = gn_round_tf32(w[i]);
}
__global__ void copy_f32(const float* __restrict__ w, int64_t total, float* __restrict__ out) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x)
        out[i] = w[i];
}
// BatchNorm eval: y = (x - mean) / sqrt(var + 1e-5) * gamma + beta  ->  y = x * scale + shift
__global__ void fold_bn(const float* g, const float* b, const float* mu, const float* var, int C, float* scale,
                        float* shift) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < C) {
        const float s = g[i] / sqrtf(var[i] + 1e-5f);
        scale[i] = s;
        shift[i] = b[i] - mu[i] * s;
    }
}

struct Arena {
    char* base; size_t off, cap;
    template <typename T> T* take(size_t n) {
        off = (off + 255) & ~size_t(255);
        T* p = reinterpret_cast<T*>(base + off);
        off += n * sizeof(T);
        return p;
    }
};

inline int conv_out(int n, int k, int s, int p) { return (n + 2 * p - k) / s + 1; }

}  // namespace

extern "C" int gn_model_create(gn_handle* h, gn_model** out, const gn_avm_state* sd, int audio_included, int S, int B,
                               void* stream) {
    if (!h || !out || !sd) return GN_E_ARG;
    *out = nullptr;
    if (S < 16 || S > 1024) return GN_E_SHAPE;   // 40 = the reference (utils.py:285); others run the tiled conv path
    const int C1 = conv_out(S, 3, 3, 3), P1 = C1 - 2, P2 = P1 - 2, P3 = P2 - 2;
    const int L1 = conv_out(B, 3, 2, 1), L2 = conv_out(L1, 3, 2, 1);
    if (audio_included && (B < 4 || L2 != 8 || B > 32)) return GN_E_SHAPE;   // tensor-core AudBl tiles: 16 frames x 8 positions
    const float* need[] = {sd->conv1_w, sd->conv1_b, sd->conv2_w, sd->conv2_b, sd->conv3_w, sd->conv3_b, sd->bn1_w,
                           sd->bn1_b, sd->bn1_m, sd->bn1_v, sd->bn2_w, sd->bn2_b, sd->bn2_m, sd->bn2_v, sd->bn3_w,
                           sd->bn3_b, sd->bn3_m, sd->bn3_v, sd->linear5_w, sd->linear5_b, sd->f0_w, sd->f0_b,
                           sd->f3_w, sd->f3_b, sd->f6_w, sd->f6_b, sd->f9_w, sd->f9_b, sd->f12_w, sd->f12_b};
    for (const float* p : need) if (!p) return GN_E_ARG;
    if (audio_included && (!sd->aconv1_w || !sd->aconv1_b || !sd->aconv2_w || !sd->aconv2_b || !sd->alinear3_w ||
                           !sd->alinear3_b)) return GN_E_ARG;

    gn_model* m = reinterpret_cast<gn_model*>(calloc(1, sizeof(gn_model)));
    if (!m) return GN_E_ARG;
    m->h = h; m->audio_included = audio_included; m->S = S; m->B = B;
    m->C1 = C1; m->P1 = P1; m->P2 = P2; m->P3 = P3; m->L1 = L1; m->L2 = L2;
    m->fin = 512 + (audio_included ? 128 : 0);
    const int64_t K5 = 512LL * P3 * P3;
    const size_t cap = (size_t)(512 * K5 + 512 * 2304 + 256 * 576) * 2 +
                       (size_t)(512 * 640 + 512 * 512 + 256 * 512 + 128 * 256) * 4 +
                       (size_t)(64 * 90 + 128 * 192 + 128 * 128 * L2) * 8 + (1 << 20);
    cudaError_t e = cudaMalloc(&m->arena, cap);
    if (e != cudaSuccess) { free(m); return (int)e; }
    Arena A{reinterpret_cast<char*>(m->arena), 0, cap};
    cudaStream_t st = gn_stream(stream);
    const int T = 256;
    auto blocks = [&](int64_t total) { int64_t b = (total + T - 1) / T; return (int)(b > 4096 ? 4096 : b); };

    m->conv1_w = A.take<float>(27 * 64);
    pack_conv1<<<(27 * 64 + T - 1) / T, T, 0, st>>>(sd->conv1_w, m->conv1_w);
    auto copyf = [&](const float* src, int64_t cnt) {
        float* d = A.take<float>(cnt);
        copy_f32<<<blocks(cnt), T, 0, st>>>(src, cnt, d);
        return d;
    };
    m->conv1_b = copyf(sd->conv1_b, 64);
    m->conv2_b = copyf(sd->conv2_b, 256);
    m->conv3_b = copyf(sd->conv3_b, 512);
    m->linear5_b = copyf(sd->linear5_b, 512);
    const float* bnp[3][4] = {{sd->bn1_w, sd->bn1_b, sd->bn1_m, sd->bn1_v},
                              {sd->bn2_w, sd->bn2_b, sd->bn2_m, sd->bn2_v},
                              {sd->bn3_w, sd->bn3_b, sd->bn3_m, sd->bn3_v}};
    const int bnc[3] = {64, 256, 512};
    for (int i = 0; i < 3; ++i) {
        m->bn_scale[i] = A.take<float>(bnc[i]);
        m->bn_shift[i] = A.take<float>(bnc[i]);
        fold_bn<<<(bnc[i] + T - 1) / T, T, 0, st>>>(bnp[i][0], bnp[i][1], bnp[i][2], bnp[i][3], bnc[i], m->bn_scale[i],
                                                    m->bn_shift[i]);
    }
    m->conv2_wk = A.take<__nv_bfloat16>(256 * 576);
    pack_conv_kmajor<<<blocks(256 * 576), T, 0, st>>>(sd->conv2_w, 256, 64, m->conv2_wk);
    m->conv3_wk = A.take<__nv_bfloat16>(512 * 2304);
    pack_conv_kmajor<<<blocks(512 * 2304), T, 0, st>>>(sd->conv3_w, 512, 256, m->conv3_wk);
    m->linear5_wk = A.take<__nv_bfloat16>(512 * K5);
    pack_linear5<<<blocks(512 * K5), T, 0, st>>>(sd->linear5_w, 512, P3 * P3, m->linear5_wk);
    if (audio_included) {
        m->aconv1_w = copyf(sd->aconv1_w, 64 * 90);
        m->aconv1_b = copyf(sd->aconv1_b, 64);
        m->aconv2_w = copyf(sd->aconv2_w, 128 * 192);
        m->aconv2_b = copyf(sd->aconv2_b, 128);
        m->alinear3_w = copyf(sd->alinear3_w, 128LL * 128 * L2);
        m->alinear3_b = copyf(sd->alinear3_b, 128);
        m->aconv2_wk = A.take<float>(128 * 192);
        pack_aconv2<<<(128 * 192 + T - 1) / T, T, 0, st>>>(sd->aconv2_w, m->aconv2_wk);
        m->alinear3_wk = A.take<float>((size_t)128 * 128 * L2);
        pack_alinear3<<<(128 * 128 * L2 + T - 1) / T, T, 0, st>>>(sd->alinear3_w, L2, m->alinear3_wk);
    }
    const float* fw[4] = {sd->f0_w, sd->f3_w, sd->f6_w, sd->f9_w};
    const float* fb[4] = {sd->f0_b, sd->f3_b, sd->f6_b, sd->f9_b};
    const int fo[4] = {512, 512, 256, 128};
    const int fi[4] = {m->fin, 512, 512, 256};
    for (int i = 0; i < 4; ++i) {
        m->f_wk[i] = A.take<float>((size_t)fo[i] * fi[i]);
        cast_tf32<<<blocks((int64_t)fo[i] * fi[i]), T, 0, st>>>(fw[i], (int64_t)fo[i] * fi[i], m->f_wk[i]);
        m->f_b[i] = copyf(fb[i], fo[i]);
    }
    m->f12_w = copyf(sd->f12_w, 128);
    m->f12_b = copyf(sd->f12_b, 1);
    h->launches += 30;
    e = cudaStreamSynchronize(st);
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess || A.off > A.cap) {
        cudaFree(m->arena);
        free(m);
        return e != cudaSuccess ? (int)e : GN_E_WORKSPACE;
    }
    *out = m;
    return GN_OK;
}

extern "C" int gn_model_set_timing(gn_model* m, int enable) {
    if (!m) return GN_E_ARG;
    if (enable && !m->ev[0])
        for (int i = 0; i <= GN_AVM_STAGES; ++i) GN_CUDA(cudaEventCreate(&m->ev[i]));
    m->timing = enable ? 1 : 0;
    return GN_OK;
}

extern "C" int gn_model_get_timing(gn_model* m, float* h_ms, int n_stages) {
    if (!m || !h_ms || n_stages < GN_AVM_STAGES || !m->ev[0]) return GN_E_ARG;
    GN_CUDA(cudaEventSynchronize(m->ev[GN_AVM_STAGES]));
    for (int i = 0; i < GN_AVM_STAGES; ++i) GN_CUDA(cudaEventElapsedTime(&h_ms[i], m->ev[i], m->ev[i + 1]));
    return GN_OK;
}

extern "C" int gn_model_destroy(gn_model* m) {
    if (!m) return GN_OK;
    if (m->ev[0])
        for (int i = 0; i <= GN_AVM_STAGES; ++i) cudaEventDestroy(m->ev[i]);
    cudaFree(m->arena);
    if (m->cls_w) cudaFree(m->cls_w);
    free(m);
    return GN_OK;
}

// Classifier variant of the fusion head (CAVM / CVM in the report's Table 1; in the reference it survives as the
// commented lines utils.py:257 `nn.Softmax(dim = 1)`, main.py:69 `nn.CrossEntropyLoss()` and main.py:97
// `torch.argmax(batch_predictions, axis = 1) + 1`): fusion.12 becomes Linear(128, 5) and Softmax replaces
// Sigmoid and 4y+1.  After this call gn_avm_forward writes [n][5] class probabilities to d_scores.
extern "C" int gn_model_set_classifier_head(gn_model* m, const float* weight, const float* bias, int n_classes,
                                            void* stream) {
    if (!m || !weight || !bias) return GN_E_ARG;
    if (n_classes != 5) return GN_E_SHAPE;
    cudaStream_t st = gn_stream(stream);
    if (!m->cls_w) {
        cudaError_t e = cudaMalloc(&m->cls_w, (size_t)(n_classes * 128 + 8) * sizeof(float));
        if (e != cudaSuccess) { m->cls_w = nullptr; return (int)e; }
        m->cls_b = m->cls_w + n_classes * 128;
    }
    cudaError_t e = cudaMemcpyAsync(m->cls_w, weight, (size_t)n_classes * 128 * sizeof(float), cudaMemcpyDeviceToDevice, st);
    if (e == cudaSuccess)
        e = cudaMemcpyAsync(m->cls_b, bias, (size_t)n_classes * sizeof(float), cudaMemcpyDeviceToDevice, st);
    if (e != cudaSuccess) return (int)e;
    m->n_cls = n_classes;
    return GN_OK;
}

// ================================================================================================
// workspace
// ================================================================================================
gn_ws gn_ws_layout(const gn_model* m, int64_t n, void* base) {
    gn_ws w{};
    size_t off = 0;
    auto take = [&](size_t elems) {
        off = (off + 1023) & ~size_t(1023);
        __nv_bfloat16* p = reinterpret_cast<__nv_bfloat16*>(reinterpret_cast<char*>(base) + off);
        off += elems * sizeof(__nv_bfloat16);
        return p;
    };
    auto takef = [&](size_t elems) { return reinterpret_cast<float*>(take(elems * 2)); };
    const size_t nn = (size_t)(n > 0 ? n : 1);
    w.a1p = take(nn * m->C1 * m->C1 * 64);
    w.a2p = take(nn * m->P1 * m->P1 * 256);
    w.a3 = take(((nn + 127) / 128 * 128) * m->P3 * m->P3 * 512);
    w.ac1 = takef(nn * (m->L1 + 2) * 64);
    w.ac2 = takef(nn * m->L2 * 128);
    w.h0 = takef(nn * m->fin);
    w.h1 = takef(nn * 512);
    w.h2 = takef(nn * 512);
    w.h3 = takef(nn * 256);
    w.part = takef((size_t)gn_linear5_splits(m) * nn * 512);
    w.bytes = (int64_t)((off + 1023) & ~size_t(1023));
    return w;
}

extern "C" int64_t gn_avm_workspace_bytes(const gn_model* m, int64_t n) {
    if (!m || n < 0) return GN_E_ARG;
    return gn_ws_layout(m, n, nullptr).bytes;
}

static int avm_forward_impl(gn_model* m, const float* d_visual, const float* d_audio, int64_t n, float* d_scores,
                            float* d_visfeat, void* d_ws, int64_t ws_bytes, void* stream, bool simt,
                            const gn_score_sink* sink = nullptr) {
    if (!m || n < 0) return GN_E_ARG;
    if (n == 0) return GN_OK;
    if (!d_visual || (!d_scores && !sink) || !d_ws) return GN_E_ARG;
    if (m->audio_included && !d_audio) return GN_E_ARG;
    if ((reinterpret_cast<uintptr_t>(d_ws) & 1023) != 0) return GN_E_ARG;
    if (n > (1LL << 31) / 256) return GN_E_RANGE;   // TMA coordinates are int32
    const gn_ws ws = gn_ws_layout(m, n, d_ws);
    if (ws.bytes > ws_bytes) return GN_E_WORKSPACE;
    cudaStream_t st = gn_stream(stream);
    GN_STAGE_MARK(m, 0, st);
    int rc = gn_launch_visbl_front(m, d_visual, n, ws, st, simt);
    if (rc) return rc;
    GN_STAGE_MARK(m, 1, st);
    rc = gn_launch_audbl(m, d_audio, n, ws, st, simt);
    if (rc) return rc;
    GN_STAGE_MARK(m, 2, st);
    return simt ? gn_launch_simt_backbone(m, n, ws, d_scores, d_visfeat, st)
                : gn_launch_tc_backbone(m, n, ws, d_scores, d_visfeat, st, sink);
}

extern "C" int gn_avm_forward_scatter(gn_model* m, const float* d_visual, const float* d_audio, int64_t n,
                                      const gn_score_sink* sink, void* d_workspace, int64_t workspace_bytes,
                                      void* stream) {
    if (!m || !sink) return GN_E_ARG;
    if (m->n_cls > 1) return GN_E_ARG;             // regression head only
    if (sink->n_dst < 1 || sink->n_dst > GN_MAX_PEERS || sink->rows_per_match < 1 || sink->dst_match_stride < 1 ||
        sink->dst_offset < 0 || sink->row0 < 0) return GN_E_ARG;
    for (int p = 0; p < sink->n_dst; ++p) if (!sink->dst[p]) return GN_E_ARG;
    return avm_forward_impl(m, d_visual, d_audio, n, nullptr, nullptr, d_workspace, workspace_bytes, stream, false, sink);
}

extern "C" int gn_avm_forward(gn_model* m, const float* d_visual, const float* d_audio, int64_t n, float* d_scores,
                              float* d_visual_features, void* d_workspace, int64_t workspace_bytes, void* stream) {
    return avm_forward_impl(m, d_visual, d_audio, n, d_scores, d_visual_features, d_workspace, workspace_bytes, stream,
                            false);
}

extern "C" int gn_dbg_avm_forward_simt(gn_model* m, const float* d_visual, const float* d_audio, int64_t n,
                                       float* d_scores, float* d_visual_features, void* d_workspace,
                                       int64_t workspace_bytes, void* stream) {
    if (m && m->n_cls > 1) return GN_E_ARG;       // the validation path has the regression head only
    return avm_forward_impl(m, d_visual, d_audio, n, d_scores, d_visual_features, d_workspace, workspace_bytes, stream,
                            true);
}

// ================================================================================================
// validation path: the same layers on CUDA cores (bf16 operands, fp32 accumulate)
// ================================================================================================
namespace {

// one CTA = (frame, 64 output channels); conv 3x3 s1 p1 on zero-bordered NHWC + ReLU + pool + BN
__global__ void __launch_bounds__(256)
simt_conv_block(const __nv_bfloat16* __restrict__ in, int OH, int Cin, const __nv_bfloat16* __restrict__ wk, int Cout,
                const float* __restrict__ bias, const float* __restrict__ sc, const float* __restrict__ sh,
                __nv_bfloat16* __restrict__ out, int out_pad, int blocked) {
    extern __shared__ float cs[];   // [OH*OH][64]
    const int f = blockIdx.x, cblk = blockIdx.y;
    const int co = cblk * 64 + (threadIdx.x & 63), pg = threadIdx.x >> 6;
    const int IW = OH + 2, K = 9 * Cin;
    const __nv_bfloat16* fin = in + (size_t)f * IW * IW * Cin;
    const __nv_bfloat16* wr = wk + (size_t)co * K;
    for (int p = pg; p < OH * OH; p += 4) {
        const int oy = p / OH, ox = p - oy * OH;
        float acc = 0.f;
        for (int tap = 0; tap < 9; ++tap) {
            const int ky = tap / 3, kx = tap - ky * 3;
            const __nv_bfloat16* ip = fin + ((size_t)(oy + ky) * IW + ox + kx) * Cin;
            const __nv_bfloat16* wp = wr + tap * Cin;
            for (int ci = 0; ci < Cin; ++ci) acc = fmaf(__bfloat162float(ip[ci]), __bfloat162float(wp[ci]), acc);
        }
        // the tensor-core epilogue rounds the raw conv sum to bf16, pools, then adds the bias and applies ReLU (max
        // commutes with the monotone x -> relu(x + b)); do the same
        cs[p * 64 + (threadIdx.x & 63)] = __bfloat162float(__float2bfloat16(acc));
    }
    __syncthreads();
    const int P = OH - 2, PW = P + 2 * out_pad;
    __nv_bfloat16* o = out + (size_t)f * PW * PW * Cout;
    for (int p = pg; p < PW * PW; p += 4) {
        const int y = p / PW - out_pad, x = p % PW - out_pad;
        float r = 0.f;
        if (y >= 0 && y < P && x >= 0 && x < P) {
            float mx = -INFINITY;
            for (int dy = 0; dy < 3; ++dy)
                for (int dx = 0; dx < 3; ++dx) mx = fmaxf(mx, cs[((y + dy) * OH + x + dx) * 64 + (threadIdx.x & 63)]);
            r = fmaf(fmaxf(mx + bias[co], 0.f), sc[co], sh[co]);
        }
        if (blocked) out[gn_a3_off(f, (int64_t)p * Cout + co, (int64_t)PW * PW * Cout / 64)] = __float2bfloat16(r);
        else o[(size_t)p * Cout + co] = __float2bfloat16(r);
    }
}

__device__ __forceinline__ float ldf(const __nv_bfloat16* p) { return __bfloat162float(*p); }
__device__ __forceinline__ float ldf(const float* p) { return *p; }

// out[r][col_off + j] = tf32(relu(A[r] . W[j] + b[j])); one warp per (row, j), lanes stride K.  T = bf16 (linear5) or
// float (head layers: TF32-rounded values, whose products are exact in fp32 -- the same products the tensor core forms)
template <typename T>
__global__ void __launch_bounds__(256)
simt_linear(const T* __restrict__ A, int64_t M, int K, int lda, const T* __restrict__ W, int N,
            const float* __restrict__ bias, float* __restrict__ out, int ld_out, int col_off,
            float* __restrict__ out_f32, int blocked) {
    const int lane = threadIdx.x & 31;
    const int64_t gw = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t o = gw; o < M * N; o += nw) {
        const int64_t r = o / N;
        const int j = (int)(o - r * N);
        const T* a = A + r * lda;
        const T* w = W + (size_t)j * K;
        float acc = 0.f;
        if (blocked) {   // linear5 operands (gn_a3_off / gn_w5_off)
            for (int k = lane; k < K; k += 32)
                acc = fmaf(ldf(A + gn_a3_off(r, k, K / 64)), ldf(W + gn_w5_off(j, k)), acc);
        } else
        for (int k = lane; k < K; k += 32) acc = fmaf(ldf(a + k), ldf(w + k), acc);
        for (int s = 16; s > 0; s >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, s);
        if (lane == 0) {
            const float v = gn_round_tf32(fmaxf(acc + bias[j], 0.f));
            out[r * ld_out + col_off + j] = v;
            if (out_f32) out_f32[r * N + j] = v;
        }
    }
}

// fusion.9 + ReLU + fusion.12 + sigmoid + 4y+1; one warp per row
__global__ void __launch_bounds__(256)
simt_score(const float* __restrict__ A, int64_t M, int K, const float* __restrict__ W,
           const float* __restrict__ bias, const float* __restrict__ wl, const float* __restrict__ bl,
           float* __restrict__ scores) {
    const int lane = threadIdx.x & 31;
    const int64_t gw = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t r = gw; r < M; r += nw) {
        float dot = 0.f;
        for (int j = 0; j < 128; ++j) {
            float acc = 0.f;
            for (int k = lane; k < K; k += 32)
                acc = fmaf(A[r * K + k], W[(size_t)j * K + k], acc);
            for (int s = 16; s > 0; s >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, s);
            dot = fmaf(fmaxf(acc + bias[j], 0.f), wl[j], dot);
        }
        if (lane == 0) scores[r] = 4.f * (1.f / (1.f + expf(-(dot + bl[0])))) + 1.f;
    }
}

// bf16 activation -> fp32, optionally NHWC(+border) -> NCHW interior
__global__ void read_act_nchw(const __nv_bfloat16* __restrict__ in, int64_t n, int C, int P, int pad, int chw_order,
                              float* __restrict__ out) {
    const int64_t total = n * C * P * P;
    const int PW = P + 2 * pad;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t f = i / ((int64_t)C * P * P);
        const int r = (int)(i - f * C * P * P);
        const int c = r / (P * P), p = r - c * P * P, y = p / P, x = p - y * P;
        if (chw_order == 2) out[i] = __bfloat162float(in[gn_a3_off(f, (int64_t)p * C + c, (int64_t)P * P * C / 64)]);   // K-blocked a3
        else out[i] = __bfloat162float(in[((f * PW + y + pad) * PW + x + pad) * C + c]);
    }
}
__global__ void read_act_flat(const float* __restrict__ in, int64_t total, float* __restrict__ out) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x)
        out[i] = in[i];
}

}  // namespace

int gn_launch_simt_backbone(gn_model* m, int64_t n, const gn_ws& ws, float* d_scores, float* d_visfeat,
                            cudaStream_t st) {
    if (n > 65535 * 16) return GN_E_SHAPE;
    if (m->P1 * m->P1 * 64 * 4 > 64 * 1024) return GN_E_SHAPE;   // validation path: one frame's conv tile in shared memory
    if (gn_attr_first(m->h, reinterpret_cast<const void*>(simt_conv_block)))
        GN_CUDA(cudaFuncSetAttribute(simt_conv_block, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
    simt_conv_block<<<dim3((unsigned)n, 4), 256, m->P1 * m->P1 * 64 * 4, st>>>(
        ws.a1p, m->P1, 64, m->conv2_wk, 256, m->conv2_b, m->bn_scale[1], m->bn_shift[1], ws.a2p, 1, 0);
    GN_LAUNCHED(m->h);
    GN_STAGE_MARK(m, 3, st);
    simt_conv_block<<<dim3((unsigned)n, 8), 256, m->P2 * m->P2 * 64 * 4, st>>>(
        ws.a2p, m->P2, 256, m->conv3_wk, 512, m->conv3_b, m->bn_scale[2], m->bn_shift[2], ws.a3, 0, 1);
    GN_LAUNCHED(m->h);
    GN_STAGE_MARK(m, 4, st);
    const int K5 = m->P3 * m->P3 * 512;
    const int aoff = m->audio_included ? 128 : 0;
    const int grid = m->h->num_sms * 8;
    simt_linear<__nv_bfloat16><<<grid, 256, 0, st>>>(ws.a3, n, K5, K5, m->linear5_wk, 512, m->linear5_b, ws.h0, m->fin, aoff, d_visfeat, 1);
    GN_LAUNCHED(m->h);
    GN_STAGE_MARK(m, 5, st);
    simt_linear<float><<<grid, 256, 0, st>>>(ws.h0, n, m->fin, m->fin, m->f_wk[0], 512, m->f_b[0], ws.h1, 512, 0, nullptr, 0);
    GN_LAUNCHED(m->h);
    simt_linear<float><<<grid, 256, 0, st>>>(ws.h1, n, 512, 512, m->f_wk[1], 512, m->f_b[1], ws.h2, 512, 0, nullptr, 0);
    GN_LAUNCHED(m->h);
    simt_linear<float><<<grid, 256, 0, st>>>(ws.h2, n, 512, 512, m->f_wk[2], 256, m->f_b[2], ws.h3, 256, 0, nullptr, 0);
    GN_LAUNCHED(m->h);
    simt_score<<<grid, 256, 0, st>>>(ws.h3, n, 256, m->f_wk[3], m->f_b[3], m->f12_w, m->f12_b, d_scores);
    GN_LAUNCHED(m->h);
    GN_STAGE_MARK(m, 6, st);
    return GN_OK;
}

extern "C" int gn_dbg_read_activation(gn_model* m, int which, int64_t n, const void* d_workspace, float* d_out,
                                      void* stream) {
    if (!m || !d_workspace || !d_out || n < 0) return GN_E_ARG;
    if (n == 0) return GN_OK;
    const gn_ws ws = gn_ws_layout(m, n, const_cast<void*>(d_workspace));
    cudaStream_t st = gn_stream(stream);
    const int grid = m->h->num_sms * 4;
    switch (which) {
        case 1: read_act_nchw<<<grid, 256, 0, st>>>(ws.a1p, n, 64, m->P1, 1, 1, d_out); break;
        case 2: read_act_nchw<<<grid, 256, 0, st>>>(ws.a2p, n, 256, m->P2, 1, 1, d_out); break;
        case 3: read_act_nchw<<<grid, 256, 0, st>>>(ws.a3, n, 512, m->P3, 0, 2, d_out); break;
        case 4: read_act_flat<<<grid, 256, 0, st>>>(ws.h0, n * m->fin, d_out); break;
        default: return GN_E_ARG;
    }
    GN_LAUNCHED(m->h);
    return GN_OK;
}
