// Context object and buffer-staging helpers shared by the C-ABI translation units.
#pragma once
#include <cuda_runtime.h>
#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <string>

#include "../../include/cgnn_b200.h"

// ---------------------------------------------------------------------------------------------
struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
};

struct cgnn_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = true;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    cudaStream_t side[2] = {nullptr, nullptr};       // ragged pairwise batches: the launches of the other kernel builds
    cudaEvent_t fork = nullptr, join[2] = {nullptr, nullptr};
    cudaDeviceProp prop{};
    std::string err;
    float last_ms = 0.f;
    int last_launches = 0;
    int pw_ctas_per_sm = 0;
    int last_cluster = 1;                            // CTAs per instance of the last graph launch
    int last_graph_mode = 0;                         // shared-memory plan of the last graph launch (cgnn_b200_last_graph_mode)
    // staging buffers (grow-only)
    DevBuf xy, nrows, inst, scores, noise, init, trace, pout, counter, dbg[5];
    DevBuf gT, gdesc, gtidx, gsub, gdata, guid;      // graph path
    DevBuf wz;                                       // Fourier parity mode
};

static inline int fail(cgnn_ctx* c, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (c) c->err = buf;
    return code;
}

#define CU(call)                                                                                         \
    do {                                                                                                 \
        cudaError_t e_ = (call);                                                                         \
        if (e_ != cudaSuccess)                                                                           \
            return fail(ctx, e_ == cudaErrorMemoryAllocation ? CGNN_ERR_NOMEM : CGNN_ERR_CUDA, "%s: %s", \
                        #call, cudaGetErrorString(e_));                                                  \
    } while (0)

static inline int ensure(cgnn_ctx* ctx, DevBuf& b, size_t bytes) {
    if (bytes <= b.cap) return 0;
    if (b.p) { CU(cudaFree(b.p)); b.p = nullptr; b.cap = 0; }
    size_t cap = std::max<size_t>(bytes, 256);
    CU(cudaMalloc(&b.p, cap));
    b.cap = cap;
    return 0;
}

static inline bool is_device_ptr(const void* p) {
    cudaPointerAttributes a{};
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

// Make `bytes` of caller memory available on the device: returns the caller's pointer if it is
// already device memory, else stages it through `b`.
static inline int stage_in(cgnn_ctx* ctx, DevBuf& b, const void* src, size_t bytes, const void** out) {
    if (!src) { *out = nullptr; return 0; }
    if (is_device_ptr(src)) { *out = src; return 0; }
    int rc = ensure(ctx, b, bytes);
    if (rc) return rc;
    CU(cudaMemcpyAsync(b.p, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
    *out = b.p;
    return 0;
}

// Device-side landing buffer for an output: the caller's own pointer if it is device memory.
static inline int stage_out(cgnn_ctx* ctx, DevBuf& b, void* dst, size_t bytes, void** out) {
    if (!dst) { *out = nullptr; return 0; }
    if (is_device_ptr(dst)) { *out = dst; return 0; }
    int rc = ensure(ctx, b, bytes);
    if (rc) return rc;
    *out = b.p;
    return 0;
}

static inline int copy_back(cgnn_ctx* ctx, void* dst, const void* dev, size_t bytes) {
    if (!dst || dst == dev) return 0;
    CU(cudaMemcpyAsync(dst, dev, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    return 0;
}


static inline int check_hp(cgnn_ctx* ctx, const cgnn_hparams* hp) {
    if (!hp) return fail(ctx, CGNN_ERR_ARG, "hparams is NULL");
    if (hp->h_layer_dim < 1 || hp->h_layer_dim > CGNN_MAX_H)
        return fail(ctx, CGNN_ERR_LIMIT, "h_layer_dim=%d outside [1,%d]", hp->h_layer_dim, CGNN_MAX_H);
    if (hp->train_epochs < 0 || hp->test_epochs < 0) return fail(ctx, CGNN_ERR_ARG, "negative epoch count");
    if (hp->exp_policy < CGNN_EXP_POLICY_AUTO || hp->exp_policy > 3) return fail(ctx, CGNN_ERR_ARG, "exp_policy=%d unknown", hp->exp_policy);
    if (hp->graph_cluster < 0 || hp->graph_cluster > 8) return fail(ctx, CGNN_ERR_ARG, "graph_cluster=%d outside [0,8]", hp->graph_cluster);
    if (hp->fast_mmd_vectors < 0 || hp->fast_mmd_vectors > CGNN_MAX_FAST_VECTORS)
        return fail(ctx, CGNN_ERR_LIMIT, "fast_mmd_vectors=%d outside [0,%d]", hp->fast_mmd_vectors, CGNN_MAX_FAST_VECTORS);
    return 0;
}
