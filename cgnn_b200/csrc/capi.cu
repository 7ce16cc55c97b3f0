// C-ABI of the cgnn_b200 engine: context, buffer staging, launches.  See include/cgnn_b200.h.
#include <cuda_runtime.h>
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <numeric>
#include <string>
#include <vector>

#include "ctx.h"
#include "common.cuh"
#include "pairwise_args.h"

using namespace cgnn;

// pairs of up to this many rows run on the 128- / 256-thread build of the pairwise kernel (measured: tools/bench_rows.py)
#ifndef CGNN_PW_ROWS_128
#define CGNN_PW_ROWS_128 256
#endif
#ifndef CGNN_PW_ROWS_256
#define CGNN_PW_ROWS_256 704
#endif

// ---------------------------------------------------------------------------------------------
// Pipe-peak micro-benchmarks: 8 independent dependency chains per thread, results kept live.
__global__ void __launch_bounds__(512) probe_mufu_kernel(float* sink, int iters) {
    float x[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) x[k] = 0.1f * float(k + 1) + 1e-3f * float(threadIdx.x & 31);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < 8; ++k) x[k] = ex2f(-x[k]);     // fixed point ~0.64: stays finite, never constant-folded
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += x[k];
    if (s == 123.456f) sink[0] = s;
}
__global__ void __launch_bounds__(512) probe_ffma_kernel(float* sink, int iters) {
    float x[8];
    const float a = 0.999f + 1e-6f * float(threadIdx.x & 31), b = 1e-3f;
#pragma unroll
    for (int k = 0; k < 8; ++k) x[k] = 0.1f * float(k + 1);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < 8; ++k) x[k] = fmaf(x[k], a, b);
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += x[k];
    if (s == 123.456f) sink[0] = s;
}

// packed FP32x2 FMA chains (fma.rn.f32x2): does the FMA pipe retire more than 128 FMA per clock per SM this way?
__global__ void __launch_bounds__(512) probe_ffma2_kernel(float* sink, int iters) {
    float2 x[8];
    const float2 a = splat2(0.999f + 1e-6f * float(threadIdx.x & 31)), b = splat2(1e-3f);
#pragma unroll
    for (int k = 0; k < 8; ++k) x[k] = make_float2(0.1f * float(k + 1), 0.05f * float(k + 1));
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < 8; ++k) x[k] = ffma2(x[k], a, b);
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += x[k].x + x[k].y;
    if (s == 123.456f) sink[0] = s;
}

// FADD2 alone, and the FMA-pipe instruction mix of the graph kernel's gradient loop (per coordinate pair: one FADD2
// difference, one FFMA2 square-accumulate, one FFMA2 with a scalar weight broadcast over the pair), all operands in
// registers: what the FMA pipe sustains on exactly these instruction forms.
__global__ void __launch_bounds__(512) probe_fadd2_kernel(float* sink, int iters) {
    float2 x[8];
    const float2 b = splat2(1e-3f + 1e-6f * float(threadIdx.x & 31));
#pragma unroll
    for (int k = 0; k < 8; ++k) x[k] = make_float2(0.1f * float(k + 1), 0.05f * float(k + 1));
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < 8; ++k) x[k] = fadd2(x[k], b);
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += x[k].x + x[k].y;
    if (s == 123.456f) sink[0] = s;
}
__global__ void __launch_bounds__(512) probe_mix_kernel(float* sink, int iters) {
    float2 p[4], acc[4], e[4];
    float w = 0.5f + 1e-3f * float(threadIdx.x & 31);
#pragma unroll
    for (int k = 0; k < 4; ++k) { p[k] = make_float2(0.1f * float(k + 1), 0.07f * float(k + 1)); acc[k] = e[k] = make_float2(0.f, 0.f); }
    float2 t = make_float2(0.3f, 0.2f);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float2 d = fadd2(t, p[k]);                 // difference
            e[k] = ffma2(d, d, e[k]);                        // squared distance
            acc[k] = ffma2(splat2(w), d, acc[k]);            // weighted accumulation
        }
        t.x = t.x * 0.999f; w = w * 0.9999f;                 // keeps the loop from being hoisted (2 scalar ops per 12 packed)
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 4; ++k) s += acc[k].x + acc[k].y + e[k].x + e[k].y;
    if (s == 123.456f) sink[0] = s;
}

// ---------------------------------------------------------------------------------------------
// The builds of the pairwise instance kernel (pairwise_inst.cu) and the rule that picks one PER PAIR, from the pair's
// row count alone - so a pair's score is the same whatever else shares its batch, and the same through every entry
// point.  Short pairs get smaller CTAs and more of them per SM: a 100-row pair has 10 tile-pair units and 25 busy
// generator threads, which 16 warps cannot share but eight 4-warp CTAs per SM can.
#define DECL_PW(TAG)                                                                                         \
    cudaError_t cgnn_pw_launch_##TAG(int, const PairwiseArgs&, int, size_t, cudaStream_t);                   \
    int cgnn_pw_occupancy_##TAG(int, size_t);                                                                \
    size_t cgnn_pw_smem_##TAG(int, int, int);
DECL_PW(128) DECL_PW(256) DECL_PW(512)
#undef DECL_PW

struct PwBuild {
    int threads, max_rows;
    cudaError_t (*launch)(int, const PairwiseArgs&, int, size_t, cudaStream_t);
    int (*occupancy)(int, size_t);
    size_t (*smem)(int, int, int);
};
static PwBuild g_pw_builds[3] = {
    {128, CGNN_PW_ROWS_128, cgnn_pw_launch_128, cgnn_pw_occupancy_128, cgnn_pw_smem_128},
    {256, CGNN_PW_ROWS_256, cgnn_pw_launch_256, cgnn_pw_occupancy_256, cgnn_pw_smem_256},
    {512, PW_MAX_ROWS, cgnn_pw_launch_512, cgnn_pw_occupancy_512, cgnn_pw_smem_512},
};
static int pw_build_of(int n_rows) {
    static bool init = false;
    if (!init) {          // tuning hook (tools/bench_rows.py): CGNN_PW_SPLIT="a,b" moves the two thresholds
        init = true;
        if (const char* e = getenv("CGNN_PW_SPLIT")) {
            int a = 0, b = 0;
            if (sscanf(e, "%d,%d", &a, &b) == 2 && a >= 0 && b >= a) { g_pw_builds[0].max_rows = a; g_pw_builds[1].max_rows = b; }
        }
    }
    for (int k = 0; k < 2; ++k)
        if (n_rows <= g_pw_builds[k].max_rows) return k;
    return 2;
}

// ---------------------------------------------------------------------------------------------
extern "C" {

int cgnn_b200_version(void) { return CGNN_B200_VERSION; }

int cgnn_b200_create(cgnn_ctx** out, int device_id) {
    if (!out) return CGNN_ERR_ARG;
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) { cudaGetLastError(); return CGNN_ERR_NODEVICE; }
    if (device_id < 0 || device_id >= n) return CGNN_ERR_ARG;
    cgnn_ctx* ctx = new cgnn_ctx();
    ctx->device = device_id;
    if (cudaSetDevice(device_id) != cudaSuccess || cudaGetDeviceProperties(&ctx->prop, device_id) != cudaSuccess) {
        cudaGetLastError();
        delete ctx;
        return CGNN_ERR_CUDA;
    }
    if (ctx->prop.major != 10) {   // the library carries sm_100a code only: fail loudly, never fall back
        delete ctx;
        return CGNN_ERR_NODEVICE;
    }
    if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreate(&ctx->ev0) != cudaSuccess || cudaEventCreate(&ctx->ev1) != cudaSuccess) {
        cudaGetLastError();
        delete ctx;
        return CGNN_ERR_CUDA;
    }
    *out = ctx;
    return CGNN_OK;
}

void cgnn_b200_destroy(cgnn_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    DevBuf* bufs[] = {&ctx->xy, &ctx->nrows, &ctx->inst, &ctx->scores, &ctx->noise, &ctx->init, &ctx->trace,
                      &ctx->pout, &ctx->counter, &ctx->dbg[0], &ctx->dbg[1], &ctx->dbg[2], &ctx->dbg[3], &ctx->dbg[4],
                      &ctx->gT, &ctx->gdesc, &ctx->gtidx, &ctx->gsub, &ctx->gdata, &ctx->guid, &ctx->wz};
    for (DevBuf* b : bufs)
        if (b->p) cudaFree(b->p);
    if (ctx->ev0) cudaEventDestroy(ctx->ev0);
    if (ctx->ev1) cudaEventDestroy(ctx->ev1);
    if (ctx->fork) cudaEventDestroy(ctx->fork);
    for (int k = 0; k < 2; ++k) {
        if (ctx->join[k]) cudaEventDestroy(ctx->join[k]);
        if (ctx->side[k]) cudaStreamDestroy(ctx->side[k]);
    }
    if (ctx->own_stream && ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

const char* cgnn_b200_last_error(const cgnn_ctx* ctx) { return ctx ? ctx->err.c_str() : "null context"; }

int cgnn_b200_set_stream(cgnn_ctx* ctx, void* stream) {
    if (!ctx) return CGNN_ERR_ARG;
    if (ctx->own_stream && ctx->stream) cudaStreamDestroy(ctx->stream);
    ctx->stream = static_cast<cudaStream_t>(stream);
    ctx->own_stream = false;
    return CGNN_OK;
}

float cgnn_b200_last_kernel_ms(const cgnn_ctx* ctx) { return ctx ? ctx->last_ms : -1.f; }
int cgnn_b200_last_launch_count(const cgnn_ctx* ctx) { return ctx ? ctx->last_launches : 0; }

int cgnn_b200_last_graph_cluster(const cgnn_ctx* ctx) { return ctx ? ctx->last_cluster : 0; }

int cgnn_b200_device_info(const cgnn_ctx* ctx, int* n_sm, int* ctas_per_sm_pairwise) {
    if (!ctx) return CGNN_ERR_ARG;
    if (n_sm) *n_sm = ctx->prop.multiProcessorCount;
    if (ctas_per_sm_pairwise) *ctas_per_sm_pairwise = ctx->pw_ctas_per_sm;
    return CGNN_OK;
}

// ---------------------------------------------------------------------------------------------
static int run_pairwise_impl(cgnn_ctx* ctx, const cgnn_hparams* hp, const float* xy, const int* n_rows, int n_pairs,
                             int ld_rows, int n_runs, unsigned long long uid_base, const float* noise,
                             const float* init_params, float* scores, float* loss_trace, float* params_out,
                             float* const dbg_host[5], int dbg_steps, const float* fourier_wz = nullptr) {
    if (!ctx) return CGNN_ERR_ARG;
    ctx->err.clear();
    int rc = check_hp(ctx, hp);
    if (rc) return rc;
    if (!xy || !n_rows) return fail(ctx, CGNN_ERR_ARG, "xy / n_rows is NULL");
    if (n_pairs < 1 || n_runs < 1 || ld_rows < 1) return fail(ctx, CGNN_ERR_ARG, "n_pairs, n_runs, ld_rows must be >= 1");
    if (!scores && !dbg_host) return fail(ctx, CGNN_ERR_ARG, "scores is NULL");
    CU(cudaSetDevice(ctx->device));

    // n_rows on the host (tiny)
    std::vector<int> nr(n_pairs);
    if (is_device_ptr(n_rows)) CU(cudaMemcpy(nr.data(), n_rows, sizeof(int) * n_pairs, cudaMemcpyDeviceToHost));
    else std::memcpy(nr.data(), n_rows, sizeof(int) * n_pairs);
    int max_n = 0;
    for (int p = 0; p < n_pairs; ++p) {
        if (nr[p] < 1 || nr[p] > ld_rows) return fail(ctx, CGNN_ERR_ARG, "n_rows[%d]=%d outside [1, ld_rows=%d]", p, nr[p], ld_rows);
        if (nr[p] > CGNN_MAX_ROWS_PAIRWISE || nr[p] > PW_MAX_ROWS)
            return fail(ctx, CGNN_ERR_LIMIT, "n_rows[%d]=%d exceeds CGNN_MAX_ROWS_PAIRWISE=%d (subsample first, GNN.py:150-153)",
                        p, nr[p], CGNN_MAX_ROWS_PAIRWISE);
        max_n = std::max(max_n, nr[p]);
    }
    const long long n_inst_ll = 2LL * n_pairs * n_runs;
    if (n_inst_ll > (1LL << 30)) return fail(ctx, CGNN_ERR_LIMIT, "too many instances in one batch");
    const int n_inst = dbg_host ? 1 : int(n_inst_ll);   // the parity probe runs (pair 0, run 0, direction 0) only
    const int h = hp->h_layer_dim, P = 4 * h + 1;
    const int epochs = hp->train_epochs + hp->test_epochs;

    // instance table: grouped by kernel build (a function of the pair's row count only), inside a group longest first
    // (cost ~ N^2) so that the tail of each persistent queue is short
    std::vector<int> order(n_pairs);
    std::iota(order.begin(), order.end(), 0);
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return nr[a] > nr[b]; });
    std::vector<int4> tab;
    tab.reserve(n_inst);
    int grp_first[4] = {0, 0, 0, 0}, grp_rows[3] = {0, 0, 0};
    for (int b = 2; b >= 0; --b) {                       // 512-thread group first: the longest instances start first
        for (int p : order) {
            if (pw_build_of(nr[p]) != b) continue;
            if (dbg_host && p != 0) continue;            // the parity probe runs (pair 0, run 0, direction 0) only
            grp_rows[b] = std::max(grp_rows[b], nr[p]);
            for (int r = 0; r < (dbg_host ? 1 : n_runs); ++r)
                for (int d = 0; d < (dbg_host ? 1 : 2); ++d) tab.push_back(make_int4(p, r, d, (p * n_runs + r) * 2 + d));
        }
        grp_first[2 - b + 1] = int(tab.size());
    }
    // grp_first[k] .. grp_first[k+1]: instances of build 2-k

    PairwiseArgs a{};
    const void* dp = nullptr;
    void* op = nullptr;
    if ((rc = stage_in(ctx, ctx->xy, xy, sizeof(float) * 2 * size_t(n_pairs) * ld_rows, &dp))) return rc;
    a.xy = static_cast<const float*>(dp);
    if ((rc = ensure(ctx, ctx->nrows, sizeof(int) * n_pairs))) return rc;
    CU(cudaMemcpyAsync(ctx->nrows.p, nr.data(), sizeof(int) * n_pairs, cudaMemcpyHostToDevice, ctx->stream));
    a.n_rows = static_cast<const int*>(ctx->nrows.p);
    if ((rc = ensure(ctx, ctx->inst, sizeof(int4) * n_inst))) return rc;
    CU(cudaMemcpyAsync(ctx->inst.p, tab.data(), sizeof(int4) * n_inst, cudaMemcpyHostToDevice, ctx->stream));
    a.inst = static_cast<const int4*>(ctx->inst.p);
    if ((rc = stage_in(ctx, ctx->noise, noise, sizeof(float) * size_t(n_inst) * epochs * ld_rows, &dp))) return rc;
    a.noise = static_cast<const float*>(dp);
    if ((rc = stage_in(ctx, ctx->init, init_params, sizeof(float) * size_t(n_inst) * P, &dp))) return rc;
    a.init_params = static_cast<const float*>(dp);
    if ((rc = stage_out(ctx, ctx->scores, scores, sizeof(float) * n_inst, &op))) return rc;
    a.scores = static_cast<float*>(op);
    if ((rc = stage_out(ctx, ctx->trace, loss_trace, sizeof(float) * size_t(n_inst) * epochs, &op))) return rc;
    a.trace = static_cast<float*>(op);
    if ((rc = stage_out(ctx, ctx->pout, params_out, sizeof(float) * size_t(n_inst) * P, &op))) return rc;
    a.params_out = static_cast<float*>(op);
    if ((rc = ensure(ctx, ctx->counter, 256))) return rc;
    a.counter = static_cast<int*>(ctx->counter.p);
    CU(cudaMemsetAsync(a.counter, 0, 256, ctx->stream));

    size_t dbg_bytes[5] = {0, 0, 0, 0, 0};
    float* dbg_dev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    if (dbg_host) {
        const size_t N0 = size_t(nr[0]);
        dbg_bytes[0] = sizeof(float) * dbg_steps;
        dbg_bytes[1] = dbg_bytes[2] = sizeof(float) * dbg_steps * N0;
        dbg_bytes[3] = dbg_bytes[4] = sizeof(float) * dbg_steps * size_t(P);
        for (int k = 0; k < 5; ++k) {
            if ((rc = stage_out(ctx, ctx->dbg[k], dbg_host[k], dbg_bytes[k], &op))) return rc;
            dbg_dev[k] = static_cast<float*>(op);
        }
        a.dbg_loss = dbg_dev[0]; a.dbg_yhat = dbg_dev[1]; a.dbg_dldy = dbg_dev[2];
        a.dbg_grad = dbg_dev[3]; a.dbg_params = dbg_dev[4];
    }

    a.fast_k = hp->fast_mmd_vectors;
    if (fourier_wz) {
        if (a.fast_k < 1) return fail(ctx, CGNN_ERR_ARG, "fourier_wz given but hparams.fast_mmd_vectors is 0");
        if ((rc = stage_in(ctx, ctx->wz, fourier_wz, sizeof(float) * size_t(n_inst) * epochs * 3 * 7 * a.fast_k, &dp))) return rc;
        a.wz_ext = static_cast<const float*>(dp);
    }
    a.ld_rows = ld_rows;
    a.h = h; a.train_epochs = hp->train_epochs; a.test_epochs = hp->test_epochs; a.trace_every = hp->trace_every;
    a.lr = hp->learning_rate; a.init_std = hp->init_std;
    a.key = make_uint2((unsigned int)(hp->seed & 0xffffffffu), (unsigned int)(hp->seed >> 32));
    a.uid_base = uid_base;
    const int pol = hp->exp_policy < 0 ? CGNN_EXP_POLICY_PAIRWISE : hp->exp_policy;

    // one launch per kernel build that has instances, back to back on the stream
    struct Launch { int build, first, count, grid; size_t smem; int np_max; };
    std::vector<Launch> launches;
    for (int k = 0; k < 3; ++k) {
        const int b = 2 - k, first = grp_first[k], count = grp_first[k + 1] - grp_first[k];
        if (count == 0) continue;
        const int np_max = 32 * ((grp_rows[b] + 31) / 32);
        const size_t smem = g_pw_builds[b].smem(np_max, h, a.fast_k);
        if (smem > size_t(ctx->prop.sharedMemPerBlockOptin))
            return fail(ctx, CGNN_ERR_LIMIT, "instance needs %zu B of shared memory, device allows %zu", smem,
                        size_t(ctx->prop.sharedMemPerBlockOptin));
        const int per_sm = g_pw_builds[b].occupancy(pol, smem);
        if (per_sm < 1) return fail(ctx, CGNN_ERR_CUDA, "pairwise kernel cannot be resident (smem %zu B)", smem);
        if (b == 2 || launches.empty()) ctx->pw_ctas_per_sm = per_sm;
        launches.push_back(Launch{b, first, count, std::min(count, per_sm * ctx->prop.multiProcessorCount), smem, np_max});
    }

    // A ragged batch has up to three launches (one per kernel build).  They are independent, so the second and third go
    // to side streams forked from the caller's stream and joined back into it: the tail of one launch no longer
    // holds the next one back.
    if (launches.size() > 1 && !ctx->fork) {
        CU(cudaEventCreateWithFlags(&ctx->fork, cudaEventDisableTiming));
        for (int k = 0; k < 2; ++k) {
            CU(cudaStreamCreateWithFlags(&ctx->side[k], cudaStreamNonBlocking));
            CU(cudaEventCreateWithFlags(&ctx->join[k], cudaEventDisableTiming));
        }
    }
    CU(cudaEventRecord(ctx->ev0, ctx->stream));
    if (launches.size() > 1) CU(cudaEventRecord(ctx->fork, ctx->stream));
    for (size_t k = 0; k < launches.size(); ++k) {
        const Launch& L = launches[k];
        PairwiseArgs ak = a;
        ak.inst = a.inst + L.first; ak.n_inst = L.count; ak.np_max = L.np_max; ak.counter = a.counter + 16 * int(k);
        cudaStream_t st = k == 0 ? ctx->stream : ctx->side[k - 1];
        if (k > 0) CU(cudaStreamWaitEvent(st, ctx->fork, 0));
        const cudaError_t le = g_pw_builds[L.build].launch(pol, ak, L.grid, L.smem, st);
        if (le != cudaSuccess) return fail(ctx, CGNN_ERR_CUDA, "pairwise kernel launch: %s", cudaGetErrorString(le));
        if (k > 0) {
            CU(cudaEventRecord(ctx->join[k - 1], st));
            CU(cudaStreamWaitEvent(ctx->stream, ctx->join[k - 1], 0));
        }
    }
    CU(cudaEventRecord(ctx->ev1, ctx->stream));
    ctx->last_launches = int(launches.size());

    if ((rc = copy_back(ctx, scores, a.scores, sizeof(float) * n_inst))) return rc;
    if ((rc = copy_back(ctx, loss_trace, a.trace, sizeof(float) * size_t(n_inst) * epochs))) return rc;
    if ((rc = copy_back(ctx, params_out, a.params_out, sizeof(float) * size_t(n_inst) * P))) return rc;
    if (dbg_host)
        for (int k = 0; k < 5; ++k)
            if ((rc = copy_back(ctx, dbg_host[k], dbg_dev[k], dbg_bytes[k]))) return rc;
    CU(cudaStreamSynchronize(ctx->stream));
    CU(cudaEventElapsedTime(&ctx->last_ms, ctx->ev0, ctx->ev1));
    return CGNN_OK;
}

int cgnn_b200_run_pairwise(cgnn_ctx* ctx, const cgnn_hparams* hp, const float* xy, const int* n_rows, int n_pairs,
                           int ld_rows, int n_runs, unsigned long long uid_base, const float* noise,
                           const float* init_params, float* scores, float* loss_trace, float* params_out) {
    return run_pairwise_impl(ctx, hp, xy, n_rows, n_pairs, ld_rows, n_runs, uid_base, noise, init_params, scores,
                             loss_trace, params_out, nullptr, 0);
}

int cgnn_b200_pairwise_step_debug(cgnn_ctx* ctx, const cgnn_hparams* hp, const float* xy, int n_rows, int n_steps,
                                  const float* noise, const float* init_params, const float* fourier_wz, float* out_loss,
                                  float* out_yhat, float* out_dldy, float* out_grad, float* out_params) {
    if (!ctx) return CGNN_ERR_ARG;
    if (!hp) return fail(ctx, CGNN_ERR_ARG, "hparams is NULL");
    if (n_steps < 1) return fail(ctx, CGNN_ERR_ARG, "n_steps must be >= 1");
    cgnn_hparams h2 = *hp;
    h2.train_epochs = n_steps;
    h2.test_epochs = 0;
    h2.trace_every = 0;
    float* dbg[5] = {out_loss, out_yhat, out_dldy, out_grad, out_params};
    return run_pairwise_impl(ctx, &h2, xy, &n_rows, 1, n_rows, 1, 0ULL, noise, init_params, nullptr, nullptr, nullptr, dbg,
                             n_steps, fourier_wz);
}


int cgnn_b200_probe_peak(cgnn_ctx* ctx, int kind, int reps, double* out_per_second) {
    if (!ctx || !out_per_second) return CGNN_ERR_ARG;
    if (kind < 0 || kind > 4) return fail(ctx, CGNN_ERR_ARG, "probe kind %d unknown", kind);
    CU(cudaSetDevice(ctx->device));
    int rc = ensure(ctx, ctx->counter, 256);
    if (rc) return rc;
    const int iters = 4096, grid = ctx->prop.multiProcessorCount * 4, nt = 512;
    double best = 0.0;
    for (int r = 0; r < std::max(reps, 1) + 1; ++r) {      // first launch is a warm-up
        CU(cudaEventRecord(ctx->ev0, ctx->stream));
        if (kind == 0)      probe_mufu_kernel<<<grid, nt, 0, ctx->stream>>>(static_cast<float*>(ctx->counter.p), iters);
        else if (kind == 1) probe_ffma_kernel<<<grid, nt, 0, ctx->stream>>>(static_cast<float*>(ctx->counter.p), iters);
        else if (kind == 2) probe_ffma2_kernel<<<grid, nt, 0, ctx->stream>>>(static_cast<float*>(ctx->counter.p), iters);
        else if (kind == 3) probe_fadd2_kernel<<<grid, nt, 0, ctx->stream>>>(static_cast<float*>(ctx->counter.p), iters);
        else                probe_mix_kernel<<<grid, nt, 0, ctx->stream>>>(static_cast<float*>(ctx->counter.p), iters);
        CU(cudaGetLastError());
        CU(cudaEventRecord(ctx->ev1, ctx->stream));
        CU(cudaStreamSynchronize(ctx->stream));
        float ms = 0.f;
        CU(cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
        // kinds 3 and 4 report packed INSTRUCTIONS x 2 lanes per second ("lane-slots": an FFMA2 = 2), i.e. half the flop/s of kind 2
        const double ops = double(grid) * nt * double(iters) * (kind == 0 ? 8.0 : kind == 1 ? 16.0 : kind == 2 ? 32.0 : kind == 3 ? 16.0 : 24.0);
        if (r > 0) best = std::max(best, ops / (double(ms) * 1e-3));
    }
    *out_per_second = best;
    return CGNN_OK;
}

}  // extern "C"

// =============================================================================================
// Entry point for callers that cannot allocate device memory (see include/cgnn_b200.h).
//
// Everything the pairwise launch needs beyond the caller's own input and output buffers - the
// work-queue head, the per-pair row counts and the instance table - lives in ONE statically
// allocated device array.  It is part of the library image: every GPU that loads the library
// carries these bytes until the process exits, whether or not this entry point is ever called.
// =============================================================================================
#define CGNN_NOALLOC_MAX_PAIRS      1024
#define CGNN_NOALLOC_MAX_INSTANCES  8192

struct NoAllocScratch {
    int  counter[64];                            // work-queue head (+ padding to 256 B); must be 0 at launch
    int  n_rows[CGNN_NOALLOC_MAX_PAIRS];
    int4 inst[CGNN_NOALLOC_MAX_INSTANCES];       // (pair, run, direction, output index)
};
static_assert(sizeof(NoAllocScratch) == 256 + 4 * CGNN_NOALLOC_MAX_PAIRS + 16 * CGNN_NOALLOC_MAX_INSTANCES, "layout");

static __device__ __align__(256) unsigned char g_noalloc_scratch[sizeof(NoAllocScratch)];   // 135,424 bytes

// Fills the scratch for one call.  Device code may name the array directly.
__global__ void noalloc_setup_kernel(int n_pairs, int n_rows, int n_runs) {
    NoAllocScratch* s = reinterpret_cast<NoAllocScratch*>(g_noalloc_scratch);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < 64) s->counter[i] = 0;               // the array keeps the previous call's contents: re-zero the queue head
    if (i < n_pairs) s->n_rows[i] = n_rows;
    if (i < 2 * n_pairs * n_runs) {
        const int d = i & 1, pr = i >> 1;
        s->inst[i] = make_int4(pr / n_runs, pr % n_runs, d, i);
    }
}

// Device address of the array, per device.  In host code the name `g_noalloc_scratch` is only a
// handle for the runtime, not a device address, so it is resolved once per device (with that
// device current) and the resulting pointer is what the kernels receive.
static void* g_noalloc_addr[64] = {nullptr};

static int na_fail(char* buf, int len, int code, const char* fmt, ...) {
    if (buf && len > 0) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(buf, size_t(len), fmt, ap);
        va_end(ap);
    }
    return code;
}

#define NA_CU(call)                                                                                      \
    do {                                                                                                 \
        cudaError_t e_ = (call);                                                                         \
        if (e_ != cudaSuccess) return na_fail(errbuf, errbuf_len, CGNN_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(e_)); \
    } while (0)

static int noalloc_impl(int device, void* stream, const cgnn_hparams* hp, const float* d_xy, int n_pairs, int n_rows,
                        int n_runs, unsigned long long uid_base, float* d_scores, char* errbuf, int errbuf_len);

extern "C" int cgnn_b200_run_pairwise_noalloc(int device, void* stream, const cgnn_hparams* hp, const float* d_xy,
                                               int n_pairs, int n_rows, int n_runs, unsigned long long uid_base,
                                               float* d_scores, char* errbuf, int errbuf_len) {
    // the call makes `device` current for its own work and hands the thread back with the device it came with
    int prev = -1;
    const bool have_prev = cudaGetDevice(&prev) == cudaSuccess;
    if (!have_prev) cudaGetLastError();
    const int rc = noalloc_impl(device, stream, hp, d_xy, n_pairs, n_rows, n_runs, uid_base, d_scores, errbuf, errbuf_len);
    if (have_prev && prev != device) cudaSetDevice(prev);
    return rc;
}

static int noalloc_impl(int device, void* stream, const cgnn_hparams* hp, const float* d_xy, int n_pairs, int n_rows,
                        int n_runs, unsigned long long uid_base, float* d_scores, char* errbuf, int errbuf_len) {
    if (errbuf && errbuf_len > 0) errbuf[0] = 0;
    if (!hp || !d_xy || !d_scores) return na_fail(errbuf, errbuf_len, CGNN_ERR_ARG, "hparams, d_xy and d_scores must not be NULL");
    if (hp->h_layer_dim < 1 || hp->h_layer_dim > CGNN_MAX_H)
        return na_fail(errbuf, errbuf_len, CGNN_ERR_LIMIT, "h_layer_dim=%d outside [1,%d]", hp->h_layer_dim, CGNN_MAX_H);
    if (hp->train_epochs < 0 || hp->test_epochs < 0 || hp->exp_policy < CGNN_EXP_POLICY_AUTO || hp->exp_policy > 3)
        return na_fail(errbuf, errbuf_len, CGNN_ERR_ARG, "bad epoch count or exp_policy");
    if (hp->fast_mmd_vectors < 0 || hp->fast_mmd_vectors > CGNN_MAX_FAST_VECTORS)
        return na_fail(errbuf, errbuf_len, CGNN_ERR_LIMIT, "fast_mmd_vectors=%d outside [0,%d]", hp->fast_mmd_vectors, CGNN_MAX_FAST_VECTORS);
    if (n_pairs < 1 || n_runs < 1 || n_rows < 1) return na_fail(errbuf, errbuf_len, CGNN_ERR_ARG, "n_pairs, n_runs, n_rows must be >= 1");
    if (n_rows > CGNN_MAX_ROWS_PAIRWISE)
        return na_fail(errbuf, errbuf_len, CGNN_ERR_LIMIT, "n_rows=%d exceeds CGNN_MAX_ROWS_PAIRWISE=%d", n_rows, CGNN_MAX_ROWS_PAIRWISE);
    if (n_pairs > CGNN_NOALLOC_MAX_PAIRS || 2LL * n_pairs * n_runs > CGNN_NOALLOC_MAX_INSTANCES)
        return na_fail(errbuf, errbuf_len, CGNN_ERR_LIMIT,
                       "batch of %d pairs x %d runs x 2 directions exceeds the static scratch (max %d pairs, %d instances): "
                       "split the batch or use cgnn_b200_run_pairwise", n_pairs, n_runs, CGNN_NOALLOC_MAX_PAIRS,
                       CGNN_NOALLOC_MAX_INSTANCES);
    if (device < 0 || device >= 64) return na_fail(errbuf, errbuf_len, CGNN_ERR_ARG, "device %d outside [0,64)", device);
    NA_CU(cudaSetDevice(device));
    if (!is_device_ptr(d_xy) || !is_device_ptr(d_scores))
        return na_fail(errbuf, errbuf_len, CGNN_ERR_ARG, "d_xy and d_scores must be device memory (staging host buffers would need an allocation)");
    int cc_major = 0, n_sm = 0, smem_optin = 0;
    NA_CU(cudaDeviceGetAttribute(&cc_major, cudaDevAttrComputeCapabilityMajor, device));
    NA_CU(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, device));
    NA_CU(cudaDeviceGetAttribute(&smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device));
    if (cc_major != 10) return na_fail(errbuf, errbuf_len, CGNN_ERR_NODEVICE, "device %d is not sm_100", device);

    // one-time set-up on this device: resolve the device address of the static array
    if (!g_noalloc_addr[device]) {
        void* p = nullptr;
        NA_CU(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        g_noalloc_addr[device] = p;
    }
    NoAllocScratch* sc = static_cast<NoAllocScratch*>(g_noalloc_addr[device]);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int n_inst = 2 * n_pairs * n_runs;

    noalloc_setup_kernel<<<(std::max(n_inst, 64) + 255) / 256, 256, 0, st>>>(n_pairs, n_rows, n_runs);
    NA_CU(cudaGetLastError());

    PairwiseArgs a{};
    a.xy = d_xy; a.n_rows = sc->n_rows; a.inst = sc->inst; a.counter = sc->counter;
    a.n_inst = n_inst; a.ld_rows = n_rows; a.np_max = 32 * ((n_rows + 31) / 32);
    a.h = hp->h_layer_dim; a.train_epochs = hp->train_epochs; a.test_epochs = hp->test_epochs; a.trace_every = 0;
    a.lr = hp->learning_rate; a.init_std = hp->init_std;
    a.key = make_uint2((unsigned int)(hp->seed & 0xffffffffu), (unsigned int)(hp->seed >> 32));
    a.uid_base = uid_base;
    a.scores = d_scores;
    a.fast_k = hp->fast_mmd_vectors;

    const PwBuild& B = g_pw_builds[pw_build_of(n_rows)];     // the same per-pair rule as cgnn_b200_run_pairwise: identical scores
    const size_t smem = B.smem(a.np_max, a.h, a.fast_k);
    if (smem > size_t(smem_optin)) return na_fail(errbuf, errbuf_len, CGNN_ERR_LIMIT, "instance needs %zu B of shared memory", smem);
    const int pol = hp->exp_policy < 0 ? CGNN_EXP_POLICY_PAIRWISE : hp->exp_policy;
    const int per_sm = B.occupancy(pol, smem);
    if (per_sm < 1) return na_fail(errbuf, errbuf_len, CGNN_ERR_CUDA, "pairwise kernel cannot be resident");
    const int grid = std::min(n_inst, per_sm * n_sm);
    const cudaError_t le = B.launch(pol, a, grid, smem, st);
    if (le != cudaSuccess) return na_fail(errbuf, errbuf_len, CGNN_ERR_CUDA, "pairwise kernel launch: %s", cudaGetErrorString(le));
    NA_CU(cudaStreamSynchronize(st));
    return CGNN_OK;
}
