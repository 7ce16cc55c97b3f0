// Pairwise GNN instance kernel (sm_100a): one persistent CTA per (pair, run, direction) instance.
//
// Replaces, fused into ONE launch for a whole batch of instances, what the reference does with
// ~150 TensorFlow launches and a host round trip per step:
//   GNN_tf.__init__  Code/cgnn/GNN.py:33-80    init, noise, 2->h->1 ReLU generator, MMD, Adam
//   GNN_tf.train     GNN.py:82-102            train_epochs x [noise, fwd, MMD fwd/bwd, bwd, Adam]
//   GNN_tf.evaluate  GNN.py:104-126           test_epochs x [noise, fwd, MMD]; mean
//   MMD_loss_tf      utils/Loss.py:12-32      block form: (PP + TT - 2 PT)/N^2, TT hoisted (constant),
//                                             PP taken once per unordered pair, direct differences
//
// Data layout (all in shared memory for the instance's whole life; HBM is touched once at
// instance start and once at the end):
//   Tc[NP] float2 = (x_j, y_j)      observed pair, cause first        NP = 32*ceil(N/32)
//   Pc[NP] float2 = (x_j, yhat_j)   generated pair (cause column copied, GNN.py:70-72)
//   es[NP] float  = noise e_j of the current step (needed again by the backward pass)
//   g64[NP] int64 = fixed-point accumulator of dLoss/d yhat_j (unscaled)
//   padded rows hold x = 1e15 so that every kernel value against them is exactly 0.
//
// MMD mapping: lane <-> row i of a 32-row block, columns streamed from shared memory.
//   PT (pred x true)  : broadcast loads of Tc[j]; row accumulator only (T has no gradient)
//   PP off-diagonal   : block pair (I,J) visited once; lane l meets column (l+s)&31 at step s
//                       (conflict-free skewed loads); the column-side gradient rides in a register
//                       that is rotated with one shuffle per evaluation; after 32 steps lane l owns
//                       column l's total, flushed with one fixed-point shared atomic
//   PP diagonal block : plain broadcast form (0.7 % of the evaluations)
// Each evaluation = dx,dy,D (4 FP32) + 7 kernels (MUFU ex2) + weights (8 FP32) + 1-2 FFMA.
#pragma once
#include "common.cuh"

namespace cgnn {

constexpr int PW_NT = 512;              // threads per CTA
constexpr int PW_NW = PW_NT / 32;       // warps per CTA
constexpr int PW_MIN_CTAS = 2;          // persistent CTAs per SM
constexpr int PW_MAX_ROWS = 4 * PW_NT;  // one 4-row quad per thread in the generator phases
constexpr float PW_PAD = 1.0e15f;

struct PairwiseArgs {
    const float* xy;          // [n_pairs][ld_rows][2]
    const int*   n_rows;      // [n_pairs]
    const int4*  inst;        // [n_inst] (pair, run, dir, out_index), sorted by decreasing cost
    int n_inst, ld_rows, np_max;
    int h, train_epochs, test_epochs, trace_every;
    float lr, init_std;
    uint2 key;
    unsigned int uid_base;
    const float* noise;       // NULL or [n_inst][epochs][ld_rows]   (indexed by out_index)
    const float* init_params; // NULL or [n_inst][4h+1]
    float* scores;            // [n_inst] by out_index
    float* trace;             // NULL or [n_inst][train+test]
    float* params_out;        // NULL or [n_inst][4h+1]
    int*   counter;           // work queue head (scratch)
    // parity probe (single instance): any may be NULL
    float *dbg_loss, *dbg_yhat, *dbg_dldy, *dbg_grad, *dbg_params;
};

struct PwSmem {
    float2* Tc; float2* Pc; float* es; long long* g64;
    float* theta; float* am; float* av; float* part; double* red;
    int part_stride;
};

__host__ __device__ inline size_t pw_smem_bytes(int np_max, int h) {
    int nchunk = (h + 7) / 8;
    int P = 4 * h + 1;
    int Pp = (P + 3) & ~3;
    size_t b = 0;
    b += size_t(np_max) * 8 * 2;                 // Tc, Pc
    b += size_t(np_max) * 8;                     // g64
    b += size_t(np_max) * 4;                     // es
    b += size_t(Pp) * 4 * 3;                     // theta, m, v
    b += size_t(PW_NW) * (nchunk * 32 + 4) * 4;  // per-warp partial gradient sums
    b += size_t(PW_NW + 4) * 8;                  // red[], TT, beta1^t, beta2^t, lr_t
    return (b + 15) & ~size_t(15);
}

// ---- unit partition -------------------------------------------------------------------------
// Row block I owns nPT broadcast tiles (PT), one diagonal tile and nPP(I) off-diagonal PP tiles
// J = (I+1+q) mod NB.  For odd NB every block has (NB-1)/2 of them; for even NB the antipodal
// pair (I, I+NB/2) is given to the lower I.  Units are flattened I-major and dealt to the warps
// in equal contiguous ranges, so a warp changes row block at most a few times per pass.
__device__ __forceinline__ int pw_npp(int I, int NB) {
    return (NB & 1) ? (NB - 1) / 2 : (NB / 2 - 1 + (I < NB / 2 ? 1 : 0));
}

// Broadcast-column tile, gradient form: returns sum_j w(i,j) * (yi - col[j].y).
template <int POL>
__device__ __forceinline__ float bc_train(float xi, float yi, const float2* __restrict__ col, int j0, int j1) {
    float acc = 0.f;
#pragma unroll 4
    for (int j = j0; j < j1; ++j) {
        const float2 c = col[j];
        const float dx = xi - c.x, dy = yi - c.y;
        const float D = fmaf(dy, dy, dx * dx);
        acc = fmaf(wsum<POL>(D), dy, acc);
    }
    return acc;
}

// Broadcast-column tile, loss form: returns sum_j sum_g k_g(i,j).
template <int POL>
__device__ __forceinline__ float bc_eval(float xi, float yi, const float2* __restrict__ col, int j0, int j1) {
    float acc = 0.f;
#pragma unroll 4
    for (int j = j0; j < j1; ++j) {
        const float2 c = col[j];
        const float dx = xi - c.x, dy = yi - c.y;
        const float D = fmaf(dy, dy, dx * dx);
        acc += ksum<POL>(D);
    }
    return acc;
}

// Off-diagonal PP tile, gradient form, each unordered pair once.  Returns the row-side sum
// sum_j w (yi - yj); colsum receives, for column J0+lane, sum_i w (yi - yj) over the 32 rows.
template <int POL>
__device__ __forceinline__ float pp_sym_train(float xi, float yi, const float2* __restrict__ Pc, int J0, int lane,
                                              float& colsum) {
    float ra = 0.f, ca = 0.f;
    const int src = (lane + 1) & 31;
#pragma unroll 8
    for (int s = 0; s < 32; ++s) {
        const float2 c = Pc[J0 + ((lane + s) & 31)];
        const float dx = xi - c.x, dy = yi - c.y;
        const float D = fmaf(dy, dy, dx * dx);
        const float t = wsum<POL>(D) * dy;
        ra += t;
        ca += t;
        ca = __shfl_sync(0xffffffffu, ca, src);
    }
    colsum = ca;
    return ra;
}

// One MMD pass over an instance.
//   TRAIN: accumulates the unscaled dLoss/d yhat into g64 (rows: +PT - PP; columns of PP tiles: +)
//   else : returns this lane's share of  PP - 2 PT  (or of the full PP when ptA == nullptr: used for TT)
template <int POL, bool TRAIN>
__device__ double mmd_pass(const float2* __restrict__ rowA, const float2* __restrict__ ptA,
                           const float2* __restrict__ ppA, long long* g64, int N, int NB, int warp, int lane) {
    const int nPT = ptA ? NB : 0;
    int total = 0;
    for (int I = 0; I < NB; ++I) total += nPT + 1 + pw_npp(I, NB);
    const int lo = int((long long)warp * total / PW_NW);
    const int hi = int((long long)(warp + 1) * total / PW_NW);
    int I = 0, base = 0;
    while (I < NB && base + nPT + 1 + pw_npp(I, NB) <= lo) { base += nPT + 1 + pw_npp(I, NB); ++I; }
    double lsum = 0.0;
    int u = lo;
    while (u < hi) {
        const int UI = nPT + 1 + pw_npp(I, NB);
        const int a0 = u - base;
        const int b0 = min(hi - base, UI);
        const int i = 32 * I + lane;
        const float2 me = rowA[i];
        const float xi = me.x, yi = me.y;
        double acc = 0.0;
        // PT tiles a0 .. min(b0, nPT)
        if (a0 < nPT) {
            const int j1 = min(32 * min(b0, nPT), N);
            for (int jc = 32 * a0; jc < j1; jc += 64) {
                const int je = min(jc + 64, j1);
                if (TRAIN) acc += double(bc_train<POL>(xi, yi, ptA, jc, je));
                else       acc -= 2.0 * double(bc_eval<POL>(xi, yi, ptA, jc, je));
            }
        }
        // diagonal tile
        if (a0 <= nPT && b0 > nPT) {
            const int je = min(32 * I + 32, N);
            if (TRAIN) acc -= double(bc_train<POL>(xi, yi, ppA, 32 * I, je));
            else       acc += double(bc_eval<POL>(xi, yi, ppA, 32 * I, je));
        }
        // off-diagonal PP tiles
        for (int q = max(a0, nPT + 1) - (nPT + 1); q < b0 - (nPT + 1); ++q) {
            int J = I + 1 + q;
            if (J >= NB) J -= NB;
            if (TRAIN) {
                float cs;
                acc -= double(pp_sym_train<POL>(xi, yi, ppA, 32 * J, lane, cs));
                if (32 * J + lane < N) fx_add(&g64[32 * J + lane], double(cs));
            } else {
                acc += 2.0 * double(bc_eval<POL>(xi, yi, ppA, 32 * J, min(32 * J + 32, N)));
            }
        }
        if (TRAIN) { if (i < N) fx_add(&g64[i], acc); }
        else       { if (i < N) lsum += acc; }
        u = base + b0;
        base += UI;
        ++I;
    }
    return lsum;
}

// Block-wide sum of per-lane doubles in a fixed order (deterministic). Result valid in thread 0.
__device__ __forceinline__ double block_sum(double v, double* red, int warp, int lane) {
    v = warp_sum(v);
    if (lane == 0) red[warp] = v;
    __syncthreads();
    double t = 0.0;
    if (warp == 0 && lane == 0) {
#pragma unroll
        for (int w = 0; w < PW_NW; ++w) t += red[w];
    }
    return t;
}

// Generator forward for this thread's 4-row quad (GNN.py:64-67), noise from Philox or the caller.
__device__ __forceinline__ void pw_forward(const PwSmem& s, const PairwiseArgs& a, int N, int h, int step,
                                           unsigned int uid, const float* noise_row, int tid) {
    const int r0 = 4 * tid;
    if (r0 < ((N + 31) & ~31)) {   // zero the gradient accumulators of this quad (padded rows too)
        s.g64[r0] = 0; s.g64[r0 + 1] = 0; s.g64[r0 + 2] = 0; s.g64[r0 + 3] = 0;
    }
    if (r0 >= N) return;
    float e[4];
    if (noise_row) {
#pragma unroll
        for (int r = 0; r < 4; ++r) e[r] = (r0 + r < N) ? noise_row[r0 + r] : 0.f;
    } else {
        const float4 n = normal4(philox4x32_10(make_uint4(tid, step, STREAM_NOISE, uid), a.key));
        e[0] = n.x; e[1] = n.y; e[2] = n.z; e[3] = n.w;
    }
    float x[4], y[4];
    const float bout = s.theta[4 * h];
#pragma unroll
    for (int r = 0; r < 4; ++r) { x[r] = s.Tc[r0 + r].x; y[r] = bout; }
    for (int k = 0; k < h; ++k) {
        const float w0 = s.theta[k], w1 = s.theta[h + k], b = s.theta[2 * h + k], wo = s.theta[3 * h + k];
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const float z = fmaf(x[r], w0, fmaf(e[r], w1, b));
            y[r] = fmaf(fmaxf(z, 0.f), wo, y[r]);
        }
    }
#pragma unroll
    for (int r = 0; r < 4; ++r)
        if (r0 + r < N) { s.Pc[r0 + r].y = y[r]; s.es[r0 + r] = e[r]; }
}

// Generator backward: per-warp partial parameter gradients into s.part (deterministic tree).
__device__ __forceinline__ void pw_backward_partials(const PwSmem& s, int N, int h, int tid, int warp, int lane) {
    const int r0 = 4 * tid;
    const float scale = 4.0f / (float(N) * float(N));
    float x[4], e[4], d[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const bool ok = r0 + r < N;
        x[r] = ok ? s.Tc[r0 + r].x : 0.f;
        e[r] = ok ? s.es[r0 + r] : 0.f;
        d[r] = ok ? float(fx_get(&s.g64[r0 + r])) * scale : 0.f;
    }
    const int nchunk = (h + 7) >> 3;
    float* mine = s.part + warp * s.part_stride;
    for (int c = 0; c < nchunk; ++c) {
        float v[32];
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
            const int k = 8 * c + kk;
            float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
            if (k < h) {
                const float w0 = s.theta[k], w1 = s.theta[h + k], b = s.theta[2 * h + k], wo = s.theta[3 * h + k];
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    const float z = fmaf(x[r], w0, fmaf(e[r], w1, b));
                    const float dz = z > 0.f ? d[r] * wo : 0.f;
                    s0 = fmaf(x[r], dz, s0);
                    s1 = fmaf(e[r], dz, s1);
                    s2 += dz;
                    s3 = fmaf(fmaxf(z, 0.f), d[r], s3);
                }
            }
            v[4 * kk] = s0; v[4 * kk + 1] = s1; v[4 * kk + 2] = s2; v[4 * kk + 3] = s3;
        }
        mine[32 * c + lane] = warp_reduce32(v, lane);
    }
    const float sd = warp_sum((d[0] + d[1]) + (d[2] + d[3]));
    if (lane == 0) mine[32 * nchunk] = sd;
}

// slot of parameter p in a per-warp partial row
__device__ __forceinline__ int pw_slot(int p, int h) {
    if (p == 4 * h) return 32 * ((h + 7) >> 3);
    const int a = p / h, k = p - a * h;       // a: 0 W_in[x], 1 W_in[e], 2 b_in, 3 W_out
    return 32 * (k >> 3) + 4 * (k & 7) + a;
}

template <int POL>
__global__ void __launch_bounds__(PW_NT, PW_MIN_CTAS) pairwise_kernel(const PairwiseArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ int s_inst;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int h = a.h, P = 4 * h + 1, Pp = (P + 3) & ~3;
    const int nchunk = (h + 7) >> 3;

    PwSmem s;
    {
        unsigned char* p = smem_raw;
        s.Tc = reinterpret_cast<float2*>(p);      p += size_t(a.np_max) * 8;
        s.Pc = reinterpret_cast<float2*>(p);      p += size_t(a.np_max) * 8;
        s.g64 = reinterpret_cast<long long*>(p);  p += size_t(a.np_max) * 8;
        s.red = reinterpret_cast<double*>(p);     p += size_t(PW_NW + 4) * 8;
        s.es = reinterpret_cast<float*>(p);       p += size_t(a.np_max) * 4;
        s.theta = reinterpret_cast<float*>(p);    p += size_t(Pp) * 4;
        s.am = reinterpret_cast<float*>(p);       p += size_t(Pp) * 4;
        s.av = reinterpret_cast<float*>(p);       p += size_t(Pp) * 4;
        s.part = reinterpret_cast<float*>(p);
        s.part_stride = nchunk * 32 + 4;
    }
    double& s_TT = s.red[PW_NW];

    for (;;) {
        __syncthreads();
        if (tid == 0) s_inst = atomicAdd(a.counter, 1);
        __syncthreads();
        const int it_idx = s_inst;
        if (it_idx >= a.n_inst) break;
        const int4 inst = a.inst[it_idx];
        const int pair = inst.x, dir = inst.z, out = inst.w;
        const unsigned int uid = a.uid_base + (unsigned int)out;
        const int N = a.n_rows[pair];
        const int NB = (N + 31) >> 5, NP = NB * 32;
        const float* src = a.xy + size_t(pair) * a.ld_rows * 2;
        const int epochs = a.train_epochs + a.test_epochs;
        const float* noise = a.noise ? a.noise + size_t(out) * epochs * a.ld_rows : nullptr;

        // ---- load the pair (cause first), pad ---------------------------------------------
        for (int i = tid; i < NP; i += PW_NT) {
            float2 t = make_float2(PW_PAD, 0.f);
            if (i < N) {
                const float2 v = reinterpret_cast<const float2*>(src)[i];
                t = dir ? make_float2(v.y, v.x) : v;
            }
            s.Tc[i] = t;
            s.Pc[i] = make_float2(t.x, 0.f);
            s.es[i] = 0.f;
        }
        // ---- parameters (GNN.py:21-29, 55-58) and Adam state -------------------------------
        if (tid < Pp) {
            float th = 0.f;
            if (tid < P) {
                if (a.init_params) th = a.init_params[size_t(out) * P + tid];
                else {
                    const float4 n = normal4(philox4x32_10(make_uint4(tid >> 2, 0u, STREAM_INIT, uid), a.key));
                    const int e = tid & 3;
                    th = a.init_std * (e == 0 ? n.x : e == 1 ? n.y : e == 2 ? n.z : n.w);
                }
            }
            s.theta[tid] = th; s.am[tid] = 0.f; s.av[tid] = 0.f;
        }
        __syncthreads();
        // ---- TT: constant per instance, fp64 accumulation ----------------------------------
        {
            const double part = mmd_pass<POL, false>(s.Tc, nullptr, s.Tc, nullptr, N, NB, warp, lane);
            const double tt = block_sum(part, s.red, warp, lane);
            if (tid == 0) s_TT = tt;
        }
        // beta^t accumulators.  TF keeps them in fp32, which puts a 6e-6 relative error on every step size
        // ((1 - 0.999f^t) inherits fp32(0.999)'s rounding); thread 0 keeps them in fp64 in shared memory.
        if (tid == 0) { s.red[PW_NW + 1] = 1.0; s.red[PW_NW + 2] = 1.0; }
        const double invN2 = 1.0 / (double(N) * double(N));

        // ---- training (GNN.py:92-96) --------------------------------------------------------
        for (int it = 0; it < a.train_epochs; ++it) {
            __syncthreads();
            if (tid == 0) {   // lr_t of this step (read after three more barriers)
                const double b1p = s.red[PW_NW + 1] * 0.9, b2p = s.red[PW_NW + 2] * 0.999;
                s.red[PW_NW + 1] = b1p; s.red[PW_NW + 2] = b2p;
                s.red[PW_NW + 3] = double(a.lr) * sqrt(1.0 - b2p) / (1.0 - b1p);
            }
            pw_forward(s, a, N, h, it, uid, noise ? noise + size_t(it) * a.ld_rows : nullptr, tid);
            __syncthreads();
            const bool traced = a.dbg_loss || (a.trace && a.trace_every > 0 && it % a.trace_every == 0);
            if (traced) {   // loss fetched with the train op is the pre-update loss (GNN.py:93-94)
                const double part = mmd_pass<POL, false>(s.Pc, s.Tc, s.Pc, nullptr, N, NB, warp, lane);
                const double sum = block_sum(part, s.red, warp, lane);
                if (tid == 0) {
                    const float loss = float((sum + s_TT) * invN2);
                    if (a.trace) a.trace[size_t(out) * epochs + it] = loss;
                    if (a.dbg_loss) a.dbg_loss[it] = loss;
                }
            } else if (a.trace && tid == 0) {
                a.trace[size_t(out) * epochs + it] = __int_as_float(0x7fc00000);
            }
            if (a.dbg_yhat) for (int i = tid; i < N; i += PW_NT) a.dbg_yhat[size_t(it) * N + i] = s.Pc[i].y;
            mmd_pass<POL, true>(s.Pc, s.Tc, s.Pc, s.g64, N, NB, warp, lane);
            __syncthreads();
            if (a.dbg_dldy) {
                const float scale = 4.0f / (float(N) * float(N));
                for (int i = tid; i < N; i += PW_NT) a.dbg_dldy[size_t(it) * N + i] = float(fx_get(&s.g64[i])) * scale;
            }
            pw_backward_partials(s, N, h, tid, warp, lane);
            __syncthreads();
            if (tid < P) {
                const int slot = pw_slot(tid, h);
                float g = 0.f;
#pragma unroll
                for (int w = 0; w < PW_NW; ++w) g += s.part[w * s.part_stride + slot];
                const float lr_t = float(s.red[PW_NW + 3]);
                float m = s.am[tid], v = s.av[tid];
                const float th = adam_update(s.theta[tid], g, m, v, lr_t);
                s.am[tid] = m; s.av[tid] = v; s.theta[tid] = th;
                if (a.dbg_grad) a.dbg_grad[size_t(it) * P + tid] = g;
                if (a.dbg_params) a.dbg_params[size_t(it) * P + tid] = th;
            }
        }
        __syncthreads();
        if (a.params_out && tid < P) a.params_out[size_t(out) * P + tid] = s.theta[tid];

        // ---- evaluation (GNN.py:115-126): mean of test_epochs fresh-noise losses ---------------
        double score_sum = 0.0;
        for (int it = 0; it < a.test_epochs; ++it) {
            __syncthreads();
            pw_forward(s, a, N, h, a.train_epochs + it, uid,
                       noise ? noise + size_t(a.train_epochs + it) * a.ld_rows : nullptr, tid);
            __syncthreads();
            const double part = mmd_pass<POL, false>(s.Pc, s.Tc, s.Pc, nullptr, N, NB, warp, lane);
            const double sum = block_sum(part, s.red, warp, lane);
            if (tid == 0) {
                const float loss = float((sum + s_TT) * invN2);    // the reference fetches an fp32 loss ...
                score_sum += double(loss);                         // ... and sums it in a Python float
                if (a.trace) a.trace[size_t(out) * epochs + a.train_epochs + it] = loss;
            }
        }
        if (tid == 0 && a.scores) a.scores[out] = a.test_epochs > 0 ? float(score_sum / double(a.test_epochs)) : 0.f;
    }
}

}  // namespace cgnn
