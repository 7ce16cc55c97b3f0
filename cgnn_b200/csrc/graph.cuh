// Graph generator instance kernel (sm_100a): one persistent CTA per (candidate graph, run) instance.
//
// Replaces, fused into ONE launch for a whole batch of candidate graphs x runs:
//   CGNN_tf.__init__ / train / evaluate            Code/cgnn/CGNN.py:35-153
//   CGNN_confounders_tf.__init__ / train / evaluate Code/cgnn/CGNN_confounders.py:37-172
//   MMD_loss_tf                                    Code/cgnn/utils/Loss.py:12-32  (block form, direct differences)
// A graph is described by its nodes in the reference's build order (CGNN.py:63-84); node k reads
// n_par generated parent columns, then its own noise column, then (confounders only) one shared
// noise column per skeleton neighbour (CGNN_confounders.py:84-96).
//
// Layout: the generated matrix Pm[NP][DP] (row-major, DP = n_var padded to a multiple of 4, pad
// columns 0) lives in shared memory for the instance's whole life; the observed matrix T[NP][DP]
// is the same for every instance of a batch (or one per row subset) and is streamed through the
// read-only path (L1/L2 resident: <= 192 KB).  Parameters, Adam moments and the fixed-point
// gradient accumulators are in shared memory too.
//
// Mapping: lane <-> row i.  A warp owns whole 32-row blocks (w, w+NW, ...): the row's generated
// values, its MMD gradient accumulators and the distance differences stay in REGISTERS across all
// 2N kernel evaluations of the row, and the row's gradient goes straight into the thread-local
// MLP backward pass - dLoss/dPred is never written to memory.  Column j (DP floats) arrives as
// DP/4 128-bit broadcast loads.  Parameter gradients are reduced across the 32 rows of a block with
// the transposing warp reduction and accumulated across warps with fixed-point shared atomics.
#pragma once
#include "common.cuh"

namespace cgnn {

constexpr int GR_NT = 512;
constexpr int GR_NW = GR_NT / 32;
constexpr int GR_MAX_VAR = 32;
constexpr int GR_MAX_H = 32;
constexpr int GR_MAX_IN = 384;           // sum over nodes of (parents + noise inputs)
constexpr uint32_t GR_STREAM_NOISE0 = 2u;

struct GraphDesc {
    int n_var, n_noise, n_params, n_in_total;
    short col[GR_MAX_VAR], n_par[GR_MAX_VAR], n_in[GR_MAX_VAR];
    int in_off[GR_MAX_VAR], par_off[GR_MAX_VAR];
    short in_idx[GR_MAX_IN];
};

struct GraphArgs {
    const float* T;            // [n_T][NP][DP]
    const int* t_index;        // NULL (all instances use T[0]) or [n_inst] by out index
    const GraphDesc* desc;     // [n_graphs]
    const int2* inst;          // [n_inst] (graph, out index)
    const unsigned long long* inst_uid;   // [n_inst] Philox stream id of each queue entry
    int n_inst, N, NP, n_noise, p_max;
    int h, train_epochs, test_epochs, trace_every;
    float lr, init_std;
    uint2 key;
    const float* noise;        // NULL or [n_inst][epochs][n_noise][N]
    const float* init_params;  // NULL or [n_inst][ld_params]
    int ld_params;
    float* scores;             // [n_inst]
    float* trace;              // NULL or [n_inst][epochs]
    float* params_out;         // NULL or [n_inst][ld_params]
    int* counter;
    float *dbg_loss, *dbg_pred, *dbg_dldp, *dbg_grad, *dbg_params;
};

struct GrSmem {
    float* Pm; float* nbuf; float* theta; float* am; float* av; long long* g64; double* red; GraphDesc* desc;
    int nn_pad;
};

__host__ __device__ inline size_t gr_smem_bytes(int NP, int DP, int n_noise, int p_max) {
    const int nn_pad = (n_noise + 3) & ~3;
    const int pp = (p_max + 3) & ~3;
    size_t b = 0;
    b += size_t(pp) * 8;                          // g64
    b += size_t(GR_NW + 4) * 8;                   // red, TT, beta powers, lr_t
    b += size_t(NP) * DP * 4;                     // Pm
    b += size_t(GR_NW) * nn_pad * 32 * 4;         // nbuf
    b += size_t(pp) * 4 * 3;                      // theta, m, v
    b += (sizeof(GraphDesc) + 15) & ~size_t(15);
    return (b + 15) & ~size_t(15);
}

// ---- noise of one row block into the warp's buffer -------------------------------------------
__device__ __forceinline__ void gr_fill_noise(const GrSmem& s, const GraphArgs& a, float* nb, int i, int lane, int step,
                                              unsigned int uid, uint2 key, const float* noise_step) {
    const int nn = s.desc->n_noise;
    if (noise_step) {
        for (int c = 0; c < nn; ++c) nb[c * 32 + lane] = (i < a.N) ? noise_step[size_t(c) * a.N + i] : 0.f;
    } else {
        for (int q = 0; 4 * q < nn; ++q) {
            const float4 n = normal4(philox4x32_10(make_uint4(i, step, GR_STREAM_NOISE0 + q, uid), key));
            nb[(4 * q + 0) * 32 + lane] = n.x; nb[(4 * q + 1) * 32 + lane] = n.y;
            nb[(4 * q + 2) * 32 + lane] = n.z; nb[(4 * q + 3) * 32 + lane] = n.w;
        }
    }
    __syncwarp();
}

// hidden pre-activations of node k for this lane's row: z[kk] = b_in[kk] + sum_p u_p W_in[p][kk]
template <int DP>
__device__ __forceinline__ void gr_node_z(const GrSmem& s, const float* nb, int k, int i, int lane, int h, float (&z)[GR_MAX_H]) {
    const GraphDesc& g = *s.desc;
    const float* th = s.theta + g.par_off[k];
    const int n_in = g.n_in[k], n_par = g.n_par[k];
    const short* idx = g.in_idx + g.in_off[k];
    const float* b_in = th + n_in * h;
#pragma unroll
    for (int kk = 0; kk < GR_MAX_H; ++kk) z[kk] = (kk < h) ? b_in[kk] : 0.f;
    for (int p = 0; p < n_in; ++p) {
        const float u = (p < n_par) ? s.Pm[i * DP + idx[p]] : nb[idx[p] * 32 + lane];
        const float* w = th + p * h;
#pragma unroll
        for (int kk = 0; kk < GR_MAX_H; ++kk)
            if (kk < h) z[kk] = fmaf(u, w[kk], z[kk]);
    }
}

// Generator forward of one row block (CGNN.py:63-84), nodes in build order.
template <int DP>
__device__ __forceinline__ void gr_forward_block(const GrSmem& s, const float* nb, int i, int lane, int h) {
    const GraphDesc& g = *s.desc;
    for (int k = 0; k < g.n_var; ++k) {
        float z[GR_MAX_H];
        gr_node_z<DP>(s, nb, k, i, lane, h, z);
        const float* w_out = s.theta + g.par_off[k] + g.n_in[k] * h + h;
        float out = w_out[h];
#pragma unroll
        for (int kk = 0; kk < GR_MAX_H; ++kk)
            if (kk < h) out = fmaf(fmaxf(z[kk], 0.f), w_out[kk], out);
        s.Pm[i * DP + g.col[k]] = out;
    }
}

template <int DP>
__device__ __forceinline__ void gr_load_col_smem(const float* __restrict__ base, int j, float (&t)[DP]) {
    const float4* r = reinterpret_cast<const float4*>(base + size_t(j) * DP);
#pragma unroll
    for (int q = 0; q < DP / 4; ++q) { const float4 v = r[q]; t[4 * q] = v.x; t[4 * q + 1] = v.y; t[4 * q + 2] = v.z; t[4 * q + 3] = v.w; }
}
template <int DP>
__device__ __forceinline__ void gr_load_col_gmem(const float* __restrict__ base, int j, float (&t)[DP]) {
    const float4* r = reinterpret_cast<const float4*>(base + size_t(j) * DP);
#pragma unroll
    for (int q = 0; q < DP / 4; ++q) { const float4 v = __ldg(r + q); t[4 * q] = v.x; t[4 * q + 1] = v.y; t[4 * q + 2] = v.z; t[4 * q + 3] = v.w; }
}

// MMD gradient of one row: tot[c] = sum_j w(P_i,T_j)(P_ic - T_jc) - sum_j w(P_i,P_j)(P_ic - P_jc)   (unscaled)
// fp32 accumulation in runs of one PT tile + one PP tile (64 evaluations), then a second-level fp32 sum
// of the (mutually cancelling, hence small) tile pairs.
template <int DP, int POL>
__device__ __forceinline__ void gr_mmd_train_row(const float* __restrict__ Pm, const float* __restrict__ T, int N,
                                                 const float (&p)[DP], float (&tot)[DP]) {
#pragma unroll
    for (int c = 0; c < DP; ++c) tot[c] = 0.f;
    for (int j0 = 0; j0 < N; j0 += 32) {
        const int j1 = min(j0 + 32, N);
        float acc[DP];
#pragma unroll
        for (int c = 0; c < DP; ++c) acc[c] = 0.f;
#pragma unroll 2
        for (int j = j0; j < j1; ++j) {
            float t[DP], df[DP];
            gr_load_col_gmem<DP>(T, j, t);
            float D = 0.f;
#pragma unroll
            for (int c = 0; c < DP; ++c) { df[c] = p[c] - t[c]; D = fmaf(df[c], df[c], D); }
            const float w = wsum<POL>(D);
#pragma unroll
            for (int c = 0; c < DP; ++c) acc[c] = fmaf(w, df[c], acc[c]);
        }
#pragma unroll 2
        for (int j = j0; j < j1; ++j) {
            float t[DP], df[DP];
            gr_load_col_smem<DP>(Pm, j, t);
            float D = 0.f;
#pragma unroll
            for (int c = 0; c < DP; ++c) { df[c] = p[c] - t[c]; D = fmaf(df[c], df[c], D); }
            const float w = -wsum<POL>(D);
#pragma unroll
            for (int c = 0; c < DP; ++c) acc[c] = fmaf(w, df[c], acc[c]);
        }
#pragma unroll
        for (int c = 0; c < DP; ++c) tot[c] += acc[c];
    }
}

// Loss share of one row: sum_j [ k(R_i, B_j) ] over a column source (smem or gmem), fp64 across tiles.
template <int DP, int POL, bool GMEM>
__device__ __forceinline__ double gr_mmd_eval_row(const float* __restrict__ B, int N, const float (&p)[DP]) {
    double tot = 0.0;
    for (int j0 = 0; j0 < N; j0 += 64) {
        const int j1 = min(j0 + 64, N);
        float acc = 0.f;
#pragma unroll 2
        for (int j = j0; j < j1; ++j) {
            float t[DP];
            if (GMEM) gr_load_col_gmem<DP>(B, j, t); else gr_load_col_smem<DP>(B, j, t);
            float D = 0.f;
#pragma unroll
            for (int c = 0; c < DP; ++c) { const float df = p[c] - t[c]; D = fmaf(df, df, D); }
            acc += ksum<POL>(D);
        }
        tot += double(acc);
    }
    return tot;
}

// Thread-local backward of one row through all nodes in reverse build order; parameter gradients are
// reduced over the 32 rows of the block (warp_reduce32) and added to the shared fixed-point accumulators.
template <int DP>
__device__ __forceinline__ void gr_backward_block(const GrSmem& s, const float* nb, int i, int lane, int h, float* gl) {
    const GraphDesc& g = *s.desc;
    for (int k = g.n_var - 1; k >= 0; --k) {
        float z[GR_MAX_H], v[32];
        gr_node_z<DP>(s, nb, k, i, lane, h, z);
        const int n_in = g.n_in[k], n_par = g.n_par[k];
        const int off = g.par_off[k];
        const short* idx = g.in_idx + g.in_off[k];
        const float* w_out = s.theta + off + n_in * h + h;
        const float delta = gl[g.col[k]];
        // b_out
        const float sb = warp_sum(delta);
        if (lane == 0) fx_add(&s.g64[off + n_in * h + 2 * h], double(sb));
        // W_out, then dz in place of z
#pragma unroll
        for (int kk = 0; kk < 32; ++kk) v[kk] = (kk < h) ? fmaxf(z[kk], 0.f) * delta : 0.f;
        float r = warp_reduce32(v, lane);
        if (lane < h) fx_add(&s.g64[off + n_in * h + h + lane], double(r));
#pragma unroll
        for (int kk = 0; kk < GR_MAX_H; ++kk) z[kk] = (kk < h && z[kk] > 0.f) ? delta * w_out[kk] : 0.f;
        // b_in
#pragma unroll
        for (int kk = 0; kk < 32; ++kk) v[kk] = z[kk];
        r = warp_reduce32(v, lane);
        if (lane < h) fx_add(&s.g64[off + n_in * h + lane], double(r));
        // W_in rows and input gradients
        for (int p = 0; p < n_in; ++p) {
            const float u = (p < n_par) ? s.Pm[i * DP + idx[p]] : nb[idx[p] * 32 + lane];
#pragma unroll
            for (int kk = 0; kk < 32; ++kk) v[kk] = u * z[kk];
            r = warp_reduce32(v, lane);
            if (lane < h) fx_add(&s.g64[off + p * h + lane], double(r));
            if (p < n_par) {
                const float* w = s.theta + off + p * h;
                float du = 0.f;
#pragma unroll
                for (int kk = 0; kk < GR_MAX_H; ++kk)
                    if (kk < h) du = fmaf(z[kk], w[kk], du);
                gl[idx[p]] += du;
            }
        }
    }
}

__device__ __forceinline__ double gr_block_sum(double v, double* red, int warp, int lane) {
    v = warp_sum(v);
    if (lane == 0) red[warp] = v;
    __syncthreads();
    double t = 0.0;
    if (warp == 0 && lane == 0) {
#pragma unroll
        for (int w = 0; w < GR_NW; ++w) t += red[w];
    }
    return t;
}

template <int DP, int POL>
__global__ void __launch_bounds__(GR_NT, 1) graph_kernel(const GraphArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ int s_inst;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int h = a.h, N = a.N, NB = (N + 31) >> 5;
    const int pp = (a.p_max + 3) & ~3;

    GrSmem s;
    {
        unsigned char* p = smem_raw;
        s.nn_pad = (a.n_noise + 3) & ~3;
        s.g64 = reinterpret_cast<long long*>(p);   p += size_t(pp) * 8;
        s.red = reinterpret_cast<double*>(p);      p += size_t(GR_NW + 4) * 8;
        s.Pm = reinterpret_cast<float*>(p);        p += size_t(a.NP) * DP * 4;
        s.nbuf = reinterpret_cast<float*>(p);      p += size_t(GR_NW) * s.nn_pad * 32 * 4;
        s.theta = reinterpret_cast<float*>(p);     p += size_t(pp) * 4;
        s.am = reinterpret_cast<float*>(p);        p += size_t(pp) * 4;
        s.av = reinterpret_cast<float*>(p);        p += size_t(pp) * 4;
        s.desc = reinterpret_cast<GraphDesc*>(p);
    }
    float* nb = s.nbuf + size_t(warp) * s.nn_pad * 32;
    const int epochs = a.train_epochs + a.test_epochs;
    const float scale = 4.0f / (float(N) * float(N));
    const double invN2 = 1.0 / (double(N) * double(N));

    for (;;) {
        __syncthreads();
        if (tid == 0) s_inst = atomicAdd(a.counter, 1);
        __syncthreads();
        if (s_inst >= a.n_inst) break;
        const int2 inst = a.inst[s_inst];
        const int out = inst.y;
        const unsigned long long uid64 = a.inst_uid[s_inst];
        const unsigned int uid = (unsigned int)(uid64 & 0xffffffffull);
        const uint2 key = make_uint2(a.key.x, a.key.y ^ (unsigned int)(uid64 >> 32));
        const float* T = a.T + size_t(a.t_index ? a.t_index[out] : 0) * a.NP * DP;
        const float* noise = a.noise ? a.noise + size_t(out) * epochs * a.n_noise * N : nullptr;

        // ---- descriptor, parameters, Adam state, generated matrix -------------------------------
        {
            const int* src = reinterpret_cast<const int*>(a.desc + inst.x);
            int* dst = reinterpret_cast<int*>(s.desc);
            for (int q = tid; q < int(sizeof(GraphDesc) / 4); q += GR_NT) dst[q] = src[q];
        }
        for (int q = tid; q < a.NP * DP; q += GR_NT) s.Pm[q] = 0.f;
        __syncthreads();
        const int P = s.desc->n_params;
        for (int q = tid; q < pp; q += GR_NT) {
            float th = 0.f;
            if (q < P) {
                if (a.init_params) th = a.init_params[size_t(out) * a.ld_params + q];
                else {
                    const float4 n = normal4(philox4x32_10(make_uint4(q >> 2, 0u, STREAM_INIT, uid), key));
                    const int e = q & 3;
                    th = a.init_std * (e == 0 ? n.x : e == 1 ? n.y : e == 2 ? n.z : n.w);
                }
            }
            s.theta[q] = th; s.am[q] = 0.f; s.av[q] = 0.f; s.g64[q] = 0;
        }
        if (tid == 0) { s.red[GR_NW + 1] = 1.0; s.red[GR_NW + 2] = 1.0; }
        // ---- TT (constant per instance) -------------------------------------------------------
        {
            double part = 0.0;
            for (int rb = warp; rb < NB; rb += GR_NW) {
                const int i = 32 * rb + lane;
                float p[DP];
                gr_load_col_gmem<DP>(T, min(i, N - 1), p);
                const double v = gr_mmd_eval_row<DP, POL, true>(T, N, p);
                if (i < N) part += v;
            }
            const double tt = gr_block_sum(part, s.red, warp, lane);
            if (tid == 0) s.red[GR_NW] = tt;
        }

        double score_sum = 0.0;
        for (int it = 0; it < epochs; ++it) {
            const bool train = it < a.train_epochs;
            const float* noise_step = noise ? noise + size_t(it) * a.n_noise * N : nullptr;
            __syncthreads();
            if (train && tid == 0) {
                const double b1p = s.red[GR_NW + 1] * 0.9, b2p = s.red[GR_NW + 2] * 0.999;
                s.red[GR_NW + 1] = b1p; s.red[GR_NW + 2] = b2p;
                s.red[GR_NW + 3] = double(a.lr) * sqrt(1.0 - b2p) / (1.0 - b1p);
            }
            // ---- generator forward, all rows ------------------------------------------------
            for (int rb = warp; rb < NB; rb += GR_NW) {
                const int i = 32 * rb + lane;
                gr_fill_noise(s, a, nb, i, lane, it, uid, key, noise_step);
                gr_forward_block<DP>(s, nb, i, lane, h);
                __syncwarp();
            }
            __syncthreads();
            if (a.dbg_pred && train)
                for (int q = tid; q < N * s.desc->n_var; q += GR_NT) {
                    const int r = q / s.desc->n_var, c = q - r * s.desc->n_var;
                    a.dbg_pred[(size_t(it) * N + r) * s.desc->n_var + c] = s.Pm[r * DP + c];
                }
            // ---- loss (evaluation steps; traced training steps) ---------------------------------
            const bool want_loss = !train || a.dbg_loss || (a.trace && a.trace_every > 0 && it % a.trace_every == 0);
            if (want_loss) {
                double part = 0.0;
                for (int rb = warp; rb < NB; rb += GR_NW) {
                    const int i = 32 * rb + lane;
                    float p[DP];
                    gr_load_col_smem<DP>(s.Pm, i, p);
                    const double v = gr_mmd_eval_row<DP, POL, false>(s.Pm, N, p) - 2.0 * gr_mmd_eval_row<DP, POL, true>(T, N, p);
                    if (i < N) part += v;
                }
                const double sum = gr_block_sum(part, s.red, warp, lane);
                if (tid == 0) {
                    const float loss = float((sum + s.red[GR_NW]) * invN2);
                    if (!train) score_sum += double(loss);
                    if (a.trace) a.trace[size_t(out) * epochs + it] = loss;
                    if (a.dbg_loss && train) a.dbg_loss[it] = loss;
                }
            } else if (a.trace && tid == 0) {
                a.trace[size_t(out) * epochs + it] = __int_as_float(0x7fc00000);
            }
            if (!train) continue;
            // ---- MMD gradient + generator backward, row block by row block ------------------------
            for (int rb = warp; rb < NB; rb += GR_NW) {
                const int i = 32 * rb + lane;
                float gl[DP];
                {
                    float p[DP], tot[DP];
                    gr_load_col_smem<DP>(s.Pm, i, p);
                    gr_mmd_train_row<DP, POL>(s.Pm, T, N, p, tot);
#pragma unroll
                    for (int c = 0; c < DP; ++c) gl[c] = (i < N) ? tot[c] * scale : 0.f;
                }
                if (a.dbg_dldp && i < N)
                    for (int c = 0; c < s.desc->n_var; ++c) a.dbg_dldp[(size_t(it) * N + i) * s.desc->n_var + c] = gl[c];
                gr_fill_noise(s, a, nb, i, lane, it, uid, key, noise_step);
                gr_backward_block<DP>(s, nb, i, lane, h, gl);
                __syncwarp();
            }
            __syncthreads();
            // ---- Adam --------------------------------------------------------------------------
            {
                const float lr_t = float(s.red[GR_NW + 3]);
                for (int q = tid; q < P; q += GR_NT) {
                    const float g = float(fx_get(&s.g64[q]));
                    s.g64[q] = 0;
                    float m = s.am[q], v = s.av[q];
                    const float th = adam_update(s.theta[q], g, m, v, lr_t);
                    s.am[q] = m; s.av[q] = v; s.theta[q] = th;
                    if (a.dbg_grad) a.dbg_grad[size_t(it) * P + q] = g;
                    if (a.dbg_params) a.dbg_params[size_t(it) * P + q] = th;
                }
            }
            if (it == a.train_epochs - 1 && a.params_out) {
                __syncthreads();
                for (int q = tid; q < P; q += GR_NT) a.params_out[size_t(out) * a.ld_params + q] = s.theta[q];
            }
        }
        if (tid == 0 && a.scores) a.scores[out] = a.test_epochs > 0 ? float(score_sum / double(a.test_epochs)) : 0.f;
    }
}

}  // namespace cgnn
