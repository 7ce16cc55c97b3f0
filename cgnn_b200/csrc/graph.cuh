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
// Layout: the generated matrix Pm[NP][DP] (row-major, DP = n_var padded to 8/16/20/22/24/32, pad
// columns 0) lives in shared memory for the instance's whole life; the observed matrix T[NP][DP]
// is the same for every instance of a batch (or one per row subset).  It is copied to shared memory next
// to Pm when both fit one SM (t_smem); otherwise every warp streams it through a cp.async ring that lives
// in its (idle) noise buffer, two halves of 16 rows (t_ring); only when neither is possible (random-feature
// mode, very narrow noise buffers) is it read through the read-only path (L1/L2).  Parameters, Adam moments
// and the fixed-point gradient accumulators are in shared memory too.  The host picks the plan from the instance
// shape (graph_api.cu); models too wide to buffer their noise columns per row block run the REGEN instantiation,
// which re-derives every noise value from its Philox counter where a node consumes it.
//
// Mapping: lane <-> row i.  A warp owns whole 32-row blocks (w, w+NW, ...): the row's generated
// values, its MMD gradient accumulators and the distance differences stay in REGISTERS across all
// 2N kernel evaluations of the row, and the row's gradient goes straight into the thread-local
// MLP backward pass - dLoss/dPred is never written to memory.  Column j (DP floats) arrives as
// broadcast loads; the gradient pass takes the observed and the generated column j in one iteration, their two
// kernel-weight chains as one FP32x2 chain (gr_grad_pair).  Parameter gradients are reduced across the 32 rows of a block with
// the transposing warp reduction and accumulated across warps with fixed-point shared atomics.
//
// Thread-block clusters: when a batch has fewer instances than the GPU has SMs (a hill-climbing tail, the
// 20-variable configuration on 8 GPUs) one instance is given a CLUSTER of 2-8 CTAs.  Every CTA keeps the whole
// generated matrix and all parameters (the generator forward is replicated - ~3 % of a step), the row blocks of
// the MMD / backward pass are split over the CTAs, and once per step the fixed-point gradient accumulators are
// summed across the cluster through distributed shared memory, so all CTAs apply the same Adam update.
#pragma once
#include <cooperative_groups.h>
#include "common.cuh"

namespace cgnn {

namespace cg = cooperative_groups;

#ifndef CGNN_GR_NT
#define CGNN_GR_NT 512
#endif
constexpr int GR_NT = CGNN_GR_NT;
constexpr int GR_MAX_CLUSTER = 8;
constexpr int GR_NW = GR_NT / 32;
constexpr int GR_MAX_VAR = 32;
constexpr int GR_MAX_H = 32;
constexpr int GR_MAX_IN = 384;           // sum over nodes of (parents + noise inputs)
constexpr uint32_t GR_STREAM_NOISE0 = 2u;

// Parameters live in shared memory in a PADDED layout: every hidden vector (a row of W_in, b_in, W_out)
// occupies HP floats (HP = 20 or 32 >= h, compile-time), the tail being zero for ever (its gradient is
// exactly 0), so the generator loops are fully unrolled 128-bit / FFMA2 code without tail predicates.
// Node k: [W_in (n_in x HP) | b_in (HP) | W_out (HP) | b_out, 3 pad] at pad_off[k]; the caller-visible flat
// layout [W_in (n_in x h) | b_in (h) | W_out (h) | b_out] starts at par_off[k].
struct GraphDesc {
    int n_var, n_noise, n_params, n_in_total, n_params_pad, pad0, pad1, pad2;
    short col[GR_MAX_VAR], n_par[GR_MAX_VAR], n_in[GR_MAX_VAR];
    int in_off[GR_MAX_VAR], par_off[GR_MAX_VAR], pad_off[GR_MAX_VAR];
    short in_idx[GR_MAX_IN];
};

// caller-visible parameter index -> index in the padded shared-memory layout
__device__ __forceinline__ int gr_pad_index(const GraphDesc& g, int q, int h, int HP) {
    int k = 0;
    while (k + 1 < g.n_var && g.par_off[k + 1] <= q) ++k;
    const int local = q - g.par_off[k], n_in = g.n_in[k];
    const int row = local / h, kk = local - row * h;           // rows 0..n_in-1: W_in, n_in: b_in, n_in+1: W_out
    if (row >= n_in + 2) return g.pad_off[k] + (n_in + 2) * HP;   // b_out
    return g.pad_off[k] + row * HP + kk;
}

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
    // random-feature ("Fourier") MMD, Loss.py:35-56: fast_k = nb_vectors_approx_MMD, 0 = Gaussian-kernel MMD
    int t_smem;                // 1: the observed matrix of the instance is copied to shared memory (it fits)
    int t_ring;                // 1: it does not fit and is streamed through the warps' idle noise buffers (GR_T_RING)
    int noise_regen;           // 1: noise values are re-derived from their Philox counter at every use instead of being
                               //    buffered per row block (wide hidden-variable models at 1500 rows: the buffer would not fit)
    int nbuf_cols;             // floats per lane of a warp's buffer: padded noise columns, or just the ring (DP), or 0
    int fast_k;
    const float* wz_ext;       // NULL or [n_inst][epochs][n_var+1][7*fast_k]: the matrix of Loss.py:35-37 per step
};

struct GrSmem {
    float* Pm; float* nbuf; float* theta; float* am; float* av; long long* g64; double* red; GraphDesc* desc;
    int nn_pad;
    float* Wc; float* bc; float* coef; long long* et; long long* ep;     // random-feature mode: one bandwidth group
    float* Ts;                                                           // copy of the observed matrix [NP][DP] (t_smem)
};

__host__ __device__ inline size_t gr_smem_bytes(int NP, int DP, int nbuf_cols, int p_max, int fast_k = 0, int t_smem = 0) {
    const int nn_pad = nbuf_cols;
    const int pp = (p_max + 3) & ~3;
    size_t b = 0;
    b += size_t(pp) * 8;                          // g64
    b += size_t(GR_NW + 4 + GR_MAX_CLUSTER) * 8;  // red, TT, beta powers, lr_t, per-CTA partials of a cluster
    b += size_t(NP) * DP * 4;                     // Pm
    b += size_t(GR_NW) * nn_pad * 32 * 4;         // nbuf
    b += size_t(pp) * 4 * 3;                      // theta, m, v
    b += (sizeof(GraphDesc) + 15) & ~size_t(15);
    const int kp = (fast_k + 3) & ~3;
    b += size_t(kp) * (8 + 8 + 4 * DP + 4 + 4);  // et, ep, Wc, bc, coef
    b = (b + 15) & ~size_t(15);
    if (t_smem) b += size_t(NP) * DP * 4;        // Ts
    return (b + 15) & ~size_t(15);
}

// ---- noise ----------------------------------------------------------------------------------------------
// Column c of row i at a step is element c & 3 of Philox(i, step, GR_STREAM_NOISE0 + c / 4, uid).  Normally a warp draws
// all columns of its row block once into its buffer (gr_fill_noise); in regeneration mode nothing is buffered and a
// value is re-derived (or re-read from the caller's noise) wherever a node consumes it.
struct GrNoise {
    const float* nb;           // the warp's buffer [column][32 lanes] (buffered mode)
    const float* ext;          // caller-supplied noise of the step [column][N], or NULL
    int i, N, lane;
    unsigned int step, uid;
    uint2 key;
};
static __device__ __noinline__ float gr_noise_regen(const float* ext, int i, int N, unsigned int step, unsigned int uid, unsigned int kx,
                                             unsigned int ky, int c) {
    if (ext) return (i < N) ? ext[size_t(c) * N + i] : 0.f;
    const float4 n = normal4(philox4x32_10(make_uint4(i, step, GR_STREAM_NOISE0 + (c >> 2), uid), make_uint2(kx, ky)));
    const int e = c & 3;
    return e == 0 ? n.x : e == 1 ? n.y : e == 2 ? n.z : n.w;
}
template <bool REGEN>
__device__ __forceinline__ float gr_noise_get(const GrNoise& z, int c) {
    if (!REGEN) return z.nb[c * 32 + z.lane];
    return gr_noise_regen(z.ext, z.i, z.N, z.step, z.uid, z.key.x, z.key.y, c);
}

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

// hidden pre-activations of node k for this lane's row: z[kk] = b_in[kk] + sum_p u_p W_in[p][kk], as HP/2 pairs
template <int DP, int HP, bool REGEN>
__device__ __forceinline__ void gr_node_z(const GrSmem& s, const GrNoise& nz, int k, int i, float2 (&z)[HP / 2]) {
    const GraphDesc& g = *s.desc;
    const float* th = s.theta + g.pad_off[k];
    const int n_in = g.n_in[k], n_par = g.n_par[k];
    const short* idx = g.in_idx + g.in_off[k];
    const float4* b4 = reinterpret_cast<const float4*>(th + n_in * HP);
#pragma unroll
    for (int q = 0; q < HP / 4; ++q) { const float4 b = b4[q]; z[2 * q] = make_float2(b.x, b.y); z[2 * q + 1] = make_float2(b.z, b.w); }
    for (int p = 0; p < n_in; ++p) {
        const float2 u = splat2((p < n_par) ? s.Pm[i * DP + idx[p]] : gr_noise_get<REGEN>(nz, idx[p]));
        const float4* w4 = reinterpret_cast<const float4*>(th + p * HP);
#pragma unroll
        for (int q = 0; q < HP / 4; ++q) {
            const float4 w = w4[q];
            z[2 * q] = ffma2(u, make_float2(w.x, w.y), z[2 * q]);
            z[2 * q + 1] = ffma2(u, make_float2(w.z, w.w), z[2 * q + 1]);
        }
    }
}

// Generator forward of one row block (CGNN.py:63-84), nodes in build order.
template <int DP, int HP, bool REGEN>
__device__ __forceinline__ void gr_forward_block(const GrSmem& s, const GrNoise& nz, int i) {
    const GraphDesc& g = *s.desc;
    for (int k = 0; k < g.n_var; ++k) {
        float2 z[HP / 2];
        gr_node_z<DP, HP, REGEN>(s, nz, k, i, z);
        const float* w_out = s.theta + g.pad_off[k] + (g.n_in[k] + 1) * HP;
        const float4* w4 = reinterpret_cast<const float4*>(w_out);
        float2 acc = make_float2(w_out[HP], 0.f);                       // b_out
#pragma unroll
        for (int q = 0; q < HP / 4; ++q) {
            const float4 w = w4[q];
            acc = ffma2(make_float2(fmaxf(z[2 * q].x, 0.f), fmaxf(z[2 * q].y, 0.f)), make_float2(w.x, w.y), acc);
            acc = ffma2(make_float2(fmaxf(z[2 * q + 1].x, 0.f), fmaxf(z[2 * q + 1].y, 0.f)), make_float2(w.z, w.w), acc);
        }
        s.Pm[i * DP + g.col[k]] = acc.x + acc.y;
    }
}

template <int DP>
__device__ __forceinline__ void gr_load_col_smem(const float* __restrict__ base, int j, float (&t)[DP]) {
    if constexpr (DP % 4 == 0) {
        const float4* r = reinterpret_cast<const float4*>(base + size_t(j) * DP);
#pragma unroll
        for (int q = 0; q < DP / 4; ++q) { const float4 v = r[q]; t[4 * q] = v.x; t[4 * q + 1] = v.y; t[4 * q + 2] = v.z; t[4 * q + 3] = v.w; }
    } else {
        const float2* r = reinterpret_cast<const float2*>(base + size_t(j) * DP);
#pragma unroll
        for (int q = 0; q < DP / 2; ++q) { const float2 v = r[q]; t[2 * q] = v.x; t[2 * q + 1] = v.y; }
    }
}
template <int DP>
__device__ __forceinline__ void gr_load_col_gmem(const float* __restrict__ base, int j, float (&t)[DP]) {
    if constexpr (DP % 4 == 0) {
        const float4* r = reinterpret_cast<const float4*>(base + size_t(j) * DP);
#pragma unroll
        for (int q = 0; q < DP / 4; ++q) { const float4 v = __ldg(r + q); t[4 * q] = v.x; t[4 * q + 1] = v.y; t[4 * q + 2] = v.z; t[4 * q + 3] = v.w; }
    } else {
        const float2* r = reinterpret_cast<const float2*>(base + size_t(j) * DP);
#pragma unroll
        for (int q = 0; q < DP / 2; ++q) { const float2 v = __ldg(r + q); t[2 * q] = v.x; t[2 * q + 1] = v.y; }
    }
}

// one row (DP floats) as DP/2 register pairs; widest aligned loads the row stride allows (DP = 22: 8-byte)
template <int DP, bool GMEM>
__device__ __forceinline__ void gr_load_pairs(const float* __restrict__ row, float2 (&t)[DP / 2]) {
    if constexpr (DP % 4 == 0) {
        const float4* r = reinterpret_cast<const float4*>(row);
#pragma unroll
        for (int q = 0; q < DP / 4; ++q) {
            const float4 v = GMEM ? __ldg(r + q) : r[q];
            t[2 * q] = make_float2(v.x, v.y); t[2 * q + 1] = make_float2(v.z, v.w);
        }
    } else {
        const float2* r = reinterpret_cast<const float2*>(row);
#pragma unroll
        for (int q = 0; q < DP / 2; ++q) t[q] = GMEM ? __ldg(r + q) : r[q];
    }
}

// where the observed rows of the MMD passes come from
enum : int { GR_T_GMEM = 0, GR_T_SMEM = 1, GR_T_RING = 2 };

// ---- gradient pass ---------------------------------------------------------------------------------------
// This lane's row is held NEGATED (np = -P_i), so a difference is one FADD2 per coordinate pair: d = X_j - P_i.
// One iteration takes the observed column j AND the generated column j: the two squared distances go through
// the kernel-weight chain as one register pair (FMUL2 / FFMA2, wsum_r2), and the two contributions
//     acc += w(P_i,T_j) (T_j - P_i) - w(P_i,P_j) (P_j - P_i)          ( = - dLoss/dP_i * N^2/4, accumulated )
// cancel as they are added - the running sum stays at the size of the (small) result for the whole sweep, which
// is what lets ONE fp32 accumulator per coordinate replace the former two-level sum (tools/expansion_error_model.py
// measures 7e-7 at the trained state of the 22-variable example).
template <int DP, int POL, bool TG>
__device__ __forceinline__ void gr_grad_pair(const float2 (&np)[DP / 2], const float* __restrict__ tj,
                                             const float* __restrict__ pj, float2 (&acc)[DP / 2]) {
    float2 d0[DP / 2], d1[DP / 2];
    gr_load_pairs<DP, TG>(tj, d0);
    gr_load_pairs<DP, false>(pj, d1);
    float2 e0a = make_float2(0.f, 0.f), e0b = e0a, e1a = e0a, e1b = e0a;
#pragma unroll
    for (int q = 0; q < DP / 2; ++q) {
        d0[q] = fadd2(d0[q], np[q]);
        d1[q] = fadd2(d1[q], np[q]);
        if (q & 1) { e0b = ffma2(d0[q], d0[q], e0b); e1b = ffma2(d1[q], d1[q], e1b); }
        else       { e0a = ffma2(d0[q], d0[q], e0a); e1a = ffma2(d1[q], d1[q], e1a); }
    }
    const float2 e0 = fadd2(e0a, e0b), e1 = fadd2(e1a, e1b);
    const float2 w = wsum_r2<POL>(make_float2(e0.x + e0.y, e1.x + e1.y));
    const float2 w0 = splat2(w.x), w1 = splat2(-w.y);
#pragma unroll
    for (int q = 0; q < DP / 2; ++q) acc[q] = ffma2(w1, d1[q], ffma2(w0, d0[q], acc[q]));
}

// ---- observed matrix streamed through a per-warp ring (shapes whose observed matrix does not fit shared memory) --
// The warp's noise buffer is idle while the warp walks the observed rows, and it holds at least 32 rows whenever
// nn_pad >= DP: two halves of GR_RING_H rows, filled with cp.async one half ahead of the arithmetic, replace
// the L1/L2 round trip of every row by a shared-memory broadcast read.
constexpr int GR_RING_H = 16;

__device__ __forceinline__ void gr_cp_async16(void* dst_smem, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(dst_smem)), "l"(src) : "memory");
}
// chunk c = rows [c*H, c*H + H) of T (padded to NP rows, so a whole chunk is always readable) -> half c & 1; one group
template <int DP>
__device__ __forceinline__ void gr_ring_issue(const float* __restrict__ T, float* ring, int c, int n_chunks, int lane) {
    if (c < n_chunks) {
        const float4* src = reinterpret_cast<const float4*>(T + size_t(c) * GR_RING_H * DP);
        float4* dst = reinterpret_cast<float4*>(ring + (c & 1) * GR_RING_H * DP);
#pragma unroll
        for (int q = lane; q < GR_RING_H * DP / 4; q += 32) gr_cp_async16(dst + q, src + q);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
}
__device__ __forceinline__ void gr_ring_wait() {            // everything but the newest group has landed, warp-wide
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    __syncwarp();
}

// dLoss/dP_i (unscaled, negated) of this lane's row over all N observed and generated columns
template <int DP, int POL, int TM>
__device__ __forceinline__ void gr_mmd_grad_row(const float* __restrict__ Pm, const float* __restrict__ T, float* ring, int N,
                                                int lane, const float2 (&np)[DP / 2], float2 (&acc)[DP / 2]) {
#pragma unroll
    for (int q = 0; q < DP / 2; ++q) acc[q] = make_float2(0.f, 0.f);
    if constexpr (TM == GR_T_RING) {
        const int n_chunks = (N + GR_RING_H - 1) / GR_RING_H;
        gr_ring_issue<DP>(T, ring, 0, n_chunks, lane);
        for (int c = 0; c < n_chunks; ++c) {
            gr_ring_issue<DP>(T, ring, c + 1, n_chunks, lane);
            gr_ring_wait();
            const float* B = ring + (c & 1) * GR_RING_H * DP;
            const float* Pc = Pm + size_t(c) * GR_RING_H * DP;
            const int n = min(GR_RING_H, N - c * GR_RING_H);
            for (int j = 0; j < n; ++j) gr_grad_pair<DP, POL, false>(np, B + j * DP, Pc + j * DP, acc);
            __syncwarp();                                   // the half may be refilled two chunks later
        }
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncwarp();
    } else {
        for (int j = 0; j < N; ++j) gr_grad_pair<DP, POL, TM == GR_T_GMEM>(np, T + size_t(j) * DP, Pm + size_t(j) * DP, acc);
    }
}

// ---- loss passes -----------------------------------------------------------------------------------------
// kernel sums of two columns at once (packed chain): (sum_g k_g(P_i, A), sum_g k_g(P_i, B))
template <int DP, int POL, bool GMEM>
__device__ __forceinline__ float2 gr_k_pair(const float2 (&np)[DP / 2], const float* __restrict__ ca, const float* __restrict__ cb) {
    float2 d0[DP / 2], d1[DP / 2];
    gr_load_pairs<DP, GMEM>(ca, d0);
    gr_load_pairs<DP, GMEM>(cb, d1);
    float2 e0a = make_float2(0.f, 0.f), e0b = e0a, e1a = e0a, e1b = e0a;
#pragma unroll
    for (int q = 0; q < DP / 2; ++q) {
        d0[q] = fadd2(d0[q], np[q]);
        d1[q] = fadd2(d1[q], np[q]);
        if (q & 1) { e0b = ffma2(d0[q], d0[q], e0b); e1b = ffma2(d1[q], d1[q], e1b); }
        else       { e0a = ffma2(d0[q], d0[q], e0a); e1a = ffma2(d1[q], d1[q], e1a); }
    }
    const float2 e0 = fadd2(e0a, e0b), e1 = fadd2(e1a, e1b);
    return ksum_r2<POL>(make_float2(e0.x + e0.y, e1.x + e1.y));
}

// sum_j k(R_i, B_j) over n consecutive rows starting at B (n <= 64): one fp32 run
template <int DP, int POL, bool GMEM>
__device__ __forceinline__ float gr_k_run(const float* __restrict__ B, int n, const float2 (&np)[DP / 2]) {
    float2 acc = make_float2(0.f, 0.f);
    int j = 0;
    for (; j + 1 < n; j += 2) acc = fadd2(acc, gr_k_pair<DP, POL, GMEM>(np, B + size_t(j) * DP, B + size_t(j + 1) * DP));
    float r = acc.x + acc.y;
    if (j < n) r += gr_k_pair<DP, POL, GMEM>(np, B + size_t(j) * DP, B + size_t(j) * DP).x;
    return r;
}

// sum_j k(R_i, T_j) over all N observed rows through the ring, fp32 runs of 64 then fp64 (as gr_mmd_eval_cols)
template <int DP, int POL>
__device__ __forceinline__ double gr_mmd_eval_row_ring(const float* __restrict__ T, float* ring, int N, int lane, const float2 (&np)[DP / 2]) {
    const int n_chunks = (N + GR_RING_H - 1) / GR_RING_H;
    double tot = 0.0;
    gr_ring_issue<DP>(T, ring, 0, n_chunks, lane);
    for (int j0 = 0; j0 < N; j0 += 64) {
        const int j1 = min(j0 + 64, N);
        float acc = 0.f;
        for (int c = j0 / GR_RING_H; c * GR_RING_H < j1; ++c) {
            gr_ring_issue<DP>(T, ring, c + 1, n_chunks, lane);
            gr_ring_wait();
            acc += gr_k_run<DP, POL, false>(ring + (c & 1) * GR_RING_H * DP, min(GR_RING_H, j1 - c * GR_RING_H), np);
            __syncwarp();
        }
        tot += double(acc);
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncwarp();
    return tot;
}

// Loss share of one row: sum_j [ k(R_i, B_j) ] over columns [c0, c1) of a source (smem or gmem), fp64 across runs of 64.
template <int DP, int POL, bool GMEM>
__device__ __forceinline__ double gr_mmd_eval_cols(const float* __restrict__ B, int c0, int c1, const float2 (&np)[DP / 2]) {
    double tot = 0.0;
    for (int j0 = c0; j0 < c1; j0 += 64)
        tot += double(gr_k_run<DP, POL, GMEM>(B + size_t(j0) * DP, min(64, c1 - j0), np));
    return tot;
}
template <int DP, int POL, bool GMEM>
__device__ __forceinline__ double gr_mmd_eval_row(const float* __restrict__ B, int N, const float2 (&np)[DP / 2]) {
    return gr_mmd_eval_cols<DP, POL, GMEM>(B, 0, N, np);
}

// Same for a block of the SAME set as the rows (pred-pred, true-true): the kernel matrix is symmetric, so
// row block rb takes its diagonal tile once and only "its" half of the other column tiles, doubled - the cyclic
// assignment J = rb+1 .. rb+npp (mod NB) covers every unordered tile pair exactly once over all row blocks.
__device__ __forceinline__ int gr_npp(int I, int NB) {
    return (NB & 1) ? (NB - 1) / 2 : (NB / 2 - 1 + (I < NB / 2 ? 1 : 0));
}
template <int DP, int POL, bool GMEM>
__device__ __forceinline__ double gr_mmd_eval_row_sym(const float* __restrict__ B, int N, int NB, int rb, const float2 (&np)[DP / 2]) {
    double tot = gr_mmd_eval_cols<DP, POL, GMEM>(B, 32 * rb, min(32 * rb + 32, N), np);
    const int npp = gr_npp(rb, NB);
    double off = 0.0;
    for (int q = 0; q < npp; ++q) {
        int J = rb + 1 + q;
        if (J >= NB) J -= NB;
        off += gr_mmd_eval_cols<DP, POL, GMEM>(B, 32 * J, min(32 * J + 32, N), np);
    }
    return tot + 2.0 * off;
}

// element kk of a hidden vector held as HP/2 register pairs (kk is a compile-time constant after unrolling)
template <int HP>
__device__ __forceinline__ float gr_zget(const float2 (&z)[HP / 2], int kk) {
    return (kk & 1) ? z[(kk >> 1) % (HP / 2)].y : z[(kk >> 1) % (HP / 2)].x;
}

// Thread-local backward of one row through all nodes in reverse build order; parameter gradients are
// reduced over the 32 rows of the block (warp_reduce32) and added to the shared fixed-point accumulators.
template <int DP, int HP, bool REGEN>
__device__ __forceinline__ void gr_backward_block(const GrSmem& s, const GrNoise& nz, int i, int lane, float* gl) {
    const GraphDesc& g = *s.desc;
    for (int k = g.n_var - 1; k >= 0; --k) {
        float2 z[HP / 2];
        gr_node_z<DP, HP, REGEN>(s, nz, k, i, z);
        const int n_in = g.n_in[k], n_par = g.n_par[k];
        const int off = g.pad_off[k];
        const short* idx = g.in_idx + g.in_off[k];
        const float* w_out = s.theta + off + (n_in + 1) * HP;
        const float delta = gl[g.col[k]];
        // b_out
        const float sb = warp_sum(delta);
        if (lane == 0) fx_add(&s.g64[off + (n_in + 2) * HP], double(sb));
        // W_out and b_in: sums over the 32 rows by the transposing reduction over the HP hidden units, the two of them
        // as one packed reduction (lane -> (slot, valid) is the same for every reduction of this width)
        float2 v[HP];
        int slot;
        bool has_slot;
#pragma unroll
        for (int q = 0; q < HP / 2; ++q) {
            const float2 zq = z[q];
            z[q].x = zq.x > 0.f ? delta * w_out[2 * q] : 0.f;                 // dz in place of z
            z[q].y = zq.y > 0.f ? delta * w_out[2 * q + 1] : 0.f;
            v[2 * q] = make_float2(fmaxf(zq.x, 0.f) * delta, z[q].x);
            v[2 * q + 1] = make_float2(fmaxf(zq.y, 0.f) * delta, z[q].y);
        }
        float2 r = warp_reduce_n2<HP>(v, lane, slot, has_slot);
        if (has_slot) {
            fx_add(&s.g64[off + (n_in + 1) * HP + slot], double(r.x));
            fx_add(&s.g64[off + n_in * HP + slot], double(r.y));
        }
        // W_in rows, two inputs per packed reduction, and the input gradients
        for (int p = 0; p < n_in; p += 2) {
            const bool two = p + 1 < n_in;
            const float u0 = (p < n_par) ? s.Pm[i * DP + idx[p]] : gr_noise_get<REGEN>(nz, idx[p]);
            const float u1 = !two ? 0.f : (p + 1 < n_par) ? s.Pm[i * DP + idx[p + 1]] : gr_noise_get<REGEN>(nz, idx[p + 1]);
#pragma unroll
            for (int kk = 0; kk < HP; ++kk) { const float dz = gr_zget<HP>(z, kk); v[kk] = make_float2(u0 * dz, u1 * dz); }
            r = warp_reduce_n2<HP>(v, lane, slot, has_slot);
            if (has_slot) {
                fx_add(&s.g64[off + p * HP + slot], double(r.x));
                if (two) fx_add(&s.g64[off + (p + 1) * HP + slot], double(r.y));
            }
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int pe = p + e;
                if (pe < n_par) {
                    const float4* w4 = reinterpret_cast<const float4*>(s.theta + off + pe * HP);
                    float2 du = make_float2(0.f, 0.f);
#pragma unroll
                    for (int q = 0; q < HP / 4; ++q) {
                        const float4 w = w4[q];
                        du = ffma2(z[2 * q], make_float2(w.x, w.y), du);
                        du = ffma2(z[2 * q + 1], make_float2(w.z, w.w), du);
                    }
                    gl[idx[pe]] += du.x + du.y;
                }
            }
        }
    }
}

// ---- random-feature ("Fourier") MMD for the graph path (Loss.py:35-56) ------------------------------
// The 7k features are processed one bandwidth group (k features) at a time: draw the group's frequencies
// and phases, column sums of cos over the rows (observed and generated), coefficients, per-row gradient.
// A warp owns exactly one row block in this mode (the host picks the cluster size accordingly), so the
// row's gradient stays in registers across the seven groups.
constexpr uint32_t GR_STREAM_FEAT = 900u;

__device__ __forceinline__ float gr_turn_reduce(float z) {
    // round-to-nearest by the 1.5*2^23 trick: two FADDs on the FMA pipe instead of an FRND on the (busy) XU pipe;
    // exact for |t| < 2^22, i.e. |z| < 2.6e7 - far beyond any frequency x coordinate product of this loss
    const float t = z * 0.15915494309189535f;
    const float r = __fsub_rn(__fadd_rn(t, 12582912.0f), 12582912.0f);
    return (t - r) * 6.283185307179586f;
}

template <int DP>
__device__ __forceinline__ void gr_fourier_features(const GrSmem& s, const GraphArgs& a, int grp, int step, unsigned int uid,
                                                    uint2 key, const float* wz_step, int tid) {
    const int k = a.fast_k, F = 7 * k, nv = s.desc->n_var;
    const float gam[7] = {G0, G1, G2, G3, G4, G5, G6};
    const float g2 = 2.0f * gam[grp];
    for (int w = tid; w < k * (DP / 4 + 1); w += GR_NT) {
        const int f = w / (DP / 4 + 1), q = w - f * (DP / 4 + 1);
        const int fid = grp * k + f;
        if (q == DP / 4) {          // phase, and clear the column sums
            float b;
            if (wz_step) b = wz_step[size_t(nv) * F + fid];
            else b = 6.283185307179586f * u01(philox4x32_10(make_uint4(fid, step, GR_STREAM_FEAT, uid), key).x);
            s.bc[f] = b; s.et[f] = 0; s.ep[f] = 0;
        } else {
            float v[4];
            if (wz_step) {
#pragma unroll
                for (int e = 0; e < 4; ++e) v[e] = (4 * q + e < nv) ? wz_step[size_t(4 * q + e) * F + fid] : 0.f;
            } else {
                const float4 n = normal4(philox4x32_10(make_uint4(fid, step, GR_STREAM_FEAT + 1 + q, uid), key));
                v[0] = g2 * n.x; v[1] = g2 * n.y; v[2] = g2 * n.z; v[3] = g2 * n.w;
#pragma unroll
                for (int e = 0; e < 4; ++e) if (4 * q + e >= nv) v[e] = 0.f;
            }
            *reinterpret_cast<float4*>(s.Wc + f * DP + 4 * q) = make_float4(v[0], v[1], v[2], v[3]);
        }
    }
}

// z = b + w . row  for a row given as DP/2 pairs
template <int DP>
__device__ __forceinline__ float gr_feature_dot(const float2 (&w)[DP / 2], const float (&r)[DP], float b) {
    float2 za = make_float2(b, 0.f), zb = make_float2(0.f, 0.f);
#pragma unroll
    for (int q = 0; q < DP / 2; ++q) {
        if (q & 1) zb = ffma2(w[q], make_float2(r[2 * q], r[2 * q + 1]), zb);
        else       za = ffma2(w[q], make_float2(r[2 * q], r[2 * q + 1]), za);
    }
    const float2 z = fadd2(za, zb);
    return z.x + z.y;
}

// column sums over the rows this CTA owns: work item = (feature, owned row block)
template <int DP>
__device__ __forceinline__ void gr_fourier_colsums(const GrSmem& s, const GraphArgs& a, const float* __restrict__ T, int N, int NB,
                                                   int crank, int csz, int tid) {
    const int k = a.fast_k;
    for (int w = tid; w < k * GR_NW; w += GR_NT) {
        const int slot = w / k, f = w - slot * k;
        const int rb = crank + csz * slot;
        if (rb >= NB) continue;
        float2 wf[DP / 2];
#pragma unroll
        for (int q = 0; q < DP / 4; ++q) {
            const float4 v = *reinterpret_cast<const float4*>(s.Wc + f * DP + 4 * q);
            wf[2 * q] = make_float2(v.x, v.y); wf[2 * q + 1] = make_float2(v.z, v.w);
        }
        const float b = s.bc[f];
        float st = 0.f, sp = 0.f;
        const int i1 = min(32 * rb + 32, N);
        for (int i = 32 * rb; i < i1; ++i) {
            float r[DP];
            gr_load_col_gmem<DP>(T, i, r);
            st += __cosf(gr_turn_reduce(gr_feature_dot<DP>(wf, r, b)));
            gr_load_col_smem<DP>(s.Pm, i, r);
            sp += __cosf(gr_turn_reduce(gr_feature_dot<DP>(wf, r, b)));
        }
        fx_add(&s.et[f], double(st));
        fx_add(&s.ep[f], double(sp));
    }
}

// per-row gradient of the current feature group, accumulated into acc (lane = row)
template <int DP>
__device__ __forceinline__ void gr_fourier_rowgrad(const GrSmem& s, int k, const float (&p)[DP], float2 (&acc)[DP / 2]) {
    for (int f = 0; f < k; ++f) {
        float2 wf[DP / 2];
#pragma unroll
        for (int q = 0; q < DP / 4; ++q) {
            const float4 v = *reinterpret_cast<const float4*>(s.Wc + f * DP + 4 * q);
            wf[2 * q] = make_float2(v.x, v.y); wf[2 * q + 1] = make_float2(v.z, v.w);
        }
        const float cs = s.coef[f] * __sinf(gr_turn_reduce(gr_feature_dot<DP>(wf, p, s.bc[f])));
        const float2 c2 = splat2(cs);
#pragma unroll
        for (int q = 0; q < DP / 2; ++q) acc[q] = ffma2(c2, wf[q], acc[q]);
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

// Sum of one double per CTA over the cluster (fixed order: deterministic); result valid in thread 0 of every CTA.
__device__ __forceinline__ double gr_cluster_sum(double v, double* red, int warp, int lane, cg::cluster_group& cl,
                                                 int csz, int crank) {
    double t = gr_block_sum(v, red, warp, lane);
    if (csz == 1) return t;
    if (warp == 0 && lane == 0)
        for (int r = 0; r < csz; ++r) cl.map_shared_rank(red, r)[GR_NW + 4 + crank] = t;
    cl.sync();
    double tot = 0.0;
    if (warp == 0 && lane == 0)
        for (int r = 0; r < csz; ++r) tot += red[GR_NW + 4 + r];
    cl.sync();
    return tot;
}

// REGEN (noise re-derived at every use, GraphArgs::noise_regen) is a separate instantiation of the whole kernel: the
// common buffered-noise kernel then carries no trace of it (a runtime switch cost it 2-5 % through register allocation).
template <int DP, int HP, int POL, bool REGEN>
__global__ void __launch_bounds__(GR_NT, 1) graph_kernel(const GraphArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ int s_inst;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int h = a.h, N = a.N, NB = (N + 31) >> 5;
    const int pp = (a.p_max + 3) & ~3;
    cg::cluster_group cl = cg::this_cluster();
    const int csz = int(cl.num_blocks()), crank = int(cl.block_rank());
    const int rb0 = crank + csz * warp, rbs = csz * GR_NW;        // this warp's row blocks of the MMD passes
    // one row block per warp, same warp in the forward and the backward pass, noise buffer not lent to the ring:
    // the noise the forward pass drew is still there when the backward pass needs it
    const bool noise_kept = csz == 1 && NB <= GR_NW && !a.t_ring;

    GrSmem s;
    {
        unsigned char* p = smem_raw;
        s.nn_pad = a.nbuf_cols;
        s.g64 = reinterpret_cast<long long*>(p);   p += size_t(pp) * 8;
        s.red = reinterpret_cast<double*>(p);      p += size_t(GR_NW + 4 + GR_MAX_CLUSTER) * 8;
        s.Pm = reinterpret_cast<float*>(p);        p += size_t(a.NP) * DP * 4;
        s.nbuf = reinterpret_cast<float*>(p);      p += size_t(GR_NW) * s.nn_pad * 32 * 4;
        s.theta = reinterpret_cast<float*>(p);     p += size_t(pp) * 4;
        s.am = reinterpret_cast<float*>(p);        p += size_t(pp) * 4;
        s.av = reinterpret_cast<float*>(p);        p += size_t(pp) * 4;
        s.desc = reinterpret_cast<GraphDesc*>(p);  p += (sizeof(GraphDesc) + 15) & ~size_t(15);
        const int kp = (a.fast_k + 3) & ~3;
        s.et = reinterpret_cast<long long*>(p);    p += size_t(kp) * 8;
        s.ep = reinterpret_cast<long long*>(p);    p += size_t(kp) * 8;
        s.Wc = reinterpret_cast<float*>(p);        p += size_t(kp) * DP * 4;
        s.bc = reinterpret_cast<float*>(p);        p += size_t(kp) * 4;
        s.coef = reinterpret_cast<float*>(p);      p += size_t(kp) * 4;
        s.Ts = reinterpret_cast<float*>(smem_raw + ((size_t(p - smem_raw) + 15) & ~size_t(15)));
    }
    float* nb = s.nbuf + size_t(warp) * s.nn_pad * 32;
    const int epochs = a.train_epochs + a.test_epochs;
    const float scale = 4.0f / (float(N) * float(N));
    const double invN2 = 1.0 / (double(N) * double(N));

    for (;;) {
        if (csz > 1) {     // CTA 0 of the cluster pulls the next instance and tells its peers
            cl.sync();
            if (crank == 0 && tid == 0) {
                const int v = atomicAdd(a.counter, 1);
                for (int r = 0; r < csz; ++r) *cl.map_shared_rank(&s_inst, r) = v;
            }
            cl.sync();
        } else {
            __syncthreads();
            if (tid == 0) s_inst = atomicAdd(a.counter, 1);
            __syncthreads();
        }
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
        if (a.t_smem) {
            const float4* src = reinterpret_cast<const float4*>(T);
            float4* dst = reinterpret_cast<float4*>(s.Ts);
            for (int q = tid; q < a.NP * DP / 4; q += GR_NT) dst[q] = __ldg(src + q);
        }
        __syncthreads();
        const int P = s.desc->n_params, PPAD = s.desc->n_params_pad;
        for (int q = tid; q < pp; q += GR_NT) { s.theta[q] = 0.f; s.am[q] = 0.f; s.av[q] = 0.f; s.g64[q] = 0; }
        __syncthreads();
        for (int q = tid; q < P; q += GR_NT) {
            float th;
            if (a.init_params) th = a.init_params[size_t(out) * a.ld_params + q];
            else {
                const float4 n = normal4(philox4x32_10(make_uint4(q >> 2, 0u, STREAM_INIT, uid), key));
                const int e = q & 3;
                th = a.init_std * (e == 0 ? n.x : e == 1 ? n.y : e == 2 ? n.z : n.w);
            }
            s.theta[gr_pad_index(*s.desc, q, h, HP)] = th;
        }
        if (tid == 0) { s.red[GR_NW + 1] = 1.0; s.red[GR_NW + 2] = 1.0; }
        // ---- TT (constant per instance) -------------------------------------------------------
        const bool fast = a.fast_k > 0;
        const float* wz = (fast && a.wz_ext) ? a.wz_ext + size_t(out) * epochs * (s.desc->n_var + 1) * 7 * a.fast_k : nullptr;
        if (!fast) {
            double part = 0.0;
            for (int rb = rb0; rb < NB; rb += rbs) {
                const int i = 32 * rb + lane;
                float2 p[DP / 2];                                // the row, negated (see gr_grad_pair)
                gr_load_pairs<DP, true>(T + size_t(min(i, N - 1)) * DP, p);
#pragma unroll
                for (int q = 0; q < DP / 2; ++q) p[q] = make_float2(-p[q].x, -p[q].y);
                const double v = gr_mmd_eval_row_sym<DP, POL, true>(T, N, NB, rb, p);
                if (i < N) part += v;
            }
            const double tt = gr_cluster_sum(part, s.red, warp, lane, cl, csz, crank);
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
                if (!REGEN) gr_fill_noise(s, a, nb, i, lane, it, uid, key, noise_step);
                gr_forward_block<DP, HP, REGEN>(s, GrNoise{nb, noise_step, i, N, lane, (unsigned)it, uid, key}, i);
                __syncwarp();
            }
            __syncthreads();
            if (a.dbg_pred && train)
                for (int q = tid; q < N * s.desc->n_var; q += GR_NT) {
                    const int r = q / s.desc->n_var, c = q - r * s.desc->n_var;
                    a.dbg_pred[(size_t(it) * N + r) * s.desc->n_var + c] = s.Pm[r * DP + c];
                }
            if (fast) {
                // ---- random-feature loss: seven bandwidth groups, the row's gradient stays in registers ----------
                const int i = 32 * rb0 + lane;
                const bool has = rb0 < NB;                   // this warp's only row block (host: NB <= csz * GR_NW)
                float pf[DP];
                float2 acc[DP / 2];
#pragma unroll
                for (int q = 0; q < DP / 2; ++q) acc[q] = make_float2(0.f, 0.f);
                if (has) gr_load_col_smem<DP>(s.Pm, i, pf);
                else {
#pragma unroll
                    for (int c = 0; c < DP; ++c) pf[c] = 0.f;
                }
                const float* wz_step = wz ? wz + size_t(it) * (s.desc->n_var + 1) * 7 * a.fast_k : nullptr;
                const double norm = sqrt(2.0 / double(a.fast_k)) / double(N);
                double lpart = 0.0;
                for (int grp = 0; grp < 7; ++grp) {
                    __syncthreads();                          // the previous group's features are no longer in use
                    gr_fourier_features<DP>(s, a, grp, it, uid, key, wz_step, tid);
                    __syncthreads();
                    gr_fourier_colsums<DP>(s, a, T, N, NB, crank, csz, tid);
                    if (csz > 1) cl.sync(); else __syncthreads();
                    for (int f = tid; f < a.fast_k; f += GR_NT) {
                        long long te = 0, pe = 0;
                        for (int r = 0; r < csz; ++r) { te += cl.map_shared_rank(s.et, r)[f]; pe += cl.map_shared_rank(s.ep, r)[f]; }
                        const double diff = norm * (double(te) - double(pe)) * CGNN_FX_INV;      // e1_f - e2_f
                        lpart += diff * diff;
                        s.coef[f] = float(2.0 * diff * norm);
                    }
                    if (csz > 1) cl.sync(); else __syncthreads();
                    if (train && has) gr_fourier_rowgrad<DP>(s, a.fast_k, pf, acc);
                }
                const double sum = gr_block_sum(lpart, s.red, warp, lane);       // the same in every CTA of the cluster
                if (tid == 0) {
                    const float loss = float(sum);
                    if (!train) score_sum += double(loss);
                    if (a.trace && crank == 0) a.trace[size_t(out) * epochs + it] = loss;
                    if (a.dbg_loss && train) a.dbg_loss[it] = loss;
                }
                if (!train) continue;
                float gl[DP];
#pragma unroll
                for (int q = 0; q < DP / 2; ++q) {
                    gl[2 * q] = (has && i < N) ? acc[q].x : 0.f;
                    gl[2 * q + 1] = (has && i < N) ? acc[q].y : 0.f;
                }
                if (a.dbg_dldp && has && i < N)
                    for (int c = 0; c < s.desc->n_var; ++c) a.dbg_dldp[(size_t(it) * N + i) * s.desc->n_var + c] = gl[c];
                if (has) {
                    if (!REGEN && !noise_kept) gr_fill_noise(s, a, nb, i, lane, it, uid, key, noise_step);
                    gr_backward_block<DP, HP, REGEN>(s, GrNoise{nb, noise_step, i, N, lane, (unsigned)it, uid, key}, i, lane, gl);
                    __syncwarp();
                }
            } else {
                // ---- loss (evaluation steps; traced training steps) ---------------------------------
                const bool want_loss = !train || a.dbg_loss || (a.trace && a.trace_every > 0 && it % a.trace_every == 0);
                if (want_loss) {
                    double part = 0.0;
                    for (int rb = rb0; rb < NB; rb += rbs) {
                        const int i = 32 * rb + lane;
                        float2 p[DP / 2];
                        gr_load_pairs<DP, false>(s.Pm + size_t(i) * DP, p);
    #pragma unroll
                        for (int q = 0; q < DP / 2; ++q) p[q] = make_float2(-p[q].x, -p[q].y);
                        const double v = gr_mmd_eval_row_sym<DP, POL, false>(s.Pm, N, NB, rb, p) -
                                         2.0 * (a.t_smem ? gr_mmd_eval_row<DP, POL, false>(s.Ts, N, p)
                                                : a.t_ring ? gr_mmd_eval_row_ring<DP, POL>(T, nb, N, lane, p) : gr_mmd_eval_row<DP, POL, true>(T, N, p));
                        if (i < N) part += v;
                    }
                    const double sum = gr_cluster_sum(part, s.red, warp, lane, cl, csz, crank);
                    if (tid == 0) {
                        const float loss = float((sum + s.red[GR_NW]) * invN2);
                        if (!train) score_sum += double(loss);
                        if (a.trace && crank == 0) a.trace[size_t(out) * epochs + it] = loss;
                        if (a.dbg_loss && train) a.dbg_loss[it] = loss;
                    }
                } else if (a.trace && tid == 0 && crank == 0) {
                    a.trace[size_t(out) * epochs + it] = __int_as_float(0x7fc00000);
                }
                if (!train) continue;
                // ---- MMD gradient + generator backward, row block by row block ------------------------
                for (int rb = rb0; rb < NB; rb += rbs) {
                    const int i = 32 * rb + lane;
                    float gl[DP];
                    {
                        float2 p[DP / 2], tot[DP / 2];
                        gr_load_pairs<DP, false>(s.Pm + size_t(i) * DP, p);
    #pragma unroll
                        for (int q = 0; q < DP / 2; ++q) p[q] = make_float2(-p[q].x, -p[q].y);
                        if (a.t_smem) gr_mmd_grad_row<DP, POL, GR_T_SMEM>(s.Pm, s.Ts, nb, N, lane, p, tot);
                        else if (a.t_ring) gr_mmd_grad_row<DP, POL, GR_T_RING>(s.Pm, T, nb, N, lane, p, tot);
                        else gr_mmd_grad_row<DP, POL, GR_T_GMEM>(s.Pm, T, nb, N, lane, p, tot);
    #pragma unroll
                        for (int q = 0; q < DP / 2; ++q) {                   // tot holds the negated sum
                            gl[2 * q] = (i < N) ? tot[q].x * -scale : 0.f;
                            gl[2 * q + 1] = (i < N) ? tot[q].y * -scale : 0.f;
                        }
                    }
                    if (a.dbg_dldp && i < N)
                        for (int c = 0; c < s.desc->n_var; ++c) a.dbg_dldp[(size_t(it) * N + i) * s.desc->n_var + c] = gl[c];
                    if (!REGEN && !noise_kept) gr_fill_noise(s, a, nb, i, lane, it, uid, key, noise_step);
                    gr_backward_block<DP, HP, REGEN>(s, GrNoise{nb, noise_step, i, N, lane, (unsigned)it, uid, key}, i, lane, gl);
                    __syncwarp();
                }
            }
            // ---- Adam --------------------------------------------------------------------------
            if (csz > 1) {
                // every CTA sums the fixed-point accumulators of the whole cluster (integer: exact, order-free)
                // and applies the same update to its own copy of the parameters
                cl.sync();
                const float lr_t = float(s.red[GR_NW + 3]);
                for (int q = tid; q < PPAD; q += GR_NT) {
                    long long tot = 0;
                    for (int r = 0; r < csz; ++r) tot += cl.map_shared_rank(s.g64, r)[q];
                    const float g = float(double(tot) * CGNN_FX_INV);
                    float m = s.am[q], v = s.av[q];
                    const float th = adam_update(s.theta[q], g, m, v, lr_t);
                    s.am[q] = m; s.av[q] = v; s.theta[q] = th;
                }
                cl.sync();                                   // peers have read this CTA's accumulators
                for (int q = tid; q < PPAD; q += GR_NT) s.g64[q] = 0;
            } else {
                __syncthreads();
                const float lr_t = float(s.red[GR_NW + 3]);
                if (a.dbg_grad)
                    for (int q = tid; q < P; q += GR_NT)
                        a.dbg_grad[size_t(it) * P + q] = float(fx_get(&s.g64[gr_pad_index(*s.desc, q, h, HP)]));
                if (a.dbg_grad) __syncthreads();
                for (int q = tid; q < PPAD; q += GR_NT) {      // pad entries: gradient 0, moments 0 -> they stay 0
                    const float g = float(fx_get(&s.g64[q]));
                    s.g64[q] = 0;
                    float m = s.am[q], v = s.av[q];
                    const float th = adam_update(s.theta[q], g, m, v, lr_t);
                    s.am[q] = m; s.av[q] = v; s.theta[q] = th;
                }
                if (a.dbg_params) {
                    __syncthreads();
                    for (int q = tid; q < P; q += GR_NT) a.dbg_params[size_t(it) * P + q] = s.theta[gr_pad_index(*s.desc, q, h, HP)];
                }
            }
        }
        if (a.params_out && crank == 0) {   // evaluation does not touch the parameters: these are the trained ones
            __syncthreads();
            for (int q = tid; q < P; q += GR_NT)
                a.params_out[size_t(out) * a.ld_params + q] = s.theta[gr_pad_index(*s.desc, q, h, HP)];
        }
        if (tid == 0 && crank == 0 && a.scores) a.scores[out] = a.test_epochs > 0 ? float(score_sum / double(a.test_epochs)) : 0.f;
    }
}

}  // namespace cgnn
