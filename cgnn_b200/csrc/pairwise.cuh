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
// instance start and once at the end).  Coordinates are pre-multiplied by S = sqrt(0.25*log2 e)
// (common.cuh, "primed" kernels) and kept as one float4 per row,
//   Q[j] = (S x_j, S y_j, S yhat_j, 0)      x = cause (copied into the generated pair, GNN.py:70-72)
// in a DOUBLED-TILE layout: rows are cut into tiles of 32 and every tile is stored twice in a row,
// Qd[64 J + t] = Q[32 J + (t & 31)], t = 0..63.  Lane l then reaches "its" column of step s,
// (l + s) & 31, at the fixed offset l + s: the skewed walk needs no index arithmetic at all.
//   xs[NP], es[NP], yh[NP]  unscaled cause, current noise and generated value (generator phases)
//   g64[NP] int64           fixed-point accumulator of dLoss/d yhat (primed, unscaled)
//   padded rows hold x = 1e15 so that every kernel value against them is exactly 0.
// Rows sit in Qd (and g64) in ascending order of the CAUSE coordinate, rank[i] = sorted position of row i; the
// per-row generator arrays xs/es/yh stay in the caller's order, so noise streams and outputs are keyed by the
// caller's row index.  The cause column is copied into the generated pair, so (x_i - x_j)^2 of any two rows
// never changes during training: a tile pair whose x ranges are further apart than CGNN_FAR_DP (primed units)
// cannot see the gamma = 50 kernel (50 e^{-50 D} < 5e-11 against the other six, < 1e-8 relative) and is walked
// with two instead of three MUFU ex2 per evaluation ("far" variant, about 58 % of the tile pairs of a
// standardised pair).
//
// MMD mapping ("trio"): for an unordered pair of DIFFERENT tiles (I, J), lane l owns row i = 32 I + l
// and meets column j = 32 J + ((l+s)&31) at step s.  The three kernel evaluations that involve the
// two rows - PP(i,j), PT(i,j) and PT(j,i) - share (x_i - x_j)^2, which is therefore computed once
// per three evaluations; the column-side gradient (from PP and from PT(j,i)) rides in ONE register
// that is rotated with one shuffle per step, so after 32 steps lane l owns column l's total, flushed
// with one fixed-point shared atomic.  Diagonal tiles use a "duo" walk (PP(i,j), PT(i,j)) over all
// ordered pairs.  Per evaluation: ~25 issue slots of which 3-7 MUFU (policy dependent).
//
// The same source is compiled three times (pairwise_inst.cu: -DCGNN_PW_NT=512|256|128 threads per CTA with 2|4|8 CTAs per
// SM, each in its own namespace): a 100-row pair has 10 tile-pair units and 25 busy generator threads - 16 warps
// cannot share that, four CTAs of 8 warps (or eight of 4) per SM can.  The host picks the build per PAIR from its row
// count alone, so a pair's score never depends on what else is in the batch.
#pragma once
#include "common.cuh"
#include "pairwise_args.h"

#ifndef CGNN_PW_NS
#define CGNN_PW_NS pw512
#endif

namespace cgnn {
namespace CGNN_PW_NS {

#ifndef CGNN_PW_NT
#define CGNN_PW_NT 512
#endif
#ifndef CGNN_PW_CTAS
#define CGNN_PW_CTAS 2
#endif
#ifndef CGNN_TRIO_UNROLL
#define CGNN_TRIO_UNROLL 2
#endif
#define CGNN_PRAGMA(x) _Pragma(#x)
#define CGNN_UNROLL(n) CGNN_PRAGMA(unroll n)
constexpr int PW_NT = CGNN_PW_NT;       // threads per CTA
constexpr int PW_NW = PW_NT / 32;       // warps per CTA
constexpr int PW_MIN_CTAS = CGNN_PW_CTAS;
constexpr float PW_PAD = 1.0e15f;


struct PwSmem {
    float4* Qd; float* xs; float* es; float* yh; long long* g64; unsigned short* rank;
    float* theta; float* am; float* av; float* part; double* red;
    int part_stride;
    float4* feat; long long* et; long long* ep;      // Fourier mode only: features (wx/S, wy/S, phase, coef), column sums
};

__host__ __device__ inline size_t pw_smem_bytes(int np_max, int h, int fast_k = 0) {
    int nchunk = (h + 7) / 8;
    int P = 4 * h + 1;
    int Pp = (P + 3) & ~3;
    size_t b = 0;
    b += size_t(np_max) * 2 * 16;                // Qd (doubled tiles)
    b += size_t(np_max) * 8;                     // g64
    b += size_t(PW_NW + 4) * 8;                  // red[], TT, beta1^t, beta2^t, lr_t
    b += size_t(np_max) * 4 * 3;                 // xs, es, yh
    b += size_t(np_max) * 2;                     // rank
    b += size_t(Pp) * 4 * 3;                     // theta, m, v
    b += size_t(PW_NW) * (nchunk * 32 + 4) * 4;  // per-warp partial gradient sums
    b = (b + 15) & ~size_t(15);
    b += size_t(7 * fast_k) * (16 + 8 + 8);      // Fourier mode: feat, et, ep
    return (b + 15) & ~size_t(15);
}

// position of row i in the doubled-tile layout (first copy; the second is +32)
__device__ __forceinline__ int pw_slot_of_row(int i) { return ((i >> 5) << 6) + (i & 31); }

// ---- unit partition -------------------------------------------------------------------------
// Row block I owns its diagonal tile and npp(I) off-diagonal tiles J = (I+1+q) mod NB.  For odd NB
// every block has (NB-1)/2 of them; for even NB the antipodal pair (I, I+NB/2) goes to the lower I.
// Units are flattened I-major and dealt to the warps in equal contiguous ranges.
__device__ __forceinline__ int pw_npp(int I, int NB) {
    return (NB & 1) ? (NB - 1) / 2 : (NB / 2 - 1 + (I < NB / 2 ? 1 : 0));
}

// Off-diagonal tile pair, gradient form.  tile = &Qd[64 J + lane].  Returns the row-side sum for row i;
// colsum receives the total for column 32 J + lane (both in primed units, unscaled).  Two steps of the
// walk are evaluated together: their six kernel evaluations run as three FP32x2-packed chains.
template <int POL, int FAR>
__device__ __forceinline__ float trio_train(float xi, float yi, float yhi, const float4* __restrict__ tile, int lane,
                                            float& colsum) {
    float2 ra = make_float2(0.f, 0.f);
    float ca = 0.f;
    const int src = (lane + 1) & 31;
    CGNN_UNROLL(CGNN_TRIO_UNROLL)
    for (int s = 0; s < 32; s += 2) {
        const float4 c0 = tile[s], c1 = tile[s + 1];
        const float dx0 = xi - c0.x, dx1 = xi - c1.x;
        const float e0 = dx0 * dx0, e1 = dx1 * dx1;
        const float2 a = make_float2(yhi - c0.z, yhi - c1.z);      // PP(i,j):  yhat_i - yhat_j
        const float2 b = make_float2(yhi - c0.y, yhi - c1.y);      // PT(i,j):  yhat_i - y_j
        const float2 d = make_float2(c0.z - yi, c1.z - yi);        // PT(j,i):  yhat_j - y_i
        const float2 wpp = wsum_p2<POL, FAR>(make_float2(fmaf(a.x, a.x, e0), fmaf(a.y, a.y, e1)));
        const float2 w1 = wsum_p2<POL, FAR>(make_float2(fmaf(b.x, b.x, e0), fmaf(b.y, b.y, e1)));
        const float2 w2 = wsum_p2<POL, FAR>(make_float2(fmaf(d.x, d.x, e0), fmaf(d.y, d.y, e1)));
        const float2 wa = fmul2(wpp, a);
        ra = fsub2(ffma2(w1, b, ra), wa);
        const float2 cc = ffma2(w2, d, wa);
        ca = __shfl_sync(0xffffffffu, ca + cc.x, src);
        ca = __shfl_sync(0xffffffffu, ca + cc.y, src);
    }
    colsum = ca;
    return ra.x + ra.y;
}

// Diagonal tile, gradient form: all ordered pairs of the tile, row side only (PT and PP of a step packed).
template <int POL>
__device__ __forceinline__ float duo_train(float xi, float yhi, const float4* __restrict__ tile) {
    float2 ra = make_float2(0.f, 0.f);
#pragma unroll 4
    for (int s = 0; s < 32; ++s) {
        const float4 c = tile[s];
        const float dx = xi - c.x;
        const float e = dx * dx;
        const float2 ab = make_float2(yhi - c.y, c.z - yhi);       // (PT: yhat_i - y_j,  PP with the sign folded in)
        const float2 w = wsum_p2<POL>(make_float2(fmaf(ab.x, ab.x, e), fmaf(ab.y, ab.y, e)));
        ra = ffma2(w, ab, ra);
    }
    return ra.x + ra.y;
}

// Loss forms: kPP - kPT(i,j) - kPT(j,i) per unordered pair (to be doubled), kPP - 2 kPT per ordered pair.
template <int POL, int FAR>
__device__ __forceinline__ float trio_eval(float xi, float yi, float yhi, const float4* __restrict__ tile) {
    float2 acc = make_float2(0.f, 0.f);
    CGNN_UNROLL(CGNN_TRIO_UNROLL)
    for (int s = 0; s < 32; s += 2) {
        const float4 c0 = tile[s], c1 = tile[s + 1];
        const float dx0 = xi - c0.x, dx1 = xi - c1.x;
        const float e0 = dx0 * dx0, e1 = dx1 * dx1;
        const float a0 = yhi - c0.z, a1 = yhi - c1.z, b0 = yhi - c0.y, b1 = yhi - c1.y, d0 = c0.z - yi, d1 = c1.z - yi;
        const float2 kpp = ksum_p2<POL, FAR>(make_float2(fmaf(a0, a0, e0), fmaf(a1, a1, e1)));
        const float2 k1 = ksum_p2<POL, FAR>(make_float2(fmaf(b0, b0, e0), fmaf(b1, b1, e1)));
        const float2 k2 = ksum_p2<POL, FAR>(make_float2(fmaf(d0, d0, e0), fmaf(d1, d1, e1)));
        acc = fadd2(acc, fsub2(kpp, fadd2(k1, k2)));
    }
    return acc.x + acc.y;
}
template <int POL>
__device__ __forceinline__ float duo_eval(float xi, float yhi, const float4* __restrict__ tile) {
    float acc = 0.f;
#pragma unroll 4
    for (int s = 0; s < 32; ++s) {
        const float4 c = tile[s];
        const float dx = xi - c.x;
        const float e = dx * dx;
        const float a = yhi - c.z, b = yhi - c.y;
        const float2 k = ksum_p2<POL>(make_float2(fmaf(a, a, e), fmaf(b, b, e)));
        acc += fmaf(-2.f, k.y, k.x);
    }
    return acc;
}

// One MMD pass over an instance.
//   TRAIN: accumulates the primed, unscaled dLoss/d yhat into g64
//   else : returns this lane's share of  PP - 2 PT
template <int POL, bool TRAIN>
__device__ double mmd_pass(const float4* __restrict__ Qd, long long* g64, int N, int NB, int warp, int lane) {
    int total = 0;
    for (int I = 0; I < NB; ++I) total += 1 + pw_npp(I, NB);
    const int lo = int((long long)warp * total / PW_NW);
    const int hi = int((long long)(warp + 1) * total / PW_NW);
    int I = 0, base = 0;
    while (I < NB && base + 1 + pw_npp(I, NB) <= lo) { base += 1 + pw_npp(I, NB); ++I; }
    double lsum = 0.0;
    int u = lo;
    while (u < hi) {
        const int UI = 1 + pw_npp(I, NB);
        const int a0 = u - base;
        const int b0 = min(hi - base, UI);
        const int i = 32 * I + lane;
        const float4 me = Qd[64 * I + lane];
        double acc = 0.0;
        for (int q = a0; q < b0; ++q) {
            if (q == 0) {
                const float4* tile = Qd + 64 * I + lane;
                if (TRAIN) acc += double(duo_train<POL>(me.x, me.z, tile));
                else       acc += double(duo_eval<POL>(me.x, me.z, tile));
            } else {
                int J = I + q;
                if (J >= NB) J -= NB;
                const float4* tile = Qd + 64 * J + lane;
                // rows are sorted by x: the gap between the x ranges of the two tiles bounds every (x_i - x_j)^2
                const float gap = (J > I) ? Qd[64 * J].x - Qd[64 * I + 31].x : Qd[64 * I].x - Qd[64 * J + 31].x;
                const float g2 = gap > 0.f ? gap * gap : 0.f;
                const int far = g2 > CGNN_FAR2_DP ? 2 : g2 > CGNN_FAR_DP ? 1 : 0;
                if (TRAIN) {
                    float cs;
                    acc += double(far == 2 ? trio_train<POL, 2>(me.x, me.y, me.z, tile, lane, cs)
                                  : far == 1 ? trio_train<POL, 1>(me.x, me.y, me.z, tile, lane, cs)
                                             : trio_train<POL, 0>(me.x, me.y, me.z, tile, lane, cs));
                    if (32 * J + lane < N) fx_add(&g64[32 * J + lane], double(cs));
                } else {
                    acc += 2.0 * double(far == 2 ? trio_eval<POL, 2>(me.x, me.y, me.z, tile)
                                        : far == 1 ? trio_eval<POL, 1>(me.x, me.y, me.z, tile) : trio_eval<POL, 0>(me.x, me.y, me.z, tile));
                }
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

// TT = sum_ij sum_g k(T_i, T_j): once per instance, plain row x all-columns walk (broadcast loads).
template <int POL>
__device__ double tt_pass(const float4* __restrict__ Qd, int N, int NB, int warp, int lane) {
    double lsum = 0.0;
    for (int I = warp; I < NB; I += PW_NW) {
        const float4 me = Qd[64 * I + lane];
        double acc = 0.0;
        for (int J = 0; J < NB; ++J) {
            float part = 0.f;
#pragma unroll 4
            for (int t = 0; t < 32; ++t) {
                const float4 c = Qd[64 * J + t];
                const float dx = me.x - c.x, dy = me.y - c.y;
                part += ksum_p<POL>(fmaf(dy, dy, dx * dx));
            }
            acc += double(part);
        }
        if (32 * I + lane < N) lsum += acc;
    }
    return lsum;
}

// ---- random-feature ("Fourier") MMD: Loss.py:35-56 ---------------------------------------------
// F = 7k features f with frequency w_f = 2 gamma_g N(0,1)^2 (g = f / k) and phase b_f = U(0, 2 pi), redrawn
// at every step (they are graph ops in the reference).  loss = sum_f (e1_f - e2_f)^2 with
// e_f = sqrt(2/k) mean_i cos(w_f . row_i + b_f) over the observed (e1) and the generated (e2) rows.
constexpr uint32_t STREAM_FEAT = 7u;
constexpr int PW_FCHUNKS = 8;            // row chunks per feature in the column-sum pass

__device__ __forceinline__ float pw_turn_reduce(float z) {       // z radians -> equivalent angle in [-pi, pi]
    // round-to-nearest by the 1.5*2^23 trick: two FADDs on the FMA pipe instead of an FRND on the (busy) XU pipe;
    // exact for |t| < 2^22, i.e. |z| < 2.6e7 - far beyond any frequency x coordinate product of this loss
    const float t = z * 0.15915494309189535f;
    const float r = __fsub_rn(__fadd_rn(t, 12582912.0f), 12582912.0f);
    return (t - r) * 6.283185307179586f;
}

// draw (or fetch) the features of this step and clear the column sums
__device__ __forceinline__ void pw_fourier_features(const PwSmem& s, const PairwiseArgs& a, int F, int step,
                                                    unsigned int uid, uint2 key, const float* wz_step, int tid) {
    const float gam[7] = {G0, G1, G2, G3, G4, G5, G6};
    for (int f = tid; f < F; f += PW_NT) {
        float wx, wy, b;
        if (wz_step) { wx = wz_step[f]; wy = wz_step[F + f]; b = wz_step[2 * F + f]; }
        else {
            const uint4 r = philox4x32_10(make_uint4(f, step, STREAM_FEAT, uid), key);
            const float4 n = normal4(r);
            const float g2 = 2.0f * gam[f / a.fast_k];
            wx = g2 * n.x; wy = g2 * n.y;
            b = 6.283185307179586f * u01(r.z ^ 0x9E3779B9u);
        }
        s.feat[f] = make_float4(wx * (1.0f / CGNN_PRESCALE), wy * (1.0f / CGNN_PRESCALE), b, wy);
        s.et[f] = 0; s.ep[f] = 0;
    }
}

// column sums of cos over the rows: work item = (feature, row chunk); fixed-point accumulation
__device__ __forceinline__ void pw_fourier_colsums(const PwSmem& s, int N, int NB, int F, int tid) {
    const int tiles_per_chunk = (NB + PW_FCHUNKS - 1) / PW_FCHUNKS;
    for (int w = tid; w < F * PW_FCHUNKS; w += PW_NT) {
        const int chunk = w / F, f = w - chunk * F;
        const float4 ft = s.feat[f];
        float st = 0.f, sp = 0.f;
        const int J0 = chunk * tiles_per_chunk, J1 = min(J0 + tiles_per_chunk, NB);
        for (int J = J0; J < J1; ++J) {
            const int nt = min(32, N - 32 * J);
#pragma unroll 4
            for (int t = 0; t < nt; ++t) {
                const float4 c = s.Qd[64 * J + t];
                const float zx = fmaf(ft.x, c.x, ft.z);
                st += __cosf(pw_turn_reduce(fmaf(ft.y, c.y, zx)));
                sp += __cosf(pw_turn_reduce(fmaf(ft.y, c.z, zx)));
            }
        }
        if (J1 > J0) { fx_add(&s.et[f], double(st)); fx_add(&s.ep[f], double(sp)); }
    }
}

// e1, e2 -> loss share of this thread; the gradient coefficient replaces feat.w
__device__ __forceinline__ double pw_fourier_coefs(const PwSmem& s, int N, int F, int k, int tid) {
    const double norm = sqrt(2.0 / double(k)) / double(N);
    double part = 0.0;
    for (int f = tid; f < F; f += PW_NT) {
        const double diff = norm * (fx_get(&s.et[f]) - fx_get(&s.ep[f]));       // e1_f - e2_f
        part += diff * diff;
        s.feat[f].w = float(2.0 * diff * norm) * s.feat[f].w;   // dLoss/d yhat_i = sum_f coef_f sin(z_if)
    }
    return part;
}

// dLoss/d yhat_i for every row: thread = row, all features broadcast from shared memory
__device__ __forceinline__ void pw_fourier_rowgrad(const PwSmem& s, int N, int F, int tid) {
    for (int i = tid; i < N; i += PW_NT) {
        const float4 c = s.Qd[pw_slot_of_row(i)];
        float g = 0.f;
#pragma unroll 4
        for (int f = 0; f < F; ++f) {
            const float4 ft = s.feat[f];
            const float z = fmaf(ft.y, c.z, fmaf(ft.x, c.x, ft.z));
            g = fmaf(ft.w, __sinf(pw_turn_reduce(z)), g);
        }
        fx_add(&s.g64[i], double(g));
    }
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

// Generator forward for 4-row quads (GNN.py:64-67), noise from Philox or the caller.
__device__ __forceinline__ void pw_forward(const PwSmem& s, const PairwiseArgs& a, int N, int NP, int h, int step,
                                           unsigned int uid, uint2 key, const float* noise_row, int tid) {
    for (int q = tid; 4 * q < NP; q += PW_NT) {
        const int r0 = 4 * q;
        s.g64[r0] = 0; s.g64[r0 + 1] = 0; s.g64[r0 + 2] = 0; s.g64[r0 + 3] = 0;
        if (r0 >= N) continue;
        float e[4];
        if (noise_row) {
#pragma unroll
            for (int r = 0; r < 4; ++r) e[r] = (r0 + r < N) ? noise_row[r0 + r] : 0.f;
        } else {
            const float4 n = normal4(philox4x32_10(make_uint4(q, step, STREAM_NOISE, uid), key));
            e[0] = n.x; e[1] = n.y; e[2] = n.z; e[3] = n.w;
        }
        float x[4], y[4];
        const float bout = s.theta[4 * h];
#pragma unroll
        for (int r = 0; r < 4; ++r) { x[r] = s.xs[r0 + r]; y[r] = bout; }
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
            if (r0 + r < N) {
                const int sl = pw_slot_of_row(s.rank[r0 + r]);
                const float yp = CGNN_PRESCALE * y[r];
                s.Qd[sl].z = yp; s.Qd[sl + 32].z = yp;
                s.es[r0 + r] = e[r]; s.yh[r0 + r] = y[r];
            }
    }
}

// Generator backward: per-warp partial parameter gradients into s.part (deterministic tree).
__device__ __forceinline__ void pw_backward_partials(const PwSmem& s, int N, int NP, int h, int tid, int warp, int lane,
                                                     float scale) {
    const int nchunk = (h + 7) >> 3;
    float* mine = s.part + warp * s.part_stride;
    for (int c = 0; c < nchunk; ++c) mine[32 * c + lane] = 0.f;
    __syncwarp();
    float sd_tot = 0.f;
    for (int q0 = 0; 4 * q0 < NP; q0 += PW_NT) {          // whole warps iterate together (uniform trip count)
        if (4 * (q0 + (tid & ~31)) >= N) continue;        // this warp's 128 rows are all padding: nothing to add
        const int r0 = 4 * (q0 + tid);
        float x[4], e[4], d[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const bool ok = r0 + r < N;
            x[r] = ok ? s.xs[r0 + r] : 0.f;
            e[r] = ok ? s.es[r0 + r] : 0.f;
            d[r] = ok ? float(fx_get(&s.g64[s.rank[r0 + r]])) * scale : 0.f;
        }
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
            mine[32 * c + lane] += warp_reduce32(v, lane);
        }
        sd_tot += (d[0] + d[1]) + (d[2] + d[3]);
    }
    const float sd = warp_sum(sd_tot);
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
        s.Qd = reinterpret_cast<float4*>(p);      p += size_t(a.np_max) * 32;
        s.g64 = reinterpret_cast<long long*>(p);  p += size_t(a.np_max) * 8;
        s.red = reinterpret_cast<double*>(p);     p += size_t(PW_NW + 4) * 8;
        s.xs = reinterpret_cast<float*>(p);       p += size_t(a.np_max) * 4;
        s.es = reinterpret_cast<float*>(p);       p += size_t(a.np_max) * 4;
        s.yh = reinterpret_cast<float*>(p);       p += size_t(a.np_max) * 4;
        s.rank = reinterpret_cast<unsigned short*>(p);  p += size_t(a.np_max) * 2;
        s.theta = reinterpret_cast<float*>(p);    p += size_t(Pp) * 4;
        s.am = reinterpret_cast<float*>(p);       p += size_t(Pp) * 4;
        s.av = reinterpret_cast<float*>(p);       p += size_t(Pp) * 4;
        s.part = reinterpret_cast<float*>(p);
        s.part_stride = nchunk * 32 + 4;
        size_t used = size_t(p - smem_raw) + size_t(PW_NW) * s.part_stride * 4;
        used = (used + 15) & ~size_t(15);
        s.feat = reinterpret_cast<float4*>(smem_raw + used);
        s.et = reinterpret_cast<long long*>(smem_raw + used + size_t(7 * a.fast_k) * 16);
        s.ep = s.et + 7 * a.fast_k;
    }
    const int F = 7 * a.fast_k;
    const bool fast = a.fast_k > 0;
    double& s_TT = s.red[PW_NW];

    for (;;) {
        __syncthreads();
        if (tid == 0) s_inst = atomicAdd(a.counter, 1);
        __syncthreads();
        const int it_idx = s_inst;
        if (it_idx >= a.n_inst) break;
        const int4 inst = a.inst[it_idx];
        const int pair = inst.x, dir = inst.z, out = inst.w;
        const unsigned long long uid64 = a.uid_base + (unsigned long long)out;
        const unsigned int uid = (unsigned int)(uid64 & 0xffffffffull);
        const uint2 key = make_uint2(a.key.x, a.key.y ^ (unsigned int)(uid64 >> 32));
        const int N = a.n_rows[pair];
        const int NB = (N + 31) >> 5, NP = NB * 32;
        const float* src = a.xy + size_t(pair) * a.ld_rows * 2;
        const int epochs = a.train_epochs + a.test_epochs;
        const float* noise = a.noise ? a.noise + size_t(out) * epochs * a.ld_rows : nullptr;

        // ---- load the pair (cause first), prescale, pad, duplicate every tile -------------------
        for (int i = tid; i < NP; i += PW_NT) {
            float x = PW_PAD;
            if (i < N) { const float2 v = reinterpret_cast<const float2*>(src)[i]; x = dir ? v.y : v.x; }
            s.xs[i] = (x != x) ? __int_as_float(0x7f800000) : x;      // sort key: NaN rows go last, every rank stays unique
            const int sl = pw_slot_of_row(i);
            const float4 pad = make_float4(PW_PAD, 0.f, 0.f, 0.f);
            s.Qd[sl] = pad; s.Qd[sl + 32] = pad;
        }
        __syncthreads();
        // rank of every row by (x, index): counting sort, N^2 / 512 comparisons per thread, once per instance
        for (int i = tid; i < NP; i += PW_NT) {
            int r = i;
            if (i < N) {
                const float xi = s.xs[i];
                r = 0;
                for (int j = 0; j < N; ++j) {
                    const float xj = s.xs[j];
                    r += (xj < xi || (xj == xi && j < i)) ? 1 : 0;
                }
            }
            s.rank[i] = (unsigned short)r;
        }
        __syncthreads();
        for (int i = tid; i < NP; i += PW_NT) {
            float x = 0.f;
            if (i < N) {
                const float2 v = reinterpret_cast<const float2*>(src)[i];
                x = dir ? v.y : v.x;
                const float y = dir ? v.x : v.y;
                const int sl = pw_slot_of_row(s.rank[i]);
                const float4 t = make_float4(CGNN_PRESCALE * x, CGNN_PRESCALE * y, 0.f, 0.f);
                s.Qd[sl] = t; s.Qd[sl + 32] = t;
            }
            s.es[i] = 0.f; s.yh[i] = 0.f;
        }
        __syncthreads();                                                     // all ranks computed from the keys
        for (int i = tid; i < NP; i += PW_NT) {                              // keys -> the cause values the generator reads
            float x = 0.f;
            if (i < N) { const float2 v = reinterpret_cast<const float2*>(src)[i]; x = dir ? v.y : v.x; }
            s.xs[i] = x;
        }
        // ---- parameters (GNN.py:21-29, 55-58) and Adam state -------------------------------
        for (int p = tid; p < Pp; p += PW_NT) {
            float th = 0.f;
            if (p < P) {
                if (a.init_params) th = a.init_params[size_t(out) * P + p];
                else {
                    const float4 n = normal4(philox4x32_10(make_uint4(p >> 2, 0u, STREAM_INIT, uid), key));
                    const int e = p & 3;
                    th = a.init_std * (e == 0 ? n.x : e == 1 ? n.y : e == 2 ? n.z : n.w);
                }
            }
            s.theta[p] = th; s.am[p] = 0.f; s.av[p] = 0.f;
        }
        __syncthreads();
        // ---- TT: constant per instance, fp64 accumulation ----------------------------------
        if (!fast) {
            const double part = tt_pass<POL>(s.Qd, N, NB, warp, lane);
            const double tt = block_sum(part, s.red, warp, lane);
            if (tid == 0) s_TT = tt;
        }
        const float* wz = (fast && a.wz_ext) ? a.wz_ext + size_t(out) * epochs * 3 * F : nullptr;
        const float gscale = fast ? 1.0f : 4.0f / (float(N) * float(N) * CGNN_PRESCALE);
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
            pw_forward(s, a, N, NP, h, it, uid, key, noise ? noise + size_t(it) * a.ld_rows : nullptr, tid);
            __syncthreads();
            const bool traced = a.dbg_loss || (a.trace && a.trace_every > 0 && it % a.trace_every == 0);
            if (a.dbg_yhat) for (int i = tid; i < N; i += PW_NT) a.dbg_yhat[size_t(it) * N + i] = s.yh[i];
            if (fast) {   // random-feature loss: the loss value is a by-product of the gradient
                pw_fourier_features(s, a, F, it, uid, key, wz ? wz + size_t(it) * 3 * F : nullptr, tid);
                __syncthreads();
                pw_fourier_colsums(s, N, NB, F, tid);
                __syncthreads();
                const double sum = block_sum(pw_fourier_coefs(s, N, F, a.fast_k, tid), s.red, warp, lane);
                if (tid == 0) {
                    const float loss = float(sum);
                    if (a.trace) a.trace[size_t(out) * epochs + it] = loss;
                    if (a.dbg_loss) a.dbg_loss[it] = loss;
                }
                __syncthreads();
                pw_fourier_rowgrad(s, N, F, tid);
            } else {
                if (traced) {   // loss fetched with the train op is the pre-update loss (GNN.py:93-94)
                    const double part = mmd_pass<POL, false>(s.Qd, nullptr, N, NB, warp, lane);
                    const double sum = block_sum(part, s.red, warp, lane);
                    if (tid == 0) {
                        const float loss = float((sum + s_TT) * invN2);
                        if (a.trace) a.trace[size_t(out) * epochs + it] = loss;
                        if (a.dbg_loss) a.dbg_loss[it] = loss;
                    }
                } else if (a.trace && tid == 0) {
                    a.trace[size_t(out) * epochs + it] = __int_as_float(0x7fc00000);
                }
                mmd_pass<POL, true>(s.Qd, s.g64, N, NB, warp, lane);
            }
            __syncthreads();
            if (a.dbg_dldy)
                for (int i = tid; i < N; i += PW_NT) a.dbg_dldy[size_t(it) * N + i] = float(fx_get(&s.g64[s.rank[i]])) * gscale;
            pw_backward_partials(s, N, NP, h, tid, warp, lane, gscale);
            __syncthreads();
            for (int p = tid; p < P; p += PW_NT) {
                const int slot = pw_slot(p, h);
                float g = 0.f;
#pragma unroll
                for (int w = 0; w < PW_NW; ++w) g += s.part[w * s.part_stride + slot];
                const float lr_t = float(s.red[PW_NW + 3]);
                float m = s.am[p], v = s.av[p];
                const float th = adam_update(s.theta[p], g, m, v, lr_t);
                s.am[p] = m; s.av[p] = v; s.theta[p] = th;
                if (a.dbg_grad) a.dbg_grad[size_t(it) * P + p] = g;
                if (a.dbg_params) a.dbg_params[size_t(it) * P + p] = th;
            }
        }
        __syncthreads();
        if (a.params_out)
            for (int p = tid; p < P; p += PW_NT) a.params_out[size_t(out) * P + p] = s.theta[p];

        // ---- evaluation (GNN.py:115-126): mean of test_epochs fresh-noise losses ---------------
        double score_sum = 0.0;
        for (int it = 0; it < a.test_epochs; ++it) {
            __syncthreads();
            pw_forward(s, a, N, NP, h, a.train_epochs + it, uid, key,
                       noise ? noise + size_t(a.train_epochs + it) * a.ld_rows : nullptr, tid);
            __syncthreads();
            double sum;
            if (fast) {
                pw_fourier_features(s, a, F, a.train_epochs + it, uid, key, wz ? wz + size_t(a.train_epochs + it) * 3 * F : nullptr, tid);
                __syncthreads();
                pw_fourier_colsums(s, N, NB, F, tid);
                __syncthreads();
                sum = block_sum(pw_fourier_coefs(s, N, F, a.fast_k, tid), s.red, warp, lane);
            } else {
                sum = block_sum(mmd_pass<POL, false>(s.Qd, nullptr, N, NB, warp, lane), s.red, warp, lane);
            }
            if (tid == 0) {
                const float loss = fast ? float(sum) : float((sum + s_TT) * invN2);    // the reference fetches an fp32 loss ...
                score_sum += double(loss);                         // ... and sums it in a Python float
                if (a.trace) a.trace[size_t(out) * epochs + a.train_epochs + it] = loss;
            }
        }
        if (tid == 0 && a.scores) a.scores[out] = a.test_epochs > 0 ? float(score_sum / double(a.test_epochs)) : 0.f;
    }
}

}  // namespace CGNN_PW_NS
}  // namespace cgnn
