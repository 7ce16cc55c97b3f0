// Shared device helpers of the cgnn_b200 engine (sm_100a).
//
// Reference math: Code/cgnn/utils/Loss.py:10-32 (MMD), GNN.py:21-29 (init), GNN.py:64 (noise),
// tf.train.AdamOptimizer (TF-1 "epsilon hat" Adam).  See DESIGN.md for the block form used here.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace cgnn {

// ---------------------------------------------------------------------------------------------
// Bandwidths (Loss.py:10).  Fixed at compile time exactly as in the reference, so that the
// multiplies by gamma become FFMA/FMUL immediates and the x2 / x4 ratios can be exploited.
// ---------------------------------------------------------------------------------------------
#define CGNN_LOG2E 1.4426950408889634f
constexpr float G0 = 0.005f, G1 = 0.05f, G2 = 0.25f, G3 = 0.5f, G4 = 1.0f, G5 = 5.0f, G6 = 50.0f;
constexpr float C0 = -G0 * CGNN_LOG2E, C1 = -G1 * CGNN_LOG2E, C2 = -G2 * CGNN_LOG2E, C3 = -G3 * CGNN_LOG2E,
                C4 = -G4 * CGNN_LOG2E, C5 = -G5 * CGNN_LOG2E, C6 = -G6 * CGNN_LOG2E;
constexpr float SUM_GAMMA_KERNELS_AT_ZERO = 7.0f;   // sum_g exp(0)

__device__ __forceinline__ float ex2f(float x) {        // one MUFU.EX2, flush-to-zero: no range fix-up code
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// The seven kernel values exp(-gamma_g * D) of one pair evaluation, reduced on the fly to
//   ksum = sum_g k_g          (loss)            wsum = sum_g gamma_g k_g   (gradient weight)
// Policies (hparams.exp_policy) trade MUFU ex2 for FP32 multiplies using the exact ratios of the
// bandwidth list (Loss.py:10): 0.25 -> 0.5 -> 1 are squarings, 0.005 -> 0.05 is a 10th power,
// 1 -> 5 a 5th power.  A power p multiplies the relative error of the ex2 it starts from by p.
//   POL 0: 7 MUFU                                          (reference-like: every kernel its own exp)
//   POL 1: 5 MUFU   gamma 0.5, 1 from 0.25                 (error x2, x4)
//   POL 2: 4 MUFU   + gamma 0.05 from 0.005                (error x10)
//   POL 3: 3 MUFU   + gamma 5 from 1                       (error x20 relative to the 0.25 ex2)
template <int POL>
__device__ __forceinline__ float pow10f(float k) { float a = k * k; float b = a * a; float c = b * k; return c * c; }

template <int POL>
__device__ __forceinline__ float wsum(float D) {
    const float k0 = ex2f(D * C0);
    const float k1 = (POL >= 2) ? pow10f<POL>(k0) : ex2f(D * C1);
    const float k2 = ex2f(D * C2);
    const float k6 = ex2f(D * C6);
    if (POL == 0) {
        const float k3 = ex2f(D * C3), k4 = ex2f(D * C4), k5 = ex2f(D * C5);
        float a = fmaf(G1, k1, G0 * k0);
        float b = fmaf(G3, k3, G2 * k2);
        float c = fmaf(G5, k5, k4);           // G4 == 1
        a = fmaf(G6, k6, a);
        return (a + b) + c;
    } else {
        const float k3 = k2 * k2;                         // gamma 0.5
        float k5;
        if (POL >= 3) { const float k4 = k3 * k3; const float a = k4 * k4; k5 = (a * a) * k4; }
        else k5 = ex2f(D * C5);
        // G2 k2 + G3 k2^2 + G4 k2^4 by Horner in k2 (k4 is never formed explicitly)
        float t = fmaf(G4, k3, G3);
        t = fmaf(t, k2, G2);
        float w = t * k2;
        float u = fmaf(G1, k1, G0 * k0);
        w = fmaf(G5, k5, w);
        u = fmaf(G6, k6, u);
        return w + u;
    }
}

template <int POL>
__device__ __forceinline__ float ksum(float D) {
    const float k0 = ex2f(D * C0);
    const float k1 = (POL >= 2) ? pow10f<POL>(k0) : ex2f(D * C1);
    const float k2 = ex2f(D * C2);
    const float k6 = ex2f(D * C6);
    float k3, k4, k5;
    if (POL == 0) { k3 = ex2f(D * C3); k4 = ex2f(D * C4); } else { k3 = k2 * k2; k4 = k3 * k3; }
    if (POL >= 3) { const float a = k4 * k4; k5 = (a * a) * k4; } else k5 = ex2f(D * C5);
    return ((k0 + k1) + (k2 + k3)) + ((k4 + k5) + k6);
}

// ---------------------------------------------------------------------------------------------
// "Primed" variants used by the pairwise kernel: coordinates are pre-multiplied by
// S = sqrt(0.25*log2(e)), so the squared distance arrives as Dp = 0.25*log2(e)*D and
//   exp(-gamma_g D) = ex2(Dp * A_g),  A_g = -gamma_g/0.25 = -0.02, -0.2, -1, -2, -4, -20, -200.
// The gamma=0.25 kernel needs no multiply at all, gamma=50 is produced already weighted
// (50*k6 = ex2(Dp*A6 + log2 50)), and the weighted sum is folded into Horner chains:
//   wsum = 50 k6 + q (G0 + G1 q^9) + p (G2 + p (G3 + G4 p^2)) + G5 p^20      q = k(0.005), p = k(0.25)
// ---------------------------------------------------------------------------------------------
#define CGNN_PRESCALE 0.6005612043932249f          /* sqrt(0.25*log2(e)) */
#define CGNN_LOG2_50  5.643856189774724f

template <int POL>
__device__ __forceinline__ float wsum_p(float Dp) {
    const float q = ex2f(Dp * -0.02f);
    const float p = ex2f(-Dp);
    float w = ex2f(fmaf(Dp, -200.0f, CGNN_LOG2_50));             // 50 * k(50)
    if (POL >= 2) {
        const float q2 = q * q, q4 = q2 * q2, q8 = q4 * q4;
        w = fmaf(fmaf(G1, q8 * q, G0), q, w);
    } else {
        w = fmaf(G1, ex2f(Dp * -0.2f), fmaf(G0, q, w));
    }
    if (POL >= 1) {
        const float p2 = p * p;
        w = fmaf(fmaf(p2 + G3, p, G2), p, w);                    // G2 p + G3 p^2 + p^4   (G4 == 1)
        if (POL >= 3) { const float p4 = p2 * p2, p8 = p4 * p4; w = fmaf(G5, (p8 * p8) * p4, w); }
        else w = fmaf(G5, ex2f(Dp * -20.0f), w);
    } else {
        w = fmaf(G2, p, w);
        w = fmaf(G3, ex2f(Dp * -2.0f), w);
        w += ex2f(Dp * -4.0f);
        w = fmaf(G5, ex2f(Dp * -20.0f), w);
    }
    return w;
}

template <int POL>
__device__ __forceinline__ float ksum_p(float Dp) {
    const float q = ex2f(Dp * -0.02f);
    const float p = ex2f(-Dp);
    float s = ex2f(Dp * -200.0f);
    if (POL >= 2) { const float q2 = q * q, q4 = q2 * q2, q5 = q4 * q; s += fmaf(q5, q5, q); }
    else s += q + ex2f(Dp * -0.2f);
    if (POL >= 1) {
        const float p2 = p * p, p4 = p2 * p2;
        float t = fmaf(p2, p2, p2) + p;                          // p + p^2 + p^4
        if (POL >= 3) { const float p8 = p4 * p4; t = fmaf(p8 * p8, p4, t); }
        else t += ex2f(Dp * -20.0f);
        s += t;
    } else {
        s += (p + ex2f(Dp * -2.0f)) + (ex2f(Dp * -4.0f) + ex2f(Dp * -20.0f));
    }
    return s;
}

// ---------------------------------------------------------------------------------------------
// Philox4x32-10 counter RNG + Box-Muller.  Engine-defined streams (the reference is unseeded):
//   key     = (seed_lo, seed_hi)
//   counter = (index, step, stream, instance uid)     -> 4 x 32 random bits -> 4 normals
// Restated in numpy in tests/philox_np.py to check the in-kernel generator.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0 = __umulhi(M0, c.x), lo0 = M0 * c.x;
        uint32_t hi1 = __umulhi(M1, c.z), lo1 = M1 * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += W0;
        k.y += W1;
    }
    return c;
}

__device__ __forceinline__ float u01(uint32_t r) {     // (0,1), exact in fp32
    return (float(r >> 9) + 0.5f) * (1.0f / 8388608.0f);
}

__device__ __forceinline__ float4 normal4(uint4 r) {
    float4 n;
    float r0 = sqrtf(-2.0f * logf(u01(r.x)));
    float r1 = sqrtf(-2.0f * logf(u01(r.z)));
    float s0, c0, s1, c1;
    sincospif(2.0f * u01(r.y), &s0, &c0);
    sincospif(2.0f * u01(r.w), &s1, &c1);
    n.x = r0 * c0; n.y = r0 * s0; n.z = r1 * c1; n.w = r1 * s1;
    return n;
}

enum : uint32_t { STREAM_NOISE = 0u, STREAM_INIT = 1u };

// ---------------------------------------------------------------------------------------------
// Packed FP32x2 arithmetic (Blackwell FFMA2 / FADD2 / FMUL2): one issue slot for two IEEE fp32
// operations on a register pair.  Same rounding as the scalar instructions; used where the kernels
// are issue-bound and the data already sits in adjacent registers (float4 column loads).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long pk(float2 a) {
    unsigned long long r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a.x), "f"(a.y));
    return r;
}
__device__ __forceinline__ float2 upk(unsigned long long r) {
    float2 a;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(a.x), "=f"(a.y) : "l"(r));
    return a;
}
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
    unsigned long long d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(pk(a)), "l"(pk(b)), "l"(pk(c)));
    return upk(d);
}
__device__ __forceinline__ float2 fadd2(float2 a, float2 b) {
    unsigned long long d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(pk(a)), "l"(pk(b)));
    return upk(d);
}
__device__ __forceinline__ float2 fsub2(float2 a, float2 b) { return fadd2(a, make_float2(-b.x, -b.y)); }
__device__ __forceinline__ float2 fmul2(float2 a, float2 b) {
    unsigned long long d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(pk(a)), "l"(pk(b)));
    return upk(d);
}
__device__ __forceinline__ float2 splat2(float w) { return make_float2(w, w); }

// ---------------------------------------------------------------------------------------------
// Two evaluations at once: the chains of wsum_p / ksum_p on register pairs (FMUL2 / FFMA2), the MUFU
// ex2 on each half.  Halves the issue slots of the FP32 part, which is what bounds exp policy 3.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float2 ex2f2(float2 x) { return make_float2(ex2f(x.x), ex2f(x.y)); }

// FAR: the caller guarantees Dp > CGNN_FAR_DP for both halves; the gamma = 50 kernel is then below 5e-11 and left out.
#define CGNN_FAR_DP 0.2f          /* primed units: D > 0.5545; 50 e^{-50 D} / (0.005 e^{-0.005 D}) < 1e-8 */

// FAR = 2 ("very far", Dp > CGNN_FAR2_DP): the gamma = 5 kernel is out of sight as well.
#define CGNN_FAR2_DP 1.85f        /* D > 5.13; 5 e^{-5 D} / (0.005 e^{-0.005 D}) < 1e-8 */

template <int POL, int FAR = 0>
__device__ __forceinline__ float2 wsum_p2(float2 Dp) {
    const float2 q = ex2f2(fmul2(Dp, splat2(-0.02f)));
    const float2 p = ex2f2(make_float2(-Dp.x, -Dp.y));
    float2 w = FAR ? make_float2(0.f, 0.f) : ex2f2(ffma2(Dp, splat2(-200.0f), splat2(CGNN_LOG2_50)));
    if (POL >= 2) {
        const float2 q2 = fmul2(q, q), q4 = fmul2(q2, q2), q8 = fmul2(q4, q4);
        w = ffma2(ffma2(splat2(G1), fmul2(q8, q), splat2(G0)), q, w);
    } else {
        w = ffma2(splat2(G1), ex2f2(fmul2(Dp, splat2(-0.2f))), ffma2(splat2(G0), q, w));
    }
    if (POL >= 1) {
        const float2 p2 = fmul2(p, p);
        w = ffma2(ffma2(fadd2(p2, splat2(G3)), p, splat2(G2)), p, w);
        if (FAR >= 2) { }
        else if (POL >= 3) { const float2 p4 = fmul2(p2, p2), p8 = fmul2(p4, p4); w = ffma2(splat2(G5), fmul2(fmul2(p8, p8), p4), w); }
        else w = ffma2(splat2(G5), ex2f2(fmul2(Dp, splat2(-20.0f))), w);
    } else {
        w = ffma2(splat2(G2), p, w);
        w = ffma2(splat2(G3), ex2f2(fmul2(Dp, splat2(-2.0f))), w);
        w = fadd2(w, ex2f2(fmul2(Dp, splat2(-4.0f))));
        w = ffma2(splat2(G5), ex2f2(fmul2(Dp, splat2(-20.0f))), w);
    }
    return w;
}

template <int POL, int FAR = 0>
__device__ __forceinline__ float2 ksum_p2(float2 Dp) {
    const float2 q = ex2f2(fmul2(Dp, splat2(-0.02f)));
    const float2 p = ex2f2(make_float2(-Dp.x, -Dp.y));
    float2 s = FAR ? make_float2(0.f, 0.f) : ex2f2(fmul2(Dp, splat2(-200.0f)));
    if (POL >= 2) { const float2 q2 = fmul2(q, q), q4 = fmul2(q2, q2), q5 = fmul2(q4, q); s = fadd2(s, ffma2(q5, q5, q)); }
    else s = fadd2(s, fadd2(q, ex2f2(fmul2(Dp, splat2(-0.2f)))));
    if (POL >= 1) {
        const float2 p2 = fmul2(p, p), p4 = fmul2(p2, p2);
        float2 t = fadd2(ffma2(p2, p2, p2), p);
        if (FAR >= 2) { }
        else if (POL >= 3) { const float2 p8 = fmul2(p4, p4); t = ffma2(fmul2(p8, p8), p4, t); }
        else t = fadd2(t, ex2f2(fmul2(Dp, splat2(-20.0f))));
        s = fadd2(s, t);
    } else {
        s = fadd2(s, fadd2(fadd2(p, ex2f2(fmul2(Dp, splat2(-2.0f)))),
                           fadd2(ex2f2(fmul2(Dp, splat2(-4.0f))), ex2f2(fmul2(Dp, splat2(-20.0f))))));
    }
    return s;
}

// ---------------------------------------------------------------------------------------------
// The same two-at-once chains on UNSCALED squared distances (graph kernel: its coordinates cannot be
// prescaled for free, one FMUL2 turns the pair of distances into ex2 arguments instead):
//   wsum = 50 k6 + G0 q + G1 k1 + p (G2 + p (G3 + p^2)) + G5 k5,   q = k(0.005), p = k(0.25), G4 = 1.
// ---------------------------------------------------------------------------------------------
template <int POL>
__device__ __forceinline__ float2 wsum_r2(float2 D) {
    const float2 q = ex2f2(fmul2(D, splat2(C0)));
    const float2 p = ex2f2(fmul2(D, splat2(C2)));
    float2 w = ex2f2(ffma2(D, splat2(C6), splat2(CGNN_LOG2_50)));               // 50 * k(50)
    if (POL >= 2) {
        const float2 q2 = fmul2(q, q), q4 = fmul2(q2, q2), q8 = fmul2(q4, q4);
        w = ffma2(ffma2(splat2(G1), fmul2(q8, q), splat2(G0)), q, w);
    } else {
        w = ffma2(splat2(G1), ex2f2(fmul2(D, splat2(C1))), ffma2(splat2(G0), q, w));
    }
    if (POL >= 1) {
        const float2 p2 = fmul2(p, p);
        w = ffma2(ffma2(fadd2(p2, splat2(G3)), p, splat2(G2)), p, w);
        if (POL >= 3) { const float2 p4 = fmul2(p2, p2), p8 = fmul2(p4, p4); w = ffma2(splat2(G5), fmul2(fmul2(p8, p8), p4), w); }
        else w = ffma2(splat2(G5), ex2f2(fmul2(D, splat2(C5))), w);
    } else {
        w = ffma2(splat2(G2), p, w);
        w = ffma2(splat2(G3), ex2f2(fmul2(D, splat2(C3))), w);
        w = fadd2(w, ex2f2(fmul2(D, splat2(C4))));
        w = ffma2(splat2(G5), ex2f2(fmul2(D, splat2(C5))), w);
    }
    return w;
}

template <int POL>
__device__ __forceinline__ float2 ksum_r2(float2 D) {
    const float2 q = ex2f2(fmul2(D, splat2(C0)));
    const float2 p = ex2f2(fmul2(D, splat2(C2)));
    float2 s = ex2f2(fmul2(D, splat2(C6)));
    if (POL >= 2) { const float2 q2 = fmul2(q, q), q4 = fmul2(q2, q2), q5 = fmul2(q4, q); s = fadd2(s, ffma2(q5, q5, q)); }
    else s = fadd2(s, fadd2(q, ex2f2(fmul2(D, splat2(C1)))));
    if (POL >= 1) {
        const float2 p2 = fmul2(p, p), p4 = fmul2(p2, p2);
        float2 t = fadd2(ffma2(p2, p2, p2), p);                                  // p + p^2 + p^4
        if (POL >= 3) { const float2 p8 = fmul2(p4, p4); t = ffma2(fmul2(p8, p8), p4, t); }
        else t = fadd2(t, ex2f2(fmul2(D, splat2(C5))));
        s = fadd2(s, t);
    } else {
        s = fadd2(s, fadd2(fadd2(p, ex2f2(fmul2(D, splat2(C3)))),
                           fadd2(ex2f2(fmul2(D, splat2(C4))), ex2f2(fmul2(D, splat2(C5))))));
    }
    return s;
}

// ---------------------------------------------------------------------------------------------
// Order-independent accumulation into shared memory: 64-bit fixed point (2^-36 resolution,
// +-2^27 range) with native integer atomics.  Integer addition is associative, so the result does
// not depend on warp scheduling: the engine is bit-reproducible run to run.
// ---------------------------------------------------------------------------------------------
#define CGNN_FX_SCALE   68719476736.0            /* 2^36 */
#define CGNN_FX_INV     (1.0 / 68719476736.0)

__device__ __forceinline__ void fx_add(long long* p, double v) {
    long long q = __double2ll_rn(v * CGNN_FX_SCALE);
    atomicAdd(reinterpret_cast<unsigned long long*>(p), static_cast<unsigned long long>(q));
}
__device__ __forceinline__ double fx_get(const long long* p) { return double(*p) * CGNN_FX_INV; }

// ---------------------------------------------------------------------------------------------
// Warp reductions
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Transposing reduction: every lane brings 32 partial sums v[0..31]; on return lane l holds the
// warp-wide total of v[l].  31 shuffles instead of 32 x 5.
__device__ __forceinline__ float warp_reduce32(float (&v)[32], int lane) {
#pragma unroll
    for (int ofs = 16, n = 32; ofs >= 1; ofs >>= 1, n >>= 1) {
        const bool up = (lane & ofs) != 0;
#pragma unroll
        for (int i = 0; i < n / 2; ++i) {
            float send = up ? v[i] : v[i + n / 2];
            float keep = up ? v[i + n / 2] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, ofs);
        }
    }
    return v[0];
}

// The same for N values per lane, N <= 32 (e.g. the 20 hidden units of the default CGNN generator): at every
// stage the lower lane of a pair keeps the first ceil(n/2) slots, the upper lane the rest (an odd count is padded
// with a zero), so N = 20 takes 10+5+3+2+1 = 21 shuffles instead of 31.  On return `slot` is the index of the
// value this lane holds and `valid` tells whether it is a real slot (N < 32 leaves some lanes without one).
template <int N, int OFS>
struct WarpReduceN {
    static __device__ __forceinline__ float run(const float (&v)[N], int lane, int& slot, int& n_valid) {
        constexpr int H = (N + 1) / 2;
        const bool up = (lane & OFS) != 0;
        float w[H];
#pragma unroll
        for (int i = 0; i < H; ++i) {
            const float hi = (H + i < N) ? v[(H + i < N) ? H + i : 0] : 0.f;
            const float send = up ? v[i] : hi;
            const float keep = up ? hi : v[i];
            w[i] = keep + __shfl_xor_sync(0xffffffffu, send, OFS);
        }
        slot += up ? H : 0;
        n_valid = up ? max(n_valid - H, 0) : min(n_valid, H);
        return WarpReduceN<H, OFS / 2>::run(w, lane, slot, n_valid);
    }
};
template <int N>
struct WarpReduceN<N, 0> {
    static __device__ __forceinline__ float run(const float (&v)[N], int, int&, int&) {
        static_assert(N == 1, "32 lanes reduce at most 32 slots");
        return v[0];
    }
};
template <int N>
__device__ __forceinline__ float warp_reduce_n(const float (&v)[N], int lane, int& slot, bool& valid) {
    static_assert(N >= 1 && N <= 32, "1..32 slots");
    slot = 0;
    int n_valid = N;
    const float r = WarpReduceN<N, 16>::run(v, lane, slot, n_valid);
    valid = n_valid >= 1;
    return r;
}

// Two such reductions at once: lane l ends with the totals of (v[l].x, v[l].y).  The same adds in the same order as
// two scalar reductions (FADD2 = two IEEE adds), but the two dependent shuffle chains overlap and half the FADD /
// select instructions disappear.
template <int N, int OFS>
struct WarpReduceN2 {
    static __device__ __forceinline__ float2 run(const float2 (&v)[N], int lane, int& slot, int& n_valid) {
        constexpr int H = (N + 1) / 2;
        const bool up = (lane & OFS) != 0;
        float2 w[H];
#pragma unroll
        for (int i = 0; i < H; ++i) {
            const float2 hi = (H + i < N) ? v[(H + i < N) ? H + i : 0] : make_float2(0.f, 0.f);
            const float2 send = up ? v[i] : hi;
            const float2 keep = up ? hi : v[i];
            w[i] = fadd2(keep, make_float2(__shfl_xor_sync(0xffffffffu, send.x, OFS), __shfl_xor_sync(0xffffffffu, send.y, OFS)));
        }
        slot += up ? H : 0;
        n_valid = up ? max(n_valid - H, 0) : min(n_valid, H);
        return WarpReduceN2<H, OFS / 2>::run(w, lane, slot, n_valid);
    }
};
template <int N>
struct WarpReduceN2<N, 0> {
    static __device__ __forceinline__ float2 run(const float2 (&v)[N], int, int&, int&) {
        static_assert(N == 1, "32 lanes reduce at most 32 slots");
        return v[0];
    }
};
template <int N>
__device__ __forceinline__ float2 warp_reduce_n2(const float2 (&v)[N], int lane, int& slot, bool& valid) {
    static_assert(N >= 1 && N <= 32, "1..32 slots");
    slot = 0;
    int n_valid = N;
    const float2 r = WarpReduceN2<N, 16>::run(v, lane, slot, n_valid);
    valid = n_valid >= 1;
    return r;
}

// ---------------------------------------------------------------------------------------------
// TF-1 Adam, fp32, one parameter per call.  lr_t = lr*sqrt(1-b2^t)/(1-b1^t) is computed by the caller.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float adam_update(float theta, float g, float& m, float& v, float lr_t) {
    // explicit rounding intrinsics: no FMA contraction choices left to the compiler, so every code path that
    // calls this (pairwise / graph kernel, single CTA / cluster) produces the same bits from the same inputs
    m = __fmaf_rn(0.9f, m, __fmul_rn(0.1f, g));                     // (1-beta1) rounds to 0.1f in fp32 as in TF
    v = __fmaf_rn(0.999f, v, __fmul_rn(0.001f, __fmul_rn(g, g)));
    return __fsub_rn(theta, __fdiv_rn(__fmul_rn(lr_t, m), __fadd_rn(__fsqrt_rn(v), 1e-8f)));
}

}  // namespace cgnn
