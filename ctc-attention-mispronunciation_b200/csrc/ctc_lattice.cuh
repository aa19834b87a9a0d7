// ctc_lattice.cuh -- lattice arithmetic shared by the linear-domain CTC kernels (ctc_pair.cu, ctc_duo.cu): geometry,
// packed alpha / beta steps for a pair of utterances, per-lane renormalisation.  See ctc_pair.cu for the scheme.
#pragma once
#include <type_traits>

#include "common.cuh"

namespace mdd {
namespace lattice {

constexpr int KF = 8;                    // frames per chunk = renormalisation period = checkpoint period
constexpr int E_EMPTY = -(1 << 20);      // exponent of a lane whose states are all zero
#ifndef MDD_SHIFT
#define MDD_SHIFT 30
#endif
#ifndef MDD_GAP8
#define MDD_GAP8 26
#endif
constexpr int SHIFT = MDD_SHIFT;         // a renormalised lane keeps its maximum in [2^SHIFT, 2^(SHIFT+1))
constexpr int OFF = 50 - SHIFT;                  // beta mantissas are kept 2^-OFF below 1 / (alpha mantissa)
constexpr int CPMAX = 64;                // staged emission row: utterance h at [h*CPMAX, +64): 32 banks x 2 levels
constexpr int RW = 2 * CPMAX;
constexpr unsigned FULL = 0xffffffffu;
template <int N>
using IC = std::integral_constant<int, N>;


template <int NSL>
struct PGeo {
    static constexpr int NL = NSL / 2;                     // label states per lane
    static constexpr int LCAP = 16 * NSL - 1;              // S = 2L+1 <= 32*NSL - 1
    static constexpr int ARS = 32 * NSL;                   // (!REGBUF) floats per frame of the alpha label-state buffer
    // mass can cross ceil(2*KF/NSL)+1 lane boundaries between two renormalisations, each multiplying the mantissa by
    // up to 2^GAP, plus 3^KF of growth, on top of the 2^SHIFT a renormalised lane starts from:
    // GAP * crossings + 14 + SHIFT < 126
    static constexpr int GAP = NSL == 2 ? 9 : NSL == 4 ? 16 : NSL == 8 ? MDD_GAP8 : 40;
    static_assert(GAP * ((2 * KF + NSL - 1) / NSL + 1) + 14 + SHIFT < 126, "alpha could overflow inside a chunk");
    static constexpr int CKW = 32 * (2 * NSL + 2);         // checkpoint words per pair and chunk
};

__device__ __forceinline__ float pow2i(int x) {            // 2^x, 0 below the normal range, clamped at 2^127
    return __int_as_float((min(max(x, -127), 127) + 127) << 23);   // branch-free: exponent field 0 is +0.0
}
constexpr float LOG2E = 1.4426950408889634f;
__device__ __forceinline__ float fast_exp2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float fast_exp(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x * 1.4426950408889634f));
    return y;
}
__device__ __forceinline__ float fmax3(float a, float b, float c) {
    float m;
    asm("max.f32 %0, %1, %2, %3;" : "=f"(m) : "f"(a), "f"(b), "f"(c));
    return m;
}
__device__ __forceinline__ float compc(const float2& v, int h) { return h ? v.y : v.x; }

// ---------------------------------------------------------------------------------------------------------------
// One lattice step for both utterances.  Lane l owns states s = NSL*l + j; even j = blank, odd j = label (j-1)/2.
//   alpha: new[s] = (a[s] + a[s-1] + ma * a[s-2]) * e[s]
//   beta : u[s]   =  b[s] + b[s+1] + mb * b[s+2];  new b[s] = u[s] * e[s]
// ep0/ep1 point at the emission of my label states inside the staged row of frame 0 (unused slots read a zero).
// ---------------------------------------------------------------------------------------------------------------
template <int NSL>
__device__ __forceinline__ void alpha_step(float2 (&a)[NSL], float2 sc, int roff, const float* __restrict__ bl0,
                                           const float* __restrict__ bl1, const float* const (&ep0)[NSL / 2],
                                           const float* const (&ep1)[NSL / 2], const float2 (&ma)[NSL / 2]) {
    float2 p = make_float2(__shfl_up_sync(FULL, a[NSL - 1].x, 1), __shfl_up_sync(FULL, a[NSL - 1].y, 1));
    p = __fmul2_rn(p, sc);                                       // sc == 0 on lane 0
    const float2 eb = make_float2(bl0[roff], bl1[roff]);
    float2 e[NSL / 2];
#pragma unroll
    for (int q = 0; q < NSL / 2; ++q) e[q] = make_float2(ep0[q][roff], ep1[q][roff]);
    float2 n[NSL];
#pragma unroll
    for (int j = 0; j < NSL; ++j) {
        const float2 m1 = j ? a[j - 1] : p;
        if (j & 1) {
            const float2 m2 = j >= 2 ? a[j - 2] : p;
            n[j] = __fmul2_rn(__ffma2_rn(ma[j >> 1], m2, __fadd2_rn(a[j], m1)), e[j >> 1]);
        } else {
            n[j] = __fmul2_rn(__fadd2_rn(a[j], m1), eb);
        }
    }
#pragma unroll
    for (int j = 0; j < NSL; ++j) a[j] = n[j];
}

template <int NSL>
__device__ __forceinline__ void beta_step(float2 (&b)[NSL], float2 (&ul)[NSL / 2], float2 sc, float2 sc2, int roff,
                                          const float* __restrict__ bl0, const float* __restrict__ bl1,
                                          const float* const (&ep0)[NSL / 2], const float* const (&ep1)[NSL / 2],
                                          const float2 (&mb)[NSL / 2]) {
    float2 n0 = make_float2(__shfl_down_sync(FULL, b[0].x, 1), __shfl_down_sync(FULL, b[0].y, 1));
    float2 n1 = make_float2(__shfl_down_sync(FULL, b[1].x, 1), __shfl_down_sync(FULL, b[1].y, 1));
    // sc * sc2 = 2^(e_beta[l+1] - e_beta[l]) in two factors: the beta mantissas are not confined to [1, 2), so a
    // product that is representable must not be lost to a clamped power of two (sc == 0 on lane 31)
    n0 = __fmul2_rn(__fmul2_rn(n0, sc), sc2);
    n1 = __fmul2_rn(__fmul2_rn(n1, sc), sc2);
    const float2 eb = make_float2(bl0[roff], bl1[roff]);
    float2 e[NSL / 2];
#pragma unroll
    for (int q = 0; q < NSL / 2; ++q) e[q] = make_float2(ep0[q][roff], ep1[q][roff]);
    float2 n[NSL];
#pragma unroll
    for (int j = 0; j < NSL; ++j) {
        const float2 p1 = (j + 1 < NSL) ? b[j + 1] : n0;
        if (j & 1) {
            const float2 p2 = (j + 2 < NSL) ? b[j + 2] : n1;
            const float2 u = __ffma2_rn(mb[j >> 1], p2, __fadd2_rn(b[j], p1));
            ul[j >> 1] = u;
            n[j] = __fmul2_rn(u, e[j >> 1]);
        } else {
            n[j] = __fmul2_rn(__fadd2_rn(b[j], p1), eb);
        }
    }
#pragma unroll
    for (int j = 0; j < NSL; ++j) b[j] = n[j];
}

// Per-lane renormalisation of alpha for both utterances (see the header).  ok[h] = false if utterance h has no
// alignment mass left or is not finite.  A lane whose own content lies far below what is about to arrive from its left
// neighbour is pulled up to 2^-GAP of it; content that underflows on the way is mass dropped from the forward sweep.
// That is harmless for the leading edge of the lattice (states far ahead of the bulk carry paths of negligible mass)
// and is checked after the fact where it is not: the backward sweep does not share the loss, so the total likelihood
// it arrives at, beta_0(0) + beta_0(1), is compared with the forward P (flag 7 in ctc_pair.cu).  `strict` restores the
// earlier a-priori criterion (any squash beyond 2^100 hands the utterance to the log-space kernel).
template <int NSL>
__device__ __forceinline__ void renorm_alpha(float2 (&a)[NSL], int (&ea)[2], float2& sca, int lane, bool (&ok)[2],
                                             bool strict = false, int decay_max = 1 << 20) {
    constexpr int GAP = PGeo<NSL>::GAP;
    float f[2], sc[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        float m = compc(a[0], h);
#pragma unroll
        for (int j = 1; j + 1 < NSL; j += 2) m = fmax3(m, compc(a[j], h), compc(a[j + 1], h));
        m = fmaxf(m, compc(a[NSL - 1], h));
        const int k = ((__float_as_int(m) >> 23) & 0xff) - 127;
        const bool some = m >= 1.17549435e-38f && m < 3.0e38f;       // normal, finite (NaN fails both)
        const int e_own = some ? ea[h] + k - SHIFT : E_EMPTY;
        int t = e_own + GAP * lane;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) t = max(t, __shfl_up_sync(FULL, t, d));   // lanes < d get their own value back
        const int e_new = t - GAP * lane;
        f[h] = pow2i(ea[h] - e_new);
        ea[h] = e_new;
        const int e_nb = __shfl_up_sync(FULL, e_new, 1);
        sc[h] = lane == 0 ? 0.f : pow2i(e_nb - e_new);
        // The lanes that hold the bulk of the mass (within 2^64 of the largest) must not have decayed by more than
        // 2^decay_max since the last renormalisation: beyond that, smaller states of such a lane have been underflowing
        // inside the chunk, and since the beta exponents are tied to alpha's, the backward sweep is blind to what a
        // lane lost once the lane's exponent has followed the loss down (the forward/backward check cannot see it).
        // Lanes far below the bulk decay faster as a matter of course (their paths are the poor ones) and are exempt.
        const int e_top = __reduce_max_sync(FULL, e_own);
        const bool fast_decay = some && e_own >= e_top - 64 && k - SHIFT < -decay_max;
        const bool bad = (m != m) || m >= 3.0e38f || fast_decay || (strict && some && e_new - e_own > 100);
        ok[h] = __any_sync(FULL, some) && !__any_sync(FULL, bad);
    }
    const float2 f2 = make_float2(f[0], f[1]);
#pragma unroll
    for (int j = 0; j < NSL; ++j) a[j] = __fmul2_rn(a[j], f2);
    sca = make_float2(sc[0], sc[1]);
}

}  // namespace lattice
}  // namespace mdd
