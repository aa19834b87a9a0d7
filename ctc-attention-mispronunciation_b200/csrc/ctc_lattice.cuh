// ctc_lattice.cuh -- lattice arithmetic shared by the linear-domain CTC kernels (ctc_pair.cu, ctc_team.cu, ctc_dual.cu): geometry,
// packed alpha / beta steps for a pair of utterances, per-lane renormalisation.  See ctc_pair.cu for the scheme.
#pragma once
#include <type_traits>

#include "common.cuh"

namespace mdd {
namespace lattice {

constexpr int KF = 8;                    // frames per chunk = renormalisation period = checkpoint period
constexpr int E_EMPTY = -(1 << 20);      // exponent of a lane whose states are all zero
constexpr int kColourMinWork = 1200;     // frames x label states per lane from which the bank colouring pays (ctc_dual.cu)
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
// earlier a-priori criterion (any squash beyond 2^100 hands the utterance to the next tier).
template <int NSL>
__device__ __forceinline__ void renorm_alpha(float2 (&a)[NSL], int (&ea)[2], float2& sca, int lane, bool (&ok)[2],
                                             bool strict = false, int decay_max = 1 << 20) {
    constexpr int GAP = PGeo<NSL>::GAP;
    // the two utterances side by side: their max-scans are two dependent chains of five shuffles each, and a warp
    // that runs them one after the other waits out ten shuffle latencies per chunk
    float m0 = a[0].x, m1 = a[0].y;
#pragma unroll
    for (int j = 1; j + 1 < NSL; j += 2) {
        m0 = fmax3(m0, a[j].x, a[j + 1].x);
        m1 = fmax3(m1, a[j].y, a[j + 1].y);
    }
    m0 = fmaxf(m0, a[NSL - 1].x);
    m1 = fmaxf(m1, a[NSL - 1].y);
    const int k0 = ((__float_as_int(m0) >> 23) & 0xff) - 127, k1 = ((__float_as_int(m1) >> 23) & 0xff) - 127;
    const bool some0 = m0 >= 1.17549435e-38f && m0 < 3.0e38f;     // normal, finite (NaN fails both)
    const bool some1 = m1 >= 1.17549435e-38f && m1 < 3.0e38f;
    const int own0 = some0 ? ea[0] + k0 - SHIFT : E_EMPTY, own1 = some1 ? ea[1] + k1 - SHIFT : E_EMPTY;
    int t0 = own0 + GAP * lane, t1 = own1 + GAP * lane;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {                            // lanes < d get their own value back
        const int s0 = __shfl_up_sync(FULL, t0, d), s1 = __shfl_up_sync(FULL, t1, d);
        t0 = max(t0, s0);
        t1 = max(t1, s1);
    }
    const int new0 = t0 - GAP * lane, new1 = t1 - GAP * lane;
    const int nb0 = __shfl_up_sync(FULL, new0, 1), nb1 = __shfl_up_sync(FULL, new1, 1);
    const int top0 = __reduce_max_sync(FULL, own0), top1 = __reduce_max_sync(FULL, own1);
    // The lanes that hold the bulk of the mass (within 2^64 of the largest) must not have decayed by more than
    // 2^decay_max since the last renormalisation: beyond that, smaller states of such a lane have been underflowing
    // inside the chunk, and since the beta exponents are tied to alpha's, the backward sweep is blind to what a
    // lane lost once the lane's exponent has followed the loss down (the forward/backward check cannot see it).
    // Lanes far below the bulk decay faster as a matter of course (their paths are the poor ones) and are exempt.
    const bool fast0 = some0 && own0 >= top0 - 64 && k0 - SHIFT < -decay_max;
    const bool fast1 = some1 && own1 >= top1 - 64 && k1 - SHIFT < -decay_max;
    const bool bad0 = (m0 != m0) || m0 >= 3.0e38f || fast0 || (strict && some0 && new0 - own0 > 100);
    const bool bad1 = (m1 != m1) || m1 >= 3.0e38f || fast1 || (strict && some1 && new1 - own1 > 100);
    const unsigned bits = __reduce_or_sync(FULL, (some0 ? 1u : 0u) | (bad0 ? 2u : 0u) | (some1 ? 4u : 0u) | (bad1 ? 8u : 0u));
    ok[0] = (bits & 3u) == 1u;
    ok[1] = (bits & 12u) == 4u;
    const float2 f2 = make_float2(pow2i(ea[0] - new0), pow2i(ea[1] - new1));
    ea[0] = new0;
    ea[1] = new1;
#pragma unroll
    for (int j = 0; j < NSL; ++j) a[j] = __fmul2_rn(a[j], f2);
    sca = make_float2(lane == 0 ? 0.f : pow2i(nb0 - new0), lane == 0 ? 0.f : pow2i(nb1 - new1));
}

// The mirror image for a beta sweep that carries its own exponents (ctc_dual.cu): mass enters a lane from its RIGHT
// neighbour, so a lane is kept within 2^GAP of lane l+1 and the max-scan runs downwards.  scb = 2^(e[l+1] - e[l])
// (0 on lane 31) is the factor beta_step applies to what it receives.  Same guards as renorm_alpha.
template <int NSL>
__device__ __forceinline__ void renorm_beta(float2 (&b)[NSL], int (&eb)[2], float2& scb, int lane, bool (&ok)[2],
                                            int decay_max = 1 << 20) {
    constexpr int GAP = PGeo<NSL>::GAP;
    float m0 = b[0].x, m1 = b[0].y;
#pragma unroll
    for (int j = 1; j + 1 < NSL; j += 2) {
        m0 = fmax3(m0, b[j].x, b[j + 1].x);
        m1 = fmax3(m1, b[j].y, b[j + 1].y);
    }
    m0 = fmaxf(m0, b[NSL - 1].x);
    m1 = fmaxf(m1, b[NSL - 1].y);
    const int k0 = ((__float_as_int(m0) >> 23) & 0xff) - 127, k1 = ((__float_as_int(m1) >> 23) & 0xff) - 127;
    const bool some0 = m0 >= 1.17549435e-38f && m0 < 3.0e38f;
    const bool some1 = m1 >= 1.17549435e-38f && m1 < 3.0e38f;
    const int own0 = some0 ? eb[0] + k0 - SHIFT : E_EMPTY, own1 = some1 ? eb[1] + k1 - SHIFT : E_EMPTY;
    int t0 = own0 - GAP * lane, t1 = own1 - GAP * lane;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {                            // lanes >= 32 - d get their own value back
        const int s0 = __shfl_down_sync(FULL, t0, d), s1 = __shfl_down_sync(FULL, t1, d);
        t0 = max(t0, s0);
        t1 = max(t1, s1);
    }
    const int new0 = t0 + GAP * lane, new1 = t1 + GAP * lane;
    const int nb0 = __shfl_down_sync(FULL, new0, 1), nb1 = __shfl_down_sync(FULL, new1, 1);
    const int top0 = __reduce_max_sync(FULL, own0), top1 = __reduce_max_sync(FULL, own1);
    const bool fast0 = some0 && own0 >= top0 - 64 && k0 - SHIFT < -decay_max;
    const bool fast1 = some1 && own1 >= top1 - 64 && k1 - SHIFT < -decay_max;
    const bool bad0 = (m0 != m0) || m0 >= 3.0e38f || fast0;
    const bool bad1 = (m1 != m1) || m1 >= 3.0e38f || fast1;
    const unsigned bits = __reduce_or_sync(FULL, (some0 ? 1u : 0u) | (bad0 ? 2u : 0u) | (some1 ? 4u : 0u) | (bad1 ? 8u : 0u));
    ok[0] = (bits & 3u) == 1u;
    ok[1] = (bits & 12u) == 4u;
    const float2 f2 = make_float2(pow2i(eb[0] - new0), pow2i(eb[1] - new1));
    eb[0] = new0;
    eb[1] = new1;
#pragma unroll
    for (int j = 0; j < NSL; ++j) b[j] = __fmul2_rn(b[j], f2);
    scb = make_float2(lane == 31 ? 0.f : pow2i(nb0 - new0), lane == 31 ? 0.f : pow2i(nb1 - new1));
}

}  // namespace lattice
}  // namespace mdd
