// ctc_fast.cu -- rescaled linear-domain CTC forward-backward: the fast path of mdd_ctc_loss.
//
// Semantics: torch aten ctc_loss / _ctc_loss_backward as reached from nn.CTCLoss at
// egs/attention_aug/steps/train_ctc.py:186,72,96 and egs/cnn-rnn-ctc/steps/train_ctc.py:148,51,67.
//
// Why not log space: a per-state log-sum-exp lattice needs ~7 MUFU ops per state-frame (S = 2L+1 = 241 states at
// L = 120), which bounds the kernel at ~10 % of the HBM roofline on the SFU pipe alone.  Here the
// emissions are exponentiated once per (t,c) and alpha/beta are propagated as plain products and sums (FADD/FMUL/
// FFMA) in a block-floating-point format: every lane carries one integer exponent for the NSL states it owns and
// renormalises its mantissas every 4 frames from its own maximum (no warp reduction).  A plain "scale the whole column"
// normalisation is not enough: with T >> L the alpha mass runs ahead to the last states and the beta mass back to the
// first ones, so at the states the posterior lives on both are > 2^126 below their column sums and fp32 loses them.
// Per-lane exponents give each 4-label stretch of the lattice its own range; neighbouring lanes exchange boundary
// states with a 2^(e_nb - e_me) factor that only changes at renormalisation time.  The loss is read off the final
// mantissa and exponent (no per-frame log), and the occupancy is self-normalised per frame
// (Z_t = sum_s alpha_t(s) u_t(s)), so no scale bookkeeping enters the gradient.  Utterances that are infeasible, empty,
// malformed or whose ranges still do not fit set flags[b] and are redone by the log-space kernel (ctc_log.cu) in the
// same stream: results never depend on this path being exact for pathological inputs.
//
// Mapping: one warp per utterance, NSL consecutive lattice states per lane held in registers; the s-1 / s-2
// neighbours across lanes come from one __shfl_up (alpha) or two __shfl_down (beta): no block barrier anywhere in the
// recursion.  Eight warps (adjacent utterances, so their 168 B rows of one frame form one contiguous run in HBM)
// share a CTA only to share the SM.
//
// Memory: the lattice never goes to HBM.  Forward sweep: log-prob rows are read KF frames ahead straight into
// registers (one 8-byte load per lane per row), exponentiated and parked in shared memory when the chunk starts;
// alpha enters each chunk as a checkpoint row written to the workspace ((S+32)*4/KF B per frame).  Backward sweep,
// chunks from the end: reload the chunk's rows (the last chunk is still in shared memory), recompute alpha for KF
// frames into shared memory, run beta; each frame's per-state occupancies overwrite the alpha row in class-sorted
// order, and a chunk-batched, deterministic class reduction (lanes = KF frames x 32/KF stretches of the sorted
// list) produces grad = (p - occupancy / Z) * g for KF rows at once.
#include <cstdlib>

#include "common.cuh"
#include "ctc_params.cuh"

namespace mdd {
namespace {

constexpr int RENORM_MASK = 3;         // renormalise before frames with (t & 3) == 0 (alpha) / == 3 (beta)
constexpr int E_EMPTY = -(1 << 20);    // exponent of a lane whose states are all zero

template <int NSL>
struct Geo {
    static constexpr int NL = NSL / 2;            // label states per lane (odd j)
    static constexpr int SLOTS = 32 * NSL;        // lattice slots per utterance
    // A lane's exponent is kept within 2^GAP of the lane it receives from.  Between two renormalisations (4 frames)
    // mass can cross ceil(8/NSL)+1 lane boundaries, each multiplying the mantissa by up to 2^GAP, so GAP * crossings
    // (+7 bits of growth) must stay below the fp32 exponent range.
    static constexpr int GAP = NSL == 2 ? 20 : NSL == 4 ? 36 : 56;
    static constexpr int RS = SLOTS + 4;          // shared-memory row stride (floats): +4 keeps 16 B alignment and
                                                  // spreads the frame rows over banks for the class walk
};

// alpha~ row <-> registers.  Layout: float4 group g of lane l at (g*128 + 4*l): conflict-free 128-bit accesses.
template <int NSL>
__device__ __forceinline__ void row_store(float* row, int lane, const float (&a)[NSL]) {
    if constexpr (NSL == 2) {
        *reinterpret_cast<float2*>(row + 2 * lane) = make_float2(a[0], a[1]);
    } else {
#pragma unroll
        for (int g = 0; g < NSL / 4; ++g)
            *reinterpret_cast<float4*>(row + g * 128 + 4 * lane) = make_float4(a[4 * g], a[4 * g + 1], a[4 * g + 2], a[4 * g + 3]);
    }
}
template <int NSL>
__device__ __forceinline__ void row_load(const float* row, int lane, float (&a)[NSL]) {
    if constexpr (NSL == 2) {
        const float2 v = *reinterpret_cast<const float2*>(row + 2 * lane);
        a[0] = v.x; a[1] = v.y;
    } else {
#pragma unroll
        for (int g = 0; g < NSL / 4; ++g) {
            const float4 v = *reinterpret_cast<const float4*>(row + g * 128 + 4 * lane);
            a[4 * g] = v.x; a[4 * g + 1] = v.y; a[4 * g + 2] = v.z; a[4 * g + 3] = v.w;
        }
    }
}

struct FastSmem {          // per-warp carve-up (floats / ints), computed on the host and in the kernel identically
    int rows_off, ctab_off, lab_off, eat_off, total_floats;
};
__host__ __device__ inline FastSmem fast_smem_layout(int NSL, int KF, int CP, int Lcap) {
    FastSmem s;
    int o = KF * (32 * NSL + 4);          // alpha chunk, then occupancy prefix sums
    s.rows_off = o; o += KF * CP;         // exp(lp) rows of the current chunk
    s.ctab_off = o; o += 128;             // per class: first slot, count in the class-sorted order
    s.lab_off = o;  o += Lcap;            // labels
    s.eat_off = o;  o += (KF / 4) * 32;   // alpha exponents of the chunk's renormalisation groups
    s.total_floats = (o + 3) & ~3;
    return s;
}

// ---------------------------------------------------------------------------------------------------------------
// One lattice step.  Lane l owns states s = NSL*l + j; even j = blank, odd j = label q = (j-1)/2.  Values are
// mantissas relative to the lane's exponent; sc_in = 2^(e_neighbour - e_me) rescales what crosses the lane boundary.
//   alpha: new[s] = (a[s] + a[s-1] + ma * a[s-2]) * e[s]        (ma = skip allowed into label s)
//   beta : u[s]   =  b[s] + b[s+1] + mb * b[s+2];  new b[s] = u[s] * e[s]   (u = beta without the emission at t)
// Label slots past the end of the target read a zero emission (row[C] = 0), which also keeps every later slot at 0.
// ---------------------------------------------------------------------------------------------------------------
template <int NSL>
__device__ __forceinline__ void alpha_step(float (&a)[NSL], float sc_in, const float* __restrict__ row, int blank,
                                           const int (&eidx)[NSL / 2], const float (&ma)[NSL / 2]) {
    const float p = __shfl_up_sync(0xffffffffu, a[NSL - 1], 1) * sc_in;     // state NSL*l - 1 (sc_in == 0 on lane 0)
    const float eb = row[blank];
    float e[NSL / 2];
#pragma unroll
    for (int q = 0; q < NSL / 2; ++q) e[q] = row[eidx[q]];
    float n[NSL];
#pragma unroll
    for (int j = 0; j < NSL; ++j) {
        const float m1 = j ? a[j - 1] : p;
        if (j & 1) {
            const float m2 = j >= 2 ? a[j - 2] : p;
            n[j] = fmaf(ma[j >> 1], m2, a[j] + m1) * e[j >> 1];
        } else {
            n[j] = (a[j] + m1) * eb;
        }
    }
#pragma unroll
    for (int j = 0; j < NSL; ++j) a[j] = n[j];
}

template <int NSL>
__device__ __forceinline__ void beta_step(float (&b)[NSL], float (&u)[NSL], float sc_in, const float* __restrict__ row,
                                          int blank, const int (&eidx)[NSL / 2], const float (&mb)[NSL / 2]) {
    const float n0 = __shfl_down_sync(0xffffffffu, b[0], 1) * sc_in;        // sc_in == 0 on lane 31
    const float n1 = __shfl_down_sync(0xffffffffu, b[1], 1) * sc_in;
    const float eb = row[blank];
    float e[NSL / 2];
#pragma unroll
    for (int q = 0; q < NSL / 2; ++q) e[q] = row[eidx[q]];
#pragma unroll
    for (int j = 0; j < NSL; ++j) {
        const float p1 = (j + 1 < NSL) ? b[j + 1] : n0;
        if (j & 1) {
            const float p2 = (j + 2 < NSL) ? b[j + 2] : n1;
            u[j] = fmaf(mb[j >> 1], p2, b[j] + p1);
        } else {
            u[j] = b[j] + p1;
        }
    }
#pragma unroll
    for (int j = 0; j < NSL; ++j) b[j] = u[j] * ((j & 1) ? e[j >> 1] : eb);
}

__device__ __forceinline__ float pow2i(int x) {            // 2^x, 0 below the normal range
    return x < -126 ? 0.f : __int_as_float((127 + min(x, 127)) << 23);
}

// Per-lane renormalisation.  v = mantissas, e = the lane's exponent.  The lane's maximum is brought to [1,2); a lane
// is never allowed more than 2^GAP below the lane it receives from (FWD: lane-1, else lane+1) so that the boundary
// factor sc_in = 2^(e_nb - e_me) stays finite; empty lanes inherit their exponent from that side.
// Returns false if every lane is empty (no alignment mass left) or real content had to be squashed by > 2^100.
template <int NSL, bool FWD>
__device__ __forceinline__ bool renorm_lane(float (&v)[NSL], int& e, float& sc_in, int lane) {
    float m = v[0];
#pragma unroll
    for (int j = 1; j < NSL; ++j) m = fmaxf(m, v[j]);
    const int k = ((__float_as_int(m) >> 23) & 0xff) - 127;
    const bool some = m >= 1.17549435e-38f && m < 3.0e38f;       // normal, finite (NaN fails both)
    const int e_own = some ? e + k : E_EMPTY;
    constexpr int GAP = Geo<NSL>::GAP;
    const int pos = FWD ? lane : 31 - lane;
    int t = e_own + GAP * pos;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int o = FWD ? __shfl_up_sync(0xffffffffu, t, d) : __shfl_down_sync(0xffffffffu, t, d);
        if (pos >= d) t = max(t, o);
    }
    const int e_new = t - GAP * pos;
    const float f = pow2i(e - e_new);
#pragma unroll
    for (int j = 0; j < NSL; ++j) v[j] *= f;
    e = e_new;
    const int e_nb = FWD ? __shfl_up_sync(0xffffffffu, e_new, 1) : __shfl_down_sync(0xffffffffu, e_new, 1);
    sc_in = (pos == 0) ? 0.f : pow2i(e_nb - e_new);
    const bool bad = (m != m) || m >= 3.0e38f || (some && e_new - e_own > 100);
    const bool alive = __any_sync(0xffffffffu, some);
    return alive && !__any_sync(0xffffffffu, bad);
}


// Two floats per lane per row.  VEC: columns (2*lane, 2*lane+1) as one 8-byte access, else columns (lane, lane+32).
// Loads are unconditional (lanes past the row re-read a clamped column) so that a chunk's loads issue back to back.
struct Cols {
    int c0, c1;         // my two columns (class ids)
    bool v0, v1;        // inside the row?
    int l0, l1;         // clamped columns for the unconditional loads
};
template <bool VEC>
__device__ __forceinline__ Cols make_cols(int lane, int C) {
    Cols k;
    if constexpr (VEC) { k.c0 = 2 * lane; k.c1 = 2 * lane + 1; k.v0 = k.v1 = (2 * lane < C); }
    else { k.c0 = lane; k.c1 = lane + 32; k.v0 = lane < C; k.v1 = lane + 32 < C; }
    k.l0 = k.v0 ? k.c0 : 0;
    k.l1 = k.v1 ? k.c1 : 0;
    return k;
}
template <bool VEC>
__device__ __forceinline__ float2 ld_row2(const float* __restrict__ src, const Cols& k) {
    if constexpr (VEC) return __ldcs(reinterpret_cast<const float2*>(src + k.l0));
    else return make_float2(__ldcs(src + k.l0), __ldcs(src + k.l1));
}
template <bool VEC>
__device__ __forceinline__ void st_row2_smem(float* __restrict__ dst, const Cols& k, float2 v) {
    if constexpr (VEC) { if (k.v0) *reinterpret_cast<float2*>(dst + k.c0) = v; }
    else { if (k.v0) dst[k.c0] = v.x; if (k.v1) dst[k.c1] = v.y; }
}
template <bool VEC>
__device__ __forceinline__ float2 ld_row2_smem(const float* __restrict__ src, const Cols& k) {
    if constexpr (VEC) return *reinterpret_cast<const float2*>(src + k.l0);
    else return make_float2(src[k.l0], src[k.l1]);
}
template <bool VEC>
__device__ __forceinline__ void st_row2_gmem(float* __restrict__ dst, const Cols& k, float2 v) {
    if constexpr (VEC) { if (k.v0) __stcs(reinterpret_cast<float2*>(dst + k.c0), v); }
    else { if (k.v0) __stcs(dst + k.c0, v.x); if (k.v1) __stcs(dst + k.c1, v.y); }
}
// exp(x) = 2^(x * log2 e) on the SFU; results below 2^-126 flush to 0 (such rows end in the log-space kernel anyway)
__device__ __forceinline__ float fast_exp(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x * 1.4426950408889634f));
    return y;
}

template <int NSL, int KF, bool VEC>
__global__ void __launch_bounds__(KF == 16 ? 256 : 512, 1)
ctc_fast_kernel(CtcParams p, float* __restrict__ ckpt_ws, int n_ckpt, int CP, int Lcap) {
    constexpr int NL = NSL / 2;
    constexpr int SLOTS = Geo<NSL>::SLOTS;
    constexpr int RS = Geo<NSL>::RS;
    constexpr int CKS = SLOTS + 32;                     // checkpoint row: mantissas + one exponent per lane
    constexpr int NPART = 32 / KF;                      // stretches of the class-sorted list walked in parallel
    // the four spare slots at the end of every chunk row
    constexpr int DUMP = SLOTS;                         // swallows the products of unused label slots
    constexpr int ZERO = SLOTS + 1;                     // always 0
    constexpr int BLK = SLOTS + 2;                      // blank-class occupancy of the frame
    constexpr int ZINV = SLOTS + 3;                     // 1 / Z_t
    static_assert(RS == SLOTS + 4, "row layout");
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int b = blockIdx.x * (blockDim.x >> 5) + warp;
    if (b >= p.B) return;

    extern __shared__ __align__(16) float smem_f[];
    const FastSmem lay = fast_smem_layout(NSL, KF, CP, Lcap);
    float* base = smem_f + (size_t)warp * lay.total_floats;
    float* chunk = base;                                         // [KF][RS]  alpha rows, then occupancy prefix sums
    float* rows = base + lay.rows_off;                           // [KF][CP] exp(lp) rows, row[C..CP) = 0
    int* ctab = reinterpret_cast<int*>(base + lay.ctab_off);     // [64] first sorted slot, [64] count per class
    int* lab = reinterpret_cast<int*>(base + lay.lab_off);       // [Lcap]
    int* eat = reinterpret_cast<int*>(base + lay.eat_off);       // [KF/4][32]

    const int C = p.C, blank = p.blank, T = p.T;
    const long long Tb64 = p.in_len[b], L64 = p.tgt_len[b];
    const bool want_grad = p.grad != nullptr;
    // anything unusual goes to the log-space kernel, which also owns the error reporting
    if (Tb64 <= 0 || Tb64 > T || L64 < 0 || L64 > p.Lmax || 2 * L64 + 1 > SLOTS) {
        if (lane == 0) p.flags[b] = 1;
        return;
    }
    const int Tb = (int)Tb64, L = (int)L64, S = 2 * L + 1;
    const int64_t* tg = p.tgt_stride ? p.targets + (int64_t)b * p.tgt_stride : p.targets + p.tgt_offsets[b];
    int flag = 0;
    for (int i = lane; i < L; i += 32) {
        const long long v = tg[i];
        if (v < 0 || v >= C || v == blank) flag = 1;
        lab[i] = (int)v;
    }
    flag = __any_sync(0xffffffffu, flag);
    if (flag) {
        if (lane == 0) p.flags[b] = 1;
        return;
    }
    // zero pads of the staged rows and the ZERO slot of the chunk rows (never overwritten)
    if (lane < KF) {
        for (int c = C; c < CP; ++c) rows[lane * CP + c] = 0.f;
        chunk[lane * RS + ZERO] = 0.f;
    }
    const Cols cols = make_cols<VEC>(lane, C);
    __syncwarp();

    // per-lane constants: emission index and skip masks of my NL label states
    int eidx[NL];
    float ma[NL], mb[NL];
#pragma unroll
    for (int q = 0; q < NL; ++q) {
        const int i = lane * NL + q;
        const bool in = i < L;
        const int c = in ? lab[i] : C;                          // row[C] == 0
        eidx[q] = c;
        ma[q] = (in && i > 0 && lab[i - 1] != c) ? 1.f : 0.f;
        mb[q] = (in && i + 1 < L && lab[i + 1] != c) ? 1.f : 0.f;
    }

    const float* lp_b = p.lp + (size_t)b * p.stride_b;
    const size_t gstride = (size_t)p.B * C;
    float* gr = want_grad ? p.grad + (size_t)b * C : nullptr;
    const float g = (p.grad_out ? p.grad_out[b] : 1.f) * p.gscalar;
    const int nch = (Tb + KF - 1) / KF;

    float2 pre[KF];                                     // the next chunk's rows, in flight
    auto load_rows = [&](int j) {                       // rows past the end of a partial chunk re-read its last row
        const int t0 = j * KF, n = min(KF, Tb - t0);
        const float* src = lp_b + (size_t)t0 * p.stride_t;
#pragma unroll
        for (int r = 0; r < KF; ++r) pre[r] = ld_row2<VEC>(src + (size_t)min(r, n - 1) * p.stride_t, cols);
    };
    auto commit_rows = [&]() {                          // p = exp(lp) into shared memory
#pragma unroll
        for (int r = 0; r < KF; ++r)
            st_row2_smem<VEC>(rows + r * CP, cols, make_float2(fast_exp(pre[r].x), fast_exp(pre[r].y)));
    };

    // ------------------------------------------------------------------------------------------ forward sweep
    float a[NSL];
#pragma unroll
    for (int j = 0; j < NSL; ++j) a[j] = 0.f;
    if (lane == 0) a[0] = 1.f;                          // virtual alpha_{-1}: one step of the recursion gives alpha_0
    int ea = 0;                                         // alpha(s) = a[j] * 2^ea
    float sca = 0.f;                                    // 2^(ea[lane-1] - ea[lane])
    float* ck = ckpt_ws + (size_t)b * n_ckpt * CKS;

    load_rows(0);
    // rows t >= T_b of the gradient are zero (torch semantics); overlaps with the first loads
    if (want_grad)
        for (int t = Tb; t < T; ++t) st_row2_gmem<VEC>(gr + (size_t)t * gstride, cols, make_float2(0.f, 0.f));

    for (int j = 0; j < nch; ++j) {
        const int n = min(KF, Tb - j * KF);
        __syncwarp();
        commit_rows();
        __syncwarp();
        if (j + 1 < nch) load_rows(j + 1);
        if (want_grad && j > 0) {                       // alpha entering chunk j (before its first renormalisation)
            row_store<NSL>(ck + (size_t)j * CKS, lane, a);
            ck[(size_t)j * CKS + SLOTS + lane] = __int_as_float(ea);
        }
#pragma unroll 4
        for (int r = 0; r < n; ++r) {
            if ((r & RENORM_MASK) == 0) {
                if (!renorm_lane<NSL, true>(a, ea, sca, lane) && !flag) flag = 2 | ((j * KF + r) << 8);
            }
            alpha_step<NSL>(a, sca, rows + r * CP, blank, eidx, ma);
        }
    }
    {   // P = alpha_{T-1}(S-1) + alpha_{T-1}(S-2), possibly in two lanes with different exponents
        float v = 0.f;
#pragma unroll
        for (int j = 0; j < NSL; ++j) {
            const int s = lane * NSL + j;
            if (s == S - 1 || s == S - 2) v += a[j];
        }
        int em = (v > 0.f) ? ea : E_EMPTY;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) em = max(em, __shfl_xor_sync(0xffffffffu, em, d));
        v = warp_sum(v * pow2i(ea - em));
        if ((em == E_EMPTY || !(v > 1e-30f) || v > 1e37f) && !flag) flag = 3;
        if (lane == 0 && !flag) {
            p.nll[b] = (float)(-((double)logf(v) + (double)em * 0.69314718055994530942));
            if (p.status) p.status[b] = MDD_ST_OK;
        }
    }
    if (flag || !want_grad) {
        if (lane == 0) p.flags[b] = flag;
        return;
    }

    // ------------------------------------------------------------------------------------------ class-sorted layout
    // Label i gets slot perm(i) of a list sorted by class (positions ascending inside a class).  split[] cuts the list
    // at class boundaries into NPART stretches of similar length; each stretch is prefix-summed independently, so the
    // occupancy of class c is P[oe] - P[os] with os = ZERO when c opens a stretch (absent classes: both ZERO).
    int perm[NL];
    int split[NPART + 1];
    int oe0, os0, oe1, os1;
    {
        int* scratch = reinterpret_cast<int*>(chunk);   // label -> slot, read back into registers below
        int cnt[2] = {0, 0};
        for (int i = 0; i < L; ++i) {
            const int v = lab[i];
            cnt[0] += (v == lane);
            cnt[1] += (v == lane + 32);
        }
        int start[2], basek = 0;
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            int inc = cnt[k];
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int o = __shfl_up_sync(0xffffffffu, inc, d);
                if (lane >= d) inc += o;
            }
            start[k] = basek + inc - cnt[k];
            basek += __shfl_sync(0xffffffffu, inc, 31);
        }
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const int c = lane + 32 * k;
            ctab[c] = start[k];
            ctab[64 + c] = cnt[k];
            if (cnt[k] > 0) {
                int w = start[k];
                for (int i = 0; i < L; ++i)
                    if (lab[i] == c) {
                        scratch[i] = w;
                        ++w;
                    }
            }
        }
        split[0] = 0;
        split[NPART] = L;
#pragma unroll
        for (int part = 1; part < NPART; ++part) {
            const int want = (L * part) / NPART;
            int best = 0x7fffffff;
#pragma unroll
            for (int k = 0; k < 2; ++k)
                if (cnt[k] > 0) best = min(best, abs(start[k] - want) * 1024 + start[k]);
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, d));
            split[part] = (best == 0x7fffffff) ? 0 : (best & 1023);
        }
#pragma unroll
        for (int part = 1; part < NPART; ++part) split[part] = max(split[part], split[part - 1]);
        __syncwarp();
#pragma unroll
        for (int q = 0; q < NL; ++q) {
            const int i = lane * NL + q;
            perm[q] = i < L ? scratch[i] : DUMP;
        }
        auto occ_slots = [&](int c, bool valid, int& oe, int& os) {
            oe = ZERO; os = ZERO;
            if (!valid) return;
            if (c == blank) { oe = BLK; return; }
            const int st = ctab[c], n = ctab[64 + c];
            if (n == 0) return;
            oe = st + n - 1;
            bool opens = false;
#pragma unroll
            for (int part = 0; part < NPART; ++part) opens |= (st == split[part]);
            os = opens ? ZERO : st - 1;
        };
        occ_slots(cols.c0, cols.v0, oe0, os0);
        occ_slots(cols.c1, cols.v1, oe1, os1);
        __syncwarp();
    }

    // ------------------------------------------------------------------------------------------ backward sweep
    float bt[NSL];
#pragma unroll
    for (int j = 0; j < NSL; ++j) bt[j] = (lane * NSL + j == S - 1) ? 1.f : 0.f;     // virtual beta_T
    int eb = 0;
    float scb = 0.f;
    const int f = lane & (KF - 1), h = lane / KF;
    int kb = 0, ke = 0;
#pragma unroll
    for (int part = 0; part < NPART; ++part)
        if (h == part) { kb = split[part]; ke = split[part + 1]; }

    for (int j = nch - 1; j >= 0; --j) {
        const int t0 = j * KF, n = min(KF, Tb - t0);
        if (j < nch - 1) {                              // the last chunk is still staged (and exponentiated)
            __syncwarp();
            commit_rows();
            __syncwarp();
        }
        if (j > 0) load_rows(j - 1);                    // prefetch the next (earlier) chunk

        // recompute alpha for the chunk from its checkpoint (bitwise the same operations as the forward sweep)
        if (j == 0) {
#pragma unroll
            for (int q = 0; q < NSL; ++q) a[q] = 0.f;
            if (lane == 0) a[0] = 1.f;
            ea = 0;
        } else {
            row_load<NSL>(ck + (size_t)j * CKS, lane, a);
            ea = __float_as_int(ck[(size_t)j * CKS + SLOTS + lane]);
        }
#pragma unroll 4
        for (int r = 0; r < n; ++r) {
            if ((r & RENORM_MASK) == 0) {
                renorm_lane<NSL, true>(a, ea, sca, lane);
                eat[(r >> 2) * 32 + lane] = ea;
            }
            alpha_step<NSL>(a, sca, rows + r * CP, blank, eidx, ma);
            row_store<NSL>(chunk + r * RS, lane, a);
        }
        __syncwarp();

        // beta, per-state occupancies alpha_t(s) * u_t(s) on a per-frame common scale
        // (labels -> class-sorted slots [0, L), blank partial sums -> [32*NL, +32))
        float scp = 0.f;                                // 2^(ea + eb - max over lanes)
        for (int r = n - 1; r >= 0; --r) {
            if (r == n - 1 || (r & RENORM_MASK) == RENORM_MASK) {
                if (!renorm_lane<NSL, false>(bt, eb, scb, lane) && !flag) flag = 4 | ((t0 + r) << 8);
                const int E = eat[(r >> 2) * 32 + lane] + eb;
                int Em = E;
#pragma unroll
                for (int d = 16; d > 0; d >>= 1) Em = max(Em, __shfl_xor_sync(0xffffffffu, Em, d));
                scp = pow2i(E - Em);
            }
            float u[NSL], al[NSL];
            beta_step<NSL>(bt, u, scb, rows + r * CP, blank, eidx, mb);
            float* cr = chunk + r * RS;
            row_load<NSL>(cr, lane, al);
            float bp = 0.f, pl[NL];
#pragma unroll
            for (int q = 0; q < NL; ++q) {
                bp = fmaf(al[2 * q], u[2 * q], bp);
                pl[q] = al[2 * q + 1] * u[2 * q + 1] * scp;
            }
            bp *= scp;
            __syncwarp();                               // every lane has read its alpha before the row is reused
#pragma unroll
            for (int q = 0; q < NL; ++q) cr[perm[q]] = pl[q];
            cr[32 * NL + lane] = bp;
        }
        __syncwarp();

        // chunk-batched class reduction: lane = (frame f, stretch h of the class-sorted list); in-place prefix sums
        {
            float* pr = chunk + f * RS;
            const bool live = f < n;
            const int k1 = live ? ke : kb;
            float acc = 0.f;
            for (int k = kb; k < k1; ++k) {
                acc += pr[k];
                pr[k] = acc;
            }
            float bs = 0.f;
            if (live) {
#pragma unroll
                for (int i = 0; i < KF; ++i) bs += pr[32 * NL + KF * h + i];
            }
#pragma unroll
            for (int d = KF; d < 32; d <<= 1) {
                bs += __shfl_xor_sync(0xffffffffu, bs, d);
                acc += __shfl_xor_sync(0xffffffffu, acc, d);
            }
            if (h == 0 && live) {
                const float Z = acc + bs;
                pr[BLK] = bs;
                pr[ZINV] = __fdividef(1.f, Z);
                if ((!(Z > 1e-30f) || Z > 1e37f) && !flag) flag = 5 | ((t0 + f) << 8);
            }
        }
        __syncwarp();

        // grad rows of the chunk: (softmax prob - occupancy / Z) * g
#pragma unroll 4
        for (int r = 0; r < n; ++r) {
            const float* pr = chunk + r * RS;
            const float zi = pr[ZINV];
            const float2 pv = ld_row2_smem<VEC>(rows + r * CP, cols);
            const float o0 = pr[oe0] - pr[os0], o1 = pr[oe1] - pr[os1];
            st_row2_gmem<VEC>(gr + (size_t)(t0 + r) * gstride, cols,
                              make_float2((pv.x - o0 * zi) * g, (pv.y - o1 * zi) * g));
        }
        __syncwarp();
    }
    {
        int fl = flag;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) fl = max(fl, __shfl_xor_sync(0xffffffffu, fl, d));
        if (lane == 0) p.flags[b] = fl;     // 0 = done here; otherwise reason | frame << 8 (diagnostic), redone in log space
    }
}

int pick_nsl(int Lmax) {
    const int S = 2 * Lmax + 1;
    if (S <= 64) return 2;
    if (S <= 128) return 4;
    if (S <= 256) return 8;
    if (S <= 512) return 16;
    return 0;
}
inline int cp_of(int C) { return (C + 2) & ~1; }
inline int lcap_of(int Lmax) { return ((Lmax > 0 ? Lmax : 1) + 3) & ~3; }

int fast_kf() {                         // frames per chunk: 8 (more warps per SM) or 16 (half the checkpoint traffic)
    static int kf = 0;
    if (!kf) {
        const char* e = getenv("MDD_CTC_FAST_KF");
        kf = (e && atoi(e) == 16) ? 16 : 8;
    }
    return kf;
}

template <int NSL, int KF, bool VEC>
int launch_cfg(const CtcParams& p, float* ws, cudaStream_t st) {
    const int CP = cp_of(p.C), Lcap = lcap_of(p.Lmax);
    const FastSmem lay = fast_smem_layout(NSL, KF, CP, Lcap);
    const size_t per_warp = (size_t)lay.total_floats * 4;
    constexpr int kMaxWarps = KF == 16 ? 8 : 16;
    int warps = (int)((224 * 1024) / per_warp);
    if (warps > kMaxWarps) warps = kMaxWarps;
    if (warps < 1) return MDD_E_UNSUPPORTED;
    // small batches: spread the utterances over the SMs instead of packing them into few CTAs
    const int spread = (p.B + kNumSM - 1) / kNumSM;
    if (warps > spread) warps = spread > 0 ? spread : 1;
    const size_t smem = per_warp * warps;
    static bool configured = false;
    if (!configured) {
        if (cudaFuncSetAttribute(ctc_fast_kernel<NSL, KF, VEC>, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024) != cudaSuccess) {
            cudaGetLastError();
            return MDD_E_CUDA;
        }
        configured = true;
    }
    const int n_ckpt = (p.T + KF - 1) / KF + 1;
    const int grid = (p.B + warps - 1) / warps;
    ctc_fast_kernel<NSL, KF, VEC><<<grid, warps * 32, smem, st>>>(p, ws, n_ckpt, CP, Lcap);
    MDD_CHECK_LAUNCH();
    return MDD_OK;
}

template <int NSL>
int launch_nsl(const CtcParams& p, float* ws, cudaStream_t st) {
    const bool vec = (p.C % 2 == 0) && (p.stride_t % 2 == 0) && (p.stride_b % 2 == 0) && ((uintptr_t)p.lp % 8 == 0) &&
                     (!p.grad || (uintptr_t)p.grad % 8 == 0);
    if (fast_kf() == 16) return vec ? launch_cfg<NSL, 16, true>(p, ws, st) : launch_cfg<NSL, 16, false>(p, ws, st);
    return vec ? launch_cfg<NSL, 8, true>(p, ws, st) : launch_cfg<NSL, 8, false>(p, ws, st);
}

}  // namespace

bool ctc_fast_supported(int C, int Lmax) { return C <= 64 && pick_nsl(Lmax) != 0; }

size_t ctc_fast_ws_bytes(int T, int B, int Lmax) {
    const int nsl = pick_nsl(Lmax);
    if (!nsl) return 0;
    const size_t n_ckpt = (size_t)(T + 7) / 8 + 1;           // sized for the smaller chunk (KF = 8)
    return (size_t)B * n_ckpt * (32 * nsl + 32) * 4;         // mantissas + per-lane exponents
}

int launch_ctc_fast(CtcParams p, void* ws, cudaStream_t st) {
    float* w = reinterpret_cast<float*>(ws);
    switch (pick_nsl(p.Lmax)) {
        case 2: return launch_nsl<2>(p, w, st);
        case 4: return launch_nsl<4>(p, w, st);
        case 8: return launch_nsl<8>(p, w, st);
        case 16: return launch_nsl<16>(p, w, st);
        default: return MDD_E_UNSUPPORTED;
    }
}

}  // namespace mdd
