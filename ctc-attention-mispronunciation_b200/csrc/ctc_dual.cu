// ctc_dual.cu -- first half of the CTC path for FEW utterances: alpha and beta swept at the same time.
//
// Semantics: torch aten ctc_loss / _ctc_loss_backward as reached from nn.CTCLoss at
// egs/attention_aug/steps/train_ctc.py:186,72,96 and egs/cnn-rnn-ctc/steps/train_ctc.py:148,51,67 -- at the batch sizes
// the reference really trains with (8-64 utterances, conf/ctc_config*.yaml:76), and for one GPU's share of a batch
// that is sharded eight ways.
//
// With a few pairs per SM the step time is ONE warp's dependent chain: forward sweep, then the backward sweep with its
// recomputation, 3 T lattice steps one after the other (ctc_team.cu).  Nothing in the mathematics orders the two
// sweeps: beta does not depend on alpha.  So here
//   * one CTA per pair runs BOTH sweeps concurrently, each with its own lattice warp and its own helper warps (rows
//     fetched by cp.async two chunks ahead, exponentiated into bank-coloured emission buffers; mbarrier hand-offs as
//     in ctc_team.cu).  alpha walks the chunks upwards and leaves a checkpoint per chunk, beta walks them downwards,
//     carries ITS OWN per-lane exponents (renorm_beta, the mirror image of renorm_alpha) and leaves a checkpoint per
//     chunk too.  The chain is T steps instead of 3 T;
//   * with both checkpoint sets in the workspace every chunk of the gradient is independent of every other: the second
//     half is the backward kernel of ctc_pair.cu started on many segments of every pair at once (SEG instantiation:
//     a segment takes beta from the checkpoint at its upper end and ties its exponents to alpha's from there on,
//     exactly as the sequential sweep does), which makes it a throughput kernel whatever the batch size.
// The two sweeps check each other: beta_0(0) + beta_0(1) from the beta sweep must agree with P from the alpha sweep
// (flag 7), mass dropped by either renormalisation is missing from exactly one of the two.
// The layout of the emission row (bank-coloured for long sweeps, one utterance per lattice warp; plain for short ones)
// is decided here and handed to the second half through the workspace.
#include <algorithm>
#include <cstdlib>

#include "common.cuh"
#include "ctc_lattice.cuh"
#include "ctc_params.cuh"

namespace mdd {
namespace {

using namespace lattice;

constexpr int DRAWW = 132;                 // floats per staged pair row: 2*C log-probs (C <= 62) + 2 row-LSE + pad
constexpr int DNSLOT = 3;                  // raw staging slots per side
constexpr int DNE = 3;                     // emission buffers per side

template <int NSL, int NH>
struct DualLayout {
    static constexpr int CKW = PGeo<NSL>::CKW;
    static constexpr int E_FLOATS = DNE * KF * RW;
    static constexpr int RAW_FLOATS = DNSLOT * KF * DRAWW;
    static constexpr int SIDE = E_FLOATS + RAW_FLOATS;           // one side: emissions, then raw rows
    static constexpr int SCR_OFF = 2 * SIDE;                     // setup tables (ints)
    static constexpr int LAB = 0, GM = 512, CPOS = 656, IPOS = 800, META = 928, SCR_INTS = 976;
    static constexpr int BAR_OFF = (SCR_OFF + SCR_INTS + 1) & ~1;
    static constexpr int NBAR = 4 * DNE;                         // per side: e_full[DNE], e_free[DNE]
    static constexpr int TOTAL = BAR_OFF + 2 * NBAR;
    static_assert(SIDE % 4 == 0, "16-byte aligned sides");
};

#ifdef MDD_DUAL_TIMING
__device__ unsigned long long g_dual_cycles[16];
#define DT(i) if (blockIdx.x == 0 && lane == 0) g_dual_cycles[i] = clock64() - dt_start;
#else
#define DT(i)
#endif

// Non-blocking look at an mbarrier phase: issued at the top of a chunk, consumed after the renormalisation, so that the
// ~100-cycle round trip of the wait that almost always succeeds (the helpers run up to three chunks ahead) is off the
// lattice warp's dependent chain.
__device__ __forceinline__ bool mbar_test(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}

__device__ __forceinline__ void set_comp(float2& v, int h, float x) { if (h) v.y = x; else v.x = x; }

template <int NSL, int NH, bool LOGITS>
__global__ void __launch_bounds__(64 * (1 + NH), NH == 1 ? 4 : 2) ctc_dual_kernel(CtcParams p, float* __restrict__ ck_alpha,
                                                                    float* __restrict__ ck_beta, int n_ckpt) {
    using G = PGeo<NSL>;
    using LY = DualLayout<NSL, NH>;
    constexpr int NL = G::NL, CKW = G::CKW, CH = KF;
    extern __shared__ __align__(16) float dual_smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int side = warp > NH ? 1 : 0;                          // 0: alpha (chunks upwards), 1: beta (downwards)
    const int wis = warp - side * (NH + 1);                      // warp within the side: 0 = lattice, 1..NH = helpers
    const bool lattice_warp = wis == 0;
    const int hid = wis - 1;
    const int pair = blockIdx.x;

    float* E = dual_smem + side * LY::SIDE;                      // [DNE][KF][RW]
    float* raw = E + LY::E_FLOATS;                               // [DNSLOT][KF][DRAWW]
    int* scr = reinterpret_cast<int*>(dual_smem + LY::SCR_OFF);
    int* lab = scr + LY::LAB;
    int* gm = scr + LY::GM;
    int* cpos = scr + LY::CPOS;
    int* ipos = scr + LY::IPOS;
    int* meta = scr + LY::META;                                  // act[2], Tb[2], L[2] | results of the two lattice warps
    uint64_t* bars = reinterpret_cast<uint64_t*>(dual_smem + LY::BAR_OFF) + side * 2 * DNE;
    uint64_t* e_full = bars;                                     // [DNE] helpers -> lattice
    uint64_t* e_free = bars + DNE;                               // [DNE] lattice -> helpers

    const int C = p.C, blank = p.blank, T = p.T;
    const bool want_grad = p.grad != nullptr;
#ifdef MDD_DUAL_TIMING
    const long long dt_start = clock64();
#endif

    // The whole grid is one wave (at most two CTAs per SM): the second half may be scheduled behind it at once and do its
    // set-up while the sweeps run (it waits for them with pdl_wait before it touches what they produce).
    pdl_launch_dependents();
    if (threadIdx.x == 0) {
        if (p.ticket && blockIdx.x == 0) *p.ticket = 0u;         // first kernel of the call (see ctc_exact.cu)
        uint64_t* b0 = reinterpret_cast<uint64_t*>(dual_smem + LY::BAR_OFF);
        for (int s = 0; s < 2; ++s)
            for (int i = 0; i < DNE; ++i) { mbar_init(&b0[s * 2 * DNE + i], NH); mbar_init(&b0[s * 2 * DNE + DNE + i], 1); }
        mbar_fence_init();
    }
    // ------------------------------------------------------------------ setup: lattice warp of side h takes utterance h
    if (lattice_warp) {
        const int h = side;
        const int b = 2 * pair + h;
        bool act = b < p.B;
        int Tb = 0, L = 0;
        if (act) {
            const long long Tb64 = p.in_len[b], L64 = p.tgt_len[b];
            bool bad = Tb64 <= 0 || Tb64 > T || L64 < 0 || L64 > p.Lmax || L64 > G::LCAP;
            if (!bad) {
                const int64_t* tg = p.tgt_stride ? p.targets + (int64_t)b * p.tgt_stride : p.targets + p.tgt_offsets[b];
                int fl = 0;
                for (int i = lane; i < (int)L64; i += 32) {
                    const long long v = tg[i];
                    if (v < 0 || v >= C || v == blank) fl = 1;
                    lab[h * 256 + i] = (int)v;
                }
                bad = __any_sync(FULL, fl);
            }
            if (bad) {
                if (lane == 0) p.flags[b] = 1;                   // the fp64 kernel owns the error reporting
                act = false;
            } else {
                Tb = (int)Tb64;
                L = (int)L64;
            }
        }
        if (lane == 0) { meta[h] = act; meta[2 + h] = Tb; meta[4 + h] = L; }
        __syncwarp();
        // Short sweeps use the plain emission row, class c at position c, here and in the second half: both are
        // latency-bound then, a two-way bank conflict in a gather costs a couple of cycles, the colouring ~12000.
        if (T * NL < kColourMinWork) {
            for (int i = lane; i < 72; i += 32) cpos[h * 72 + i] = i;
        } else {
        // bank colouring of this utterance's emission row (see ctc_pair.cu): items = classes 0..C-1 and the zero item C
        for (int i = lane; i < 72; i += 32) gm[h * 72 + i] = 0;
        __syncwarp();
#pragma unroll
        for (int q = 0; q < NL; ++q) {
            const int i = lane * NL + q;
            atomicOr(&gm[h * 72 + (i < L ? lab[h * 256 + i] : C)], 1 << q);
        }
        __syncwarp();
        {
            int Gm[3], posk[3];
            bool valid[3];
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const int c = lane + 32 * k;
                valid[k] = c <= C;
                Gm[k] = valid[k] ? gm[h * 72 + c] : 0;
                posk[k] = 0;
            }
            int bankmask = 0, cnt = 0;                           // lane = bank
            for (int pass = 0; pass < 3; ++pass) {
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    const int pc = __popc(Gm[k]);
                    const bool sel = valid[k] && (pass == 0 ? pc >= 2 : pass == 1 ? pc == 1 : pc == 0);
                    unsigned m = __ballot_sync(FULL, sel);
                    while (m) {
                        const int src = __ffs(m) - 1;
                        m &= m - 1;
                        const int gq = __shfl_sync(FULL, Gm[k], src);
                        const bool ok = ((bankmask & gq) == 0) && cnt < 2;
                        unsigned e = __ballot_sync(FULL, ok && cnt == 0);
                        if (!e) e = __ballot_sync(FULL, ok);
                        if (!e) e = __ballot_sync(FULL, cnt < 2);
                        const int bsel = __ffs(e) - 1;
                        const int pcur = __shfl_sync(FULL, cnt, bsel) * 32 + bsel;
                        if (lane == bsel) { bankmask |= gq; ++cnt; }
                        if (lane == src) posk[k] = pcur;
                    }
                }
            }
#pragma unroll
            for (int k = 0; k < 3; ++k)
                if (valid[k]) cpos[h * 72 + lane + 32 * k] = posk[k];
        }
        }
        __syncwarp();
        for (int i = lane; i < 64; i += 32) ipos[h * 64 + i] = -1;
        __syncwarp();
        for (int c = lane; c <= C; c += 32) ipos[h * 64 + cpos[h * 72 + c]] = c;
        if (p.cpos_ws)                                           // the second half reads the layout instead of redoing it
            for (int i = lane; i < 72; i += 32) p.cpos_ws[(size_t)pair * 144 + h * 72 + i] = cpos[h * 72 + i];
    }
    __syncthreads();
    const bool act[2] = {meta[0] != 0, meta[1] != 0};
    if (!act[0] && !act[1]) return;
    const int Tb[2] = {meta[2], meta[3]}, L[2] = {meta[4], meta[5]};
    const int bidx[2] = {2 * pair, 2 * pair + 1};
    const int S[2] = {2 * L[0] + 1, 2 * L[1] + 1};

    // the zero item's emissions in every buffer of both sides (set once, never written again)
    for (int i = threadIdx.x; i < 2 * DNE * KF; i += blockDim.x) {
        float* Ei = dual_smem + (i / (DNE * KF)) * LY::SIDE + (i % (DNE * KF)) * RW;
        Ei[cpos[C]] = 0.f;
        Ei[CPMAX + cpos[72 + C]] = 0.f;
    }

    // geometry (as ctc_pair.cu): joint phase [0, Tmin), then the longer utterance alone; full chunks and single frames
    const int Tmin = min(Tb[0], Tb[1]), Tmax = max(Tb[0], Tb[1]);
    const int nJ8 = Tmin / CH, cJ = nJ8 + Tmin % CH;
    const int nS8 = (Tmax - Tmin) / CH, cS = cJ + nS8;
    const int nch = cS + (Tmax - Tmin) % CH;
    auto chunk_geom = [&](int j, int& t0) -> bool {
        if (j < nJ8) { t0 = j * CH; return true; }
        if (j < cJ) { t0 = nJ8 * CH + (j - nJ8); return false; }
        if (j < cS) { t0 = Tmin + (j - cJ) * CH; return true; }
        t0 = Tmin + nS8 * CH + (j - cS);
        return false;
    };
    auto seq2j = [&](int k) { return side ? nch - 1 - k : k; };  // k-th chunk of this side's sweep
    const int bsrc0 = act[0] ? bidx[0] : bidx[1], bsrc1 = act[1] ? bidx[1] : bidx[0];
    __syncthreads();
    if (warp == 0) { DT(0) }

    if (!lattice_warp) {
        // ================================================================================================ helpers
        int oc[2][2], wp[4];
        bool zs[4];                                      // my own position of (utterance, level) holds the zero item
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const unsigned free0 = __ballot_sync(FULL, ipos[h * 64 + lane] < 0), free1 = __ballot_sync(FULL, ipos[h * 64 + 32 + lane] < 0);
            const int fpos = free0 ? __ffs(free0) - 1 : 32 + __ffs(free1) - 1;
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                const int it = ipos[h * 64 + k * 32 + lane];     // class, C = the zero item, -1 = nobody
                oc[h][k] = (it >= 0 && it < C) ? it : 0;
                zs[2 * h + k] = it == C;
                wp[2 * h + k] = h * CPMAX + (it == C ? fpos : k * 32 + lane);
            }
        }
        const int ro00 = oc[0][0], ro01 = oc[0][1], ro10 = C + oc[1][0], ro11 = C + oc[1][1];
        const float* lpb0 = p.lp + (size_t)bsrc0 * p.stride_b;
        const float* lpb1 = p.lp + (size_t)bsrc1 * p.stride_b;
        const bool vec8 = ((C & 1) == 0) && ((p.stride_t & 1) == 0) && ((p.stride_b & 1) == 0) &&
                          ((reinterpret_cast<uintptr_t>(p.lp) & 7) == 0);
        const int hC = C >> 1;
        const bool pv0 = lane < 2 * hC, pv1 = lane + 32 < 2 * hC;
        const int ph0 = lane >= hC ? 1 : 0, ph1 = lane + 32 >= hC ? 1 : 0;
        const int pd0 = 2 * lane, pd1 = 2 * (lane + 32);
        const float* ps0 = (ph0 ? lpb1 : lpb0) + 2 * (lane - ph0 * hC);
        const float* ps1 = (ph1 ? lpb1 : lpb0) + 2 * (lane + 32 - ph1 * hC);
        const unsigned raw_s = smem_u32(raw);
        float* erow_pair = (want_grad && p.erow) ? p.erow + (size_t)pair * p.T * RW : nullptr;
        static_assert(CH % NH == 0, "a full chunk must split evenly over the helpers");
        const unsigned tstride_b = (unsigned)(p.stride_t * 4);   // row pitch in bytes (< 4 GB, checked by the launcher)
        auto issue_rows = [&](int k) {                   // my rows of this side's k-th chunk -> slot k % DNSLOT
            if (k >= 0 && k < nch) {
                int t0;
                const bool full = chunk_geom(seq2j(k), t0);
                const int so = (k % DNSLOT) * (KF * DRAWW);
                auto body = [&](auto nrc, int r0, int rs) {
                    constexpr int NR = decltype(nrc)::value;
#pragma unroll
                    for (int kk = 0; kk < NR; ++kk) {
                        const int r = r0 + kk * rs;
                        const unsigned long long o = (unsigned long long)(unsigned)(t0 + r) * tstride_b;
                        if (vec8) {
                            if (pv0) asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(raw_s + (so + r * DRAWW + pd0) * 4),
                                                  "l"(reinterpret_cast<const char*>(ps0) + o));
                            if (pv1) asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(raw_s + (so + r * DRAWW + pd1) * 4),
                                                  "l"(reinterpret_cast<const char*>(ps1) + o));
                        } else {
                            for (int i = lane; i < 2 * C; i += 32) {
                                const int h = i >= C ? 1 : 0;
                                raw[so + r * DRAWW + i] = (h ? lpb1 : lpb0)[(size_t)(t0 + r) * p.stride_t + i - h * C];
                            }
                        }
                        if constexpr (LOGITS) {
                            if (lane < 2)
                                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(raw_s + (so + r * DRAWW + 2 * C + lane) * 4),
                                             "l"(p.lse + (size_t)(t0 + r) * p.B + (lane ? bsrc1 : bsrc0)));
                        }
                    }
                };
                if (full) body(IC<CH / NH>{}, hid, NH);
                else if (hid == 0) body(IC<1>{}, 0, 1);
            }
            asm volatile("cp.async.commit_group;");
        };
        auto commit_rows = [&](int k) {                  // my rows of the k-th chunk: exp(lp), bank-coloured, into E[k % DNE]
            int t0;
            const bool full = chunk_geom(seq2j(k), t0);
            const bool kill0 = !act[0] || t0 >= Tb[0], kill1 = !act[1] || t0 >= Tb[1];
            const float* slotp = raw + (k % DNSLOT) * (KF * DRAWW);
            float* Eb = E + (k % DNE) * (KF * RW);
            auto body = [&](auto nrc, int r0, int rs) {
                constexpr int NR = decltype(nrc)::value;
                float x[NR][4], nl[NR][2];
#pragma unroll
                for (int kk = 0; kk < NR; ++kk) {        // all loads of my rows before the first exponential
                    const float* s = slotp + (r0 + kk * rs) * DRAWW;
                    x[kk][0] = s[ro00]; x[kk][1] = s[ro01]; x[kk][2] = s[ro10]; x[kk][3] = s[ro11];
                    if constexpr (LOGITS) { nl[kk][0] = s[2 * C]; nl[kk][1] = s[2 * C + 1]; }
                }
#pragma unroll
                for (int kk = 0; kk < NR; ++kk) {
                    float2 v0, v1;
                    if constexpr (LOGITS) {
                        const float n0 = nl[kk][0] * -LOG2E, n1 = nl[kk][1] * -LOG2E;
                        v0 = make_float2(fast_exp2(fmaf(x[kk][0], LOG2E, n0)), fast_exp2(fmaf(x[kk][1], LOG2E, n0)));
                        v1 = make_float2(fast_exp2(fmaf(x[kk][2], LOG2E, n1)), fast_exp2(fmaf(x[kk][3], LOG2E, n1)));
                    } else {
                        v0 = make_float2(fast_exp(x[kk][0]), fast_exp(x[kk][1]));
                        v1 = make_float2(fast_exp(x[kk][2]), fast_exp(x[kk][3]));
                    }
                    if (kill0) v0 = make_float2(0.f, 0.f);
                    if (kill1) v1 = make_float2(0.f, 0.f);
                    float* d = Eb + (r0 + kk * rs) * RW;
                    d[wp[0]] = v0.x;
                    d[wp[1]] = v0.y;
                    d[wp[2]] = v1.x;
                    d[wp[3]] = v1.y;
                    // the beta side (its sweep has the lighter steps and ends first) also leaves the row, as it lies in
                    // shared memory (position l + 32 k from lane l), to the segments of the second half: they neither
                    // read the log-probs nor exponentiate them again
                    if (side == 1 && erow_pair) {
                        float* eg = erow_pair + (size_t)(t0 + r0 + kk * rs) * RW + lane;
                        eg[0] = zs[0] ? 0.f : v0.x;
                        eg[32] = zs[1] ? 0.f : v0.y;
                        eg[64] = zs[2] ? 0.f : v1.x;
                        eg[96] = zs[3] ? 0.f : v1.y;
                    }
                }
            };
            if (full) body(IC<CH / NH>{}, hid, NH);
            else if (hid == 0) body(IC<1>{}, 0, 1);
        };
        unsigned ph_free = 0u;                           // parity to wait for next, per e_free barrier
        issue_rows(0);
        issue_rows(1);
        for (int k = 0; k < nch; ++k) {
            asm volatile("cp.async.wait_group 1;" ::: "memory");
            __syncwarp();
            if (k >= DNE) {                              // the lattice warp has left the buffer (chunk k - DNE of the sweep)
                mbar_wait_idle(&e_free[k % DNE], (ph_free >> (k % DNE)) & 1u);
                ph_free ^= 1u << (k % DNE);
            }
            commit_rows(k);
            __syncwarp();
            if (lane == 0) mbar_arrive(&e_full[k % DNE]);
            issue_rows(k + 2);
        }
        asm volatile("cp.async.wait_all;" ::: "memory");
        __syncthreads();                                 // (A) both sweeps are done
        return;
    }

    // ==================================================================================================== lattice warps
    const float* ep0[NL];
    const float* ep1[NL];
    float2 mk[NL];                                       // skip masks: alpha side looks down (s-2), beta side up (s+2)
#pragma unroll
    for (int q = 0; q < NL; ++q) {
        const int i = lane * NL + q;
        int eo[2];
        float m[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const bool in = i < L[h];
            const int c = in ? lab[h * 256 + i] : -1;
            eo[h] = h * CPMAX + cpos[h * 72 + (in ? c : C)];
            if (side == 0) m[h] = (in && i > 0 && lab[h * 256 + i - 1] != c) ? 1.f : 0.f;
            else m[h] = (in && i + 1 < L[h] && lab[h * 256 + i + 1] != c) ? 1.f : 0.f;
        }
        ep0[q] = E + eo[0];
        ep1[q] = E + eo[1];
        mk[q] = make_float2(m[0], m[1]);
    }
    const float* bl0 = E + cpos[blank];
    const float* bl1 = E + CPMAX + cpos[72 + blank];
    float* res = reinterpret_cast<float*>(meta + 8);     // [side][2][4]: alpha: flag, 1/mantissa(P), exponent, mantissa; beta: flag, b0+b1, exponent
    unsigned ph_full = 0u;
    int flag[2] = {0, 0};

    if (side == 0) {
        // ---------------------------------------------------------------------------------------- alpha, upwards
        float2 a[NSL];
#pragma unroll
        for (int j = 0; j < NSL; ++j) a[j] = make_float2(0.f, 0.f);
        if (lane == 0) a[0] = make_float2(act[0] ? 1.f : 0.f, act[1] ? 1.f : 0.f);   // virtual alpha_{-1}
        int ea[2] = {0, 0};
        float2 sca = make_float2(0.f, 0.f);
        float zi[2] = {0.f, 0.f}, vm[2] = {1.f, 1.f};
        int koff[2] = {0, 0};
        float* ck = ck_alpha + (size_t)pair * n_ckpt * CKW;
        auto fwd_chunk = [&](auto nfc, int j, int t0) {
            constexpr int NF = decltype(nfc)::value;
            const bool kill0 = !act[0] || t0 >= Tb[0], kill1 = !act[1] || t0 >= Tb[1];
            const bool ready = mbar_test(&e_full[j % DNE], (ph_full >> (j % DNE)) & 1u);
            {
                bool ok[2];
                renorm_alpha<NSL>(a, ea, sca, lane, ok, p.strict_squash != 0, p.decay_max);
                if (!ok[0] && !kill0 && !flag[0]) flag[0] = 2 | (t0 << 8);
                if (!ok[1] && !kill1 && !flag[1]) flag[1] = 2 | (t0 << 8);
            }
            if (want_grad) {                            // alpha entering chunk j, after its renormalisation
                float* c = ck + (size_t)j * CKW;
#pragma unroll
                for (int q = 0; q < NSL; ++q) reinterpret_cast<float2*>(c)[q * 32 + lane] = a[q];
                reinterpret_cast<int2*>(c + 64 * NSL)[lane] = make_int2(ea[0], ea[1]);
            }
            const int eb = (j % DNE) * (KF * RW);
            if (!ready) mbar_wait(&e_full[j % DNE], (ph_full >> (j % DNE)) & 1u);
            ph_full ^= 1u << (j % DNE);
#pragma unroll
            for (int r = 0; r < NF; ++r) alpha_step<NSL>(a, sca, eb + r * RW, bl0, bl1, ep0, ep1, mk);
            __syncwarp();
            if (lane == 0) mbar_arrive(&e_free[j % DNE]);
            const int tend = t0 + NF;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                if (act[h] && tend == Tb[h]) {
                    float v = 0.f;
#pragma unroll
                    for (int jj = 0; jj < NSL; ++jj) {
                        const int s = lane * NSL + jj;
                        if (s == S[h] - 1 || s == S[h] - 2) v += compc(a[jj], h);
                    }
                    int em = (v > 0.f) ? ea[h] : E_EMPTY;
#pragma unroll
                    for (int d = 16; d > 0; d >>= 1) em = max(em, __shfl_xor_sync(FULL, em, d));
                    v = warp_sum(v * pow2i(ea[h] - em));
                    if ((em == E_EMPTY || !(v > 1e-30f) || v > 1e37f) && !flag[h]) flag[h] = 3;
                    if (!flag[h]) {
                        const int kx = ((__float_as_int(v) >> 23) & 0xff) - 127;
                        v *= pow2i(-kx);
                        em += kx;
                        if (lane == 0) {
                            p.nll[bidx[h]] = (float)(-((double)logf(v) + (double)em * 0.69314718055994530942));
                            if (p.status) p.status[bidx[h]] = MDD_ST_OK;
                        }
                        zi[h] = pow2i(OFF) / v;
                        vm[h] = v;
                        koff[h] = em + OFF;
                    }
                }
            }
        };
        DT(1)
        for (int j = 0; j < nch; ++j) {
            int t0;
            if (chunk_geom(j, t0)) fwd_chunk(IC<CH>{}, j, t0);
            else fwd_chunk(IC<1>{}, j, t0);
            if (j == 0) { DT(2) }
            if (j == 1) { DT(3) }
            if (j == 2) { DT(8) }
        }
        DT(4)
        if (lane == 0) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                res[h * 4 + 0] = __int_as_float(flag[h]);
                res[h * 4 + 1] = zi[h];
                res[h * 4 + 2] = __int_as_float(koff[h]);
                res[h * 4 + 3] = vm[h];
            }
        }
    } else {
        // ---------------------------------------------------------------------------------------- beta, downwards
        float2 bt[NSL];
#pragma unroll
        for (int j = 0; j < NSL; ++j) bt[j] = make_float2(0.f, 0.f);
        int eb[2] = {0, 0};
        float2 scb = make_float2(0.f, 0.f);
        const float2 one2 = make_float2(1.f, 1.f);
        float* ck = ck_beta + (size_t)pair * n_ckpt * CKW;
        auto beta_chunk = [&](auto nfc, int k, int j, int t0) {
            constexpr int NF = decltype(nfc)::value;
            const bool kill0 = !act[0] || t0 >= Tb[0], kill1 = !act[1] || t0 >= Tb[1];
            const int tend = t0 + NF;
            const bool ready = mbar_test(&e_full[k % DNE], (ph_full >> (k % DNE)) & 1u);
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                if (act[h] && tend == Tb[h]) {          // beta_T enters: 1 at the last state, nothing anywhere else
                    eb[h] = 0;
#pragma unroll
                    for (int jj = 0; jj < NSL; ++jj) set_comp(bt[jj], h, (lane * NSL + jj == S[h] - 1) ? 1.f : 0.f);
                }
            }
            {
                bool ok[2];
                renorm_beta<NSL>(bt, eb, scb, lane, ok, p.decay_max);
                if (!ok[0] && !kill0 && !flag[0]) flag[0] = 2 | (t0 << 8);
                if (!ok[1] && !kill1 && !flag[1]) flag[1] = 2 | (t0 << 8);
            }
            if (want_grad) {                            // beta entering chunk j from above, after its renormalisation
                float* c = ck + (size_t)j * CKW;
#pragma unroll
                for (int q = 0; q < NSL; ++q) reinterpret_cast<float2*>(c)[q * 32 + lane] = bt[q];
                reinterpret_cast<int2*>(c + 64 * NSL)[lane] = make_int2(eb[0], eb[1]);
            }
            const int ebuf = (k % DNE) * (KF * RW);
            if (!ready) mbar_wait(&e_full[k % DNE], (ph_full >> (k % DNE)) & 1u);
            ph_full ^= 1u << (k % DNE);
#pragma unroll
            for (int r = NF - 1; r >= 0; --r) {
                float2 ul[NL];
                beta_step<NSL>(bt, ul, scb, one2, ebuf + r * RW, bl0, bl1, ep0, ep1, mk);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&e_free[k % DNE]);
        };
        DT(5)
        for (int k = 0; k < nch; ++k) {
            const int j = nch - 1 - k;
            int t0;
            if (chunk_geom(j, t0)) beta_chunk(IC<CH>{}, k, j, t0);
            else beta_chunk(IC<1>{}, k, j, t0);
            if (k == 0) { DT(6) }
        }
        DT(7)
        if (lane == 0) {                                // the beta sweep's own total likelihood: beta_0(0) + beta_0(1)
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                res[8 + h * 4 + 0] = __int_as_float(flag[h]);
                res[8 + h * 4 + 1] = compc(bt[0], h) + compc(bt[1], h);
                res[8 + h * 4 + 2] = __int_as_float(eb[h]);
            }
        }
    }
    __syncthreads();                                     // (A) both sweeps are done
    if (side == 0 && lane < 2 && act[lane]) {
        const int h = lane;
        int fl = __float_as_int(res[h * 4]);
        const float zi = res[h * 4 + 1], v = res[h * 4 + 3];
        const int koff = __float_as_int(res[h * 4 + 2]);
        const int flb = __float_as_int(res[8 + h * 4]);
        if (!fl && flb) fl = flb;
        if (!fl) {
            // mass dropped by either sweep is missing from exactly one of the two totals
            const float r = res[8 + h * 4 + 1] * pow2i(__float_as_int(res[8 + h * 4 + 2]) - (koff - OFF)) / v;
            if (!(fabsf(r - 1.f) <= p.pbeta_tol)) fl = 7;
        }
        if (want_grad) p.rec[bidx[h]] = make_float4(zi, __int_as_float(koff), __int_as_float(fl), 0.f);
        p.flags[bidx[h]] = fl;
    }
}

template <int NSL, int NH, bool LOGITS>
int launch_dual1(const CtcParams& p, float* cka, float* ckb, cudaStream_t st, int npairs) {
    using LY = DualLayout<NSL, NH>;
    const size_t smem = (size_t)LY::TOTAL * 4;
    if (cudaFuncSetAttribute(ctc_dual_kernel<NSL, NH, LOGITS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
        cudaGetLastError();
        return MDD_E_CUDA;
    }
    const int n_ckpt = p.T / KF + 2 * KF;
    ctc_dual_kernel<NSL, NH, LOGITS><<<npairs, 64 * (1 + NH), smem, st>>>(p, cka, ckb, n_ckpt);
    MDD_CHECK_LAUNCH();
    return MDD_OK;
}

template <int NSL, bool LOGITS>
int launch_dual_nsl(const CtcParams& p, float* cka, float* ckb, cudaStream_t st, int npairs) {
    // 4 helpers per side while every pair has an SM's worth of registers to itself, 2 when two CTAs share one, 1 (four
    // warps per CTA) when four do: the whole grid stays one wave up to four pairs per SM
    if (npairs <= sm_count()) return launch_dual1<NSL, 4, LOGITS>(p, cka, ckb, st, npairs);
    if (npairs <= 2 * sm_count()) return launch_dual1<NSL, 2, LOGITS>(p, cka, ckb, st, npairs);
    return launch_dual1<NSL, 1, LOGITS>(p, cka, ckb, st, npairs);
}

}  // namespace

#ifdef MDD_DUAL_TIMING
extern "C" void mdd_debug_dual_cycles(unsigned long long* out) {
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out, g_dual_cycles, sizeof(unsigned long long) * 16);
}
#endif

// alpha sweep + beta sweep of every pair, checkpoints of both into the workspace (cka / ckb: [npairs][T/8 + 16][CKW])
int launch_ctc_dual(const CtcParams& p, float* cka, float* ckb, cudaStream_t st, int npairs) {
    const int nsl = p.Lmax <= PGeo<2>::LCAP ? 2 : p.Lmax <= PGeo<4>::LCAP ? 4 : 8;
    const bool lg = p.lse != nullptr;
    switch (nsl) {
        case 2: return lg ? launch_dual_nsl<2, true>(p, cka, ckb, st, npairs) : launch_dual_nsl<2, false>(p, cka, ckb, st, npairs);
        case 4: return lg ? launch_dual_nsl<4, true>(p, cka, ckb, st, npairs) : launch_dual_nsl<4, false>(p, cka, ckb, st, npairs);
        default: return lg ? launch_dual_nsl<8, true>(p, cka, ckb, st, npairs) : launch_dual_nsl<8, false>(p, cka, ckb, st, npairs);
    }
}

}  // namespace mdd
