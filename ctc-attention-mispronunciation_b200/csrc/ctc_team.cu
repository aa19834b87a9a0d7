// ctc_team.cu -- CTC loss + gradient for FEW utterances: one CTA per pair, a lattice warp and NH helper warps.
//
// Semantics and arithmetic are those of ctc_pair.cu (packed fp32, block floating point, beta exponents tied to
// alpha's, fixed-point class cells; see there) -- this file changes who does what.  In ctc_pair.cu one warp owns a pair
// for the whole kernel and the SM hides its latencies with other pairs; with fewer than a few pairs per SM (the
// reference's training batches of 8-64 utterances, egs/attention_aug/steps/train_ctc.py:72 with conf/ctc_config*.yaml;
// a 4096-utterance batch sharded over 8 GPUs; the utterances the first pass flagged) there is nothing to hide them
// with, and the kernel runs at the speed of ONE warp's dependent chain: 8300 cycles per chunk of 8 frames, of which
// only ~1400 are the lattice recursion.  Here
//   * warp 0 (the lattice warp) runs the recursion and nothing else: renormalisation, checkpoints, alpha, beta, the
//     products into the class cells;
//   * warps 1..NH (helpers) do everything that is parallel over frames: row r of a chunk belongs to helper r % NH, which
//     fetches it (cp.async, two chunks ahead), exponentiates it into the bank-coloured emission buffer, and later turns
//     its class cells into the gradient row and stores it.
// Hand-offs are mbarriers in shared memory: emission buffers (three, indexed by chunk % 3) full / free, class-cell
// buffers (two) full / free.  No helper ever touches another helper's rows, so they need no synchronisation among
// themselves.
#include <algorithm>
#include <cstdlib>

#include "common.cuh"
#include "ctc_lattice.cuh"
#include "ctc_params.cuh"

namespace mdd {
namespace {

using namespace lattice;

constexpr int TRAWW = 132;                 // floats per staged pair row: 2*C log-probs (C <= 62) + 2 row-LSE + pad
constexpr int TNSLOT = 3;                  // raw staging slots (chunk % 3)
constexpr int TNE = 3;                     // emission buffers (chunk % 3)

#ifdef MDD_TEAM_TIMING
__device__ unsigned long long g_team_cycles[32];
#define TT_DECL long long tt_last = clock64(); long long tt_acc[16] = {0,0,0,0,0,0,0,0,0,0,0,0,0,0,0,0};
#define TT(i) { asm volatile("" ::: "memory"); const long long tt_now = clock64(); tt_acc[i] += tt_now - tt_last; tt_last = tt_now; }
#define TT_FLUSH(base) if (lane == 0 && blockIdx.x == 0) { for (int i = 0; i < 16; ++i) atomicAdd(&g_team_cycles[(base) + i], (unsigned long long)tt_acc[i]); }
#else
#define TT_DECL
#define TT(i)
#define TT_FLUSH(base)
#endif

__device__ __forceinline__ void mbar_inval(uint64_t* bar) {
    asm volatile("mbarrier.inval.shared::cta.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

template <int NSL, int NH>
struct TeamLayout {
    static constexpr int RPH = (KF + NH - 1) / NH;              // rows per helper and chunk (at most)
    static constexpr int CKW = PGeo<NSL>::CKW;
    static constexpr int E_OFF = 0;                             // [TNE][KF][RW] emissions
    static constexpr int OCC_OFF = E_OFF + TNE * KF * RW;       // [2][KF][RW] class cells
    static constexpr int CKST_OFF = OCC_OFF + 2 * KF * RW;      // [CKW]
    static constexpr int RAW_OFF = (CKST_OFF + CKW + 3) & ~3;   // [TNSLOT][KF][TRAWW]
    static constexpr int GST_OFF = RAW_OFF + TNSLOT * KF * TRAWW;   // [NH][RPH][TRAWW] gradient rows on their way out
    static constexpr int SCR_OFF = GST_OFF + NH * RPH * TRAWW;  // setup tables (ints)
    static constexpr int LAB = 0, HIST = 512, GM = 656, CPOS = 800, IPOS = 944, META = 1072, SCR_INTS = 1104;
    static constexpr int BAR_OFF = (SCR_OFF + SCR_INTS + 1) & ~1;   // mbarriers (8 B each)
    static constexpr int NBAR = 2 * TNE + TNE + 4;              // e_full, e_free (fwd), eb_full (bwd), occ_full[2], occ_free[2]
    static constexpr int TOTAL = BAR_OFF + 2 * NBAR;
};

template <int NSL, int NH, bool LOGITS, int CH, bool LIST>
__device__ __forceinline__ void ctc_team_body(const CtcParams& p, float* __restrict__ ckpt_ws, int n_ckpt, int pair,
                                              int slot, int n_utt, float* smem_f, bool first) {
    using G = PGeo<NSL>;
    using LY = TeamLayout<NSL, NH>;
    constexpr int NL = G::NL, CKW = G::CKW, RPH = LY::RPH;
    static_assert(CH >= 1 && CH <= KF && (CH & (CH - 1)) == 0, "chunk length: a power of two up to the staged rows");
    static_assert(NSL <= 8, "the chunk's alpha label states live in registers");
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool lattice_warp = warp == 0;
    const int hid = warp - 1;                                    // helper index (rows r % NH == hid)

    float* E = smem_f + LY::E_OFF;
    unsigned* occ = reinterpret_cast<unsigned*>(smem_f + LY::OCC_OFF);
    float* ckst = smem_f + LY::CKST_OFF;
    float* raw = smem_f + LY::RAW_OFF;
    float* gst = smem_f + LY::GST_OFF;
    int* scr = reinterpret_cast<int*>(smem_f + LY::SCR_OFF);
    int* lab = scr + LY::LAB;
    int* hist = scr + LY::HIST;
    int* gm = scr + LY::GM;
    int* cpos = scr + LY::CPOS;
    int* ipos = scr + LY::IPOS;
    int* meta = scr + LY::META;                                  // act[2], Tb[2], L[2], bidx[2], go, occ_hi[NH]
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem_f + LY::BAR_OFF);
    uint64_t* e_full = bars;                                     // [TNE] forward: helpers -> lattice
    uint64_t* e_free = bars + TNE;                               // [TNE] forward: lattice -> helpers
    uint64_t* eb_full = bars + 2 * TNE;                          // [TNE] backward: helpers -> lattice
    uint64_t* occ_full = bars + 3 * TNE;                         // [2]   lattice -> helpers
    uint64_t* occ_free = bars + 3 * TNE + 2;                     // [2]   helpers -> lattice
    constexpr int OCC_DELTA = LY::OCC_OFF - LY::E_OFF;           // class cell of an emission slot (buffer 0 of each)

    const int C = p.C, blank = p.blank, T = p.T;
    const bool want_grad = p.grad != nullptr;

    __syncthreads();                                             // the previous pair of this CTA is done with shared memory
    if (threadIdx.x == 0) {
        if (!first)
            for (int i = 0; i < LY::NBAR; ++i) mbar_inval(&bars[i]);
        for (int i = 0; i < TNE; ++i) { mbar_init(&e_full[i], NH); mbar_init(&e_free[i], 1); mbar_init(&eb_full[i], NH); }
        for (int i = 0; i < 2; ++i) { mbar_init(&occ_full[i], 1); mbar_init(&occ_free[i], NH); }
        mbar_fence_init();
    }
    // ------------------------------------------------------------------------------------------ setup (lattice warp)
    if (lattice_warp) {
        int bidx[2], Tb[2], L[2];
        bool act[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const bool have = 2 * pair + h < n_utt;
            const int b = LIST ? (have ? p.list[2 * pair + h] : p.list[2 * pair]) : 2 * pair + h;
            bidx[h] = b;
            act[h] = have;
            Tb[h] = 0; L[h] = 0;
            if (act[h]) {
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
                    if (lane == 0) p.flags[b] = 1;               // the fp64 kernel owns the error reporting
                    act[h] = false;
                } else {
                    Tb[h] = (int)Tb64;
                    L[h] = (int)L64;
                }
            }
        }
        if (lane == 0) {
            meta[0] = act[0]; meta[1] = act[1]; meta[2] = Tb[0]; meta[3] = Tb[1]; meta[4] = L[0]; meta[5] = L[1];
            meta[6] = bidx[0]; meta[7] = bidx[1]; meta[8] = 0;
        }
        __syncwarp();
        if (act[0] || act[1]) {
            // bank colouring of the emission row (see ctc_pair.cu): items = classes 0..C-1 and the zero item C
            for (int i = lane; i < 144; i += 32) { gm[i] = 0; hist[i] = 0; }
            __syncwarp();
#pragma unroll
            for (int h = 0; h < 2; ++h)
#pragma unroll
                for (int q = 0; q < NL; ++q) {
                    const int i = lane * NL + q;
                    const int it = h * 72 + (i < L[h] ? lab[h * 256 + i] : C);
                    atomicOr(&gm[it], 1 << q);
                    atomicAdd(&hist[it], 1);
                }
            __syncwarp();
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                int Gm[3], posk[3];
                bool valid[3];
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    const int c = lane + 32 * k;
                    valid[k] = c <= C;
                    Gm[k] = valid[k] ? gm[h * 72 + c] : 0;
                    posk[k] = 0;
                }
                int bankmask = 0, cnt = 0;                       // lane = bank
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
            for (int i = lane; i < 128; i += 32) ipos[i] = -1;
            __syncwarp();
            for (int c = lane; c <= C; c += 32) {
                ipos[cpos[c]] = c;
                ipos[64 + cpos[72 + c]] = c;
            }
        }
    }
    __syncthreads();
    const bool act[2] = {meta[0] != 0, meta[1] != 0};
    if (!act[0] && !act[1]) return;
    const int Tb[2] = {meta[2], meta[3]}, L[2] = {meta[4], meta[5]}, bidx[2] = {meta[6], meta[7]};
    const int S[2] = {2 * L[0] + 1, 2 * L[1] + 1};

    // cells start at zero; so do the zero item's emissions in every buffer (set once, never written again)
    for (int i = threadIdx.x; i < 2 * KF * RW; i += blockDim.x) occ[i] = 0u;
    for (int i = threadIdx.x; i < TNE * KF; i += blockDim.x) {
        E[i * RW + cpos[C]] = 0.f;
        E[i * RW + CPMAX + cpos[72 + C]] = 0.f;
    }

    // geometry: joint phase [0, Tmin), then the longer utterance alone [Tmin, Tmax); full chunks and single frames
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
    const int bsrc0 = act[0] ? bidx[0] : bidx[1], bsrc1 = act[1] ? bidx[1] : bidx[0];
    const size_t gstride = (size_t)p.B * C;
    float* ck = ckpt_ws + (size_t)slot * n_ckpt * CKW;
    __syncthreads();

    if (!lattice_warp) {
        // ================================================================================================ helpers
        int oc[2][2];
        bool ov[2][2];
        unsigned nbias[4];
        int wp[4], wg[4];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const unsigned free0 = __ballot_sync(FULL, ipos[h * 64 + lane] < 0), free1 = __ballot_sync(FULL, ipos[h * 64 + 32 + lane] < 0);
            const int fpos = free0 ? __ffs(free0) - 1 : 32 + __ffs(free1) - 1;
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                const int it = ipos[h * 64 + k * 32 + lane];     // class, C = the zero item, -1 = nobody
                ov[h][k] = it >= 0 && it < C;
                oc[h][k] = ov[h][k] ? it : 0;
                nbias[2 * h + k] = it >= 0 ? (unsigned)hist[h * 72 + it] * 0x4B000000u : 0u;
                wp[2 * h + k] = h * CPMAX + (it == C ? fpos : k * 32 + lane);
                wg[2 * h + k] = ov[h][k] ? h * C + oc[h][k] : 2 * C + 2;
            }
        }
        const bool isb00 = ov[0][0] && oc[0][0] == blank, isb01 = ov[0][1] && oc[0][1] == blank;
        const bool isb10 = ov[1][0] && oc[1][0] == blank, isb11 = ov[1][1] && oc[1][1] == blank;
        const int ro00 = oc[0][0], ro01 = oc[0][1], ro10 = C + oc[1][0], ro11 = C + oc[1][1];
        const float* lpb0 = p.lp + (size_t)bsrc0 * p.stride_b;
        const float* lpb1 = p.lp + (size_t)bsrc1 * p.stride_b;
        float* gr0 = want_grad ? p.grad + (size_t)bidx[0] * C : nullptr;
        float* gr1 = want_grad ? p.grad + (size_t)bidx[1] * C : nullptr;
        float g[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) g[h] = act[h] ? (p.grad_out ? p.grad_out[bidx[h]] : 1.f) * p.gscalar : 0.f;
        const bool vec8 = ((C & 1) == 0) && ((p.stride_t & 1) == 0) && ((p.stride_b & 1) == 0) &&
                          ((reinterpret_cast<uintptr_t>(p.lp) & 7) == 0);
        const int hC = C >> 1;
        const bool pv0 = lane < 2 * hC, pv1 = lane + 32 < 2 * hC;
        const int ph0 = lane >= hC ? 1 : 0, ph1 = lane + 32 >= hC ? 1 : 0;
        const int pd0 = 2 * lane, pd1 = 2 * (lane + 32);
        const float* ps0 = (ph0 ? lpb1 : lpb0) + 2 * (lane - ph0 * hC);
        const float* ps1 = (ph1 ? lpb1 : lpb0) + 2 * (lane + 32 - ph1 * hC);
        const unsigned raw_s = smem_u32(raw);

        // A full chunk gives every helper CH / NH rows (r = hid, hid + NH, ...), handled as straight-line code so that
        // the rows' latencies overlap; the single frames at the end of a phase belong to helper 0.
        static_assert(CH % NH == 0, "a full chunk must split evenly over the helpers");
        const unsigned tstride_b = (unsigned)(p.stride_t * 4);   // row pitch in bytes (< 4 GB, checked by the launcher)
        auto issue_rows = [&](int j) {                   // my rows of chunk j -> slot j % TNSLOT (one cp.async group)
            if (j >= 0 && j < nch) {
                int t0;
                const bool full = chunk_geom(j, t0);
                const int so = (j % TNSLOT) * (KF * TRAWW);
                auto body = [&](auto nrc, int r0, int rs) {
                    constexpr int NR = decltype(nrc)::value;
#pragma unroll
                    for (int k = 0; k < NR; ++k) {
                        const int r = r0 + k * rs;
                        const unsigned long long o = (unsigned long long)(unsigned)(t0 + r) * tstride_b;
                        if (vec8) {
                            if (pv0) asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(raw_s + (so + r * TRAWW + pd0) * 4),
                                                  "l"(reinterpret_cast<const char*>(ps0) + o));
                            if (pv1) asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(raw_s + (so + r * TRAWW + pd1) * 4),
                                                  "l"(reinterpret_cast<const char*>(ps1) + o));
                        } else {
                            for (int i = lane; i < 2 * C; i += 32) {
                                const int h = i >= C ? 1 : 0;
                                raw[so + r * TRAWW + i] = (h ? lpb1 : lpb0)[(size_t)(t0 + r) * p.stride_t + i - h * C];
                            }
                        }
                        if constexpr (LOGITS) {
                            if (lane < 2)
                                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(raw_s + (so + r * TRAWW + 2 * C + lane) * 4),
                                             "l"(p.lse + (size_t)(t0 + r) * p.B + (lane ? bsrc1 : bsrc0)));
                        }
                    }
                };
                if (full) body(IC<CH / NH>{}, hid, NH);
                else if (hid == 0) body(IC<1>{}, 0, 1);
            }
            asm volatile("cp.async.commit_group;");
        };
        auto commit_rows = [&](int j) {                  // my rows of chunk j: exp(lp), bank-coloured, into E[j % TNE]
            int t0;
            const bool full = chunk_geom(j, t0);
            const bool kill0 = !act[0] || t0 >= Tb[0], kill1 = !act[1] || t0 >= Tb[1];
            const float* slotp = raw + (j % TNSLOT) * (KF * TRAWW);
            float* Eb = E + (j % TNE) * (KF * RW);
            auto body = [&](auto nrc, int r0, int rs) {
                constexpr int NR = decltype(nrc)::value;
#pragma unroll
                for (int k = 0; k < NR; ++k) {
                    const int r = r0 + k * rs;
                    const float* s = slotp + r * TRAWW;
                    float2 v0, v1;
                    if constexpr (LOGITS) {
                        const float n0 = s[2 * C] * -LOG2E, n1 = s[2 * C + 1] * -LOG2E;
                        v0 = make_float2(fast_exp2(fmaf(s[ro00], LOG2E, n0)), fast_exp2(fmaf(s[ro01], LOG2E, n0)));
                        v1 = make_float2(fast_exp2(fmaf(s[ro10], LOG2E, n1)), fast_exp2(fmaf(s[ro11], LOG2E, n1)));
                    } else {
                        v0 = make_float2(fast_exp(s[ro00]), fast_exp(s[ro01]));
                        v1 = make_float2(fast_exp(s[ro10]), fast_exp(s[ro11]));
                    }
                    if (kill0) v0 = make_float2(0.f, 0.f);
                    if (kill1) v1 = make_float2(0.f, 0.f);
                    float* d = Eb + r * RW;
                    d[wp[0]] = v0.x;
                    d[wp[1]] = v0.y;
                    d[wp[2]] = v1.x;
                    d[wp[3]] = v1.y;
                }
            };
            if (full) body(IC<CH / NH>{}, hid, NH);
            else if (hid == 0) body(IC<1>{}, 0, 1);
        };
        auto arrive = [&](uint64_t* bar) {
            __syncwarp();
            if (lane == 0) mbar_arrive(bar);
        };

        // rows t >= T_max of the gradient are zero (torch semantics)
        if (want_grad)
            for (int t = Tmax + hid; t < T; t += NH)
                for (int i = lane; i < C; i += 32) {
                    if (act[0]) __stcs(gr0 + (size_t)t * gstride + i, 0.f);
                    if (act[1]) __stcs(gr1 + (size_t)t * gstride + i, 0.f);
                }

        // ------------------------------------------------------------------------------------------ forward
        unsigned ph_free = 0u;                           // parity to wait for next, per e_free barrier
        issue_rows(0);
        issue_rows(1);
        TT_DECL
        for (int j = 0; j < nch; ++j) {
            asm volatile("cp.async.wait_group 1;" ::: "memory");
            __syncwarp();
            TT(0)                                        // helper fwd: wait rows
            if (j >= TNE) {                              // the lattice warp has left the buffer (chunk j - TNE)
                mbar_wait_idle(&e_free[j % TNE], (ph_free >> (j % TNE)) & 1u);
                ph_free ^= 1u << (j % TNE);
            }
            TT(1)                                        // helper fwd: wait e_free
            commit_rows(j);
            arrive(&e_full[j % TNE]);
            issue_rows(j + 2);
            TT(2)                                        // helper fwd: commit + issue
        }
        asm volatile("cp.async.wait_all;" ::: "memory");
        __syncthreads();                                 // (A) forward done; the lattice warp decides whether to go on
        __syncthreads();                                 // (B) meta[8] = go
        if (!meta[8]) return;

        // ------------------------------------------------------------------------------------------ backward
        constexpr float FIX = 8388608.f * 0.999f;
        constexpr int fixi = (int)FIX;
        unsigned occ_hi = 0u;
        unsigned ph_occ = 0u;                            // parity per occ_full barrier
        float* mygst = gst + hid * (RPH * TRAWW);
        const bool gvec = (C & 1) == 0 && ((reinterpret_cast<uintptr_t>(p.grad) & 7) == 0);   // 8-byte stores of a row
        const unsigned gstride_b = (unsigned)(gstride * 4);
        auto gradient_rows = [&](int j, int nb) {        // my rows of chunk j (the nb-th chunk of the backward sweep)
            int t0;
            const bool full = chunk_geom(j, t0);
            const bool kill0 = !act[0] || t0 >= Tb[0], kill1 = !act[1] || t0 >= Tb[1];
            const float g0 = kill0 ? 0.f : g[0], g1 = kill1 ? 0.f : g[1];
            const float og0 = g0 * (1.f / FIX), og1 = g1 * (1.f / FIX);
            TT(3)                                        // helper bwd: commit / issue
            mbar_wait_idle(&occ_full[nb & 1], (ph_occ >> (nb & 1)) & 1u);
            ph_occ ^= 1u << (nb & 1);
            TT(4)                                        // helper bwd: wait occ_full
            const float* Eb = E + (j % TNE) * (KF * RW);
            unsigned* ob = occ + (nb & 1) * (KF * RW);
            auto cells = [&](auto nrc, int r0, int rs) {
                constexpr int NR = decltype(nrc)::value;
#pragma unroll
                for (int k = 0; k < NR; ++k) {
                    const int r = r0 + k * rs;
                    const float* d = Eb + r * RW;
                    unsigned* oq = ob + r * RW;
                    int o00 = (int)(oq[lane] - nbias[0]), o01 = (int)(oq[32 + lane] - nbias[1]);
                    int o10 = (int)(oq[CPMAX + lane] - nbias[2]), o11 = (int)(oq[CPMAX + 32 + lane] - nbias[3]);
                    oq[lane] = 0u; oq[32 + lane] = 0u; oq[CPMAX + lane] = 0u; oq[CPMAX + 32 + lane] = 0u;
                    const float2 pv0 = make_float2(d[lane], d[32 + lane]), pv1 = make_float2(d[CPMAX + lane], d[CPMAX + 32 + lane]);
                    const int ts0 = __reduce_add_sync(FULL, o00 + o01);      // blank: what the label classes leave of Z
                    const int ts1 = __reduce_add_sync(FULL, o10 + o11);
                    if (isb00) o00 = fixi - ts0;
                    if (isb01) o01 = fixi - ts0;
                    if (isb10) o10 = fixi - ts1;
                    if (isb11) o11 = fixi - ts1;
                    occ_hi = max(occ_hi, max((unsigned)(ts0 + 8192), (unsigned)(ts1 + 8192)));
                    float* w = mygst + k * TRAWW;
                    w[wg[0]] = fmaf(-(float)o00, og0, pv0.x * g0);
                    w[wg[1]] = fmaf(-(float)o01, og0, pv0.y * g0);
                    w[wg[2]] = fmaf(-(float)o10, og1, pv1.x * g1);
                    w[wg[3]] = fmaf(-(float)o11, og1, pv1.y * g1);
                }
            };
            auto stores = [&](auto nrc, int r0, int rs) {
                constexpr int NR = decltype(nrc)::value;
#pragma unroll
                for (int k = 0; k < NR; ++k) {
                    const int r = r0 + k * rs;
                    const unsigned long long go = (unsigned long long)(unsigned)(t0 + r) * gstride_b;
                    const float* w = mygst + k * TRAWW;
                    if (gvec) {
                        if (lane < C / 2) {
                            if (act[0]) __stcs(reinterpret_cast<float2*>(reinterpret_cast<char*>(gr0) + go) + lane, reinterpret_cast<const float2*>(w)[lane]);
                            if (act[1]) __stcs(reinterpret_cast<float2*>(reinterpret_cast<char*>(gr1) + go) + lane, reinterpret_cast<const float2*>(w + C)[lane]);
                        }
                    } else {
#pragma unroll
                        for (int i0 = 0; i0 < 64; i0 += 32) {
                            const int i = i0 + lane;
                            if (i < C) {
                                if (act[0]) __stcs(reinterpret_cast<float*>(reinterpret_cast<char*>(gr0) + go) + i, w[i]);
                                if (act[1]) __stcs(reinterpret_cast<float*>(reinterpret_cast<char*>(gr1) + go) + i, w[C + i]);
                            }
                        }
                    }
                }
            };
            if (full) cells(IC<CH / NH>{}, hid, NH);
            else if (hid == 0) cells(IC<1>{}, 0, 1);
            arrive(&occ_free[nb & 1]);                   // cells are zero again (and my staged rows are complete)
            if (full) stores(IC<CH / NH>{}, hid, NH);
            else if (hid == 0) stores(IC<1>{}, 0, 1);
            __syncwarp();                                // staged rows have been read before the next chunk overwrites them
            TT(5)                                        // helper bwd: gradient rows
        };
        // Emission buffers in use while the lattice warp works on chunk j: j (being read), j + 1 (its gradient rows are
        // next) and j - 1 (committed ahead, so that the lattice warp never waits for its emissions).
        arrive(&eb_full[(nch - 1) % TNE]);               // the last chunk is still in its buffer from the forward sweep
        issue_rows(nch - 2);
        issue_rows(nch - 3);
        for (int j = nch - 1; j >= 0; --j) {
            if (j >= 1) {
                asm volatile("cp.async.wait_group 1;" ::: "memory");   // rows(j - 1) have landed (j - 2 may be in flight)
                __syncwarp();
                commit_rows(j - 1);                      // into the buffer chunk j + 2 has left (its gradient is done)
                arrive(&eb_full[(j - 1) % TNE]);
            }
            issue_rows(j - 3);
            gradient_rows(j, nch - 1 - j);
        }
        asm volatile("cp.async.wait_all;" ::: "memory");
        if (lane == 0) meta[9 + hid] = (int)occ_hi;
        if (hid == 0) { TT_FLUSH(16) }
        __syncthreads();                                 // (C) helpers' range checks are visible
        return;
    }

    // ==================================================================================================== lattice warp
    const float* ep0[NL];
    const float* ep1[NL];
    float2 ma[NL], mb[NL];
#pragma unroll
    for (int q = 0; q < NL; ++q) {
        const int i = lane * NL + q;
        int eo[2];
        float am[2], bm[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const bool in = i < L[h];
            const int c = in ? lab[h * 256 + i] : -1;
            eo[h] = h * CPMAX + cpos[h * 72 + (in ? c : C)];
            am[h] = (in && i > 0 && lab[h * 256 + i - 1] != c) ? 1.f : 0.f;
            bm[h] = (in && i + 1 < L[h] && lab[h * 256 + i + 1] != c) ? 1.f : 0.f;
        }
        ep0[q] = E + eo[0];
        ep1[q] = E + eo[1];
        ma[q] = make_float2(am[0], am[1]);
        mb[q] = make_float2(bm[0], bm[1]);
    }
    const float* bl0 = E + cpos[blank];
    const float* bl1 = E + CPMAX + cpos[72 + blank];

    float2 a[NSL];
#pragma unroll
    for (int j = 0; j < NSL; ++j) a[j] = make_float2(0.f, 0.f);
    if (lane == 0) a[0] = make_float2(act[0] ? 1.f : 0.f, act[1] ? 1.f : 0.f);   // virtual alpha_{-1}
    int ea[2] = {0, 0};
    float2 sca = make_float2(0.f, 0.f);
    int flag[2] = {0, 0};
    float zi[2] = {0.f, 0.f};
    int koff[2] = {0, 0};
    unsigned ph_full = 0u;
    auto arrive = [&](uint64_t* bar) {
        __syncwarp();
        if (lane == 0) mbar_arrive(bar);
    };

    TT_DECL
    auto fwd_chunk = [&](auto nfc, int j, int t0) {
        constexpr int NF = decltype(nfc)::value;
        const bool kill0 = !act[0] || t0 >= Tb[0], kill1 = !act[1] || t0 >= Tb[1];
        TT(0)
        {
            bool ok[2];
            renorm_alpha<NSL>(a, ea, sca, lane, ok, p.strict_squash != 0, p.decay_max);
            if (!ok[0] && !kill0 && !flag[0]) flag[0] = 2 | (t0 << 8);
            if (!ok[1] && !kill1 && !flag[1]) flag[1] = 2 | (t0 << 8);
        }
        if (want_grad) {                                // alpha entering chunk j, after its renormalisation
            float* c = ck + (size_t)j * CKW;
#pragma unroll
            for (int q = 0; q < NSL; ++q) reinterpret_cast<float2*>(c)[q * 32 + lane] = a[q];
            reinterpret_cast<int2*>(c + 64 * NSL)[lane] = make_int2(ea[0], ea[1]);
        }
        const int eb = (j % TNE) * (KF * RW);
        TT(1)                                           // lattice fwd: renorm + ckpt store
        mbar_wait(&e_full[j % TNE], (ph_full >> (j % TNE)) & 1u);
        ph_full ^= 1u << (j % TNE);
        TT(2)                                           // lattice fwd: wait e_full
#pragma unroll
        for (int r = 0; r < NF; ++r) alpha_step<NSL>(a, sca, eb + r * RW, bl0, bl1, ep0, ep1, ma);
        arrive(&e_free[j % TNE]);
        TT(3)                                           // lattice fwd: alpha steps
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
                    koff[h] = em + OFF;
                }
            }
        }
    };
    for (int j = 0; j < nch; ++j) {
        int t0;
        if (chunk_geom(j, t0)) fwd_chunk(IC<CH>{}, j, t0);
        else fwd_chunk(IC<1>{}, j, t0);
    }
    __syncthreads();                                     // (A)
    const bool go = want_grad && !((flag[0] || !act[0]) && (flag[1] || !act[1]));
    if (lane == 0) meta[8] = go ? 1 : 0;
    __syncthreads();                                     // (B)
    if (!go) {
#pragma unroll
        for (int h = 0; h < 2; ++h)
            if (act[h] && lane == 0) p.flags[bidx[h]] = flag[h];
        return;
    }

    // ------------------------------------------------------------------------------------------ backward sweep
    float2 bt[NSL];
#pragma unroll
    for (int j = 0; j < NSL; ++j) bt[j] = make_float2(0.f, 0.f);
    int eb_[2] = {0, 0};
    constexpr float FIX = 8388608.f * 0.999f;
    const float2 zs2 = make_float2(zi[0] * FIX, zi[1] * FIX);
    const float2 magic2 = make_float2(8388608.f, 8388608.f);
    unsigned ph_ebf = 0u, ph_ofree = 0u;
    const unsigned ckst_s = smem_u32(ckst);
    auto ckpt_issue = [&](int j) {
        if (j >= 0) {
            const float* c = ck + (size_t)j * CKW;
#pragma unroll
            for (int q = 0; q < NSL / 2; ++q)
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(ckst_s + (q * 32 + lane) * 16), "l"(c + (q * 32 + lane) * 4));
            asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(ckst_s + (64 * NSL + lane * 2) * 4), "l"(c + 64 * NSL + lane * 2));
        }
        asm volatile("cp.async.commit_group;");
    };
    auto bwd_chunk = [&](auto nfc, int j, int t0, int nb) {
        constexpr int NF = decltype(nfc)::value;
        TT(4)
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncwarp();
        TT(5)                                           // lattice bwd: wait checkpoint
#pragma unroll
        for (int q = 0; q < NSL; ++q) a[q] = reinterpret_cast<const float2*>(ckst)[q * 32 + lane];
        {
            const int2 e = reinterpret_cast<const int2*>(ckst + 64 * NSL)[lane];
            ea[0] = e.x; ea[1] = e.y;
        }
        __syncwarp();
        ckpt_issue(j - 1);
        float2 scb, scb2;
        {
            float fb[2], fb2[2], sa[2], sb[2], sb2[2], ini[2];
            const int tend = t0 + NF;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int el = __shfl_up_sync(FULL, ea[h], 1), er = __shfl_down_sync(FULL, ea[h], 1);
                sa[h] = lane == 0 ? 0.f : pow2i(el - ea[h]);
                const int dx = max(ea[h] - er, -252);
                sb[h] = lane == 31 ? 0.f : pow2i(dx >> 1);
                sb2[h] = pow2i(dx - (dx >> 1));
                const int ebn = koff[h] - ea[h];
                const bool starts = act[h] && tend == Tb[h];
                const int df = min(max(eb_[h] - ebn, -252), 252);
                fb[h] = starts ? 0.f : pow2i(df >> 1);
                fb2[h] = pow2i(df - (df >> 1));
                ini[h] = starts ? pow2i(-ebn) : 0.f;
                if (starts) {
                    const bool owner = (S[h] - 1) / NSL == lane;
                    if (__any_sync(FULL, owner && (ebn > 120 || ebn < -120)) && !flag[h]) flag[h] = 6;
                }
                eb_[h] = ebn;
            }
            sca = make_float2(sa[0], sa[1]);
            scb = make_float2(sb[0], sb[1]);
            scb2 = make_float2(sb2[0], sb2[1]);
            const float2 f2 = make_float2(fb[0], fb[1]), f22 = make_float2(fb2[0], fb2[1]);
#pragma unroll
            for (int jj = 0; jj < NSL; ++jj) bt[jj] = __fmul2_rn(__fmul2_rn(bt[jj], f2), f22);
            if ((act[0] && tend == Tb[0]) || (act[1] && tend == Tb[1])) {
#pragma unroll
                for (int jj = 0; jj < NSL; ++jj) {
                    const int s = lane * NSL + jj;
                    if (s == S[0] - 1) bt[jj].x += ini[0];
                    if (s == S[1] - 1) bt[jj].y += ini[1];
                }
            }
        }
        const int eb = (j % TNE) * (KF * RW);
        TT(6)                                           // lattice bwd: take + bookkeeping
        mbar_wait(&eb_full[j % TNE], (ph_ebf >> (j % TNE)) & 1u);
        ph_ebf ^= 1u << (j % TNE);
        TT(7)                                           // lattice bwd: wait eb_full
        float2 A[NF][NL];
#pragma unroll
        for (int r = 0; r < NF; ++r) {
            alpha_step<NSL>(a, sca, eb + r * RW, bl0, bl1, ep0, ep1, ma);
#pragma unroll
            for (int q = 0; q < NL; ++q) A[r][q] = a[2 * q + 1];
        }
        TT(8)                                           // lattice bwd: alpha recompute
        if (nb >= 2) {                                  // the helpers have read (and zeroed) the cells of chunk nb - 2
            mbar_wait(&occ_free[nb & 1], (ph_ofree >> (nb & 1)) & 1u);
            ph_ofree ^= 1u << (nb & 1);
        }
        TT(9)                                           // lattice bwd: wait occ_free
        const int ob = OCC_DELTA + (nb & 1) * (KF * RW);
#pragma unroll
        for (int r = NF - 1; r >= 0; --r) {
            float2 ul[NL];
            beta_step<NSL>(bt, ul, scb, scb2, eb + r * RW, bl0, bl1, ep0, ep1, mb);
#pragma unroll
            for (int q = 0; q < NL; ++q) {
                const float2 fx = __ffma2_rn(__fmul2_rn(A[r][q], ul[q]), zs2, magic2);
                atomicAdd(reinterpret_cast<unsigned*>(const_cast<float*>(ep0[q])) + (ob + r * RW), __float_as_uint(fx.x));
                atomicAdd(reinterpret_cast<unsigned*>(const_cast<float*>(ep1[q])) + (ob + r * RW), __float_as_uint(fx.y));
            }
        }
        arrive(&occ_full[nb & 1]);
        TT(10)                                          // lattice bwd: beta + atomics
    };
    ckpt_issue(nch - 1);
    for (int j = nch - 1; j >= 0; --j) {
        int t0;
        if (chunk_geom(j, t0)) bwd_chunk(IC<CH>{}, j, t0, nch - 1 - j);
        else bwd_chunk(IC<1>{}, j, t0, nch - 1 - j);
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
    TT_FLUSH(0)
    __syncthreads();                                     // (C)
    {
        constexpr unsigned OMAX = (unsigned)(8388608.0 * 1.002) + 8192u;
        unsigned hi = 0u;
        for (int h = 0; h < NH; ++h) hi = max(hi, (unsigned)meta[9 + h]);
        int bad_occ = hi > OMAX ? 3 : 0;
#pragma unroll
        for (int jj = 0; jj < NSL; ++jj) {
            if (!(bt[jj].x < 3.0e38f)) bad_occ |= 1;
            if (!(bt[jj].y < 3.0e38f)) bad_occ |= 2;
        }
        bad_occ = __reduce_or_sync(FULL, bad_occ);
        if ((bad_occ & 1) && !flag[0]) flag[0] = 5;
        if ((bad_occ & 2) && !flag[1]) flag[1] = 5;
        const float r0 = (bt[0].x + bt[1].x) * pow2i(-ea[0]) * zi[0];
        const float r1 = (bt[0].y + bt[1].y) * pow2i(-ea[1]) * zi[1];
        const int disagree = (!(fabsf(r0 - 1.f) <= p.pbeta_tol) ? 1 : 0) | (!(fabsf(r1 - 1.f) <= p.pbeta_tol) ? 2 : 0);
        const int d0 = __shfl_sync(FULL, disagree, 0);
        if ((d0 & 1) && act[0] && !flag[0]) flag[0] = 7;
        if ((d0 & 2) && act[1] && !flag[1]) flag[1] = 7;
#pragma unroll
        for (int h = 0; h < 2; ++h)
            if (act[h] && lane == 0) p.flags[bidx[h]] = flag[h];
    }
}

template <int NSL, int NH, bool LOGITS, int CH, bool LIST>
__global__ void __launch_bounds__(32 * (1 + NH), 2) ctc_team_kernel(CtcParams p, float* __restrict__ ckpt_ws, int n_ckpt) {
    extern __shared__ __align__(16) float team_smem[];
    if constexpr (!LIST) {
        if (p.ticket && blockIdx.x == 0 && threadIdx.x == 0) *p.ticket = 0u;   // first kernel of the call (see ctc_exact.cu)
        ctc_team_body<NSL, NH, LOGITS, CH, false>(p, ckpt_ws, n_ckpt, blockIdx.x, blockIdx.x, p.B, team_smem, true);
    } else {
        // the flagged utterances: this CTA owns checkpoint block blockIdx.x and walks the pairs blockIdx.x, + grid, ...
        pdl_wait();                                              // (the next tier is released when this grid ends)
        const int n_utt = *p.list_count;
        if (n_utt > p.list_team_max) return;                     // many flagged utterances: the pair kernel's second tier takes them
        for (int pr = blockIdx.x; 2 * pr < n_utt; pr += gridDim.x)
            ctc_team_body<NSL, NH, LOGITS, CH, true>(p, ckpt_ws, n_ckpt, pr, blockIdx.x, n_utt, team_smem, pr == (int)blockIdx.x);
    }
}

template <int NSL, int NH, bool LOGITS, int CH, bool LIST>
int launch_team1(const CtcParams& p, float* ws, cudaStream_t st, int nblocks) {
    using LY = TeamLayout<NSL, NH>;
    const size_t smem = (size_t)LY::TOTAL * 4;
    if (cudaFuncSetAttribute(ctc_team_kernel<NSL, NH, LOGITS, CH, LIST>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)smem) != cudaSuccess) {
        cudaGetLastError();
        return MDD_E_CUDA;
    }
    const int n_ckpt = p.T / CH + 2 * CH;
    if constexpr (LIST) {
        if (launch_dependent(ctc_team_kernel<NSL, NH, LOGITS, CH, LIST>, dim3(nblocks), dim3(32 * (1 + NH)), smem, st, p, ws, n_ckpt) !=
            cudaSuccess) {
            cudaGetLastError();
            return MDD_E_CUDA;
        }
    } else {
        ctc_team_kernel<NSL, NH, LOGITS, CH, LIST><<<nblocks, 32 * (1 + NH), smem, st>>>(p, ws, n_ckpt);
    }
    MDD_CHECK_LAUNCH();
    return MDD_OK;
}

template <int NSL, bool LOGITS>
int launch_team_nsl(const CtcParams& p, float* ws, cudaStream_t st, int ch, bool list, int nblocks) {
    (void)ch;
    if (!list) return launch_team1<NSL, 4, LOGITS, KF, false>(p, ws, st, nblocks);   // 8 frames per chunk, 4 helpers x 2 rows
    return launch_team1<NSL, 2, LOGITS, 2, true>(p, ws, st, nblocks);                // 2 frames per chunk, 2 helpers x 1 row
}

}  // namespace

#ifdef MDD_TEAM_TIMING
extern "C" void mdd_debug_team_cycles(unsigned long long* out, int reset) {
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out, g_team_cycles, sizeof(unsigned long long) * 32);
    if (reset) {
        unsigned long long z[32] = {0};
        cudaMemcpyToSymbol(g_team_cycles, z, sizeof(z));
    }
}
#endif

bool ctc_team_supported(int C, int Lmax) { return C <= 62 && Lmax <= PGeo<8>::LCAP; }   // (see ctc_pair_supported)

// first pass over the whole batch (list == false) or second tier over the flagged list at 2 frames per chunk
int launch_ctc_team(const CtcParams& p, void* ws, cudaStream_t st, bool list, int nblocks) {
    float* w = reinterpret_cast<float*>(ws);
    const int nsl = p.Lmax <= PGeo<2>::LCAP ? 2 : p.Lmax <= PGeo<4>::LCAP ? 4 : 8;
    const bool lg = p.lse != nullptr;
    switch (nsl) {
        case 2: return lg ? launch_team_nsl<2, true>(p, w, st, 0, list, nblocks) : launch_team_nsl<2, false>(p, w, st, 0, list, nblocks);
        case 4: return lg ? launch_team_nsl<4, true>(p, w, st, 0, list, nblocks) : launch_team_nsl<4, false>(p, w, st, 0, list, nblocks);
        default: return lg ? launch_team_nsl<8, true>(p, w, st, 0, list, nblocks) : launch_team_nsl<8, false>(p, w, st, 0, list, nblocks);
    }
}

}  // namespace mdd
