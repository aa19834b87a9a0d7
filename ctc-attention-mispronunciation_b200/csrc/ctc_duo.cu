// ctc_duo.cu -- CTC loss + gradient: the pair kernel's arithmetic (ctc_pair.cu) run by TWO warps per pair of utterances.
//
// Semantics: torch aten ctc_loss / _ctc_loss_backward as reached from nn.CTCLoss at
// egs/attention_aug/steps/train_ctc.py:186,72,96 and egs/cnn-rnn-ctc/steps/train_ctc.py:148,51,67.
//
// ctc_pair_kernel is bound by the latency of one warp's dependent instruction chain (7 warps per SM, 255 registers
// each: DESIGN.md 5.1).  Here the work of a pair is split into two instruction streams, one CTA = one pair = two warps:
//   * the LATTICE warp keeps alpha / beta and the chunk's alpha label states in registers (per-utterance setup,
//     renormalisation, checkpoints, alpha and beta steps, occupancy atomics, P, flags);
//   * the ROW warp moves data (log-prob rows HBM -> registers -> exp -> staged emission rows in both sweeps; gradient
//     rows from the occupancy cells back to HBM).
// Emission rows and occupancy cells are double buffered (chunk j lives in buffer j & 1 in both sweeps); the two warps
// hand buffers over through mbarriers in shared memory (one arrival per phase by lane 0 after __syncwarp; every lane
// of the waiting warp polls, which gives each lane its own acquire; phase numbers follow from the chunk index).
//
// STATUS (round 1): correct (same parity tests as ctc_pair.cu, MDD_CTC_DUO=1) and 27 % faster for a pair that has an SM
// to itself (0.375 vs 0.513 ms at T = 1000), but NOT the default: a scheduler's register file holds three warps of
// 129..168 registers, so six pairs fit on an SM and config 3 (13.84 pairs per SM) needs three rounds of 0.53 ms where
// ctc_pair_kernel needs two of 0.66 ms.  DESIGN.md 6 has the numbers and what would have to change.
//
// Shapes: NSL = 8 (L in 64..127), C <= 63.  Everything else, and every flagged utterance, goes the usual way
// (ctc_pair.cu / ctc_log.cu).
#include <cstdio>
#include <cstdlib>

#include "common.cuh"
#include "ctc_lattice.cuh"
#include "ctc_params.cuh"

namespace mdd {
namespace {

using namespace lattice;

constexpr int NSL = 8, NL = 4;
constexpr int CKW = PGeo<NSL>::CKW;
#ifndef MDD_DUO_SLEEP
#define MDD_DUO_SLEEP 100
#endif
constexpr int BUF = KF * RW;                 // one buffer of emission rows / occupancy cells (floats)

// per-pair shared memory (float offsets)
constexpr int O_OCC = 0;                     // [2][KF][RW] u32 (setup scratch: occ + rows)
constexpr int O_ROWS = O_OCC + 2 * BUF;      // [2][KF][RW] f32
constexpr int O_CKST = O_ROWS + 2 * BUF;     // [CKW] checkpoint landing by cp.async
constexpr int O_LTAB = O_CKST + CKW;         // [32][12] i32 per-lane table for the row warp
constexpr int O_MISC = O_LTAB + 32 * 12;     // [16] i32
constexpr int O_BARS = O_MISC + 16;          // [12] u64 mbarriers
constexpr int PAIR_FLOATS = O_BARS + 24 + 8; // 16-byte multiple
static_assert(PAIR_FLOATS % 4 == 0, "pair block must keep 16-byte alignment");
constexpr int OCCD = O_OCC - O_ROWS;         // occupancy cell of an emission slot (same buffer): rows -> occ

enum { M_INACTIVE = 0, M_T0 = 1, M_T1 = 2, M_ACT0 = 3, M_ACT1 = 4, M_GO_BWD = 5, M_BAD_ROWS = 6 };
enum { B_SETUP = 0, B_LDONE = 1, B_FW_FULL = 2, B_FW_FREE = 4, B_BW_FULL = 6, B_OCC_FULL = 8, B_OCC_FREE = 10 };

#ifdef MDD_PAIR_TIMING
__device__ unsigned long long g_duo_cycles[16];
#define DT_DECL long long dt_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0}; long long dt_t0 = clock64();
#define DT_WAIT(i, stmt) { const long long w0 = clock64(); stmt; dt_acc[i] += clock64() - w0; }
#define DT_MARK(i) { const long long n = clock64(); dt_acc[i] += n - dt_t0; dt_t0 = n; }
#define DT_FLUSH(off) if (lane == 0) { for (int i = 0; i < 8; ++i) atomicAdd(&g_duo_cycles[off + i], (unsigned long long)dt_acc[i]); }
#else
#define DT_DECL
#define DT_WAIT(i, stmt) stmt;
#define DT_MARK(i)
#define DT_FLUSH(off)
#endif

__device__ __forceinline__ void mbar_init(unsigned bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(unsigned bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, int phase) {   // returns when phase number `phase` has completed
    unsigned done;
    for (;;) {
        asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                     : "=r"(done) : "r"(bar), "r"(phase & 1) : "memory");
        if (done) break;
        __nanosleep(MDD_DUO_SLEEP);                      // a polling warp must not eat the issue slots and LSU cycles of its partner
    }
}
// one arrival per phase: the warp's shared-memory writes first, then lane 0 signals
__device__ __forceinline__ void warp_signal(unsigned bar, int lane) {
    __syncwarp();
    if (lane == 0) mbar_arrive(bar);
}

struct Geo {                                 // chunk geometry of a pair (see ctc_pair.cu)
    int Tmin, Tmax, nJ8, cJ, cS, nS8, nch;
    __device__ __forceinline__ void init(int T0, int T1) {
        Tmin = min(T0, T1); Tmax = max(T0, T1);
        nJ8 = Tmin / KF; cJ = nJ8 + Tmin % KF;
        nS8 = (Tmax - Tmin) / KF; cS = cJ + nS8;
        nch = cS + (Tmax - Tmin) % KF;
    }
    __device__ __forceinline__ bool chunk(int j, int& t0) const {   // true: full chunk
        if (j < nJ8) { t0 = j * KF; return true; }
        if (j < cJ) { t0 = nJ8 * KF + (j - nJ8); return false; }
        if (j < cS) { t0 = Tmin + (j - cJ) * KF; return true; }
        t0 = Tmin + nS8 * KF + (j - cS);
        return false;
    }
};

// ------------------------------------------------------------------------------------------------------------------
// ROW warp: log-prob rows HBM -> registers -> exp -> staged emission rows (both sweeps); gradient rows (backward).
// Lane l owns the classes placed in bank l of each level of the coloured row (table written by the lattice warp).
// ------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void row_role(const CtcParams& p, float* base, unsigned bars, int lane, int pair) {
    unsigned* occ = reinterpret_cast<unsigned*>(base + O_OCC);
    float* rows = base + O_ROWS;
    const int* ltab = reinterpret_cast<const int*>(base + O_LTAB) + lane * 12;
    volatile int* misc = reinterpret_cast<volatile int*>(base + O_MISC);

    mbar_wait(bars + 8 * B_SETUP, 0);
    if (misc[M_INACTIVE]) return;
    const int C = p.C, blank = p.blank, T = p.T;
    const bool want_grad = p.grad != nullptr;
    int oc[2][2];
    bool ov[2][2];
    unsigned nbias[4];
#pragma unroll
    for (int h = 0; h < 2; ++h)
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            oc[h][k] = ltab[2 * h + k];
            ov[h][k] = (ltab[4] >> (2 * h + k)) & 1;
            nbias[2 * h + k] = (unsigned)ltab[5 + 2 * h + k];
        }
    const int Tb[2] = {misc[M_T0], misc[M_T1]};
    const bool act[2] = {misc[M_ACT0] != 0, misc[M_ACT1] != 0};
    const int bidx[2] = {2 * pair, 2 * pair + 1};
    Geo geo;
    geo.init(Tb[0], Tb[1]);
    const int nch = geo.nch, Tmax = geo.Tmax;
    DT_DECL

    // rows are always read from a live utterance's memory; an idle half re-reads its partner's rows
    const int bsrc0 = act[0] ? bidx[0] : bidx[1], bsrc1 = act[1] ? bidx[1] : bidx[0];
    const float* lpb0 = p.lp + (size_t)bsrc0 * p.stride_b;
    const float* lpb1 = p.lp + (size_t)bsrc1 * p.stride_b;
    const size_t gstride = (size_t)p.B * C;
    float* gr0 = want_grad ? p.grad + (size_t)bidx[0] * C : nullptr;
    float* gr1 = want_grad ? p.grad + (size_t)bidx[1] * C : nullptr;
    float g[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) g[h] = act[h] ? (p.grad_out ? p.grad_out[bidx[h]] : 1.f) * p.gscalar : 0.f;

    const float* lp00 = lpb0 + oc[0][0];                         // my level-0 / level-1 class of each utterance
    const float* lp01 = lpb0 + oc[0][1];
    const float* lp10 = lpb1 + oc[1][0];
    const float* lp11 = lpb1 + oc[1][1];
    float2 pre0[KF], pre1[KF];                          // the next chunk's rows, in flight
    auto load_rows = [&](auto nfc, int t0) {
        constexpr int NF = decltype(nfc)::value;
        size_t o = (size_t)t0 * p.stride_t;
#pragma unroll
        for (int r = 0; r < NF; ++r) {
            asm volatile("ld.global.cs.f32 %0, [%1];" : "=f"(pre0[r].x) : "l"(lp00 + o));
            asm volatile("ld.global.cs.f32 %0, [%1];" : "=f"(pre0[r].y) : "l"(lp01 + o));
            asm volatile("ld.global.cs.f32 %0, [%1];" : "=f"(pre1[r].x) : "l"(lp10 + o));
            asm volatile("ld.global.cs.f32 %0, [%1];" : "=f"(pre1[r].y) : "l"(lp11 + o));
            o += p.stride_t;
        }
    };
    auto load_chunk = [&](int j) {
        int t0;
        DT_WAIT(5, if (geo.chunk(j, t0)) load_rows(IC<KF>{}, t0); else load_rows(IC<1>{}, t0))
    };
    auto prefetch_chunk = [&](int j) {                  // a full chunk's rows -> L2
        int tq;
        if (j >= 0 && j < nch && geo.chunk(j, tq)) {
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                const int idx = lane + 32 * k;          // 48 tasks: row, utterance, 128-byte part of its 4*C bytes
                const int r = idx / 6, u = (idx % 6) / 3, part = idx % 3;
                if (idx < 6 * KF) {
                    const float* q = (u ? lpb1 : lpb0) + (size_t)(tq + r) * p.stride_t + min(part * 32, C - 1);
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(q));
                }
            }
        }
    };
    auto commit_rows = [&](auto nfc, float* rb, int t0) {   // p = exp(lp) into shared memory; an idle half emits zeros
        constexpr int NF = decltype(nfc)::value;
        const bool kill0 = !act[0] || t0 >= Tb[0], kill1 = !act[1] || t0 >= Tb[1];
#pragma unroll
        for (int r = 0; r < NF; ++r) {
            float2 v0 = make_float2(fast_exp(pre0[r].x), fast_exp(pre0[r].y));
            float2 v1 = make_float2(fast_exp(pre1[r].x), fast_exp(pre1[r].y));
            if (kill0) v0 = make_float2(0.f, 0.f);
            if (kill1) v1 = make_float2(0.f, 0.f);
            float* d = rb + r * RW;
            if (ov[0][0]) d[lane] = v0.x;
            if (ov[0][1]) d[32 + lane] = v0.y;
            if (ov[1][0]) d[CPMAX + lane] = v1.x;
            if (ov[1][1]) d[CPMAX + 32 + lane] = v1.y;
        }
    };
    auto commit_chunk = [&](int j) {
        int t0;
        float* rb = rows + (j & 1) * BUF;
        if (geo.chunk(j, t0)) commit_rows(IC<KF>{}, rb, t0);
        else commit_rows(IC<1>{}, rb, t0);
    };

    // ------------------------------------------------------------------------------------------ forward sweep
    load_chunk(0);
    prefetch_chunk(1);
    prefetch_chunk(2);
    if (want_grad) {                                    // rows t >= T_max of the gradient are zero (torch semantics)
        for (int t = Tmax; t < T; ++t) {
            float* z0 = gr0 + (size_t)t * gstride;
            float* z1 = gr1 + (size_t)t * gstride;
            if (act[0] && ov[0][0]) __stcs(z0 + oc[0][0], 0.f);
            if (act[0] && ov[0][1]) __stcs(z0 + oc[0][1], 0.f);
            if (act[1] && ov[1][0]) __stcs(z1 + oc[1][0], 0.f);
            if (act[1] && ov[1][1]) __stcs(z1 + oc[1][1], 0.f);
        }
    }
    DT_MARK(7)
    for (int j = 0; j < nch; ++j) {
        if (j >= 2) DT_WAIT(0, mbar_wait(bars + 8 * (B_FW_FREE + (j & 1)), (j >> 1) - 1))   // the lattice warp is through chunk j - 2
        DT_WAIT(4, commit_chunk(j))
        warp_signal(bars + 8 * (B_FW_FULL + (j & 1)), lane);
        if (j + 1 < nch) load_chunk(j + 1);
        prefetch_chunk(j + 3);
    }
    if (!want_grad) return;
    DT_MARK(2)                                          // row warp: forward sweep
    // the lattice warp has finished the forward sweep (and decided whether a backward sweep follows)
    // (arrival number u on the barrier of buffer b is that of its u-th chunk: chunk j is arrival j >> 1)
    if (nch >= 2) mbar_wait(bars + 8 * (B_FW_FREE + ((nch - 2) & 1)), (nch - 2) >> 1);
    mbar_wait(bars + 8 * (B_FW_FREE + ((nch - 1) & 1)), (nch - 1) >> 1);
    if (!misc[M_GO_BWD]) return;

    // ------------------------------------------------------------------------------------------ backward sweep
    const bool isb00 = ov[0][0] && oc[0][0] == blank, isb01 = ov[0][1] && oc[0][1] == blank;
    const bool isb10 = ov[1][0] && oc[1][0] == blank, isb11 = ov[1][1] && oc[1][1] == blank;
    const bool s00 = act[0] && ov[0][0], s01 = act[0] && ov[0][1], s10 = act[1] && ov[1][0], s11 = act[1] && ov[1][1];
    constexpr float FIX = 8388608.f * 0.999f;           // fixed-point scale of an occupancy fraction (ctc_pair.cu)
    constexpr int OMAX = (int)(8388608.0 * 1.002);
    int bad_occ = 0;
    auto grad_rows = [&](auto nfc, int b, int t0) {      // (softmax prob - occupancy / Z) * g; cells zeroed for reuse
        constexpr int NF = decltype(nfc)::value;
        const bool kill0 = !act[0] || t0 >= Tb[0], kill1 = !act[1] || t0 >= Tb[1];
        const float g0 = kill0 ? 0.f : g[0], g1 = kill1 ? 0.f : g[1];
        const float og0 = g0 * (1.f / FIX), og1 = g1 * (1.f / FIX);
        float* w00 = gr0 + (size_t)t0 * gstride + oc[0][0];
        float* w01 = gr0 + (size_t)t0 * gstride + oc[0][1];
        float* w10 = gr1 + (size_t)t0 * gstride + oc[1][0];
        float* w11 = gr1 + (size_t)t0 * gstride + oc[1][1];
#pragma unroll
        for (int r = 0; r < NF; ++r) {
            const float* d = rows + b * BUF + r * RW;
            unsigned* oq = occ + b * BUF + r * RW;
            const float2 pv0 = make_float2(d[lane], d[32 + lane]), pv1 = make_float2(d[CPMAX + lane], d[CPMAX + 32 + lane]);
            int o00 = (int)(oq[lane] - nbias[0]), o01 = (int)(oq[32 + lane] - nbias[1]);
            int o10 = (int)(oq[CPMAX + lane] - nbias[2]), o11 = (int)(oq[CPMAX + 32 + lane] - nbias[3]);
            oq[lane] = 0u; oq[32 + lane] = 0u; oq[CPMAX + lane] = 0u; oq[CPMAX + 32 + lane] = 0u;
            const int t0s = __reduce_add_sync(FULL, (ov[0][0] ? o00 : 0) + (ov[0][1] ? o01 : 0));
            const int t1s = __reduce_add_sync(FULL, (ov[1][0] ? o10 : 0) + (ov[1][1] ? o11 : 0));
            const int fixi = (int)FIX;
            if (isb00) o00 = fixi - t0s;                 // blank: what the label classes leave of Z
            if (isb01) o01 = fixi - t0s;
            if (isb10) o10 = fixi - t1s;
            if (isb11) o11 = fixi - t1s;
            if (!kill0 && ((unsigned)(t0s + 8192) > (unsigned)(OMAX + 8192))) bad_occ |= 1;
            if (!kill1 && ((unsigned)(t1s + 8192) > (unsigned)(OMAX + 8192))) bad_occ |= 2;
            if (s00) __stcs(w00, fmaf(-(float)o00, og0, pv0.x * g0));
            if (s01) __stcs(w01, fmaf(-(float)o01, og0, pv0.y * g0));
            if (s10) __stcs(w10, fmaf(-(float)o10, og1, pv1.x * g1));
            if (s11) __stcs(w11, fmaf(-(float)o11, og1, pv1.y * g1));
            w00 += gstride; w01 += gstride; w10 += gstride; w11 += gstride;
#ifdef MDD_DUO_GRAD_FENCE
            if ((r % MDD_DUO_GRAD_FENCE) == MDD_DUO_GRAD_FENCE - 1) asm volatile("" ::: "memory");   // bound the live range of the hoisted loads
#endif
        }
    };
    auto grad_chunk = [&](int j) {
        const int b = j & 1;
        int t0;
        DT_WAIT(1, mbar_wait(bars + 8 * (B_OCC_FULL + b), (nch - 1 - j) >> 1))   // the lattice warp has accumulated chunk j
        DT_WAIT(6, if (geo.chunk(j, t0)) grad_rows(IC<KF>{}, b, t0); else grad_rows(IC<1>{}, b, t0))
        warp_signal(bars + 8 * (B_OCC_FREE + b), lane);
    };
    if (nch >= 2) load_chunk(nch - 2);
    prefetch_chunk(nch - 3);
    for (int j = nch - 1; j >= 0; --j) {
        if (j < nch - 1) DT_WAIT(4, commit_chunk(j))    // the last chunk is still staged from the forward sweep
        warp_signal(bars + 8 * (B_BW_FULL + (j & 1)), lane);
        if (j < nch - 1 && j > 0) load_chunk(j - 1);    // in flight while the gradient rows below are written
        prefetch_chunk(j - 3);
        if (j < nch - 1) grad_chunk(j + 1);
    }
    grad_chunk(0);
    DT_MARK(3)                                          // row warp: backward sweep (from the end of the forward one)
    DT_FLUSH(8)
    bad_occ = __reduce_or_sync(FULL, bad_occ);
    if (lane == 0) misc[M_BAD_ROWS] = bad_occ;
    warp_signal(bars + 8 * B_LDONE, lane);
}

// ------------------------------------------------------------------------------------------------------------------
// LATTICE warp: per-utterance setup, alpha sweep with checkpoints, P, beta sweep with occupancy accumulation, flags.
// ------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void lattice_role(const CtcParams& p, float* ckpt_ws, int n_ckpt, float* base, unsigned bars,
                                             int lane, int pair) {
    unsigned* occ = reinterpret_cast<unsigned*>(base + O_OCC);
    float* rows = base + O_ROWS;
    float* ckst = base + O_CKST;
    int* ltab = reinterpret_cast<int*>(base + O_LTAB) + lane * 12;
    volatile int* misc = reinterpret_cast<volatile int*>(base + O_MISC);
    const int C = p.C, blank = p.blank, T = p.T;
    const bool want_grad = p.grad != nullptr;
    if (lane < 16) misc[lane] = 0;
    __syncwarp();

    // ------------------------------------------------------------------------------------------ per-utterance setup
    int bidx[2], Tb[2], L[2], S[2];
    bool act[2];
    int* scratch = reinterpret_cast<int*>(occ);                  // lab[2][256] | hist[2][64] | ...
    int* lab = scratch;
    int* hist = scratch + 512;
    static_assert(2 * KF * RW >= 1280 + 288 + 128, "setup scratch must fit in the occupancy + emission rows");
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int b = 2 * pair + h;
        bidx[h] = b;
        act[h] = b < p.B;
        Tb[h] = 0; L[h] = 0;
        if (act[h]) {
            const long long Tb64 = p.in_len[b], L64 = p.tgt_len[b];
            bool bad = Tb64 <= 0 || Tb64 > T || L64 < 0 || L64 > p.Lmax || L64 > PGeo<NSL>::LCAP;
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
                if (lane == 0) p.flags[b] = 1;                   // the log-space kernel owns the error reporting
                act[h] = false;
            } else {
                Tb[h] = (int)Tb64;
                L[h] = (int)L64;
            }
        }
        S[h] = 2 * L[h] + 1;
    }
    if (!act[0] && !act[1]) {
        if (lane == 0) misc[M_INACTIVE] = 1;
        warp_signal(bars + 8 * B_SETUP, lane);
        return;
    }
    __syncwarp();

    // Emission layout.  The lattice gathers e[t][label] with one LDS per label state; with C > 32 classes a plain row
    // makes two of the requested classes share a bank in almost every gather.  Per utterance, classes (and the zero
    // that unused label slots read, item C) are therefore placed at position bank + 32*level such that two classes
    // requested by the same gather instruction (same label slot q of different lanes) never share a bank: items are
    // placed greedily, those present in most gather groups first, each into a bank whose residents are requested by
    // none of its groups.
    int* gm = scratch + 1280;                                    // [2][72] gather-group mask per item
    int* cpos = scratch + 1280 + 144;                            // [2][72] position per item
    for (int i = lane; i < 144; i += 32) gm[i] = 0;
    __syncwarp();
#pragma unroll
    for (int h = 0; h < 2; ++h) {
#pragma unroll
        for (int q = 0; q < NL; ++q) {
            const int i = lane * NL + q;
            atomicOr(&gm[h * 72 + (i < L[h] ? lab[h * 256 + i] : C)], 1 << q);
        }
    }
    __syncwarp();
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        int G[3], posk[3];
        bool valid[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const int c = lane + 32 * k;
            valid[k] = c <= C;
            G[k] = valid[k] ? gm[h * 72 + c] : 0;
            posk[k] = 0;
        }
        int bankmask = 0, cnt = 0;                               // lane = bank
        for (int pass = 0; pass < 3; ++pass) {
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const int pc = __popc(G[k]);
                const bool sel = valid[k] && (pass == 0 ? pc >= 2 : pass == 1 ? pc == 1 : pc == 0);
                unsigned m = __ballot_sync(FULL, sel);
                while (m) {
                    const int src = __ffs(m) - 1;
                    m &= m - 1;
                    const int gq = __shfl_sync(FULL, G[k], src);
                    const bool ok = ((bankmask & gq) == 0) && cnt < 2;
                    unsigned e = __ballot_sync(FULL, ok && cnt == 0);
                    if (!e) e = __ballot_sync(FULL, ok);
                    if (!e) e = __ballot_sync(FULL, cnt < 2);   // no conflict-free bank left: accept a conflict
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
    __syncwarp();
    // Row ownership follows the layout: lane l commits (and later differentiates) the class placed in bank l of each
    // level, so that the row commit, its read-back and the staged row itself are conflict-free by construction.
    int* ipos = scratch + 1280 + 288;                            // [2][64] class at a position, -1 = none / zero slot
    for (int i = lane; i < 128; i += 32) ipos[i] = -1;
    __syncwarp();
    for (int c = lane; c < C; c += 32) {
        ipos[cpos[c]] = c;
        ipos[64 + cpos[72 + c]] = c;
    }
    __syncwarp();
    int oc[2][2];                                                // my class of utterance h at level k (clamped to 0)
    bool ov[2][2];
#pragma unroll
    for (int h = 0; h < 2; ++h)
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const int c = ipos[h * 64 + k * 32 + lane];
            ov[h][k] = c >= 0;
            oc[h][k] = c >= 0 ? c : 0;
        }

    // per-lane constants: emission pointers and skip masks of my label states
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
            eo[h] = h * CPMAX + cpos[h * 72 + (in ? c : C)];     // item C: a slot that always holds 0
            am[h] = (in && i > 0 && lab[h * 256 + i - 1] != c) ? 1.f : 0.f;
            bm[h] = (in && i + 1 < L[h] && lab[h * 256 + i + 1] != c) ? 1.f : 0.f;
        }
        ep0[q] = rows + eo[0];
        ep1[q] = rows + eo[1];
        ma[q] = make_float2(am[0], am[1]);
        mb[q] = make_float2(bm[0], bm[1]);
    }
    const float* bl0 = rows + cpos[blank];
    const float* bl1 = rows + CPMAX + cpos[72 + blank];
    const int zslot0 = cpos[C], zslot1 = CPMAX + cpos[72 + C];

    // Class reduction.  occupancy_c(t) = sum of alpha*u over the label states of class c is accumulated with
    // shared-memory integer atomics: a product, scaled to a fraction of Z (<= 1) times 2^23 * 0.999, is rounded to an
    // integer by adding 2^23 in fp32, and the BIT PATTERN is added to the class cell (the n_c copies of the bias
    // 0x4B000000 are subtracted by the reader).  Integer addition is associative, so the result does not depend on the
    // order in which conflicting lanes are served: bitwise reproducible.  The cell of a class sits at the same
    // position as its emission (bank-coloured), so the emission pointers address it and the owner lane reads it back
    // without conflicts.  Resolution 1.2e-7 of Z per product.
    unsigned nbias[4];                                           // n_c * bias of my classes: [2*h + level]
    for (int i = lane; i < 128; i += 32) hist[i] = 0;
    __syncwarp();
#pragma unroll
    for (int h = 0; h < 2; ++h)
#pragma unroll
        for (int q = 0; q < NL; ++q) {
            const int i = lane * NL + q;
            if (i < L[h]) atomicAdd(&hist[h * 64 + lab[h * 256 + i]], 1);
        }
    __syncwarp();
#pragma unroll
    for (int h = 0; h < 2; ++h)
#pragma unroll
        for (int k = 0; k < 2; ++k) nbias[2 * h + k] = (unsigned)hist[h * 64 + oc[h][k]] * 0x4B000000u;
    const bool isb00 = ov[0][0] && oc[0][0] == blank, isb01 = ov[0][1] && oc[0][1] == blank;
    const bool isb10 = ov[1][0] && oc[1][0] == blank, isb11 = ov[1][1] && oc[1][1] == blank;
    __syncwarp();
    // the occupancy cells start at zero; zero pads of the staged rows (emission of unused label slots)
    for (int i = lane; i < 2 * BUF; i += 32) occ[i] = 0u;
    if (lane < 2 * KF) {
        rows[lane * RW + zslot0] = 0.f;
        rows[lane * RW + zslot1] = 0.f;
    }
    __syncwarp();

    // hand the row warp its table
#pragma unroll
    for (int h = 0; h < 2; ++h)
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            ltab[2 * h + k] = oc[h][k];
            ltab[5 + 2 * h + k] = (int)nbias[2 * h + k];
        }
    ltab[4] = (int)ov[0][0] | ((int)ov[0][1] << 1) | ((int)ov[1][0] << 2) | ((int)ov[1][1] << 3);
    if (lane == 0) {
        misc[M_T0] = Tb[0]; misc[M_T1] = Tb[1];
        misc[M_ACT0] = act[0]; misc[M_ACT1] = act[1];
    }
    warp_signal(bars + 8 * B_SETUP, lane);
    (void)T; (void)isb00; (void)isb01; (void)isb10; (void)isb11;

    Geo geo;
    geo.init(Tb[0], Tb[1]);
    const int nch = geo.nch;
    // chunk j stages its emissions in buffer j & 1: the emission pointers follow
    int curb = 0;
    auto set_buf = [&](int b) {
        if (b != curb) {
            const int d = b ? BUF : -BUF;
#pragma unroll
            for (int q = 0; q < NL; ++q) { ep0[q] += d; ep1[q] += d; }
            bl0 += d; bl1 += d;
            curb = b;
        }
    };

    // ------------------------------------------------------------------------------------------ forward sweep
    float2 a[NSL];
#pragma unroll
    for (int j = 0; j < NSL; ++j) a[j] = make_float2(0.f, 0.f);
    if (lane == 0) a[0] = make_float2(act[0] ? 1.f : 0.f, act[1] ? 1.f : 0.f);   // virtual alpha_{-1}
    int ea[2] = {0, 0};
    float2 sca = make_float2(0.f, 0.f);
    int flag[2] = {0, 0};
    float zi[2] = {0.f, 0.f};
    int koff[2] = {0, 0};
    float* ck = ckpt_ws + (size_t)pair * n_ckpt * CKW;
    DT_DECL

    auto fwd_chunk = [&](auto nfc, int j, int t0) {
        constexpr int NF = decltype(nfc)::value;
        const bool kill0 = !act[0] || t0 >= Tb[0], kill1 = !act[1] || t0 >= Tb[1];
        const int b = j & 1;
        set_buf(b);
        {
            bool ok[2];
            renorm_alpha<NSL>(a, ea, sca, lane, ok, true);
            if (!ok[0] && !kill0 && !flag[0]) flag[0] = 2 | (t0 << 8);
            if (!ok[1] && !kill1 && !flag[1]) flag[1] = 2 | (t0 << 8);
        }
        if (want_grad) {                                // alpha entering chunk j, after its renormalisation
            float* c = ck + (size_t)j * CKW;
#pragma unroll
            for (int q = 0; q < NSL / 2; ++q)
                reinterpret_cast<float4*>(c)[q * 32 + lane] = make_float4(a[2 * q].x, a[2 * q].y, a[2 * q + 1].x, a[2 * q + 1].y);
            reinterpret_cast<int2*>(c + 64 * NSL)[lane] = make_int2(ea[0], ea[1]);
        }
        DT_WAIT(0, mbar_wait(bars + 8 * (B_FW_FULL + b), j >> 1))      // the row warp has staged chunk j
#pragma unroll
        for (int r = 0; r < NF; ++r) alpha_step<NSL>(a, sca, r * RW, bl0, bl1, ep0, ep1, ma);
        const int tend = t0 + NF;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            if (act[h] && tend == Tb[h]) {
                // P = alpha_{T-1}(S-1) + alpha_{T-1}(S-2), possibly in two lanes with different exponents
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
                    const int kx = ((__float_as_int(v) >> 23) & 0xff) - 127;     // P = v * 2^em with v in [1, 2)
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
        if (j == nch - 1) {                             // the row warp reads this after the last hand-back
            const bool go = want_grad && !((flag[0] || !act[0]) && (flag[1] || !act[1]));
            if (lane == 0) misc[M_GO_BWD] = go;
        }
        warp_signal(bars + 8 * (B_FW_FREE + b), lane);
    };
    for (int j = 0; j < nch; ++j) {
        int t0;
        if (geo.chunk(j, t0)) fwd_chunk(IC<KF>{}, j, t0);
        else fwd_chunk(IC<1>{}, j, t0);
    }
    if (!want_grad || ((flag[0] || !act[0]) && (flag[1] || !act[1]))) {
#pragma unroll
        for (int h = 0; h < 2; ++h)
            if (act[h] && lane == 0) p.flags[bidx[h]] = flag[h];
        return;
    }

    DT_MARK(4)                                          // lattice warp: forward sweep
    // ------------------------------------------------------------------------------------------ backward sweep
    float2 bt[NSL];
#pragma unroll
    for (int j = 0; j < NSL; ++j) bt[j] = make_float2(0.f, 0.f);
    int eb[2] = {0, 0};
    constexpr float FIX = 8388608.f * 0.999f;           // fixed-point scale of an occupancy fraction (ctc_pair.cu)
    const float2 zs2 = make_float2(zi[0] * FIX, zi[1] * FIX);
    const float2 magic2 = make_float2(8388608.f, 8388608.f);
    const unsigned ckst_s = (unsigned)__cvta_generic_to_shared(ckst);
    auto ckpt_issue = [&](int j) {
        const float* c = ck + (size_t)j * CKW;
#pragma unroll
        for (int q = 0; q < NSL / 2; ++q)
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(ckst_s + (q * 32 + lane) * 16), "l"(c + (q * 32 + lane) * 4));
        asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(ckst_s + (64 * NSL + lane * 2) * 4), "l"(c + 64 * NSL + lane * 2));
        asm volatile("cp.async.commit_group;");
    };
    auto ckpt_take = [&]() {
        asm volatile("cp.async.wait_group 0;" ::: "memory");
#pragma unroll
        for (int q = 0; q < NSL / 2; ++q) {
            const float4 v = reinterpret_cast<const float4*>(ckst)[q * 32 + lane];
            a[2 * q] = make_float2(v.x, v.y);
            a[2 * q + 1] = make_float2(v.z, v.w);
        }
        const int2 e = reinterpret_cast<const int2*>(ckst + 64 * NSL)[lane];
        ea[0] = e.x; ea[1] = e.y;
    };

    auto bwd_chunk = [&](auto nfc, int j, int t0) {
        constexpr int NF = decltype(nfc)::value;
        const int b = j & 1;
        set_buf(b);
        ckpt_take();
        // a / ea hold the checkpoint of this chunk: boundary factors and the beta exponents that go with it
        float2 scb;
        {
            float fb[2], sa[2], sb[2], ini[2];
            const int tend = t0 + NF;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int el = __shfl_up_sync(FULL, ea[h], 1), er = __shfl_down_sync(FULL, ea[h], 1);
                sa[h] = lane == 0 ? 0.f : pow2i(el - ea[h]);
                sb[h] = lane == 31 ? 0.f : pow2i(ea[h] - er);            // 2^(e_beta[l+1] - e_beta[l])
                const int ebn = koff[h] - ea[h];
                const bool starts = act[h] && tend == Tb[h];             // beta_T enters here
                fb[h] = starts ? 0.f : pow2i(eb[h] - ebn);
                ini[h] = starts ? pow2i(-ebn) : 0.f;
                if (starts) {                            // beta_T = 1 must be representable in the lane that owns S-1
                    const bool owner = (S[h] - 1) / NSL == lane;
                    if (__any_sync(FULL, owner && (ebn > 120 || ebn < -120)) && !flag[h]) flag[h] = 6;
                }
                eb[h] = ebn;
            }
            sca = make_float2(sa[0], sa[1]);
            scb = make_float2(sb[0], sb[1]);
            const float2 f2 = make_float2(fb[0], fb[1]);
#pragma unroll
            for (int jj = 0; jj < NSL; ++jj) {
                bt[jj] = __fmul2_rn(bt[jj], f2);
                const int s = lane * NSL + jj;
                if (s == S[0] - 1) bt[jj].x += ini[0];
                if (s == S[1] - 1) bt[jj].y += ini[1];
            }
        }
        DT_WAIT(1, mbar_wait(bars + 8 * (B_BW_FULL + b), (nch - 1 - j) >> 1))        // the row warp has staged chunk j
        // recompute alpha over the chunk, keeping the label states
        float2 A[NF][NL];
#pragma unroll
        for (int r = 0; r < NF; ++r) {
            alpha_step<NSL>(a, sca, r * RW, bl0, bl1, ep0, ep1, ma);
#pragma unroll
            for (int q = 0; q < NL; ++q) A[r][q] = a[2 * q + 1];
        }
        if (j > 0) ckpt_issue(j - 1);                   // the alpha registers are free: the next checkpoint sets off
        if (nch - 1 - j >= 2) DT_WAIT(2, mbar_wait(bars + 8 * (B_OCC_FREE + b), ((nch - 1 - j) >> 1) - 1))   // the cells of chunk j + 2 are read and reset
        // beta over the chunk; label products alpha_t(s) * u_t(s) into the class cells
#pragma unroll
        for (int r = NF - 1; r >= 0; --r) {
            float2 ul[NL];
            beta_step<NSL>(bt, ul, scb, make_float2(1.f, 1.f), r * RW, bl0, bl1, ep0, ep1, mb);
#pragma unroll
            for (int q = 0; q < NL; ++q) {
                const float2 fx = __ffma2_rn(__fmul2_rn(A[r][q], ul[q]), zs2, magic2);
                atomicAdd(reinterpret_cast<unsigned*>(const_cast<float*>(ep0[q])) + (OCCD + r * RW), __float_as_uint(fx.x));
                atomicAdd(reinterpret_cast<unsigned*>(const_cast<float*>(ep1[q])) + (OCCD + r * RW), __float_as_uint(fx.y));
            }
        }
        warp_signal(bars + 8 * (B_OCC_FULL + b), lane);
    };

    ckpt_issue(nch - 1);
    for (int j = nch - 1; j >= 0; --j) {
        int t0;
        if (geo.chunk(j, t0)) bwd_chunk(IC<KF>{}, j, t0);
        else bwd_chunk(IC<1>{}, j, t0);
    }
    DT_MARK(5)                                          // lattice warp: backward sweep
    DT_FLUSH(0)
    {   // beta must have stayed finite to the very first frame (a blank state's overflow shows nowhere else)
        int bad = 0;
#pragma unroll
        for (int jj = 0; jj < NSL; ++jj) {
            if (!(bt[jj].x < 3.0e38f)) bad |= 1;
            if (!(bt[jj].y < 3.0e38f)) bad |= 2;
        }
        bad = __reduce_or_sync(FULL, bad);
        mbar_wait(bars + 8 * B_LDONE, 0);               // the row warp's occupancy checks
        bad |= misc[M_BAD_ROWS];
        if ((bad & 1) && !flag[0]) flag[0] = 5;
        if ((bad & 2) && !flag[1]) flag[1] = 5;
#pragma unroll
        for (int h = 0; h < 2; ++h)
            if (act[h] && lane == 0) p.flags[bidx[h]] = flag[h];   // 0 = done here; else redone in log space
    }
}

#ifndef MDD_DUO_CTAS
#define MDD_DUO_CTAS 6                                   // pairs (= CTAs of two warps) per SM
#endif
// three warps per scheduler: 16384 / 3 / 32 = 170 registers
__global__ void __maxnreg__(168)
ctc_duo_kernel(CtcParams p, float* __restrict__ ckpt_ws, int n_ckpt) {
    extern __shared__ __align__(16) float smem_f[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int pair = blockIdx.x;
    float* base = smem_f;
    const unsigned bars = (unsigned)__cvta_generic_to_shared(base + O_BARS);
    if (warp == 0 && lane < 12) mbar_init(bars + 8 * lane, 1);
    __syncthreads();                                    // barriers initialised before either warp touches them
    if (warp) row_role(p, base, bars, lane, pair);
    else lattice_role(p, ckpt_ws, n_ckpt, base, bars, lane, pair);
}

// Four pairs per CTA: warps 0..3 (one warpgroup) are the lattice warps, warps 4..7 the row warps of the same four
// pairs.  The CTA starts with 128 registers per thread (two CTAs = 16 warps = 8 pairs per SM); the row warpgroup hands
// 40 registers per thread back to the CTA's pool and the lattice warpgroup takes them (setmaxnreg is a warpgroup-wide
// instruction, hence four pairs per CTA).
#ifndef MDD_QUAD_LREGS
#define MDD_QUAD_LREGS 168
#define MDD_QUAD_RREGS 88
#endif
#define MDD_STR2(x) #x
#define MDD_STR(x) MDD_STR2(x)
__global__ void __launch_bounds__(256, 2)
ctc_quad_kernel(CtcParams p, float* __restrict__ ckpt_ws, int n_ckpt) {
    extern __shared__ __align__(16) float smem_f[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int slot = warp & 3;
    const int pair = blockIdx.x * 4 + slot;
    float* base = smem_f + slot * PAIR_FLOATS;
    const unsigned bars = (unsigned)__cvta_generic_to_shared(base + O_BARS);
    if (warp < 4 && lane < 12) mbar_init(bars + 8 * lane, 1);
    __syncthreads();                                    // barriers initialised before either warp touches them
    if (warp >= 4) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 " MDD_STR(MDD_QUAD_RREGS) ";");
        row_role(p, base, bars, lane, pair);
    } else {
        asm volatile("setmaxnreg.inc.sync.aligned.u32 " MDD_STR(MDD_QUAD_LREGS) ";");
        lattice_role(p, ckpt_ws, n_ckpt, base, bars, lane, pair);
    }
}

#ifdef MDD_ROWONLY_TEST                                 // register census of each role on its own (compile-only)
__global__ void __maxnreg__(MDD_ROW_TEST_REGS) row_only_kernel(CtcParams p) {
    extern __shared__ __align__(16) float smem_f[];
    row_role(p, smem_f, (unsigned)__cvta_generic_to_shared(smem_f + O_BARS), threadIdx.x & 31, blockIdx.x);
}
__global__ void lattice_only_kernel(CtcParams p, float* __restrict__ ckpt_ws, int n_ckpt) {
    extern __shared__ __align__(16) float smem_f[];
    lattice_role(p, ckpt_ws, n_ckpt, smem_f, (unsigned)__cvta_generic_to_shared(smem_f + O_BARS), threadIdx.x & 31, blockIdx.x);
}
#endif

}  // namespace

#ifdef MDD_PAIR_TIMING
extern "C" void mdd_debug_duo_cycles(unsigned long long* out, int reset) {
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out, g_duo_cycles, sizeof(unsigned long long) * 16);
    if (reset) {
        unsigned long long z[16] = {0};
        cudaMemcpyToSymbol(g_duo_cycles, z, sizeof(z));
    }
}
#endif

bool ctc_duo_supported(int C, int Lmax) { return C <= 63 && Lmax <= PGeo<NSL>::LCAP && Lmax > PGeo<4>::LCAP; }

int launch_ctc_duo(CtcParams p, void* ws, cudaStream_t st) {
    // one CTA = one pair = two warps; shared memory padded so that exactly MDD_DUO_CTAS of them share an SM
    size_t smem = (size_t)PAIR_FLOATS * 4;
    const size_t want = ((233472 / (size_t)MDD_DUO_CTAS) - 1024 - 64) & ~(size_t)15;
    if (want > smem) smem = want;
    static bool configured = false;
    if (!configured) {
        if (cudaFuncSetAttribute(ctc_duo_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
            cudaGetLastError();
            return MDD_E_CUDA;
        }
        cudaFuncSetAttribute(ctc_duo_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
        configured = true;
        if (getenv("MDD_CTC_DUO_DEBUG")) {
            int nb = 0;
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, ctc_duo_kernel, 64, smem);
            cudaFuncAttributes fa;
            cudaFuncGetAttributes(&fa, ctc_duo_kernel);
            fprintf(stderr, "ctc_duo: %d CTAs per SM, %d registers, %zu B dynamic + %zu B static smem\n", nb, fa.numRegs, smem,
                    fa.sharedSizeBytes);
        }
    }
    const int npairs = (p.B + 1) / 2;
    const int n_ckpt = p.T / KF + 2 * KF;
    static const bool quad = getenv("MDD_CTC_DUO") && atoi(getenv("MDD_CTC_DUO")) == 2;
    if (quad) {
        const size_t qsmem = (size_t)4 * PAIR_FLOATS * 4;
        static bool qconf = false;
        if (!qconf) {
            if (cudaFuncSetAttribute(ctc_quad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)qsmem) != cudaSuccess) {
                cudaGetLastError();
                return MDD_E_CUDA;
            }
            cudaFuncSetAttribute(ctc_quad_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
            qconf = true;
            if (getenv("MDD_CTC_DUO_DEBUG")) {
                int nb = 0;
                cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, ctc_quad_kernel, 256, qsmem);
                cudaFuncAttributes fa;
                cudaFuncGetAttributes(&fa, ctc_quad_kernel);
                fprintf(stderr, "ctc_quad: %d CTAs per SM, %d registers, %zu B dynamic smem\n", nb, fa.numRegs, qsmem);
            }
        }
        ctc_quad_kernel<<<(npairs + 3) / 4, 256, qsmem, st>>>(p, reinterpret_cast<float*>(ws), n_ckpt);
        MDD_CHECK_LAUNCH();
        return MDD_OK;
    }
    ctc_duo_kernel<<<npairs, 64, smem, st>>>(p, reinterpret_cast<float*>(ws), n_ckpt);
    MDD_CHECK_LAUNCH();
    return MDD_OK;
}

}  // namespace mdd
