// ctc_pair.cu -- CTC loss + gradient, default fast path of mdd_ctc_loss: two utterances per warp, packed FP32.
//
// Semantics: torch aten ctc_loss / _ctc_loss_backward as reached from nn.CTCLoss at
// egs/attention_aug/steps/train_ctc.py:186,72,96 and egs/cnn-rnn-ctc/steps/train_ctc.py:148,51,67.
//
// The lattice recursion is bound by per-warp instruction latency and by shared-memory wavefronts, not by HBM (DESIGN.md
// 5.1), so this kernel is organised around issuing as few instructions per lattice state as possible:
//
//  * One warp owns a PAIR of utterances.  Lane l holds NSL consecutive extended-label states of both utterances as
//    float2 registers (.x = first utterance, .y = second); every lattice update is a packed sm_100 instruction
//    (FADD2 / FFMA2 / FMUL2), i.e. half the FP32 issue slots of a scalar kernel.  Neighbour states across lanes come
//    from warp shuffles; there is no block barrier anywhere.
//  * Linear domain with block floating point (as ctc_fast.cu): emissions are exponentiated once per (t,c); alpha is
//    renormalised once per chunk of KF = 8 frames from each lane's own maximum, a max-scan keeps a lane within 2^GAP
//    of the lane it receives from.  The checkpoint of a chunk is taken AFTER its renormalisation, so the recomputation
//    in the backward sweep needs no renormalisation at all and is bitwise the forward's.
//  * beta has no normalisation of its own: its per-lane exponent is DEFINED as  e_beta = e_P + OFF - e_alpha  (e_P the
//    exponent of the total likelihood P).  Because alpha_t(s) * u_t(s) <= P for every state (u = beta without the
//    emission at t), the beta mantissas are bounded by 2^-OFF / (alpha mantissa), and whatever underflows belongs to
//    paths of relative mass < 2^-75.  The products alpha * u of all lanes are then on ONE common scale, sum_s = P, so
//    neither a per-frame normaliser nor the blank states' products are needed:
//        occupancy_c / Z = (sum_{s: l'_s = c} alpha u) * (2^OFF / mantissa(P)),  occupancy_blank = Z - sum of labels.
//  * Ragged pairs: frames [0, min(T0,T1)) are processed jointly, then the longer utterance alone (the other half runs
//    on zero emissions, stays finite, and its gradient rows are written as zeros).  Chunks never straddle the boundary.
//  * Class reduction: the product of a label state, as a fixed-point fraction of Z, is added with a shared-memory
//    integer atomic to the cell of its class, which sits at the class's (bank-coloured) emission position; the lane
//    that owns the position reads it back.  Integer addition is associative: results are bitwise reproducible.
//  * Nothing that travels far touches a register: log-prob rows go global -> shared by cp.async (8-byte pieces, two
//    chunks ahead, three staging slots), checkpoints likewise one chunk ahead.  Gradient values are stored straight
//    from the lanes that own their classes (four 4-byte streaming stores per row; the sectors fill up in L2): staging
//    the row in shared memory for wide stores costs more on the shared-memory pipe, which binds the kernel.
//  * Dynamic range.  One exponent per lane and chunk is a finite window, and peaky posteriors move 20-35 bits per
//    frame.  A renormalised lane keeps its maximum at 2^SHIFT; a lane far below its left neighbour is pulled up to
//    2^-GAP of it and may lose its own content, which the backward sweep does not share: its total likelihood
//    beta_0(0) + beta_0(1) must agree with the forward P (flag 7).  What that check cannot see (beta's exponents are
//    tied to alpha's, so beta is blind to what a lane lost once the lane's exponent followed the loss down) is
//    excluded in the forward sweep: the lanes holding the bulk of the mass must not decay by more than 2^decay_max
//    inside one chunk (flag 2).  DESIGN.md 5.1 has the measured envelope.
//  * Three tiers.  Anything unusual (infeasible, empty, malformed, over/underflow, the guards above) sets flags[b].
//    The flagged utterances are compacted into a list and run again by this same kernel at CH = 2 frames per chunk
//    (LIST instantiation, persistent warps over the first pass's checkpoint blocks; few flagged utterances go to the
//    team kernel instead); the fp64 kernel of ctc_exact.cu then redoes whatever is still flagged in the same stream
//    with torch's inf / NaN / zero_infinity semantics.
//  * Two kernels per step (PHASE 1: forward sweep at 128 registers, PHASE 2: backward sweep), and for few utterances the
//    backward sweep on independent segments of every pair (SEG) behind the concurrent alpha / beta sweeps of
//    ctc_dual.cu.  Kernels that follow another kernel of the call are launched as programmatic dependents.
//  * LOGITS instantiation: lp holds the output layer's logits, lse their row log-sum-exp (row_lse.cu); the emission
//    is 2^(x log2 e - lse log2 e) and the gradient written is the one with respect to the logits.
// This file is compiled twice so that the two halves of its instantiations build in parallel: as itself (log-prob
// input, MDD_PAIR_TU_LOGITS 0, plus the public entry points) and through ctc_pair_logits.cu (logit input).
#ifndef MDD_PAIR_TU_LOGITS
#define MDD_PAIR_TU_LOGITS 0
#endif
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <type_traits>

#include "common.cuh"
#include "ctc_lattice.cuh"
#include "ctc_params.cuh"

namespace mdd {
namespace {

using namespace lattice;

constexpr int RAWW = 132;                  // floats per staged pair row: 2*C log-probs (C <= 62) + 2 row-LSE + pad (16 B multiple)
// staging slots: the chunk being consumed and two in flight; the forward-only kernel (more warps per SM) keeps one in flight
__host__ __device__ constexpr int nslot(int phase) { return phase == 1 ? 2 : 3; }

struct PairSmem {
    int occ_off, rows_off, abuf_off, ckst_off, raw_off, total_floats;
};
__host__ __device__ inline PairSmem pair_smem_layout(int NSL, bool regbuf, int phase = 0, bool seg = false) {
    PairSmem s;
    int o = 0;
    // class occupancies of the chunk, fixed point.  The setup tables (2*KF*RW ints) live in this block and the next;
    // the forward-only kernel has no occupancies: its tables lie over the emission rows and the staging slots.
    s.occ_off = o; if (phase != 1) o += KF * RW;
    // exp(lp) rows of the current chunk, both utterances (segments: a ring of three chunks of the sweep kernel's rows)
    s.rows_off = o; o += (seg ? nslot(phase) : 1) * KF * RW;
    s.abuf_off = o; if (!regbuf && phase != 1) o += KF * 32 * NSL;   // alpha label states of the chunk when they do not fit in registers
    s.ckst_off = o; if (phase != 1) o += 32 * (2 * NSL + 2);   // the next checkpoint, landing by cp.async (backward sweep)
    o = (o + 3) & ~3;
    s.raw_off = o; if (!seg) o += nslot(phase) * KF * RAWW;  // log-prob rows as they arrive (class order)
    s.total_floats = (o + 3) & ~3;
    return s;
}

// CH = frames per chunk (renormalisation / checkpoint period): 8 for the first pass over the whole batch, 2 for the
// second tier, which (LIST) walks the compacted list of the utterances the first pass flagged: four times the dynamic
// range per frame for about twice the time per frame, in front of the fp64 kernel.
// One pair of utterances, start to finish.  `pair` selects the utterances (2*pair, 2*pair+1 of the batch, or of the
// compacted list), `slot` the checkpoint block of the workspace this warp may use.
// PHASE 0: both sweeps in one kernel.  PHASE 1: the forward sweep alone (loss, checkpoints, a record of P per
// utterance) -- it needs a third of the registers, so two to three times as many warps share an SM; PHASE 2: the
// backward sweep alone, from the checkpoints and the record.
// SEG (with PHASE 2): the backward sweep of ONE SEGMENT of the pair's chunks, [seg * nch / nseg, (seg + 1) * nch / nseg):
// beta enters from the checkpoint the concurrent beta sweep of ctc_dual.cu left at the segment's upper end (its own
// exponents; from there on they are tied to alpha's as everywhere else), the emission-row layout comes from the
// workspace.  Segments of a pair are independent of each other, so few utterances still fill the machine.
template <int NSL, bool REGBUF, bool LOGITS, int CH, bool LIST, int PHASE = 0, bool SEG = false>
__device__ __forceinline__ void ctc_pair_body(const CtcParams& p, float* __restrict__ ckpt_ws, int n_ckpt, int pair,
                                              int slot, int n_utt, int seg = 0) {
    static_assert(!SEG || (PHASE == 2 && !LIST), "segments: backward kernel of the first pass only");
    // Segments do not read the log-probs at all: the sweep kernel's helpers have stored the exponentiated rows they
    // committed, exactly as they lie in shared memory (512 B per pair-frame), and a segment lands them by cp.async where
    // its lattice gathers them -- no second exponentiation, no second commit (13 of the backward sweep's 90 shared-
    // memory wavefronts per pair-frame).  (At full batch size the extra 1 GB of writes makes the forward kernel
    // DRAM-bound and the step no faster; a few hundred utterances are nowhere near that.)
    constexpr bool ELOAD = SEG;
    using G = PGeo<NSL>;
    constexpr int NL = G::NL, ARS = G::ARS, CKW = G::CKW;
    static_assert(CH >= 1 && CH <= KF && (CH & (CH - 1)) == 0, "chunk length: a power of two up to the staged rows");
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    extern __shared__ __align__(16) float smem_f[];
    constexpr int NSLOT = nslot(PHASE);
    const PairSmem lay = pair_smem_layout(NSL, REGBUF, PHASE, SEG);
    float* base = smem_f + (size_t)warp * lay.total_floats;
    unsigned* occ = reinterpret_cast<unsigned*>(base + lay.occ_off);   // [KF][RW], laid out like the emission rows
    float* rows = base + lay.rows_off;                           // [KF][RW]
    float* abuf = base + lay.abuf_off;                           // [KF][ARS] (!REGBUF)
    float* ckst = base + lay.ckst_off;                           // [CKW]
    float* raw = base + lay.raw_off;                             // [NSLOT][KF][RAWW]
    constexpr int OCCD = -KF * RW;                               // occupancy cell of an emission slot: rows -> occ

    const int C = p.C, blank = p.blank, T = p.T;
    const bool want_grad = p.grad != nullptr;

    // ------------------------------------------------------------------------------------------ per-utterance setup
    int bidx[2], Tb[2], L[2], S[2];
    bool act[2];
    int* scratch = reinterpret_cast<int*>(occ);                  // lab[2][256] | hist[2][64] | ...
    int* lab = scratch;
    int* hist = scratch + 512;
    static_assert(2 * KF * RW >= 1280 + 288 + 128, "setup scratch must fit in the occupancy + emission rows");
    static_assert(KF * RW + 2 * KF * RAWW >= 2 * KF * RW, "forward-only kernel: setup scratch over the rows + staging slots");
    static_assert(512 + 144 <= 1280, "label histogram [2][72] behind the labels");
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
                if (lane == 0) p.flags[b] = 1;                   // the fp64 kernel owns the error reporting
                act[h] = false;
            } else {
                Tb[h] = (int)Tb64;
                L[h] = (int)L64;
            }
        }
        S[h] = 2 * L[h] + 1;
    }
    if (!act[0] && !act[1]) return;
    __syncwarp();

    // Emission layout.  The lattice gathers e[t][label] with one LDS per label state; with C > 32 classes a plain row
    // makes two of the requested classes share a bank in almost every gather.  Per utterance, classes (and the zero
    // that unused label slots read, item C) are therefore placed at position bank + 32*level such that two classes
    // requested by the same gather instruction (same label slot q of different lanes) never share a bank: items are
    // placed greedily, those present in most gather groups first, each into a bank whose residents are requested by
    // none of its groups.
    int* gm = scratch + 1280;                                    // [2][72] gather-group mask per item
    int* cpos = scratch + 1280 + 144;                            // [2][72] position per item
    if constexpr (PHASE == 2 && !LIST && !SEG) {                 // the forward kernel has coloured the row of this pair
        pdl_wait();
        for (int i = lane; i < 144; i += 32) cpos[i] = p.cpos_ws[(size_t)pair * 144 + i];
    } else if constexpr (SEG) {
        if (p.T * NL < kColourMinWork) {                         // short sweeps: the plain row, nothing to wait for yet
            for (int i = lane; i < 144; i += 32) cpos[i] = i % 72;
        } else {
            pdl_wait();                                          // the layout comes from the sweep kernel
            for (int i = lane; i < 144; i += 32) cpos[i] = p.cpos_ws[(size_t)pair * 144 + i];
        }
    } else {
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
    if constexpr (PHASE == 1 && !LIST) {                         // ... and hands the layout to the backward kernel
        __syncwarp();
        if (want_grad && p.cpos_ws)
            for (int i = lane; i < 144; i += 32) p.cpos_ws[(size_t)pair * 144 + i] = cpos[i];
    }
    }
    __syncwarp();
    // Row ownership follows the layout: lane l commits (and later differentiates) the class placed in bank l of each
    // level, so that the row commit, its read-back and the staged row itself are conflict-free by construction.
    int* ipos = scratch + 1280 + 288;                            // [2][64] class at a position, -1 = none / zero slot
    for (int i = lane; i < 128; i += 32) ipos[i] = -1;
    __syncwarp();
    for (int c = lane; c <= C; c += 32) {                         // item C = the zero that unused label slots read
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
            ov[h][k] = c >= 0 && c < C;
            oc[h][k] = ov[h][k] ? c : 0;
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
    // without conflicts.  Resolution 1.2e-7 of Z per product.  Unused label slots add the bare bias to the cell of
    // the zero item; the lane that reads that position subtracts their count like a class's, positions nobody is
    // placed at never change, so every cell a lane reads comes out as a plain integer and needs no masking.
    unsigned nbias[4];                                           // n * bias of the item at my positions: [2*h + level]
    for (int i = lane; i < 144; i += 32) hist[i] = 0;            // [2][72]: classes 0..C-1, the zero item C
    __syncwarp();
#pragma unroll
    for (int h = 0; h < 2; ++h)
#pragma unroll
        for (int q = 0; q < NL; ++q) {
            const int i = lane * NL + q;
            atomicAdd(&hist[h * 72 + (i < L[h] ? lab[h * 256 + i] : C)], 1);
        }
    __syncwarp();
#pragma unroll
    for (int h = 0; h < 2; ++h)
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const int it = ipos[h * 64 + k * 32 + lane];         // class, C = the zero item, -1 = nobody
            nbias[2 * h + k] = it >= 0 ? (unsigned)hist[h * 72 + it] * 0x4B000000u : 0u;
        }
    const bool isb00 = ov[0][0] && oc[0][0] == blank, isb01 = ov[0][1] && oc[0][1] == blank;
    const bool isb10 = ov[1][0] && oc[1][0] == blank, isb11 = ov[1][1] && oc[1][1] == blank;
    __syncwarp();
    // the occupancy cells start at zero; zero pads of the staged rows (emission of unused label slots)
    if constexpr (PHASE != 1)
        for (int i = lane; i < KF * RW; i += 32) occ[i] = 0u;
    __syncwarp();

    // ------------------------------------------------------------------------------------------ geometry
    // Chunks are either full (KF frames) or single frames (the remainders of the two phases), so that every chunk
    // body is straight-line code: joint phase [0, Tmin), then the longer utterance alone [Tmin, Tmax).
    const int Tmin = min(Tb[0], Tb[1]), Tmax = max(Tb[0], Tb[1]);
    const int nJ8 = Tmin / CH, cJ = nJ8 + Tmin % CH;
    const int nS8 = (Tmax - Tmin) / CH, cS = cJ + nS8;
    const int nch = cS + (Tmax - Tmin) % CH;
    auto chunk_geom = [&](int j, int& t0) -> bool {               // true: full chunk
        if (j < nJ8) { t0 = j * CH; return true; }
        if (j < cJ) { t0 = nJ8 * CH + (j - nJ8); return false; }
        if (j < cS) { t0 = Tmin + (j - cJ) * CH; return true; }
        t0 = Tmin + nS8 * CH + (j - cS);
        return false;
    };
    // the chunks this warp sweeps backwards: all of them, or its segment
    const int jlo = SEG ? (int)((long long)seg * nch / p.nseg) : 0;
    const int jhi = SEG ? (int)((long long)(seg + 1) * nch / p.nseg) : nch;
    if (SEG && jlo >= jhi) return;
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

    // Log-prob rows travel global -> shared by cp.async in 8-byte pieces (a row of C floats is 8-byte aligned when C
    // and the strides are even): the pair's row is 2*C/2 pieces, lane l moves pieces l and l + 32.  Staged in class
    // order: utterance h at [h*C, h*C + C) of the slot's row, the two row-LSE values (LOGITS) behind them.
    const bool vec8 = ((C & 1) == 0) && ((p.stride_t & 1) == 0) && ((p.stride_b & 1) == 0) &&
                      ((reinterpret_cast<uintptr_t>(p.lp) & 7) == 0);
    const int hC = C >> 1;
    const bool pv0 = lane < 2 * hC, pv1 = lane + 32 < 2 * hC;
    const int ph0 = lane >= hC ? 1 : 0, ph1 = lane + 32 >= hC ? 1 : 0;
    const int pd0 = 2 * lane, pd1 = 2 * (lane + 32);             // destination (floats): h*C + 2*(u - h*hC) = 2*u
    const float* ps0 = (ph0 ? lpb1 : lpb0) + 2 * (lane - ph0 * hC);
    const float* ps1 = (ph1 ? lpb1 : lpb0) + 2 * (lane + 32 - ph1 * hC);
    const float* lsep = nullptr;
    if constexpr (LOGITS) lsep = p.lse + ((lane & KF) ? bsrc1 : bsrc0);
    const unsigned raw_s = (unsigned)__cvta_generic_to_shared(raw);
    const unsigned tstride_b = (unsigned)(p.stride_t * 4);       // row pitch in bytes (< 4 GB, checked by the launcher)
    const float* er = ELOAD ? p.erow + (size_t)slot * p.T * RW : nullptr;   // [T][RW] emission rows of this pair
    const unsigned rows_s = (unsigned)__cvta_generic_to_shared(rows);
    auto issue_rows = [&](int j) {                       // rows of chunk j -> slot j % NSLOT (its own cp.async group)
        if constexpr (ELOAD) {                           // 16 bytes per lane and frame, straight into the ring
            if (j >= jlo && j < nch) {
                int t0;
                const bool full = chunk_geom(j, t0);
                const unsigned d = rows_s + ((j % NSLOT) * (KF * RW) + lane * 4) * 4;
                const float* q = er + (size_t)t0 * RW + lane * 4;
                if (full) {
#pragma unroll
                    for (int r = 0; r < CH; ++r)
                        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d + r * (RW * 4)), "l"(q + r * RW));
                } else {
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(q));
                }
            }
            asm volatile("cp.async.commit_group;");
            return;
        }
        if (j >= jlo && j < nch) {
            int t0;
            const bool full = chunk_geom(j, t0);
            const int so = (j % NSLOT) * (KF * RAWW);
            auto body = [&](auto nfc) {
                constexpr int NF = decltype(nfc)::value;
                if (vec8) {
                    const unsigned d0 = raw_s + (so + pd0) * 4, d1 = raw_s + (so + pd1) * 4;
                    const unsigned long long o = (unsigned long long)(unsigned)t0 * tstride_b;
                    const char* q0 = reinterpret_cast<const char*>(ps0) + o;
                    const char* q1 = reinterpret_cast<const char*>(ps1) + o;
#pragma unroll
                    for (int r = 0; r < NF; ++r) {
                        if (pv0) asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d0 + r * (RAWW * 4)), "l"(q0));
                        if (pv1) asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d1 + r * (RAWW * 4)), "l"(q1));
                        q0 += tstride_b;
                        q1 += tstride_b;
                    }
                } else {                                 // odd C / odd strides: element-wise, through registers
                    size_t o = (size_t)t0 * p.stride_t;
                    for (int r = 0; r < NF; ++r) {
                        for (int i = lane; i < 2 * C; i += 32) {
                            const int h = i >= C ? 1 : 0;
                            raw[so + r * RAWW + i] = (h ? lpb1 : lpb0)[o + i - h * C];
                        }
                        o += p.stride_t;
                    }
                }
                if constexpr (LOGITS) {
                    const int r = lane & (KF - 1);
                    if (lane < 2 * KF && r < NF)
                        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(raw_s + (so + r * RAWW + 2 * C + (lane >> 3)) * 4),
                                     "l"(lsep + (size_t)(t0 + r) * p.B));
                }
            };
            if (full) body(IC<CH>{});
            else body(IC<1>{});
        }
        asm volatile("cp.async.commit_group;");
    };
    // my four (utterance, level) items inside a staged row in class order (a position without a class reads class 0)
    const int ro00 = oc[0][0], ro01 = oc[0][1], ro10 = C + oc[1][0], ro11 = C + oc[1][1];
    // ... and where their exponentials / gradients go.  Stores are unconditional: a position nobody is placed at takes
    // whatever its lane computes (nobody reads it); the lane at the zero item's position writes to a free one instead;
    // gradient values without a class land in the pad of the staging row.
    int wp[4], wg[4];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const unsigned free0 = __ballot_sync(FULL, ipos[h * 64 + lane] < 0), free1 = __ballot_sync(FULL, ipos[h * 64 + 32 + lane] < 0);
        const int fpos = free0 ? __ffs(free0) - 1 : 32 + __ffs(free1) - 1;       // 64 slots, C + 1 <= 64 items: one is free
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const bool zero_item = ipos[h * 64 + k * 32 + lane] == C;
            wp[2 * h + k] = h * CPMAX + (zero_item ? fpos : k * 32 + lane);
            wg[2 * h + k] = ov[h][k] ? h * C + oc[h][k] : 2 * C + 2;
        }
    }
    // Only now, with the last of the set-up tables read: the emission of unused label slots.  (The tables lie over the
    // emission rows; zeroing these slots before `ipos` had been read for the last time could turn the one free position
    // of a 62-class row into an occupied one -- the zero item's owner then parked its store on a real class -- or hide
    // the partner's zero item from its owner.  Found by the randomised sweep of round 2.)
    __syncwarp();
    if (lane < KF) {
        rows[lane * RW + zslot0] = 0.f;
        rows[lane * RW + zslot1] = 0.f;
    }
    __syncwarp();
    auto commit_rows = [&](auto nfc, int j, bool kill0, bool kill1) {   // p = exp(lp), bank-coloured; an idle half emits zeros
        constexpr int NF = decltype(nfc)::value;
        const float* slot = raw + (j % NSLOT) * (KF * RAWW);
        auto body = [&](auto killc) {
            constexpr bool KILL = decltype(killc)::value != 0;
            // rows in batches of up to four: all of a batch's loads are issued before the first exponential, so that
            // a batch costs one shared-memory latency instead of one per row
            constexpr int RB = NF < 4 ? NF : 4;
#pragma unroll
            for (int r0 = 0; r0 < NF; r0 += RB) {
                float x[RB][4], nl[RB][2];
#pragma unroll
                for (int i = 0; i < RB; ++i) {
                    const float* s = slot + (r0 + i) * RAWW;
                    x[i][0] = s[ro00]; x[i][1] = s[ro01]; x[i][2] = s[ro10]; x[i][3] = s[ro11];
                    if constexpr (LOGITS) { nl[i][0] = s[2 * C]; nl[i][1] = s[2 * C + 1]; }
                }
#pragma unroll
                for (int i = 0; i < RB; ++i) {
                    float2 v0, v1;
                    if constexpr (LOGITS) {
                        const float n0 = nl[i][0] * -LOG2E, n1 = nl[i][1] * -LOG2E;
                        v0 = make_float2(fast_exp2(fmaf(x[i][0], LOG2E, n0)), fast_exp2(fmaf(x[i][1], LOG2E, n0)));
                        v1 = make_float2(fast_exp2(fmaf(x[i][2], LOG2E, n1)), fast_exp2(fmaf(x[i][3], LOG2E, n1)));
                    } else {
                        v0 = make_float2(fast_exp(x[i][0]), fast_exp(x[i][1]));
                        v1 = make_float2(fast_exp(x[i][2]), fast_exp(x[i][3]));
                    }
                    if constexpr (KILL) {
                        if (kill0) v0 = make_float2(0.f, 0.f);
                        if (kill1) v1 = make_float2(0.f, 0.f);
                    }
                    float* d = rows + (r0 + i) * RW;
                    d[wp[0]] = v0.x;
                    d[wp[1]] = v0.y;
                    d[wp[2]] = v1.x;
                    d[wp[3]] = v1.y;
                }
            }
        };
        if (kill0 || kill1) body(IC<1>{});
        else body(IC<0>{});
    };

    // ------------------------------------------------------------------------------------------ forward sweep
    float2 a[NSL];
#pragma unroll
    for (int j = 0; j < NSL; ++j) a[j] = make_float2(0.f, 0.f);
    if (lane == 0) a[0] = make_float2(act[0] ? 1.f : 0.f, act[1] ? 1.f : 0.f);   // virtual alpha_{-1}
    int ea[2] = {0, 0};
    float2 sca = make_float2(0.f, 0.f);
    int flag[2] = {0, 0};
    float zi[2] = {0.f, 0.f}, zc[2] = {0.f, 0.f};
    int koff[2] = {0, 0};
    float* ck = ckpt_ws + (size_t)slot * n_ckpt * CKW;

    if constexpr (PHASE != 2) {
        issue_rows(0);
        if constexpr (NSLOT == 3) issue_rows(1);
    }
    if (want_grad && PHASE != 1 && jhi == nch) {        // rows t >= T_max of the gradient are zero (torch semantics)
        for (int t = Tmax; t < T; ++t)
            for (int i = lane; i < C; i += 32) {
                if (act[0]) __stcs(gr0 + (size_t)t * gstride + i, 0.f);
                if (act[1]) __stcs(gr1 + (size_t)t * gstride + i, 0.f);
            }
    }

    auto fwd_chunk = [&](auto nfc, int j, int t0) {
        constexpr int NF = decltype(nfc)::value;
        const bool kill0 = !act[0] || t0 >= Tb[0], kill1 = !act[1] || t0 >= Tb[1];
        if constexpr (NSLOT == 3) asm volatile("cp.async.wait_group 1;" ::: "memory");   // rows of chunk j have landed (j + 1 may be in flight)
        else asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncwarp();
        commit_rows(nfc, j, kill0, kill1);
        __syncwarp();
        issue_rows(j + NSLOT - 1);                      // into the slot chunk j - 1 was consumed from
        {
            bool ok[2];
            renorm_alpha<NSL>(a, ea, sca, lane, ok, p.strict_squash != 0, p.decay_max);
            if (!ok[0] && !kill0 && !flag[0]) flag[0] = 2 | (t0 << 8);
            if (!ok[1] && !kill1 && !flag[1]) flag[1] = 2 | (t0 << 8);
        }
        if (want_grad) {                                // alpha entering chunk j, after its renormalisation
            float* c = ck + (size_t)j * CKW;
#pragma unroll
            for (int q = 0; q < NSL; ++q) reinterpret_cast<float2*>(c)[q * 32 + lane] = a[q];   // register pairs as they are
            reinterpret_cast<int2*>(c + 64 * NSL)[lane] = make_int2(ea[0], ea[1]);
        }
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
                    // P = v * 2^em with v brought into [1, 2): the beta exponents below are tied to P's own exponent,
                    // so the products alpha*u have mantissas of about 2^-OFF whatever the final lanes' scale was
                    const int kx = ((__float_as_int(v) >> 23) & 0xff) - 127;
                    v *= pow2i(-kx);
                    em += kx;
                    if (lane == 0) {
                        p.nll[bidx[h]] = (float)(-((double)logf(v) + (double)em * 0.69314718055994530942));
                        if (p.status) p.status[bidx[h]] = MDD_ST_OK;
                    }
                    zc[h] = v * pow2i(-OFF);
                    zi[h] = pow2i(OFF) / v;
                    koff[h] = em + OFF;
                }
            }
        }
    };
    if constexpr (PHASE != 2) {
        for (int j = 0; j < nch; ++j) {
            int t0;
            if (chunk_geom(j, t0)) fwd_chunk(IC<CH>{}, j, t0);
            else fwd_chunk(IC<1>{}, j, t0);
        }
        asm volatile("cp.async.wait_all;" ::: "memory");
    }
    if constexpr (PHASE == 1) {                         // hand P over to the backward kernel: 1 / mantissa, exponent, flag
        if (want_grad && lane == 0) {
#pragma unroll
            for (int h = 0; h < 2; ++h)
                if (act[h]) p.rec[bidx[h]] = make_float4(zi[h], __int_as_float(koff[h]), __int_as_float(flag[h]), 0.f);
        }
#pragma unroll
        for (int h = 0; h < 2; ++h)
            if (act[h] && lane == 0 && (!want_grad || flag[h])) p.flags[bidx[h]] = flag[h];
        return;
    }
    if constexpr (PHASE == 2) {
        pdl_wait();                                     // set-up done; from here on the forward kernel's results are read
#pragma unroll
        for (int h = 0; h < 2; ++h)
            if (act[h]) {
                const float4 r = p.rec[bidx[h]];
                zi[h] = r.x; koff[h] = __float_as_int(r.y); flag[h] = __float_as_int(r.z);
            }
    }
    if (!want_grad || ((flag[0] || !act[0]) && (flag[1] || !act[1]))) {
#pragma unroll
        for (int h = 0; h < 2; ++h)
            if (act[h] && lane == 0) p.flags[bidx[h]] = flag[h];
        return;
    }

    // ------------------------------------------------------------------------------------------ backward sweep
    float2 bt[NSL];
#pragma unroll
    for (int j = 0; j < NSL; ++j) bt[j] = make_float2(0.f, 0.f);
    int eb[2] = {0, 0};
    constexpr float FIX = 8388608.f * 0.999f;           // fixed-point scale of an occupancy fraction
    const float2 zs2 = make_float2(zi[0] * FIX, zi[1] * FIX);
    const float2 magic2 = make_float2(8388608.f, 8388608.f);
    unsigned occ_hi = 0u;                               // largest (row total + 8192) seen: range check of the occupancies

    // The gradient rows of a chunk leave as 16-byte streaming stores of the staged row when the pair's two rows are
    // adjacent and 16-byte aligned in the gradient array; otherwise element-wise from the same staging.
    const bool gvec = !LIST && act[0] && act[1] && bidx[1] == bidx[0] + 1 && (C & 1) == 0 && ((p.B * C) & 3) == 0 &&
                      (bidx[0] & 1) == 0 && ((reinterpret_cast<uintptr_t>(p.grad) & 15) == 0);
    const unsigned gstride_b = (unsigned)(gstride * 4);           // gradient row pitch in bytes (< 4 GB, checked by the launcher)

    // alpha entering chunk j (post-renormalisation) + exponents travel global -> shared by cp.async, one chunk ahead,
    // and are picked up with LDS.
    const unsigned ckst_s = (unsigned)__cvta_generic_to_shared(ckst);
    auto ckpt_issue = [&](int j) {
        if (j >= jlo) {
            const float* c = ck + (size_t)j * CKW;
#pragma unroll
            for (int q = 0; q < NSL / 2; ++q)
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(ckst_s + (q * 32 + lane) * 16), "l"(c + (q * 32 + lane) * 4));
            asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(ckst_s + (64 * NSL + lane * 2) * 4), "l"(c + 64 * NSL + lane * 2));
        }
        asm volatile("cp.async.commit_group;");
    };
    auto ckpt_take = [&]() {
#pragma unroll
        for (int q = 0; q < NSL; ++q) a[q] = reinterpret_cast<const float2*>(ckst)[q * 32 + lane];
        const int2 e = reinterpret_cast<const int2*>(ckst + 64 * NSL)[lane];
        ea[0] = e.x; ea[1] = e.y;
    };

    auto bwd_chunk = [&](auto nfc, int j, int t0) {
        constexpr int NF = decltype(nfc)::value;
        const bool kill0 = !act[0] || t0 >= Tb[0], kill1 = !act[1] || t0 >= Tb[1];
        // in flight, oldest first: rows(j), checkpoint(j), rows(j-1): everything but the newest has landed
        asm volatile("cp.async.wait_group 1;" ::: "memory");
        __syncwarp();
        ckpt_take();
        if constexpr (!ELOAD)
            if (PHASE == 2 || j < nch - 1) commit_rows(nfc, j, kill0, kill1);   // (one kernel: the last chunk is still staged)
        const int eo = ELOAD ? (j % NSLOT) * (KF * RW) : 0;     // where this chunk's emission rows lie
        __syncwarp();
        ckpt_issue(j - 1);                              // the staging block has been read
        // a / ea hold the checkpoint of this chunk: boundary factors and the beta exponents that go with it
        float2 scb, scb2;
        {
            float fb[2], fb2[2], sa[2], sb[2], sb2[2], ini[2];
            const int tend = t0 + NF;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int el = __shfl_up_sync(FULL, ea[h], 1), er = __shfl_down_sync(FULL, ea[h], 1);
                sa[h] = lane == 0 ? 0.f : pow2i(el - ea[h]);
                const int dx = max(ea[h] - er, -252);                    // 2^(e_beta[l+1] - e_beta[l]), in two factors
                sb[h] = lane == 31 ? 0.f : pow2i(dx >> 1);
                sb2[h] = pow2i(dx - (dx >> 1));
                const int ebn = koff[h] - ea[h];
                const bool starts = act[h] && tend == Tb[h];             // beta_T enters here
                const int df = min(max(eb[h] - ebn, -252), 252);
                fb[h] = starts ? 0.f : pow2i(df >> 1);
                fb2[h] = pow2i(df - (df >> 1));
                ini[h] = starts ? pow2i(-ebn) : 0.f;
                if (starts) {                            // beta_T = 1 must be representable in the lane that owns S-1
                    const bool owner = (S[h] - 1) / NSL == lane;
                    if (__any_sync(FULL, owner && (ebn > 120 || ebn < -120)) && !flag[h]) flag[h] = 6;
                }
                eb[h] = ebn;
            }
            sca = make_float2(sa[0], sa[1]);
            scb = make_float2(sb[0], sb[1]);
            scb2 = make_float2(sb2[0], sb2[1]);
            const float2 f2 = make_float2(fb[0], fb[1]), f22 = make_float2(fb2[0], fb2[1]);
#pragma unroll
            for (int jj = 0; jj < NSL; ++jj) bt[jj] = __fmul2_rn(__fmul2_rn(bt[jj], f2), f22);
            if ((act[0] && tend == Tb[0]) || (act[1] && tend == Tb[1])) {   // once per utterance: beta_T enters
#pragma unroll
                for (int jj = 0; jj < NSL; ++jj) {
                    const int s = lane * NSL + jj;
                    if (s == S[0] - 1) bt[jj].x += ini[0];
                    if (s == S[1] - 1) bt[jj].y += ini[1];
                }
            }
        }
        // recompute alpha over the chunk, keeping the label states
        float2 A[REGBUF ? NF : 1][NL];
#pragma unroll
        for (int r = 0; r < NF; ++r) {
            alpha_step<NSL>(a, sca, eo + r * RW, bl0, bl1, ep0, ep1, ma);
#pragma unroll
            for (int q = 0; q < NL; ++q) {
                if constexpr (REGBUF) A[r][q] = a[2 * q + 1];
                else reinterpret_cast<float2*>(abuf + r * ARS)[q * 32 + lane] = a[2 * q + 1];
            }
        }
        issue_rows(j - 2);                              // into the slot the gradient rows of chunk j + 1 have left
        __syncwarp();                                   // the previous chunk's gradient rows have been read
        // beta over the chunk; label products alpha_t(s) * u_t(s) into the class cells
#pragma unroll
        for (int r = NF - 1; r >= 0; --r) {
            float2 ul[NL], al[NL];
#pragma unroll
            for (int q = 0; q < NL; ++q) {
                if constexpr (REGBUF) al[q] = A[r][q];
                else al[q] = reinterpret_cast<const float2*>(abuf + r * ARS)[q * 32 + lane];
            }
            beta_step<NSL>(bt, ul, scb, scb2, eo + r * RW, bl0, bl1, ep0, ep1, mb);
#pragma unroll
            for (int q = 0; q < NL; ++q) {
                const float2 fx = __ffma2_rn(__fmul2_rn(al[q], ul[q]), zs2, magic2);
                atomicAdd(reinterpret_cast<unsigned*>(const_cast<float*>(ep0[q])) + (OCCD + r * RW), __float_as_uint(fx.x));
                atomicAdd(reinterpret_cast<unsigned*>(const_cast<float*>(ep1[q])) + (OCCD + r * RW), __float_as_uint(fx.y));
            }
        }
        __syncwarp();
        // gradient rows of the chunk: (softmax prob - occupancy / Z) * g, assembled in class order in the staging slot
        // the chunk's log-probs arrived in; the cells are zeroed for the next chunk
        {
            const float g0 = kill0 ? 0.f : g[0], g1 = kill1 ? 0.f : g[1];
            const float og0 = g0 * (1.f / FIX), og1 = g1 * (1.f / FIX);
            float* gs = raw + (j % NSLOT) * (KF * RAWW);
            constexpr int fixi = (int)FIX;
#ifndef MDD_PAIR_DIRECT_GRAD
#define MDD_PAIR_DIRECT_GRAD 1
#endif
            // Where my four values of a row go.  Direct: straight to the gradient array, four 4-byte streaming stores per
            // row (every float of the pair's row is written by exactly one lane; the sectors fill up in L2) -- the
            // shared-memory pipe is this kernel's bound, and staging the row for wide stores costs it 9 wavefronts a frame.
            constexpr bool DIRECT = MDD_PAIR_DIRECT_GRAD != 0;
            char* dq[4];
            if constexpr (DIRECT) {
                const unsigned long long go = (unsigned long long)(unsigned)t0 * gstride_b;
                dq[0] = reinterpret_cast<char*>(gr0 + oc[0][0]) + go;
                dq[1] = reinterpret_cast<char*>(gr0 + oc[0][1]) + go;
                dq[2] = reinterpret_cast<char*>(gr1 + oc[1][0]) + go;
                dq[3] = reinterpret_cast<char*>(gr1 + oc[1][1]) + go;
            }
            const bool sv0 = act[0] && ov[0][0], sv1 = act[0] && ov[0][1], sv2 = act[1] && ov[1][0], sv3 = act[1] && ov[1][1];
            // one row ahead: cells and their warp totals (REDUX) of row r + 1 are on their way while row r is finished
            int oc4[4], on4[4] = {0, 0, 0, 0}, ts0, ts1, tn0 = 0, tn1 = 0;
            auto cells = [&](int r, int (&o)[4], int& s0, int& s1) {
                unsigned* oq = occ + r * RW;
                o[0] = (int)(oq[lane] - nbias[0]);
                o[1] = (int)(oq[32 + lane] - nbias[1]);
                o[2] = (int)(oq[CPMAX + lane] - nbias[2]);
                o[3] = (int)(oq[CPMAX + 32 + lane] - nbias[3]);
                oq[lane] = 0u; oq[32 + lane] = 0u; oq[CPMAX + lane] = 0u; oq[CPMAX + 32 + lane] = 0u;
                s0 = __reduce_add_sync(FULL, o[0] + o[1]);       // blank: what the label classes leave of Z
                s1 = __reduce_add_sync(FULL, o[2] + o[3]);
            };
            cells(0, oc4, ts0, ts1);
#pragma unroll
            for (int r = 0; r < NF; ++r) {
                if (r + 1 < NF) cells(r + 1, on4, tn0, tn1);
                const float* d = rows + eo + r * RW;
                const float2 pv0 = make_float2(d[lane], d[32 + lane]), pv1 = make_float2(d[CPMAX + lane], d[CPMAX + 32 + lane]);
                if (isb00) oc4[0] = fixi - ts0;
                if (isb01) oc4[1] = fixi - ts0;
                if (isb10) oc4[2] = fixi - ts1;
                if (isb11) oc4[3] = fixi - ts1;
                // every row total must be a fraction of Z in [0, 1] up to rounding: catches overflow / NaN upstream
                occ_hi = max(occ_hi, max((unsigned)(ts0 + 8192), (unsigned)(ts1 + 8192)));
                const float v0 = fmaf(-(float)oc4[0], og0, pv0.x * g0), v1 = fmaf(-(float)oc4[1], og0, pv0.y * g0);
                const float v2 = fmaf(-(float)oc4[2], og1, pv1.x * g1), v3 = fmaf(-(float)oc4[3], og1, pv1.y * g1);
                if constexpr (DIRECT) {
                    if (sv0) __stcs(reinterpret_cast<float*>(dq[0]), v0);
                    if (sv1) __stcs(reinterpret_cast<float*>(dq[1]), v1);
                    if (sv2) __stcs(reinterpret_cast<float*>(dq[2]), v2);
                    if (sv3) __stcs(reinterpret_cast<float*>(dq[3]), v3);
#pragma unroll
                    for (int k = 0; k < 4; ++k) dq[k] += gstride_b;
                } else {
                    float* w = gs + r * RAWW;
                    w[wg[0]] = v0;
                    w[wg[1]] = v1;
                    w[wg[2]] = v2;
                    w[wg[3]] = v3;
                }
#pragma unroll
                for (int k = 0; k < 4; ++k) oc4[k] = on4[k];
                ts0 = tn0; ts1 = tn1;
            }
            __syncwarp();
            if constexpr (DIRECT) {
            } else if (gvec) {                                 // the pair's two rows are one 16-byte aligned run of 2*C floats
                char* gq = reinterpret_cast<char*>(gr0) + (unsigned long long)(unsigned)t0 * gstride_b + lane * 16;
#pragma unroll
                for (int r = 0; r < NF; ++r) {
                    if (lane < C / 2) __stcs(reinterpret_cast<float4*>(gq), reinterpret_cast<const float4*>(gs + r * RAWW)[lane]);
                    gq += gstride_b;
                }
            } else {
#pragma unroll
                for (int r = 0; r < NF; ++r) {
                    const size_t go = (size_t)(t0 + r) * gstride;
                    for (int i = lane; i < C; i += 32) {
                        if (act[0]) __stcs(gr0 + go + i, gs[r * RAWW + i]);
                        if (act[1]) __stcs(gr1 + go + i, gs[r * RAWW + C + i]);
                    }
                }
            }
        }
    };

    // in flight when chunk j starts, oldest first: rows(j), checkpoint(j), rows(j - 1); in the single kernel the last
    // chunk's rows are still staged from the forward sweep
    if constexpr (SEG) {
        if (jhi < nch) {                                // beta entering chunk jhi - 1 from above, with the beta sweep's own exponents
            const float* c = p.ckb + ((size_t)slot * n_ckpt + (jhi - 1)) * CKW;
#pragma unroll
            for (int q = 0; q < NSL; ++q) bt[q] = reinterpret_cast<const float2*>(c)[q * 32 + lane];
            const int2 e = reinterpret_cast<const int2*>(c + 64 * NSL)[lane];
            eb[0] = e.x; eb[1] = e.y;
        }
    }
    if constexpr (PHASE == 2) issue_rows(jhi - 1);
    ckpt_issue(jhi - 1);
    issue_rows(jhi - 2);
    for (int j = jhi - 1; j >= jlo; --j) {
        int t0;
        if (chunk_geom(j, t0)) bwd_chunk(IC<CH>{}, j, t0);
        else bwd_chunk(IC<1>{}, j, t0);
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
    {   // beta must have stayed finite to the very first frame (a blank state's overflow shows nowhere else)
        constexpr unsigned OMAX = (unsigned)(8388608.0 * 1.002) + 8192u;   // a row total above Z (+ rounding): upstream overflow
        int bad_occ = occ_hi > OMAX ? 3 : 0;            // which of the two is decided below by the likelihood check
#pragma unroll
        for (int jj = 0; jj < NSL; ++jj) {
            if (!(bt[jj].x < 3.0e38f)) bad_occ |= 1;
            if (!(bt[jj].y < 3.0e38f)) bad_occ |= 2;
        }
        bad_occ = __reduce_or_sync(FULL, bad_occ);
        if ((bad_occ & 1) && !flag[0]) flag[0] = 5;
        if ((bad_occ & 2) && !flag[1]) flag[1] = 5;
        // The backward sweep's own total likelihood, beta_0(0) + beta_0(1), against the forward P.  Mass dropped by
        // either sweep (alpha content squashed in a renormalisation, beta content below its tied exponent) is missing
        // from exactly one of the two, so their ratio bounds what the occupancies can be off by.
        if (!SEG || jlo == 0) {
            const float r0 = (bt[0].x + bt[1].x) * pow2i(-ea[0]) * zi[0];
            const float r1 = (bt[0].y + bt[1].y) * pow2i(-ea[1]) * zi[1];
            const int disagree = (!(fabsf(r0 - 1.f) <= p.pbeta_tol) ? 1 : 0) | (!(fabsf(r1 - 1.f) <= p.pbeta_tol) ? 2 : 0);
#ifdef MDD_PAIR_DEBUG
            if (lane == 0) printf("PB b=%d r=%.7f flag=%d | b=%d r=%.7f flag=%d\n", bidx[0], r0, flag[0], bidx[1], r1, flag[1]);
#endif
            const int d0 = __shfl_sync(FULL, disagree, 0);
            if ((d0 & 1) && act[0] && !flag[0]) flag[0] = 7;
            if ((d0 & 2) && act[1] && !flag[1]) flag[1] = 7;
        }
#pragma unroll
        for (int h = 0; h < 2; ++h)                          // 0 = done here; else redone by the next tier.  Segments share the
            if (act[h] && lane == 0 && (!SEG || flag[h])) p.flags[bidx[h]] = flag[h];   // word: each only ever raises it
    }
}

template <int NSL, bool REGBUF, int MAXR, bool LOGITS, int CH, bool LIST, int PHASE = 0, bool SEG = false>
__global__ void __maxnreg__(MAXR)
ctc_pair_kernel(CtcParams p, float* __restrict__ ckpt_ws, int n_ckpt) {
    const int w = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if constexpr (!LIST && !SEG && PHASE != 2) {                 // first kernel of the call (see ctc_exact.cu)
        if (p.ticket && blockIdx.x == 0 && threadIdx.x == 0) *p.ticket = 0u;
    }
    if constexpr (SEG) {
        const int pr = w / p.nseg;                               // segments of a pair sit in neighbouring CTAs
        if (pr < ((p.B + 1) >> 1)) ctc_pair_body<NSL, REGBUF, LOGITS, CH, false, PHASE, true>(p, ckpt_ws, n_ckpt, pr, pr, p.B, w - pr * p.nseg);
        pdl_launch_dependents();                                 // (at the end: the tier kernels must not take this grid's SMs)
    } else if constexpr (!LIST) {
        if (w < ((p.B + 1) >> 1)) ctc_pair_body<NSL, REGBUF, LOGITS, CH, false, PHASE>(p, ckpt_ws, n_ckpt, w, w, p.B);
        pdl_launch_dependents();
    } else {
        pdl_wait();                                              // (the fp64 sweep behind this grid is released when it ends)
        // second tier: this warp owns checkpoint block w and walks the pairs w, w + #warps, ... of the flagged list
        const int n_utt = *p.list_count, stride = gridDim.x * (blockDim.x >> 5);
        if (n_utt <= p.list_team_max) return;                    // few flagged utterances: the team kernel has taken them
        for (int pr = w; 2 * pr < n_utt; pr += stride) {
            ctc_pair_body<NSL, REGBUF, LOGITS, CH, true>(p, ckpt_ws, n_ckpt, pr, w, n_utt);
            asm volatile("cp.async.wait_all;" ::: "memory");
            __syncwarp();
        }
    }
}

constexpr int kDualMaxPairs = 4 * kNumSM;                        // the concurrent-sweep path serves up to this many pairs (four CTAs per SM)
// floats from the start of the pair workspace to the beta checkpoints: alpha checkpoints, one record per utterance
size_t dual_beta_offset(size_t ck_floats, int B) { return (ck_floats + ((size_t)B + 2) * 4 + 63) & ~(size_t)63; }

int pick_nsl(int Lmax) {
    if (Lmax <= PGeo<2>::LCAP) return 2;
    if (Lmax <= PGeo<4>::LCAP) return 4;
    if (Lmax <= PGeo<8>::LCAP) return 8;
    if (Lmax <= PGeo<16>::LCAP) return 16;
    return 0;
}

int env_int(const char* name, int dflt) {
    const char* e = getenv(name);
    return e ? atoi(e) : dflt;
}

// ordered compaction of the flagged utterances: list[0..count) = { b : flags[b] != 0 }, ascending (one CTA)
__global__ void __launch_bounds__(1024) compact_flagged_kernel(const int32_t* __restrict__ flags, int B,
                                                               int32_t* __restrict__ list, int32_t* __restrict__ count) {
    __shared__ int wsum[32];
    __shared__ int base;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    pdl_launch_dependents();
    pdl_wait();
    if (tid == 0) base = 0;
    __syncthreads();
    for (int b0 = 0; b0 < B; b0 += 1024) {
        const int b = b0 + tid;
        const bool f = b < B && flags[b] != 0;
        const unsigned m = __ballot_sync(0xffffffffu, f);
        if (lane == 0) wsum[warp] = __popc(m);
        __syncthreads();
        int off = base;
        for (int w = 0; w < warp; ++w) off += wsum[w];
        if (f) list[off + __popc(m & ((1u << lane) - 1))] = b;
        __syncthreads();
        if (tid == 0) {
            int t = 0;
            for (int w = 0; w < 32; ++w) t += wsum[w];
            base += t;
        }
        __syncthreads();
    }
    if (tid == 0) *count = base;
}

template <int NSL, bool REGBUF, int MAXR, bool LOGITS, int CH, bool LIST, int PHASE = 0, bool SEG = false>
int launch_cfg1(const CtcParams& p, float* ws, cudaStream_t st, int npairs) {
    const PairSmem lay = pair_smem_layout(NSL, REGBUF, PHASE, SEG);
    size_t smem = (size_t)lay.total_floats * 4;                  // one warp (= one pair) per CTA
    // Warps per SM: the register file allows 8; fewer when that packs the pairs into full waves (a pair runs for the
    // whole kernel, so 13.8 pairs per SM are better served by 7 warps x 2 waves than by 8 + 5.8).
    static int wps_env = env_int("MDD_CTC_PAIR_WPS", 0);
    int wps = wps_env;
    if (wps <= 0) {
        double best = 0.0;
        int wmax = 65536 / (32 * ((MAXR + 7) & ~7));             // register file
        wmax = std::min(wmax, (int)((227 * 1024) / (smem + 1024)));   // shared memory
        for (int w = wmax; w >= std::max(1, wmax - 2); --w) {
            const long long slots = (long long)kNumSM * w;
            const double eff = (double)npairs / (double)(((npairs + slots - 1) / slots) * slots);
            if (eff > best + 0.04) { best = eff; wps = w; }
        }
    }
    if (wps > 0 && wps < 32) {
        const size_t want = ((233472 / (size_t)wps) - 1024 - 64) & ~(size_t)15;
        if (want > smem && want <= 227 * 1024) smem = want;
    }
    if (cudaFuncSetAttribute(ctc_pair_kernel<NSL, REGBUF, MAXR, LOGITS, CH, LIST, PHASE, SEG>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)smem) != cudaSuccess) {                // per device: set on every call (cheap)
        cudaGetLastError();
        return MDD_E_CUDA;
    }
    const int n_ckpt = p.T / CH + 2 * CH;
    if constexpr (PHASE == 2 || LIST) {                          // behind another kernel of this call: dependent launch
        if (launch_dependent(ctc_pair_kernel<NSL, REGBUF, MAXR, LOGITS, CH, LIST, PHASE, SEG>, dim3(npairs), dim3(32), smem, st, p, ws,
                             n_ckpt) != cudaSuccess) {
            cudaGetLastError();
            return MDD_E_CUDA;
        }
    } else {
        ctc_pair_kernel<NSL, REGBUF, MAXR, LOGITS, CH, LIST, PHASE, SEG><<<npairs, 32, smem, st>>>(p, ws, n_ckpt);
    }
    MDD_CHECK_LAUNCH();
    return MDD_OK;
}

constexpr int CH2 = 2;                                           // chunk length of the second tier

template <int NSL, bool REGBUF, int MAXR, bool LOGITS>
int launch_tiers(CtcParams p, float* ws, cudaStream_t st) {
    const int npairs = (p.B + 1) / 2;
    // Few pairs per SM: one CTA per pair with helper warps (ctc_team.cu) -- nothing hides a lone warp's latencies here.
    static int team_max = env_int("MDD_CTC_TEAM_MAX_PAIRS", 2 * kNumSM);
    const bool team_ok = ctc_team_supported(p.C, p.Lmax);
    int rc;
    static int split = env_int("MDD_CTC_SPLIT", 1);
    const size_t ck_floats = (size_t)npairs * (p.T / KF + 2 * KF) * PGeo<NSL>::CKW;
    p.rec = reinterpret_cast<float4*>(ws + ck_floats);
    // Fewer still (and a gradient wanted): alpha and beta swept concurrently by ctc_dual.cu, then the backward kernel on
    // independent segments of every pair -- the dependent chain is T lattice steps instead of 3 T.
    // (measured at T = 1000: 0.37 / 0.43 / 0.47 ms at 320 / 384 / 448 pairs against 0.47 / 0.48 / 0.50 for the kernels
    // below; level at 512 pairs, slower beyond)
    static int dual_max = std::min(env_int("MDD_CTC_DUAL_MAX_PAIRS", 3 * kNumSM), kDualMaxPairs);
    static int nseg_env = env_int("MDD_CTC_DUAL_NSEG", 0);
    if (team_ok && p.grad && npairs <= dual_max) {
        float* ckb = ws + dual_beta_offset(ck_floats, p.B);
        p.ckb = ckb;
        p.cpos_ws = reinterpret_cast<int*>(ckb + ck_floats);
        p.erow = ckb + ck_floats + (((size_t)npairs * 144 + 63) & ~(size_t)63);
        rc = launch_ctc_dual(p, ws, ckb, st, npairs);
        if (!rc) {
            const int want = nseg_env > 0 ? nseg_env : (7 * sm_count()) / npairs;     // one wave of 7 warps per SM
            p.nseg = std::max(1, std::min(want, std::max(1, p.T / KF)));
            rc = launch_cfg1<NSL, REGBUF, MAXR, LOGITS, KF, false, 2, true>(p, ws, st, npairs * p.nseg);
        }
    } else if (team_ok && npairs <= team_max) rc = launch_ctc_team(p, ws, st, false, npairs);
    else if (split || !p.grad) {
        // forward sweep at high occupancy (a third of the registers), then the backward sweep from its checkpoints
        p.cpos_ws = reinterpret_cast<int*>(ws + dual_beta_offset(ck_floats, p.B));
        rc = launch_cfg1<NSL, REGBUF, 128, LOGITS, KF, false, 1>(p, ws, st, npairs);
        if (!rc && p.grad) rc = launch_cfg1<NSL, REGBUF, MAXR, LOGITS, KF, false, 2>(p, ws, st, npairs);
    } else rc = launch_cfg1<NSL, REGBUF, MAXR, LOGITS, KF, false>(p, ws, st, npairs);
    if (rc || !p.list) return rc;
    // second tier: the flagged utterances again, renormalised every CH2 frames, in the same checkpoint workspace
    const size_t ws_floats = (size_t)npairs * (p.T / KF + 2 * KF) * PGeo<NSL>::CKW;
    const size_t per_pair = (size_t)(p.T / CH2 + 2 * CH2) * PGeo<NSL>::CKW;
    const int cap = (int)std::min<size_t>(ws_floats / per_pair, (size_t)npairs);
    if (cap <= 0) return MDD_OK;
    if (launch_dependent(compact_flagged_kernel, dim3(1), dim3(1024), 0, st, (const int32_t*)p.flags, p.B, p.list, p.list_count) != cudaSuccess) {
        cudaGetLastError();
        return MDD_E_CUDA;
    }
    MDD_CHECK_LAUNCH();
    p.list_cap_pairs = cap;
    // Few flagged utterances: one CTA per pair with helper warps (each pair alone on its SM would crawl in the pair
    // kernel); many: the pair kernel's own occupancy wins.  The count is on the device, so both are launched and each
    // leaves at once when the other is in charge.
    static int team_tier2 = env_int("MDD_CTC_TEAM_TIER2", 1);
    p.list_team_max = (team_ok && team_tier2) ? 2 * std::min(cap, 2 * sm_count()) : 0;
    if (p.list_team_max > 0) {
        rc = launch_ctc_team(p, ws, st, true, std::min(cap, 2 * sm_count()));
        if (rc) return rc;
        if (p.B <= p.list_team_max) return MDD_OK;               // the whole batch fits the team kernel's share: nothing can be left
    }
    return launch_cfg1<NSL, REGBUF, MAXR, LOGITS, CH2, true>(p, ws, st, cap);
}

template <int NSL, bool REGBUF, int MAXR>
int launch_cfg(const CtcParams& p, float* ws, cudaStream_t st) {
    return launch_tiers<NSL, REGBUF, MAXR, MDD_PAIR_TU_LOGITS != 0>(p, ws, st);
}

template <int NSL>
int launch_nsl(const CtcParams& p, float* ws, cudaStream_t st) {
    if constexpr (NSL == 8) {
#ifdef MDD_PAIR_EXPERIMENTS                                      // A/B builds only (scratch/variants.sh): alpha label states
        static int regbuf = env_int("MDD_CTC_PAIR_REGBUF", 1), maxr = env_int("MDD_CTC_PAIR_MAXR", 255);   // in shared memory,
        if (!regbuf) {                                           // optionally with a register cap
            if (maxr >= 255) return launch_cfg<NSL, false, 255>(p, ws, st);
            return launch_cfg<NSL, false, 168>(p, ws, st);
        }
#endif
#ifndef MDD_PAIR_MAXR
#define MDD_PAIR_MAXR 255
#endif
        return launch_cfg<NSL, true, MDD_PAIR_MAXR>(p, ws, st);
    }
    if constexpr (NSL < 8) return launch_cfg<NSL, true, 255>(p, ws, st);
    return launch_cfg<NSL, false, 255>(p, ws, st);               // NSL = 16: the alpha buffer does not fit in registers
}

}  // namespace

#if !MDD_PAIR_TU_LOGITS
// C classes + the zero item that unused label slots read + one free slot (where the lane that owns the zero item's
// position parks its own store) must fit the 64 positions of an utterance's row
bool ctc_pair_supported(int C, int Lmax) { return C <= 62 && pick_nsl(Lmax) != 0; }

size_t ctc_pair_ws_bytes(int T, int B, int Lmax) {
    const int nsl = pick_nsl(Lmax);
    if (!nsl) return 0;
    const size_t n_ckpt = (size_t)T / KF + 2 * KF;
    const size_t npairs = (size_t)((B + 1) / 2), ck_floats = npairs * n_ckpt * 32 * (2 * nsl + 2);
    if ((int)npairs <= kDualMaxPairs)                            // + beta checkpoints and the emission-row layouts (ctc_dual.cu)
        return (dual_beta_offset(ck_floats, B) + ck_floats + npairs * 144 + 128 + npairs * (size_t)T * RW) * 4;   // + emission rows
    return (dual_beta_offset(ck_floats, B) + npairs * 144 + 64) * 4;   // + one record per utterance, one row layout per pair
}
#endif

// the instantiations of this translation unit (log-prob input here, logit input in ctc_pair_logits.cu)
#if MDD_PAIR_TU_LOGITS
int launch_ctc_pair_tu_logits(const CtcParams& p, float* w, cudaStream_t st) {
#else
int launch_ctc_pair_tu_logits(const CtcParams& p, float* w, cudaStream_t st);
static int launch_ctc_pair_tu(const CtcParams& p, float* w, cudaStream_t st) {
#endif
    switch (pick_nsl(p.Lmax)) {
#ifndef MDD_PAIR_ONLY8
        case 2: return launch_nsl<2>(p, w, st);
        case 4: return launch_nsl<4>(p, w, st);
        case 16: return launch_nsl<16>(p, w, st);
#endif
        case 8: return launch_nsl<8>(p, w, st);
        default: return MDD_E_UNSUPPORTED;
    }
}

#if !MDD_PAIR_TU_LOGITS
int launch_ctc_pair(CtcParams p, void* ws, cudaStream_t st) {
    float* w = reinterpret_cast<float*>(ws);
    static const char* tol_env = getenv("MDD_CTC_PBETA_TOL");
    static int strict = env_int("MDD_CTC_STRICT_SQUASH", 0);
    p.pbeta_tol = tol_env ? (float)atof(tol_env) : 3e-5f;
    p.strict_squash = strict;
    static int decay = env_int("MDD_CTC_DECAY_MAX", 76 + SHIFT);
    p.decay_max = decay;
    return p.lse ? launch_ctc_pair_tu_logits(p, w, st) : launch_ctc_pair_tu(p, w, st);
}
#endif

}  // namespace mdd
