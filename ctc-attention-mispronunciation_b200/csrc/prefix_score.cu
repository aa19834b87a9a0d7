// prefix_score.cu -- CTC prefix scorer + beam selection for CTC/attention joint beam decoding.
//
// Named by BASELINE.json's north_star; the reference has no such function (its nearest code is the
// (prBlank, prNonBlank, prTotal) recursion of egs/attention_aug/utils/BeamSearch.py:52-66,99-125).  Formulation:
// Watanabe et al. 2017 / ESPnet CTCPrefixScore; parity anchor = the fp64 CPU restatement used by the tests and the
// psi(l.eos) == -nll invariant against torch CPU ctc_loss.
//
// Arithmetic is fp64 log space: the scores are log-likelihoods of magnitude |nll| (thousands of nats at T = 500), and
// the beam keeps the W best of W*C candidates by DIFFERENCES of such numbers; in fp32 (ulp 1.2e-4 at 1024) near-ties
// at the beam boundary flip against an fp64 restatement a few times per thousand decisions, which changes decoded
// tokens.  In fp64 the selection is the restatement's.
//
//  * prefix_score_kernel<false>: one CTA per (utterance, hypothesis); thread c scans t once for candidate label c
//    with (r_n, r_b, psi) in registers.  The parent state r_prev[b,w,:,:] and the rows x[t,b,:] are shared by all
//    candidates and are served from L1/L2.
//  * prefix_select_kernel: one CTA per utterance; forms score(g) + w_ctc (psi(g.c) - psi(g)) + (1 - w_ctc) att[c] for
//    all W*C candidates (eos scored with psi_eos, finished hypotheses carried unchanged), takes the W best by W rounds
//    of block arg-max (ties: smaller w*C + c first) and writes the new beam: parents, labels, scores, psi, lengths,
//    finished flags and the token histories.  No host round trip, no torch.topk / gather between the steps.
//  * prefix_score_kernel<true>: recomputes and stores the state of the W chosen (parent, label) pairs, so state stays
//    O(W*T) per utterance.
#include "common.cuh"

namespace mdd {
namespace {

#define NEGD (-CUDART_INF)

__device__ __forceinline__ double lse2d(double a, double b) {
    const double m = fmax(a, b);
    if (m == NEGD) return NEGD;
    return m + log(exp(a - m) + exp(b - m));
}

__global__ void prefix_init_kernel(const float* __restrict__ x, int T, int B, int C, const int64_t* __restrict__ in_len,
                                   int blank, int W, double* __restrict__ r0) {
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= B) return;
    const int b = warp;
    const int Tb = in_len ? (int)max(0ll, min((long long)in_len[b], (long long)T)) : T;
    double run = 0.0;
    for (int base = 0; base < T; base += 32) {
        const int t = base + lane;
        double v = (t < Tb) ? (double)x[((size_t)t * B + b) * C + blank] : 0.0;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const double o = __shfl_up_sync(0xffffffffu, v, d);
            if (lane >= d) v += o;
        }
        v += run;
        run = __shfl_sync(0xffffffffu, v, 31);
        if (t < T) {
            const double rb = (t < Tb) ? v : NEGD;
            for (int w = 0; w < W; ++w) {
                double2* dst = reinterpret_cast<double2*>(r0 + (((size_t)b * W + w) * T + t) * 2);
                *dst = make_double2(NEGD, rb);
            }
        }
    }
}

// ADVANCE = false: thread = candidate label of hypothesis (b, w); psi out.
// ADVANCE = true : thread = new hypothesis slot (b, k) with parent sel_hyp and label sel_label; r_new out.
template <bool ADVANCE>
__global__ void prefix_score_kernel(const float* __restrict__ x, int T, int B, int C, const int64_t* __restrict__ in_len,
                                    int blank, int W, const double* __restrict__ r_prev, const int32_t* __restrict__ last,
                                    const int32_t* __restrict__ glen, const int32_t* __restrict__ skip,
                                    const int32_t* __restrict__ sel_hyp, const int32_t* __restrict__ sel_label,
                                    double* __restrict__ out, double* __restrict__ psi_eos) {
    int bw, c, src_w;
    if (ADVANCE) {
        bw = blockIdx.x * blockDim.x + threadIdx.x;
        if (bw >= B * W) return;
        src_w = sel_hyp[bw];
        c = sel_label[bw];
    } else {
        bw = blockIdx.x;                     // b * W + w
        c = threadIdx.x;
        if (c >= C) return;
        src_w = bw % W;
    }
    const int b = bw / W;
    const int Tb = in_len ? (int)max(0ll, min((long long)in_len[b], (long long)T)) : T;
    if (ADVANCE) {
        // a slot that is dead (no parent), finished (skip flag of the NEW slot) or mislabelled keeps no CTC state
        const bool valid = src_w >= 0 && src_w < W && c >= 0 && c < C && c != blank && !(skip && skip[bw]);
        if (!valid) return;
    } else if (skip && skip[bw]) {           // finished / dead hypotheses are not extended
        out[(size_t)bw * C + c] = NEGD;
        if (c == blank && psi_eos) psi_eos[bw] = NEGD;
        return;
    }
    double2* rout = ADVANCE ? reinterpret_cast<double2*>(out + (size_t)bw * T * 2) : nullptr;
    const double2* rp = reinterpret_cast<const double2*>(r_prev + ((size_t)b * W + src_w) * T * 2);
    const int lastl = last[b * W + src_w], gl = glen[b * W + src_w];
    const int start = gl > 1 ? gl : 1;
    if (!ADVANCE && c == blank) {
        out[(size_t)bw * C + c] = NEGD;
        if (psi_eos) {
            double e = NEGD;
            if (Tb > 0) { const double2 r = rp[Tb - 1]; e = lse2d(r.x, r.y); }
            psi_eos[bw] = e;
        }
        return;
    }
    double rn = NEGD, rb = NEGD;
    if (start - 1 < Tb && gl == 0) rn = (double)x[(size_t)b * C + c];
    if (ADVANCE) {
        for (int t = 0; t < start - 1 && t < T; ++t) rout[t] = make_double2(NEGD, NEGD);
        if (start - 1 < T) rout[start - 1] = (start - 1 < Tb) ? make_double2(rn, rb) : make_double2(NEGD, NEGD);
    }
    double ps = rn;
    for (int t = start; t < Tb; ++t) {
        const double2 r = rp[t - 1];
        const double rsum = lse2d(r.x, r.y);
        const double phi = (c == lastl) ? r.y : rsum;
        const float* row = x + ((size_t)t * B + b) * C;
        const double xc = (double)row[c], xb = (double)row[blank];
        const double nrn = lse2d(rn, phi) + xc;
        const double nrb = lse2d(rn, rb) + xb;
        ps = lse2d(ps, phi + xc);
        rn = nrn; rb = nrb;
        if (ADVANCE) rout[t] = make_double2(rn, rb);
    }
    if (ADVANCE) {
        for (int t = max(Tb, start); t < T; ++t) rout[t] = make_double2(NEGD, NEGD);
    } else {
        out[(size_t)bw * C + c] = ps;
    }
}

// One CTA per utterance.  Candidate (w, c) -> value; the W best (ties: smaller w*C + c) become the new beam.
__global__ void __launch_bounds__(256) prefix_select_kernel(
    int B, int W, int C, int blank, int eos, double ctc_weight, int max_len,
    const double* __restrict__ psi, const double* __restrict__ psi_eos, const float* __restrict__ att,
    const double* __restrict__ score, const double* __restrict__ psi_prev, const int32_t* __restrict__ glen,
    const int32_t* __restrict__ finished, const int32_t* __restrict__ tokens,
    int32_t* __restrict__ sel_hyp, int32_t* __restrict__ sel_label, double* __restrict__ score_new,
    double* __restrict__ psi_new, int32_t* __restrict__ glen_new, int32_t* __restrict__ finished_new,
    int32_t* __restrict__ tokens_new) {
    extern __shared__ __align__(16) unsigned char sm_raw[];
    double* cand = reinterpret_cast<double*>(sm_raw);            // [W*C]
    __shared__ double s_val[8];
    __shared__ int s_idx[8];
    __shared__ int s_pick;
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nw = blockDim.x >> 5;
    const int n = W * C;
    for (int i = tid; i < n; i += blockDim.x) {
        const int w = i / C, c = i - w * C;
        const int bw = b * W + w;
        const double sc = score[bw];
        double v = NEGD;
        if (finished[bw]) {
            if (c == (eos >= 0 ? eos : blank)) v = sc;           // a finished hypothesis is carried unchanged
        } else if (sc > NEGD && c != blank) {
            const double pp = psi_prev[bw];
            const double pc = (c == eos) ? psi_eos[bw] : psi[(size_t)bw * C + c];
            if (pc > NEGD && pp > NEGD) v = sc + ctc_weight * (pc - pp) + (1.0 - ctc_weight) * (double)att[(size_t)bw * C + c];
            if (!(v == v)) v = NEGD;                             // NaN never enters the beam
        }
        cand[i] = v;
    }
    __syncthreads();
    for (int k = 0; k < W; ++k) {
        double bv = NEGD;
        int bi = 0x7fffffff;
        for (int i = tid; i < n; i += blockDim.x) {
            const double v = cand[i];
            if (v > bv || (v == bv && i < bi && v > NEGD)) { bv = v; bi = i; }
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            const double ov = __shfl_xor_sync(0xffffffffu, bv, d);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, d);
            if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
        }
        if (lane == 0) { s_val[warp] = bv; s_idx[warp] = bi; }
        __syncthreads();
        if (tid == 0) {
            for (int w2 = 1; w2 < nw; ++w2)
                if (s_val[w2] > bv || (s_val[w2] == bv && s_idx[w2] < bi)) { bv = s_val[w2]; bi = s_idx[w2]; }
            const int slot = b * W + k;
            if (bv > NEGD) {
                const int w = bi / C, c = bi - w * C;
                const int pw = b * W + w;
                const int pf = finished[pw] != 0;
                const int fin = pf || (c == eos);
                sel_hyp[slot] = w;
                sel_label[slot] = c;
                score_new[slot] = bv;
                psi_new[slot] = pf ? psi_prev[pw] : (c == eos ? psi_eos[pw] : psi[(size_t)pw * C + c]);
                glen_new[slot] = glen[pw] + (pf ? 0 : 1);
                finished_new[slot] = fin;
                cand[bi] = NEGD;
                s_pick = pw;
            } else {                                             // fewer than W live candidates: a dead slot
                sel_hyp[slot] = -1;
                sel_label[slot] = -1;
                score_new[slot] = NEGD;
                psi_new[slot] = NEGD;
                glen_new[slot] = 0;
                finished_new[slot] = 2;                          // 0 live, 1 finished, 2 dead
                s_pick = -1;
            }
        }
        __syncthreads();
        {   // token history of the new slot: the parent's, plus the new label
            const int pw = s_pick, slot = b * W + k;
            int32_t* dst = tokens_new + (size_t)slot * max_len;
            if (pw >= 0) {
                const int32_t* src = tokens + (size_t)pw * max_len;
                const int gl = glen[pw], pf = finished[pw] != 0;
                const int lbl = sel_label[slot];
                for (int i = tid; i < max_len; i += blockDim.x) dst[i] = (!pf && i == gl) ? lbl : src[i];
            } else {
                for (int i = tid; i < max_len; i += blockDim.x) dst[i] = -1;
            }
        }
        __syncthreads();
    }
}

}  // namespace
}  // namespace mdd

using namespace mdd;

extern "C" int mdd_prefix_init(const float* x, int T, int B, int C, const int64_t* input_lengths, int blank, int W,
                               double* r0, mdd_stream_t stream) {
    if (!x || !r0 || T <= 0 || B < 0 || C <= 0 || W <= 0 || blank < 0 || blank >= C) return MDD_E_ARG;
    if (B == 0) return MDD_OK;
    prefix_init_kernel<<<ceil_div(B, 4), 128, 0, (cudaStream_t)stream>>>(x, T, B, C, input_lengths, blank, W, r0);
    MDD_CHECK_LAUNCH();
    return MDD_OK;
}

extern "C" int mdd_prefix_score(const float* x, int T, int B, int C, const int64_t* input_lengths, int blank, int W,
                                const double* r_prev, const int32_t* last, const int32_t* glen, const int32_t* skip,
                                double* psi, double* psi_eos, mdd_stream_t stream) {
    if (!x || !r_prev || !last || !glen || !psi || T <= 0 || B < 0 || C <= 0 || C > MDD_MAX_CLASSES || W <= 0 ||
        blank < 0 || blank >= C)
        return MDD_E_ARG;
    if (B == 0) return MDD_OK;
    const int nt = ceil_div(C, 32) * 32;
    prefix_score_kernel<false><<<B * W, nt, 0, (cudaStream_t)stream>>>(x, T, B, C, input_lengths, blank, W, r_prev, last,
                                                                       glen, skip, nullptr, nullptr, psi, psi_eos);
    MDD_CHECK_LAUNCH();
    return MDD_OK;
}

extern "C" int mdd_prefix_advance(const float* x, int T, int B, int C, const int64_t* input_lengths, int blank, int W,
                                  const double* r_prev, const int32_t* last, const int32_t* glen, const int32_t* sel_hyp,
                                  const int32_t* sel_label, const int32_t* skip_new, double* r_new, mdd_stream_t stream) {
    if (!x || !r_prev || !last || !glen || !sel_hyp || !sel_label || !r_new || T <= 0 || B < 0 || C <= 0 || W <= 0 ||
        blank < 0 || blank >= C)
        return MDD_E_ARG;
    if (B == 0) return MDD_OK;
    prefix_score_kernel<true><<<ceil_div(B * W, 32), 32, 0, (cudaStream_t)stream>>>(
        x, T, B, C, input_lengths, blank, W, r_prev, last, glen, skip_new, sel_hyp, sel_label, r_new, nullptr);
    MDD_CHECK_LAUNCH();
    return MDD_OK;
}

extern "C" int mdd_prefix_select(int B, int W, int C, int blank, int eos, double ctc_weight, int max_len,
                                 const double* psi, const double* psi_eos, const float* att, const double* score,
                                 const double* psi_prev, const int32_t* glen, const int32_t* finished,
                                 const int32_t* tokens, int32_t* sel_hyp, int32_t* sel_label, double* score_new,
                                 double* psi_new, int32_t* glen_new, int32_t* finished_new, int32_t* tokens_new,
                                 mdd_stream_t stream) {
    if (!psi || !psi_eos || !att || !score || !psi_prev || !glen || !finished || !tokens || !sel_hyp || !sel_label ||
        !score_new || !psi_new || !glen_new || !finished_new || !tokens_new || B < 0 || W <= 0 || C <= 0 ||
        C > MDD_MAX_CLASSES || blank < 0 || blank >= C || eos >= C || max_len <= 0)
        return MDD_E_ARG;
    if (B == 0) return MDD_OK;
    const size_t smem = (size_t)W * C * sizeof(double);
    if (smem > 48 * 1024) return MDD_E_UNSUPPORTED;
    prefix_select_kernel<<<B, 256, smem, (cudaStream_t)stream>>>(B, W, C, blank, eos, ctc_weight, max_len, psi, psi_eos,
                                                                 att, score, psi_prev, glen, finished, tokens, sel_hyp,
                                                                 sel_label, score_new, psi_new, glen_new, finished_new,
                                                                 tokens_new);
    MDD_CHECK_LAUNCH();
    return MDD_OK;
}
