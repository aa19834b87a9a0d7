// prefix_score.cu -- CTC prefix scorer + beam selection for CTC/attention joint beam decoding.
//
// Named by BASELINE.json's north_star; the reference has no such function (its nearest code is the
// (prBlank, prNonBlank, prTotal) recursion of egs/attention_aug/utils/BeamSearch.py:52-66,99-125).  Formulation:
// Watanabe et al. 2017 / ESPnet CTCPrefixScore; parity anchor = the fp64 CPU restatement used by the tests and the
// psi(l.eos) == -nll invariant against torch CPU ctc_loss.
//
// Arithmetic is fp64: the scores are log-likelihoods of magnitude |nll| (thousands of nats at T = 500), and the beam
// keeps the W best of W*C candidates by DIFFERENCES of such numbers; in fp32 (ulp 1.2e-4 at 1024) near-ties at the
// beam boundary flip against an fp64 restatement a few times per thousand decisions, which changes decoded tokens.
//
// The state r[b,w,t] = (log r_n, log r_b) is kept in log space between steps, but the scans over t run in the LINEAR
// domain relative to the hypothesis' own maximum M = max_t max(r_n, r_b): with phi_c[t] = the mass that may take label
// c at frame t (r_b[t-1] alone when c repeats the last label, r_n[t-1] + r_b[t-1] otherwise, and 1 at t = 0 for the
// empty prefix)
//     psi(g.c)   = M + log sum_t phi_c[t] p_t[c]                     -- a dot product, no recursion
//     r_n'[t]    = (r_n'[t-1] + phi_c[t]) p_t[c],   r_b'[t] = (r_n'[t-1] + r_b'[t-1]) p_t[blank]
// so a step costs 2 exp per (hypothesis, frame) instead of 6 exp + 3 log per (hypothesis, candidate, frame); the
// emissions are exponentiated once per utterance (prefix_probs_kernel).  Entries more than e^-708 below M flush to
// zero (log: -inf); their weight in anything downstream is below fp64 resolution.
//
//  * prefix_probs_kernel : p[b,t,c] = exp(x[t,b,c]) in fp64, utterance-major.
//  * prefix_score_kernel : one CTA per utterance; the warps linearise the hypotheses' states into shared memory, then
//    thread (w, c) takes its dot product over t (p rows coalesced over c, phi broadcast from shared memory).
//  * prefix_select_kernel: one CTA per utterance; forms score(g) + w_ctc (psi(g.c) - psi(g)) + (1 - w_ctc) att[c] for
//    all W*C candidates (eos scored with psi_eos, finished hypotheses carried unchanged), takes the W best by W rounds
//    of block arg-max (ties: smaller w*C + c first) and writes the new beam: parents, labels, scores, psi, lengths,
//    finished flags and the token histories.  No host round trip, no torch.topk / gather between the steps.
//  * prefix_advance_kernel: one warp per new slot; the lanes linearise the parent's state and gather p_t[c], p_t[blank]
//    into shared memory, lane 0 runs the two-term recursion (T dependent FMAs), the lanes take the logs and store.
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

__global__ void prefix_probs_kernel(const float* __restrict__ x, int T, int B, int C, double* __restrict__ p) {
    const size_t n = (size_t)T * B * C;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const int c = (int)(i % C);
        const size_t tb = i / C;
        const int b = (int)(tb % B), t = (int)(tb / B);
        p[((size_t)b * T + t) * C + c] = exp((double)x[i]);
    }
}

// The warp's view of hypothesis (b, w): M and, through fill(t, phis, phib), the linear masses entering frame t.
__device__ __forceinline__ double warp_state_max(const double2* __restrict__ rp, int Tb, int gl, int lane) {
    double m = (gl == 0) ? 0.0 : NEGD;                           // the empty prefix enters frame 0 with mass 1
    for (int t = lane; t < Tb; t += 32) {
        const double2 r = rp[t];
        m = fmax(m, fmax(r.x, r.y));
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, d));
    return (m == NEGD) ? 0.0 : m;                                // a dead state: every mass is zero, any scale does
}

// grid (B, ceil(W / WG)); shared memory: WG * T double2 (phis, phib) + WG doubles (M) + WG ints (skip).
__global__ void __launch_bounds__(256) prefix_score_kernel(
    const double* __restrict__ p, int T, int B, int C, const int64_t* __restrict__ in_len, int blank, int W, int WG,
    const double* __restrict__ r_prev, const int32_t* __restrict__ last, const int32_t* __restrict__ glen,
    const int32_t* __restrict__ skip, double* __restrict__ psi, double* __restrict__ psi_eos) {
    extern __shared__ __align__(16) unsigned char sm_raw[];
    double2* ph = reinterpret_cast<double2*>(sm_raw);            // [WG][T]: masses entering frame t
    double* sM = reinterpret_cast<double*>(ph + (size_t)WG * T);
    int* sSkip = reinterpret_cast<int*>(sM + WG);
    const int b = blockIdx.x, w0 = blockIdx.y * WG, nw = min(WG, W - w0);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    const int Tb = in_len ? (int)max(0ll, min((long long)in_len[b], (long long)T)) : T;
    for (int j = warp; j < nw; j += nwarps) {
        const int bw = b * W + w0 + j;
        const bool sk = skip && skip[bw];
        if (lane == 0) sSkip[j] = sk;
        if (sk) {
            if (lane == 0) { sM[j] = 0.0; if (psi_eos) psi_eos[bw] = NEGD; }
            continue;
        }
        const double2* rp = reinterpret_cast<const double2*>(r_prev + (size_t)bw * T * 2);
        const int gl = glen[bw];
        const double M = warp_state_max(rp, Tb, gl, lane);
        double2* dst = ph + (size_t)j * T;
        for (int t = lane; t < Tb; t += 32) {
            double2 v;
            if (t == 0) {
                const double e = (gl == 0) ? exp(-M) : 0.0;
                v = make_double2(e, e);
            } else {
                const double2 r = rp[t - 1];
                const double en = exp(r.x - M), eb = exp(r.y - M);
                v = make_double2(en + eb, eb);
            }
            dst[t] = v;
        }
        if (lane == 0) {
            sM[j] = M;
            if (psi_eos) {
                double e = NEGD;
                if (Tb > 0) { const double2 r = rp[Tb - 1]; e = lse2d(r.x, r.y); }
                psi_eos[bw] = e;
            }
        }
    }
    __syncthreads();
    const double* pb = p + (size_t)b * T * C;
    for (int i = tid; i < nw * C; i += blockDim.x) {
        const int j = i / C, c = i - j * C;
        const int bw = b * W + w0 + j;
        double out = NEGD;
        if (!sSkip[j] && c != blank) {
            const bool rep = c == last[bw];
            const double2* f = ph + (size_t)j * T;
            const double* pc = pb + c;
            double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
            int t = 0;
            for (; t + 4 <= Tb; t += 4) {
                const double2 f0 = f[t], f1 = f[t + 1], f2 = f[t + 2], f3 = f[t + 3];
                const double p0 = pc[(size_t)t * C], p1 = pc[(size_t)(t + 1) * C], p2 = pc[(size_t)(t + 2) * C],
                             p3 = pc[(size_t)(t + 3) * C];
                a0 = fma(rep ? f0.y : f0.x, p0, a0);
                a1 = fma(rep ? f1.y : f1.x, p1, a1);
                a2 = fma(rep ? f2.y : f2.x, p2, a2);
                a3 = fma(rep ? f3.y : f3.x, p3, a3);
            }
            for (; t < Tb; ++t) {
                const double2 f0 = f[t];
                a0 = fma(rep ? f0.y : f0.x, pc[(size_t)t * C], a0);
            }
            const double s = (a0 + a1) + (a2 + a3);
            out = (s > 0.0) ? sM[j] + log(s) : NEGD;
        }
        psi[(size_t)bw * C + c] = out;
    }
}

// One warp per new slot (b, k) with parent sel_hyp and label sel_label; shared memory: 3 * T doubles per warp.
__global__ void prefix_advance_kernel(const double* __restrict__ p, int T, int B, int C,
                                      const int64_t* __restrict__ in_len, int blank, int W,
                                      const double* __restrict__ r_prev, const int32_t* __restrict__ last,
                                      const int32_t* __restrict__ glen, const int32_t* __restrict__ sel_hyp,
                                      const int32_t* __restrict__ sel_label, const int32_t* __restrict__ skip,
                                      double* __restrict__ r_new) {
    extern __shared__ __align__(16) unsigned char sm_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
    const int slot = blockIdx.x * wpb + warp;
    if (slot >= B * W) return;
    const int src_w = sel_hyp[slot], c = sel_label[slot];
    // a slot that is dead (no parent), finished (skip flag of the NEW slot) or mislabelled keeps no CTC state
    if (!(src_w >= 0 && src_w < W && c >= 0 && c < C && c != blank && !(skip && skip[slot]))) return;
    const int b = slot / W;
    const int Tb = in_len ? (int)max(0ll, min((long long)in_len[b], (long long)T)) : T;
    double* A = reinterpret_cast<double*>(sm_raw) + (size_t)warp * 3 * T;   // phi_c[t] in, r_n'[t] out
    double* PC = A + T;                                                     // p_t[c]
    double* PB = PC + T;                                                    // p_t[blank] in, r_b'[t] out
    const int pw = b * W + src_w;
    const double2* rp = reinterpret_cast<const double2*>(r_prev + (size_t)pw * T * 2);
    const int gl = glen[pw];
    const bool rep = c == last[pw];
    const double M = warp_state_max(rp, Tb, gl, lane);
    const double* pb = p + (size_t)b * T * C;
    for (int t = lane; t < Tb; t += 32) {
        double f;
        if (t == 0) {
            f = (gl == 0) ? exp(-M) : 0.0;
        } else {
            const double2 r = rp[t - 1];
            const double eb = exp(r.y - M);
            f = rep ? eb : exp(r.x - M) + eb;
        }
        A[t] = f;
        PC[t] = pb[(size_t)t * C + c];
        PB[t] = pb[(size_t)t * C + blank];
    }
    __syncwarp();
    if (lane == 0) {
        double a = 0.0, bl = 0.0;
#pragma unroll 4
        for (int t = 0; t < Tb; ++t) {
            const double na = (a + A[t]) * PC[t];
            const double nb = (a + bl) * PB[t];
            A[t] = na;
            PB[t] = nb;
            a = na;
            bl = nb;
        }
    }
    __syncwarp();
    double2* rout = reinterpret_cast<double2*>(r_new + (size_t)slot * T * 2);
    for (int t = lane; t < T; t += 32) {
        double2 v = make_double2(NEGD, NEGD);
        if (t < Tb) {
            const double a = A[t], bl = PB[t];
            v = make_double2(a > 0.0 ? M + log(a) : NEGD, bl > 0.0 ? M + log(bl) : NEGD);
        }
        rout[t] = v;
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

extern "C" int mdd_prefix_probs(const float* x, int T, int B, int C, double* p_lin, mdd_stream_t stream) {
    if (!x || !p_lin || T <= 0 || B < 0 || C <= 0) return MDD_E_ARG;
    if (B == 0) return MDD_OK;
    const size_t n = (size_t)T * B * C;
    const int nb = (int)std::min<size_t>((n + 255) / 256, (size_t)sm_count() * 16);
    prefix_probs_kernel<<<nb, 256, 0, (cudaStream_t)stream>>>(x, T, B, C, p_lin);
    MDD_CHECK_LAUNCH();
    return MDD_OK;
}

static constexpr size_t PREFIX_SMEM_MAX = 200 * 1024;

extern "C" int mdd_prefix_score(const double* p_lin, int T, int B, int C, const int64_t* input_lengths, int blank, int W,
                                const double* r_prev, const int32_t* last, const int32_t* glen, const int32_t* skip,
                                double* psi, double* psi_eos, mdd_stream_t stream) {
    if (!p_lin || !r_prev || !last || !glen || !psi || T <= 0 || B < 0 || C <= 0 || C > MDD_MAX_CLASSES || W <= 0 ||
        blank < 0 || blank >= C)
        return MDD_E_ARG;
    if (B == 0) return MDD_OK;
    // hypotheses per CTA: as many as fit in half an SM's shared memory (two CTAs per SM), at least one
    const size_t per_hyp = (size_t)T * sizeof(double2) + sizeof(double) + sizeof(int);
    if (per_hyp + 16 > PREFIX_SMEM_MAX) return MDD_E_UNSUPPORTED;
    int WG = (int)std::min<size_t>((size_t)W, std::max<size_t>(1, (100 * 1024) / per_hyp));
    const size_t smem = (size_t)WG * per_hyp + 16;
    if (smem > 48 * 1024 &&
        cudaFuncSetAttribute(prefix_score_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
        cudaGetLastError();
        return MDD_E_CUDA;
    }
    const dim3 grid(B, ceil_div(W, WG));
    prefix_score_kernel<<<grid, 256, smem, (cudaStream_t)stream>>>(p_lin, T, B, C, input_lengths, blank, W, WG, r_prev,
                                                                   last, glen, skip, psi, psi_eos);
    MDD_CHECK_LAUNCH();
    return MDD_OK;
}

extern "C" int mdd_prefix_advance(const double* p_lin, int T, int B, int C, const int64_t* input_lengths, int blank,
                                  int W, const double* r_prev, const int32_t* last, const int32_t* glen,
                                  const int32_t* sel_hyp, const int32_t* sel_label, const int32_t* skip_new,
                                  double* r_new, mdd_stream_t stream) {
    if (!p_lin || !r_prev || !last || !glen || !sel_hyp || !sel_label || !r_new || T <= 0 || B < 0 || C <= 0 ||
        W <= 0 || blank < 0 || blank >= C)
        return MDD_E_ARG;
    if (B == 0) return MDD_OK;
    const size_t per_warp = (size_t)3 * T * sizeof(double);
    if (per_warp > PREFIX_SMEM_MAX) return MDD_E_UNSUPPORTED;
    const int wpb = (int)std::min<size_t>(4, std::max<size_t>(1, (48 * 1024) / per_warp));
    const size_t smem = wpb * per_warp;
    if (smem > 48 * 1024 &&
        cudaFuncSetAttribute(prefix_advance_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
        cudaGetLastError();
        return MDD_E_CUDA;
    }
    prefix_advance_kernel<<<ceil_div(B * W, wpb), 32 * wpb, smem, (cudaStream_t)stream>>>(
        p_lin, T, B, C, input_lengths, blank, W, r_prev, last, glen, sel_hyp, sel_label, skip_new, r_new);
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
