// prefix_score.cu -- CTC prefix scorer for CTC/attention joint beam decoding.
//
// Named by BASELINE.json's north_star; the reference has no such function (its nearest code is the
// (prBlank, prNonBlank, prTotal) recursion of egs/attention_aug/utils/BeamSearch.py:52-66,99-125).  Formulation:
// Watanabe et al. 2017 / ESPnet CTCPrefixScore; parity anchor = the fp64 CPU restatement used by the tests and the
// psi(l.eos) == -nll invariant against torch CPU ctc_loss.
//
// One CTA per (utterance, hypothesis); thread c scans t once for candidate label c with (r_n, r_b, psi) in
// registers.  The previous state r_prev[b,w,:,:] and the rows x[t,b,:] are shared by all candidates and are served
// from L1/L2 (every row is re-read W times by the W CTAs of the utterance).
#include "common.cuh"

namespace mdd {
namespace {

__global__ void prefix_init_kernel(const float* __restrict__ x, int T, int B, int C, const int64_t* __restrict__ in_len,
                                   int blank, int W, float* __restrict__ r0) {
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= B) return;
    const int b = warp;
    const int Tb = in_len ? (int)max(0ll, min((long long)in_len[b], (long long)T)) : T;
    float run = 0.f;
    for (int base = 0; base < T; base += 32) {
        const int t = base + lane;
        float v = (t < Tb) ? x[((size_t)t * B + b) * C + blank] : 0.f;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const float o = __shfl_up_sync(0xffffffffu, v, d);
            if (lane >= d) v += o;
        }
        v += run;
        run = __shfl_sync(0xffffffffu, v, 31);
        if (t < T) {
            const float rb = (t < Tb) ? v : -CUDART_INF_F;
            for (int w = 0; w < W; ++w) {
                float2* dst = reinterpret_cast<float2*>(r0 + (((size_t)b * W + w) * T + t) * 2);
                *dst = make_float2(-CUDART_INF_F, rb);
            }
        }
    }
}

template <bool ADVANCE>
__global__ void prefix_score_kernel(const float* __restrict__ x, int T, int B, int C, const int64_t* __restrict__ in_len,
                                    int blank, int W, const float* __restrict__ r_prev, const int32_t* __restrict__ last,
                                    const int32_t* __restrict__ glen, const int32_t* __restrict__ sel_hyp,
                                    const int32_t* __restrict__ sel_label, float* __restrict__ out,
                                    float* __restrict__ psi_eos) {
    const int bw = blockIdx.x;               // b * W + w
    const int b = bw / W;
    const int Tb = in_len ? (int)max(0ll, min((long long)in_len[b], (long long)T)) : T;
    int src_w = bw - b * W;
    int c = threadIdx.x;
    if (ADVANCE) {
        src_w = sel_hyp[bw];
        c = sel_label[bw];
        if (threadIdx.x != 0) return;
    } else if (c >= C) return;
    const bool valid = !ADVANCE || (src_w >= 0 && src_w < W && c >= 0 && c < C && c != blank);
    float2* rout = ADVANCE ? reinterpret_cast<float2*>(out + (size_t)bw * T * 2) : nullptr;
    if (ADVANCE && !valid) {
        for (int t = 0; t < T; ++t) rout[t] = make_float2(-CUDART_INF_F, -CUDART_INF_F);
        return;
    }
    const float2* rp = reinterpret_cast<const float2*>(r_prev + ((size_t)b * W + src_w) * T * 2);
    const int lastl = last[b * W + src_w], gl = glen[b * W + src_w];
    const int start = gl > 1 ? gl : 1;
    const float NEG = -CUDART_INF_F;
    if (!ADVANCE && c == blank) {
        out[(size_t)bw * C + c] = NEG;
        if (psi_eos) {
            float e = NEG;
            if (Tb > 0) { const float2 r = rp[Tb - 1]; e = lse2_acc(r.x, r.y); }
            psi_eos[bw] = e;
        }
        return;
    }
    float rn = NEG, rb = NEG;
    if (start - 1 < Tb && gl == 0) rn = x[(size_t)b * C + c];
    if (ADVANCE) {
        for (int t = 0; t < start - 1 && t < T; ++t) rout[t] = make_float2(NEG, NEG);
        if (start - 1 < T) rout[start - 1] = (start - 1 < Tb) ? make_float2(rn, rb) : make_float2(NEG, NEG);
    }
    float ps = rn;
    for (int t = start; t < Tb; ++t) {
        const float2 r = rp[t - 1];
        const float rsum = lse2_acc(r.x, r.y);
        const float phi = (c == lastl) ? r.y : rsum;
        const float* row = x + ((size_t)t * B + b) * C;
        const float xc = row[c], xb = row[blank];
        const float nrn = lse2_acc(rn, phi) + xc;
        const float nrb = lse2_acc(rn, rb) + xb;
        ps = lse2_acc(ps, phi + xc);
        rn = nrn; rb = nrb;
        if (ADVANCE) rout[t] = make_float2(rn, rb);
    }
    if (ADVANCE) {
        for (int t = max(Tb, start); t < T; ++t) rout[t] = make_float2(NEG, NEG);
    } else {
        out[(size_t)bw * C + c] = ps;
    }
}

}  // namespace
}  // namespace mdd

using namespace mdd;

extern "C" int mdd_prefix_init(const float* x, int T, int B, int C, const int64_t* input_lengths, int blank, int W,
                               float* r0, mdd_stream_t stream) {
    if (!x || !r0 || T <= 0 || B < 0 || C <= 0 || W <= 0 || blank < 0 || blank >= C) return MDD_E_ARG;
    if (B == 0) return MDD_OK;
    prefix_init_kernel<<<ceil_div(B, 4), 128, 0, (cudaStream_t)stream>>>(x, T, B, C, input_lengths, blank, W, r0);
    MDD_CHECK_LAUNCH();
    return MDD_OK;
}

extern "C" int mdd_prefix_score(const float* x, int T, int B, int C, const int64_t* input_lengths, int blank, int W,
                                const float* r_prev, const int32_t* last, const int32_t* glen, float* psi,
                                float* psi_eos, mdd_stream_t stream) {
    if (!x || !r_prev || !last || !glen || !psi || T <= 0 || B < 0 || C <= 0 || C > MDD_MAX_CLASSES || W <= 0 ||
        blank < 0 || blank >= C)
        return MDD_E_ARG;
    if (B == 0) return MDD_OK;
    const int nt = ceil_div(C, 32) * 32;
    prefix_score_kernel<false><<<B * W, nt, 0, (cudaStream_t)stream>>>(x, T, B, C, input_lengths, blank, W, r_prev, last,
                                                                       glen, nullptr, nullptr, psi, psi_eos);
    MDD_CHECK_LAUNCH();
    return MDD_OK;
}

extern "C" int mdd_prefix_advance(const float* x, int T, int B, int C, const int64_t* input_lengths, int blank, int W,
                                  const float* r_prev, const int32_t* last, const int32_t* glen, const int32_t* sel_hyp,
                                  const int32_t* sel_label, float* r_new, mdd_stream_t stream) {
    if (!x || !r_prev || !last || !glen || !sel_hyp || !sel_label || !r_new || T <= 0 || B < 0 || C <= 0 || W <= 0 ||
        blank < 0 || blank >= C)
        return MDD_E_ARG;
    if (B == 0) return MDD_OK;
    prefix_score_kernel<true><<<B * W, 32, 0, (cudaStream_t)stream>>>(x, T, B, C, input_lengths, blank, W, r_prev, last,
                                                                      glen, sel_hyp, sel_label, r_new, nullptr);
    MDD_CHECK_LAUNCH();
    return MDD_OK;
}
