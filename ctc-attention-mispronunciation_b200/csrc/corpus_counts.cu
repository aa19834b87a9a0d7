// corpus_counts.cu -- corpus MDD evaluation counters (TA / FR / FA / TR-correct / TR-wrong, PER numerators), one warp
// per utterance.
//
// Replaces the per-utterance body of egs/attention_aug/steps/test_ctc_nosil.py:218-297 together with
// print_align_space_canonical_origin (:33-82): given the op paths of wer(labels, canonical) and
// wer(decoded, canonical) (mdd_edit_align) and the human / recognised token rows, every canonical position j gets
// d1[j], d2[j] in {'-', 'D', 'S'+phone}; the positions are classified (:246-264) and the two insertion lists are
// matched with the reference's remove-while-iterating loop (:265-287), restated literally on one lane because its
// result depends on the iteration order (Python skips the element that slides into a freed index).
//
// Integer, on-chip: the paths are read once (coalesced bytes), the d maps live in shared memory; per-utterance
// counters go to counts[B,12], corpus totals are accumulated with 64-bit atomics (integer: order-independent).
#include "common.cuh"

namespace mdd {
namespace {

constexpr int WARPS = 4;
constexpr unsigned FULLM = 0xffffffffu;

struct CorpusParams {
    const int32_t *lab, *lab_len, *dec, *dec_len;
    int lab_stride, dec_stride, B, max_lab, max_dec, max_can;
    const int32_t *lab_map, *dec_map;
    int n_map;
    const uint8_t *lc_ops, *dc_ops;
    int lc_stride, dc_stride;
    const int32_t *lc_len, *dc_len;
    const int32_t* dist_dl;
    const uint8_t* group;
    int32_t* counts;
    unsigned long long* totals;
    int32_t* status;
    int words_per_warp;
};

// token row -> optional map -> compaction into shared memory (same rule as mdd_edit_align); returns the kept length
__device__ int load_row(const int32_t* src, int n, const int32_t* map, int n_map, int32_t* dst, int cap, int lane) {
    int out = 0;
    for (int base = 0; base < n; base += 32) {
        const int k = base + lane;
        int v = -1;
        bool keep = false;
        if (k < n) {
            v = src[k];
            if (map) v = (v >= 0 && v < n_map) ? map[v] : -1;
            keep = map ? (v >= 0) : true;
        }
        const unsigned m = __ballot_sync(FULLM, keep);
        const int at = out + __popc(m & ((1u << lane) - 1));
        if (keep && at < cap) dst[at] = v;
        out += __popc(m);
    }
    return min(out, cap);
}

// d[j] for every canonical position of one path (0 '-', 1 'D', 2 + token 'S'+token) and the insertion list (the
// canonical position each 'I' precedes).  Returns the number of insertions; n_can = canonical positions seen.
__device__ int build_dmap(const uint8_t* ops, int n, const int32_t* hyp, int Lh, int* d, int cap_d, int* ilist,
                          int cap_i, int lane, int& n_can) {
    int i_base = 0, j_base = 0, ni = 0;
    const unsigned lt = (1u << lane) - 1;
    for (int base = 0; base < n; base += 32) {
        const int k = base + lane;
        const int op = k < n ? ops[k] : 0;
        const bool is_i = op == 'I', is_d = op == 'D', is_m = op == '-' || op == 'S';
        const unsigned mh = __ballot_sync(FULLM, is_i || is_m), mc = __ballot_sync(FULLM, is_d || is_m);
        const unsigned mi = __ballot_sync(FULLM, is_i);
        const int i = i_base + __popc(mh & lt), j = j_base + __popc(mc & lt);
        if ((is_d || is_m) && j < cap_d) d[j] = op == '-' ? 0 : is_d ? 1 : 2 + (i < Lh ? hyp[i] : -1);
        const int at = ni + __popc(mi & lt);
        if (is_i && at < cap_i) ilist[at] = j;
        i_base += __popc(mh);
        j_base += __popc(mc);
        ni += __popc(mi);
    }
    n_can = j_base;
    return min(ni, cap_i);
}

__global__ void __launch_bounds__(WARPS * 32) corpus_counts_kernel(CorpusParams p) {
    extern __shared__ int32_t smem_i[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int b = blockIdx.x * WARPS + warp;
    if (b >= p.B) return;
    int32_t* w = smem_i + (size_t)warp * p.words_per_warp;
    int32_t* labf = w;                                   // [max_lab] filtered human labels
    int32_t* decf = labf + p.max_lab;                    // [max_dec] filtered recognised phones
    int* d1 = decf + p.max_dec;                          // [max_can]
    int* d2 = d1 + p.max_can;                            // [max_can]
    int* i1 = d2 + p.max_can;                            // [max_lab] insertions of the human transcription
    int* i2 = i1 + p.max_lab;                            // [max_dec] insertions of the recogniser

    const int Ll = load_row(p.lab + (size_t)b * p.lab_stride, max(0, min(p.lab_len[b], p.max_lab)), p.lab_map, p.n_map,
                            labf, p.max_lab, lane);
    const int Ld = load_row(p.dec + (size_t)b * p.dec_stride, max(0, min(p.dec_len[b], p.max_dec)), p.dec_map, p.n_map,
                            decf, p.max_dec, lane);
    __syncwarp();
    int c1 = 0, c2 = 0;
    int n1 = build_dmap(p.lc_ops + (size_t)b * p.lc_stride, max(0, min(p.lc_len[b], p.lc_stride)), labf, Ll, d1, p.max_can,
                        i1, p.max_lab, lane, c1);
    int n2 = build_dmap(p.dc_ops + (size_t)b * p.dc_stride, max(0, min(p.dc_len[b], p.dc_stride)), decf, Ld, d2, p.max_can,
                        i2, p.max_dec, lane, c2);
    __syncwarp();
    int32_t* out = p.counts + (size_t)b * 12;
    int st = MDD_ST_OK;
    if (Ll == 0 || Ld == 0 || c1 == 0) st = MDD_ST_EMPTY;            // decoder.wer raises TypeError (ctcDecoder.py:125,137-138)
    else if (c1 != c2 || c1 > p.max_can) st = MDD_ST_BADLEN;         // the two paths are not against one canonical row
    if (st != MDD_ST_OK) {
        if (lane < 12) out[lane] = 0;
        if (lane == 0 && p.status) p.status[b] = st;
        return;
    }
    int ta = 0, fr = 0, fa = 0, trc = 0, trw = 0;
    for (int k = lane; k < c1; k += 32) {
        const int a = d1[k], r = d2[k];
        if (a == 0 && r == 0) ++ta;
        else if (a == 0) ++fr;
        else if (r == 0) ++fa;
        else if (a == r) ++trc;
        else ++trw;
    }
    ta = __reduce_add_sync(FULLM, ta);
    fr = __reduce_add_sync(FULLM, fr);
    fa = __reduce_add_sync(FULLM, fa);
    trc = __reduce_add_sync(FULLM, trc);
    trw = __reduce_add_sync(FULLM, trw);
    if (lane == 0) {
        int m_fr = fr, m_fa = fa;
        if (n1 && !n2) { fa += n1; m_fa += 1; }
        else if (!n1 && n2) { fr += n2; m_fr += 1; }
        else if (n1 && n2) {
            for (int idx = 0; idx < n1; ++idx) {                     // `for e in d1['I']` over the live list
                const int e = i1[idx];
                int at2 = -1;
                for (int q = 0; q < n2; ++q) if (i2[q] == e) { at2 = q; break; }
                if (at2 < 0) continue;
                int at1 = 0;
                while (i1[at1] != e) ++at1;                          // list.remove: first occurrence
                for (int q = at1; q + 1 < n1; ++q) i1[q] = i1[q + 1];
                --n1;
                for (int q = at2; q + 1 < n2; ++q) i2[q] = i2[q + 1];
                --n2;
                ++trc;
            }
            fa += n1; fr += n2; m_fa += n1; m_fr += n2;
        }
        const int dl = p.dist_dl ? p.dist_dl[b] : 0;
        out[0] = ta; out[1] = fr; out[2] = fa; out[3] = trc; out[4] = trw; out[5] = c1;
        out[6] = m_fr; out[7] = m_fa; out[8] = Ll; out[9] = dl; out[10] = 0; out[11] = 0;
        if (p.status) p.status[b] = MDD_ST_OK;
        if (p.totals) {
            const int all[10] = {ta, fr, fa, trc, trw, c1, m_fr, m_fa, Ll, dl};
            for (int k = 0; k < 10; ++k) if (all[k]) atomicAdd(p.totals + k, (unsigned long long)(long long)all[k]);
            if (p.group && p.group[b]) {                             // the m_speaker counters (:236-297, 318-343)
                const int grp[10] = {ta, m_fr, m_fa, trc, trw, c1, m_fr, m_fa, Ll, dl};
                for (int k = 0; k < 10; ++k) if (grp[k]) atomicAdd(p.totals + 12 + k, (unsigned long long)(long long)grp[k]);
            }
        }
    }
}

}  // namespace
}  // namespace mdd

extern "C" int mdd_corpus_counts(const int32_t* lab, int lab_stride, const int32_t* lab_len, const int32_t* dec,
                                 int dec_stride, const int32_t* dec_len, int B, int max_lab_len, int max_dec_len,
                                 int max_can_len, const int32_t* lab_map, const int32_t* dec_map, int n_map,
                                 const uint8_t* lc_ops, int lc_stride, const int32_t* lc_len, const uint8_t* dc_ops,
                                 int dc_stride, const int32_t* dc_len, const int32_t* dist_dl, const uint8_t* group,
                                 int32_t* counts, int64_t* totals, int32_t* status, mdd_stream_t stream) {
    using namespace mdd;
    if (B == 0) return MDD_OK;
    if (!lab || !lab_len || !dec || !dec_len || !lc_ops || !lc_len || !dc_ops || !dc_len || !counts || B < 0 ||
        max_lab_len < 0 || max_dec_len < 0 || max_can_len < 0 || lc_stride < 0 || dc_stride < 0)
        return MDD_E_ARG;
    if ((lab_map || dec_map) && n_map <= 0) return MDD_E_ARG;
    CorpusParams p{};
    p.lab = lab; p.lab_len = lab_len; p.dec = dec; p.dec_len = dec_len;
    p.lab_stride = lab_stride; p.dec_stride = dec_stride; p.B = B;
    p.max_lab = max_lab_len > 0 ? max_lab_len : 1; p.max_dec = max_dec_len > 0 ? max_dec_len : 1;
    p.max_can = max_can_len > 0 ? max_can_len : 1;
    p.lab_map = lab_map; p.dec_map = dec_map; p.n_map = n_map;
    p.lc_ops = lc_ops; p.dc_ops = dc_ops; p.lc_stride = lc_stride; p.dc_stride = dc_stride;
    p.lc_len = lc_len; p.dc_len = dc_len; p.dist_dl = dist_dl; p.group = group;
    p.counts = counts; p.totals = reinterpret_cast<unsigned long long*>(totals); p.status = status;
    p.words_per_warp = 2 * (p.max_lab + p.max_dec + p.max_can);
    const size_t smem = (size_t)WARPS * p.words_per_warp * sizeof(int32_t);
    if (smem > 200 * 1024) return MDD_E_UNSUPPORTED;
    if (smem > 48 * 1024) {
        if (cudaFuncSetAttribute(corpus_counts_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
            cudaGetLastError();
            return MDD_E_CUDA;
        }
    }
    corpus_counts_kernel<<<ceil_div(B, WARPS), WARPS * 32, smem, (cudaStream_t)stream>>>(p);
    MDD_CHECK_LAUNCH();
    return MDD_OK;
}
