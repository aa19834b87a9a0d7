/*
 * oracle/mdd_oracle.c -- CPU restatement of the reference's CTC hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it.
 * The product path (ctc-attention-mispronunciation_b200/) never links, imports or calls it.
 *
 * Every function restates one piece of dyustc/CTC-Attention-Mispronunciation (paths relative to
 * the reference root) or of the third-party arithmetic its call sites reach:
 *
 *   orc_ctc_loss_f64 / _f32   nn.CTCLoss(reduction='sum') call site
 *                             egs/attention_aug/steps/train_ctc.py:186,72,96 and
 *                             egs/cnn-rnn-ctc/steps/train_ctc.py:148,51,67.  The arithmetic is torch's
 *                             aten ctc_loss (third-party, un-vendored, version unpinned by the
 *                             reference; pinned here to torch 2.11.0 CPU by tests/test_oracle_pin.py).
 *   orc_greedy                GreedyDecoder.decode   egs/attention_aug/utils/ctcDecoder.py:186-200,80-116
 *   orc_compute_wer           CTC_Model.compute_wer  egs/attention_aug/models/model_ctc.py:227-244
 *   orc_edit_align            Decoder._edit_distance / printChanges  utils/ctcDecoder.py:134-184
 *   orc_beam_search           ctcBeamSearch.decode   egs/attention_aug/utils/BeamSearch.py:73-153
 *                             (+ calcExtPr :52-66, log_add_prob :43-50, BeamState.sort/norm :23-33)
 *   orc_mdd_post              align_canonical_decoded / stastics / score
 *                             egs/attention_aug/infer.py:162-209, 405-433, 338-342
 *   orc_corpus_counts         TA/FR/FA/TR counters of the corpus MDD evaluation
 *                             egs/attention_aug/steps/test_ctc_nosil.py:218-293 (+ :33-82)
 *   orc_prefix_score          CTC prefix scorer for CTC/attention joint decoding.  NOT in the reference
 *                             (Watanabe et al. 2017 formulation): parity unpinned
 *                             by the reference; pinned by the psi(l.eos) == -nll invariant against
 *                             orc_ctc_loss_f64 / torch.
 *
 * Build: see oracle/Makefile (gcc -O2 -shared -fPIC -lm).  No -ffast-math: the beam search must
 * reproduce CPython's math.log / math.exp (glibc libm) double arithmetic bit for bit.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_API __attribute__((visibility("default")))

/* ------------------------------------------------------------------------------------------------
 * 1. CTC loss + gradient (torch semantics)
 * ---------------------------------------------------------------------------------------------- */

#define DEFINE_CTC(NAME, real, EXP, LOG)                                                          \
static real NAME##_lse3(real a, real b, real c) {                                                 \
    real m = a > b ? a : b; if (c > m) m = c;                                                     \
    if (m == -INFINITY) m = 0;                                                                    \
    return LOG(EXP(a - m) + EXP(b - m) + EXP(c - m)) + m;                                         \
}                                                                                                 \
static real NAME##_lse2(real a, real b) {                                                         \
    real m = a > b ? a : b;                                                                       \
    if (m == -INFINITY) m = 0;                                                                    \
    return LOG(EXP(a - m) + EXP(b - m)) + m;                                                      \
}                                                                                                 \
/* lp: (T,B,C) contiguous. targets: row b at targets + b*tgt_stride (tgt_stride==0: 1-D           \
 * concatenated, offsets = prefix sums of tgt_len).  grad_out: per-utterance upstream gradient    \
 * (NULL = ones).  grad may be NULL (forward only).  alpha_out (optional): (B,T,2*Lmax+1).  */    \
ORC_API int NAME(const real* lp, int T, int B, int C, const int64_t* targets, int64_t tgt_stride, \
                 const int64_t* in_len, const int64_t* tgt_len, int blank, int zero_infinity,     \
                 const real* grad_out, real* nll, real* grad, real* alpha_out, int Lmax) {        \
    int64_t off = 0;                                                                              \
    int Smax_all = 1;                                                                             \
    for (int b = 0; b < B; ++b) { int s = 2 * (int)tgt_len[b] + 1; if (s > Smax_all) Smax_all = s; } \
    real* A = (real*)malloc(sizeof(real) * (size_t)(T > 0 ? T : 1) * Smax_all);                   \
    real* Bt = (real*)malloc(sizeof(real) * 2 * Smax_all);                                        \
    real* occ = (real*)malloc(sizeof(real) * C);                                                  \
    for (int b = 0; b < B; ++b) {                                                                 \
        const int Tb = (int)in_len[b], L = (int)tgt_len[b], S = 2 * L + 1;                        \
        const int64_t* tg = tgt_stride ? targets + (int64_t)b * tgt_stride : targets + off;       \
        off += L;                                                                                 \
        if (Tb < 0 || Tb > T || L < 0) { free(A); free(Bt); free(occ); return -1; }               \
        if (grad) for (int t = 0; t < T; ++t) for (int c = 0; c < C; ++c)                         \
            grad[((size_t)t * B + b) * C + c] = 0;                                                \
        if (Tb == 0) {                                                                            \
            nll[b] = (L == 0) ? (real)0 : (real)INFINITY;                                         \
            if (zero_infinity && nll[b] == (real)INFINITY) nll[b] = 0;                            \
            continue;                                                                             \
        }                                                                                         \
        for (int s = 0; s < S; ++s) A[s] = -INFINITY;                                             \
        A[0] = lp[((size_t)0 * B + b) * C + blank];                                               \
        if (L > 0) A[1] = lp[((size_t)0 * B + b) * C + tg[0]];                                    \
        for (int t = 1; t < Tb; ++t) {                                                            \
            const real* row = lp + ((size_t)t * B + b) * C;                                       \
            const real* p = A + (size_t)(t - 1) * S; real* q = A + (size_t)t * S;                 \
            for (int s = 0; s < S; ++s) {                                                         \
                int64_t cur = (s & 1) ? tg[s >> 1] : blank;                                       \
                real a1 = p[s], a2 = s > 0 ? p[s - 1] : (real)-INFINITY, a3 = -INFINITY;          \
                if (s > 1 && (s & 1) && tg[(s - 2) >> 1] != cur) a3 = p[s - 2];                   \
                q[s] = NAME##_lse3(a1, a2, a3) + row[cur];                                        \
            }                                                                                     \
        }                                                                                         \
        const real* last = A + (size_t)(Tb - 1) * S;                                              \
        real ll = S > 1 ? NAME##_lse2(last[S - 1], last[S - 2]) : last[S - 1];                    \
        real nl = -ll;                                                                            \
        if (alpha_out) for (int t = 0; t < Tb; ++t) for (int s = 0; s < S; ++s)                   \
            alpha_out[((size_t)b * T + t) * (2 * Lmax + 1) + s] = A[(size_t)t * S + s];           \
        if (!grad) { nll[b] = (zero_infinity && nl == (real)INFINITY) ? 0 : nl; continue; }       \
        if (zero_infinity && nl == (real)INFINITY) { nll[b] = 0; continue; }                      \
        nll[b] = nl;                                                                              \
        const real g = grad_out ? grad_out[b] : (real)1;                                          \
        real* bn = Bt; real* bo = Bt + Smax_all;                                                  \
        for (int t = Tb - 1; t >= 0; --t) {                                                       \
            const real* row = lp + ((size_t)t * B + b) * C;                                       \
            if (t == Tb - 1) {                                                                    \
                for (int s = 0; s < S; ++s) bn[s] = -INFINITY;                                    \
                bn[S - 1] = row[blank];                                                           \
                if (L > 0) bn[S - 2] = row[tg[L - 1]];                                            \
            } else {                                                                              \
                for (int s = 0; s < S; ++s) {                                                     \
                    int64_t cur = (s & 1) ? tg[s >> 1] : blank;                                   \
                    real b1 = bo[s], b2 = s + 1 < S ? bo[s + 1] : (real)-INFINITY, b3 = -INFINITY;\
                    if (s + 2 < S && (s & 1) && tg[(s + 2) >> 1] != cur) b3 = bo[s + 2];          \
                    bn[s] = NAME##_lse3(b1, b2, b3) + row[cur];                                   \
                }                                                                                 \
            }                                                                                     \
            for (int c = 0; c < C; ++c) occ[c] = -INFINITY;                                       \
            const real* al = A + (size_t)t * S;                                                   \
            for (int s = 0; s < S; ++s) {                                                         \
                int64_t cur = (s & 1) ? tg[s >> 1] : blank;                                       \
                occ[cur] = NAME##_lse2(occ[cur], al[s] + bn[s]);                                  \
            }                                                                                     \
            real* gr = grad + ((size_t)t * B + b) * C;                                            \
            for (int c = 0; c < C; ++c)                                                           \
                gr[c] = (EXP(row[c]) - EXP(occ[c] + nl - row[c])) * g;                            \
            real* tmp = bn; bn = bo; bo = tmp;                                                    \
        }                                                                                         \
    }                                                                                             \
    free(A); free(Bt); free(occ);                                                                 \
    return 0;                                                                                     \
}

DEFINE_CTC(orc_ctc_loss_f64, double, exp, log)
DEFINE_CTC(orc_ctc_loss_f32, float, expf, logf)

/* ------------------------------------------------------------------------------------------------
 * 2. Greedy decode: frame argmax (first max wins, like torch.max) + collapse.
 *    ctcDecoder.py:186-200 -> _convert_to_strings :94-116 -> _process_string(remove_rep=True) :80-92.
 *    The reference compares the previous FRAME's word string; word_id[c] gives an id per word string
 *    so that two classes sharing a string compare equal exactly as there (NULL = identity).
 *    tokens: (B, T) int32, out_len: (B).  lp is (T,B,C) contiguous.
 * ---------------------------------------------------------------------------------------------- */
ORC_API int orc_greedy(const float* lp, int T, int B, int C, const int64_t* in_len, int blank,
                       const int32_t* word_id, int32_t* tokens, int32_t* out_len, int32_t* argmax_out) {
    for (int b = 0; b < B; ++b) {
        int n = 0, prev = -1;
        int Tb = (int)in_len[b];
        if (Tb > T || Tb < 0) return -1;
        for (int t = 0; t < T; ++t) {
            const float* row = lp + ((size_t)t * B + b) * C;
            int best = 0; float bv = row[0];
            for (int c = 1; c < C; ++c) if (row[c] > bv) { bv = row[c]; best = c; }
            /* torch.max propagates NaN: the first NaN wins */
            for (int c = 0; c < C; ++c) if (row[c] != row[c]) { best = c; break; }
            if (argmax_out) argmax_out[(size_t)b * T + t] = best;
            if (t < Tb) {
                int w = word_id ? word_id[best] : best;
                int wb = word_id ? word_id[blank] : blank;
                if (w != wb && !(t != 0 && w == prev)) tokens[(size_t)b * T + n++] = best;
                prev = w;
            }
        }
        out_len[b] = n;
    }
    return 0;
}

/* ------------------------------------------------------------------------------------------------
 * 3. Levenshtein distance + op-path traceback (ctcDecoder.py:134-184).
 *    hyp = s1 (recognised), ref = s2 (canonical).  ops: '-','S','I','D' bytes, forward order.
 *    status: 0 ok, 1 = an empty side (the reference's _edit_distance returns a bare int there and
 *    wer() raises TypeError when unpacking, ctcDecoder.py:125,137-138); dist is still filled.
 * ---------------------------------------------------------------------------------------------- */
ORC_API int orc_edit_align_one(const int32_t* a, int L1, const int32_t* b, int L2,
                               int32_t* dist_out, uint8_t* ops, int32_t* ops_len) {
    if (L1 == 0 || L2 == 0) { *dist_out = L1 == 0 ? L2 : L1; if (ops_len) *ops_len = 0; return 1; }
    int W = L2 + 1;
    int32_t* d = (int32_t*)malloc(sizeof(int32_t) * (size_t)(L1 + 1) * W);
    for (int j = 0; j <= L2; ++j) d[j] = j;
    for (int i = 1; i <= L1; ++i) d[(size_t)i * W] = i;
    for (int i = 1; i <= L1; ++i)
        for (int j = 1; j <= L2; ++j) {
            int cost = a[i - 1] == b[j - 1] ? 0 : 1;
            int x = d[(size_t)i * W + j - 1] + 1, y = d[(size_t)(i - 1) * W + j] + 1,
                z = d[(size_t)(i - 1) * W + j - 1] + cost;
            int m = x < y ? x : y; if (z < m) m = z;
            d[(size_t)i * W + j] = m;
        }
    *dist_out = d[(size_t)L1 * W + L2];
    if (ops) {
        int i = L1, j = L2, n = 0;
        uint8_t* tmp = (uint8_t*)malloc((size_t)L1 + L2 + 1);
        while (i > 0 || j > 0) {
            if (i == 0) { tmp[n++] = 'D'; --j; }
            else if (j == 0) { tmp[n++] = 'I'; --i; }
            else if (a[i - 1] == b[j - 1]) { --i; --j; tmp[n++] = '-'; }
            else if (d[(size_t)i * W + j] == d[(size_t)(i - 1) * W + j - 1] + 1) { --i; --j; tmp[n++] = 'S'; }
            else if (d[(size_t)i * W + j] == d[(size_t)(i - 1) * W + j] + 1) { --i; tmp[n++] = 'I'; }
            else if (d[(size_t)i * W + j] == d[(size_t)i * W + j - 1] + 1) { --j; tmp[n++] = 'D'; }
        }
        for (int k = 0; k < n; ++k) ops[k] = tmp[n - 1 - k];
        *ops_len = n;
        free(tmp);
    }
    free(d);
    return 0;
}

/* batched: hyp (B, hyp_stride), ref (B, ref_stride); ops (B, ops_stride) */
ORC_API int orc_edit_align(const int32_t* hyp, int hyp_stride, const int32_t* hyp_len,
                           const int32_t* ref, int ref_stride, const int32_t* ref_len, int B,
                           int32_t* dist, uint8_t* ops, int ops_stride, int32_t* ops_len, int32_t* status) {
    for (int b = 0; b < B; ++b) {
        int32_t ol = 0;
        status[b] = orc_edit_align_one(hyp + (size_t)b * hyp_stride, hyp_len[b], ref + (size_t)b * ref_stride,
                                       ref_len[b], dist + b, ops ? ops + (size_t)b * ops_stride : NULL, &ol);
        if (ops_len) ops_len[b] = ol;
    }
    return 0;
}

/* model_ctc.py:227-244: index (B,T) argmax ids; collapse (drop 0, drop equal-to-previous-frame),
 * errs += edit distance(label[:L_b], pred), tokens += L_b.  The reference hard-codes blank == 0. */
ORC_API int orc_compute_wer(const int64_t* index, int B, int T, const int64_t* in_len, const int64_t* targets,
                            int Lmax, const int64_t* tgt_len, int64_t* errs, int64_t* tokens) {
    int32_t* pred = (int32_t*)malloc(sizeof(int32_t) * (T + 1));
    int32_t* lab = (int32_t*)malloc(sizeof(int32_t) * (Lmax + 1));
    int64_t e = 0, n = 0;
    for (int b = 0; b < B; ++b) {
        int np = 0;
        for (int j = 0; j < (int)in_len[b] && j < T; ++j) {
            int64_t v = index[(size_t)b * T + j];
            if (v == 0) continue;
            if (j == 0 || v != index[(size_t)b * T + j - 1]) pred[np++] = (int32_t)v;
        }
        int L = (int)tgt_len[b];
        for (int i = 0; i < L; ++i) lab[i] = (int32_t)targets[(size_t)b * Lmax + i];
        int32_t d; orc_edit_align_one(lab, L, pred, np, &d, NULL, NULL);
        e += d; n += L;
    }
    *errs = e; *tokens = n;
    free(pred); free(lab);
    return 0;
}

/* ------------------------------------------------------------------------------------------------
 * 4. CTC prefix beam search with bigram-LM fusion (BeamSearch.py:73-153), float64 scores on float32
 *    probabilities, dict-insertion-order tie-breaks, numpy-2 float32 threshold semantics.
 *    probs: (B,T,C) float32 PROBABILITIES (ctcDecoder.py:223-225 applies exp before the call).
 *    lm: (C+1)x(C+1) double table, lm[c1*(C+1)+c2] = get_bi_prob(classes[c1], classes[c2]) with index C
 *    standing for '' (<s> as c1, </s> as c2) (NgramLM.py:65-78); NaN = the lookup raises KeyError.
 *    status: 0 ok, 1 = IndexError (empty prefix in the final beam, BeamSearch.py:135),
 *            2 = ValueError (math.log(0), :64/66/103/106), 3 = KeyError (LM vocabulary miss).
 * ---------------------------------------------------------------------------------------------- */
#define ORC_LOG_ZERO (-99999999.0)

typedef struct { int parent, label, len; } orc_node;
typedef struct { int node; double pb, pnb, pt; } orc_entry;

typedef struct {
    orc_node* nodes; int n_nodes, cap_nodes;
    int* htab; int hcap;             /* (parent,label) -> node */
} orc_trie;

static unsigned orc_hash(int parent, int label, int cap) {
    uint64_t x = ((uint64_t)(uint32_t)parent << 20) ^ (uint32_t)label;
    x *= 0x9E3779B97F4A7C15ull; x ^= x >> 29;
    return (unsigned)(x & (uint64_t)(cap - 1));
}
static void orc_trie_rehash(orc_trie* tr) {
    int ncap = tr->hcap * 2;
    int* nt = (int*)malloc(sizeof(int) * ncap);
    for (int i = 0; i < ncap; ++i) nt[i] = -1;
    for (int n = 1; n < tr->n_nodes; ++n) {
        unsigned h = orc_hash(tr->nodes[n].parent, tr->nodes[n].label, ncap);
        while (nt[h] >= 0) h = (h + 1) & (ncap - 1);
        nt[h] = n;
    }
    free(tr->htab); tr->htab = nt; tr->hcap = ncap;
}
static int orc_trie_child(orc_trie* tr, int parent, int label) {
    unsigned h = orc_hash(parent, label, tr->hcap);
    while (tr->htab[h] >= 0) {
        orc_node* nd = &tr->nodes[tr->htab[h]];
        if (nd->parent == parent && nd->label == label) return tr->htab[h];
        h = (h + 1) & (tr->hcap - 1);
    }
    if (tr->n_nodes == tr->cap_nodes) {
        tr->cap_nodes *= 2;
        tr->nodes = (orc_node*)realloc(tr->nodes, sizeof(orc_node) * tr->cap_nodes);
    }
    int id = tr->n_nodes++;
    tr->nodes[id].parent = parent; tr->nodes[id].label = label; tr->nodes[id].len = tr->nodes[parent].len + 1;
    tr->htab[h] = id;
    if (tr->n_nodes * 2 > tr->hcap) orc_trie_rehash(tr);
    return id;
}

static double orc_log_add(double x, double y) {       /* BeamSearch.py:43-50 */
    if (x <= ORC_LOG_ZERO) return y;
    if (y <= ORC_LOG_ZERO) return x;
    if ((y - x) > 0.0) { double t = x; x = y; y = t; }
    return x + log(1 + exp(y - x));
}

/* stable descending sort of entry indices by pt (Python sorted(reverse=True) keeps ties in order) */
static void orc_top(const orc_entry* e, int n, int W, int* out, int* n_out) {
    int* idx = (int*)malloc(sizeof(int) * (n > 0 ? n : 1));
    for (int i = 0; i < n; ++i) idx[i] = i;
    /* insertion sort is fine for the oracle: n <= W*C + W */
    for (int i = 1; i < n; ++i) {
        int k = idx[i]; int j = i - 1;
        while (j >= 0 && e[idx[j]].pt < e[k].pt) { idx[j + 1] = idx[j]; --j; }
        idx[j + 1] = k;
    }
    int m = n < W ? n : W;
    for (int i = 0; i < m; ++i) out[i] = idx[i];
    *n_out = m;
    free(idx);
}

static int orc_find(const orc_entry* e, const int* slot_of_node, int node) { (void)e; return slot_of_node[node]; }

ORC_API int orc_beam_search(const float* probs, int B, int T, int C, const int64_t* in_len, int W, int blank,
                            const double* lm, double lm_alpha, int32_t* tokens, int tok_stride,
                            int32_t* out_len, int32_t* status, double* best_score) {
    const int cap_e = W * C + W + 4;
    orc_entry* last = (orc_entry*)malloc(sizeof(orc_entry) * cap_e);
    orc_entry* curr = (orc_entry*)malloc(sizeof(orc_entry) * cap_e);
    int* bhat = (int*)malloc(sizeof(int) * (W + 1));
    const float thr_lo = 0.1f, thr_hi = 0.9f;       /* numpy>=2: python float is weak -> float32 compare */
    for (int b = 0; b < B; ++b) {
        const float* mat = probs + (size_t)b * T * C;
        int Tb = in_len ? (int)in_len[b] : T;
        orc_trie tr; tr.cap_nodes = 1024; tr.nodes = (orc_node*)malloc(sizeof(orc_node) * tr.cap_nodes);
        tr.n_nodes = 1; tr.nodes[0].parent = -1; tr.nodes[0].label = -1; tr.nodes[0].len = 0;
        tr.hcap = 4096; tr.htab = (int*)malloc(sizeof(int) * tr.hcap);
        for (int i = 0; i < tr.hcap; ++i) tr.htab[i] = -1;
        int slot_cap = 1024; int* slot = (int*)malloc(sizeof(int) * slot_cap);   /* node -> index in curr */
        int n_last = 1, n_curr = 0, st = 0;
        last[0].node = 0; last[0].pb = 0.0; last[0].pt = 0.0; last[0].pnb = ORC_LOG_ZERO;
        for (int t = 0; t < Tb && !st; ++t) {
            const float* row = mat + (size_t)t * C;
            if ((float)(1.0f - row[blank]) < thr_lo) continue;                 /* BeamSearch.py:93 */
            int nb; orc_top(last, n_last, W, bhat, &nb);
            n_curr = 0;
            /* slot map must cover every node that can exist after this frame */
            while (tr.n_nodes + nb * C + 8 > slot_cap) { slot_cap *= 2; slot = (int*)realloc(slot, sizeof(int) * slot_cap); }
            for (int i = 0; i < slot_cap && i < tr.n_nodes + nb * C + 8; ++i) slot[i] = -1;
            for (int r = 0; r < nb && !st; ++r) {
                const orc_entry* y = &last[bhat[r]];
                const orc_node* yn = &tr.nodes[y->node];
                double pnb = ORC_LOG_ZERO;
                if (yn->len > 0) {
                    if (!(row[yn->label] > 0.0f)) { st = 2; break; }
                    pnb = y->pnb + log((double)row[yn->label]);
                }
                if (!(row[blank] > 0.0f)) { st = 2; break; }
                double pb = y->pt + log((double)row[blank]);
                int k_self = orc_find(curr, slot, y->node);
                if (k_self < 0) { k_self = n_curr++; slot[y->node] = k_self; curr[k_self].node = y->node;
                                  curr[k_self].pb = curr[k_self].pnb = curr[k_self].pt = ORC_LOG_ZERO; }
                curr[k_self].pnb = orc_log_add(curr[k_self].pnb, pnb);
                curr[k_self].pb = orc_log_add(curr[k_self].pb, pb);
                curr[k_self].pt = orc_log_add(curr[k_self].pt, orc_log_add(pb, pnb));
                for (int k = 0; k < C; ++k) {
                    if (k == blank) continue;
                    double bg = 0.0;
                    if (lm) {
                        int c1 = yn->len ? yn->label : C;
                        double v = lm[(size_t)c1 * (C + 1) + k];
                        if (v != v) { st = 3; break; }
                        bg = v * lm_alpha;
                    }
                    if (!(row[k] > 0.0f)) { st = 2; break; }
                    double base;
                    if (yn->len && yn->label == k && (mat + (size_t)(t - 1) * C)[blank] < thr_hi) base = y->pb;
                    else base = y->pt;
                    double ext = log((double)row[k]) + bg + base;              /* left-to-right adds */
                    int child = orc_trie_child(&tr, y->node, k);
                    yn = &tr.nodes[y->node];                                    /* realloc-safe */
                    if (child >= slot_cap) { int oc = slot_cap; slot_cap *= 2; slot = (int*)realloc(slot, sizeof(int) * slot_cap);
                                             for (int i = oc; i < slot_cap; ++i) slot[i] = -1; }
                    int kc = slot[child];
                    if (kc < 0) { kc = n_curr++; slot[child] = kc; curr[kc].node = child;
                                  curr[kc].pb = curr[kc].pnb = curr[kc].pt = ORC_LOG_ZERO; }
                    curr[kc].pnb = orc_log_add(curr[kc].pnb, ext);
                    curr[kc].pt = orc_log_add(curr[kc].pt, ext);
                }
            }
            orc_entry* tmp = last; last = curr; curr = tmp; n_last = n_curr;
        }
        int n_tok = 0; double best = 0.0;
        if (!st) {
            int nb; orc_top(last, n_last, W, bhat, &nb);
            /* EOS term + length norm (BeamSearch.py:130-145); first-best under a stable descending sort */
            int best_i = -1;
            for (int r = 0; r < nb; ++r) {
                const orc_entry* y = &last[bhat[r]];
                const orc_node* yn = &tr.nodes[y->node];
                if (yn->len == 0) { st = 1; break; }
                double v = lm ? lm[(size_t)yn->label * (C + 1) + C] : NAN;
                if (v != v) { st = 3; break; }
                double p = y->pt + v * lm_alpha;
                double tot = orc_log_add(ORC_LOG_ZERO, p);
                tot = tot * (1.0 / (yn->len ? yn->len : 1));
                if (best_i < 0 || tot > best) { best = tot; best_i = bhat[r]; }
            }
            if (!st && best_i >= 0) {
                int node = last[best_i].node; n_tok = tr.nodes[node].len;
                for (int i = n_tok - 1; i >= 0; --i) { if (i < tok_stride) tokens[(size_t)b * tok_stride + i] = tr.nodes[node].label; node = tr.nodes[node].parent; }
            }
        }
        out_len[b] = st ? 0 : n_tok; status[b] = st;
        if (best_score) best_score[b] = best;
        free(tr.nodes); free(tr.htab); free(slot);
    }
    free(last); free(curr); free(bhat);
    return 0;
}

/* ------------------------------------------------------------------------------------------------
 * 5. MDD post-processing on integer tokens (infer.py:162-209 align_canonical_decoded,
 *    :405-433 stastics, :338-342 counts).  Inputs: dec[Ld], can[Lc], ops[n] from orc_edit_align
 *    (hyp = decoded, ref = canonical).  Outputs are the three equal-length aligned lists with
 *    pad ids (PAD_D in the decoded row where the path says 'D', PAD_I in the canonical row where it
 *    says 'I'), after the leading-insertion trims; fault lists; counts.
 *    counts[0..5] = n_aligned, n_ins, n_sub, n_del, correct_cnt, del_sub_cnt.
 * ---------------------------------------------------------------------------------------------- */
#define ORC_PAD_D (-2)
#define ORC_PAD_I (-3)
ORC_API int orc_mdd_post(const int32_t* dec, int Ld, const int32_t* can, int Lc, const uint8_t* ops, int n,
                         int32_t* a_dec, int32_t* a_can, uint8_t* a_ops,
                         int32_t* ins, int32_t* sub, int32_t* del, int32_t* counts) {
    /* build the padded triples: every op consumes one slot in each row */
    int i = 0, j = 0;
    for (int k = 0; k < n; ++k) {
        if (ops[k] == 'D') { a_dec[k] = ORC_PAD_D; a_can[k] = j < Lc ? can[j] : -1; ++j; }
        else if (ops[k] == 'I') { a_dec[k] = i < Ld ? dec[i] : -1; a_can[k] = ORC_PAD_I; ++i; }
        else { a_dec[k] = i < Ld ? dec[i] : -1; a_can[k] = j < Lc ? can[j] : -1; ++i; ++j; }
        a_ops[k] = ops[k];
    }
    /* leading insertions: d['I'][i] == i holds exactly for the leading run of 'I' */
    int lead = 0; while (lead < n && ops[lead] == 'I') ++lead;
    int start = lead > 0 ? lead - 1 : 0;
    int m = n - start;
    /* drop the kept leading insertion if it duplicates the next decoded phone (infer.py:203-207) */
    if (m >= 1 && a_ops[start] == 'I' && a_can[start] == ORC_PAD_I && m >= 2 && a_dec[start] == a_dec[start + 1]) {
        ++start; --m;
    }
    int ni = 0, ns = 0, nd = 0, cc = 0, ds = 0;
    for (int k = 0; k < m; ++k) {
        a_dec[k] = a_dec[start + k]; a_can[k] = a_can[start + k]; a_ops[k] = a_ops[start + k];
        switch (a_ops[k]) {
            case '-': ++cc; break;
            case 'S': sub[ns++] = a_can[k]; ++ds; break;
            case 'I': ins[ni++] = a_dec[k]; break;
            default:  del[nd++] = a_can[k]; ++ds; break;
        }
    }
    counts[0] = m; counts[1] = ni; counts[2] = ns; counts[3] = nd; counts[4] = cc; counts[5] = ds;
    return 0;
}

/* ------------------------------------------------------------------------------------------------
 * 5b. Corpus MDD evaluation: the per-utterance body of egs/attention_aug/steps/test_ctc_nosil.py:218-293
 *     over integer tokens.  lc_ops = path of wer(labels, canonical), dc_ops = path of wer(decoded, canonical)
 *     (both against the same canonical row, so both hold the same number of non-'I' ops = canonical phones).
 *     print_align_space_canonical_origin (:33-82) turns a path into d[j] for every canonical position j
 *     ('-' | 'S'+hyp phone | 'D') and d['I'] = [str(j-1)+str(j) per insertion]; for consecutive integers that
 *     string determines j, so the lists hold j itself.  The 'I' branch (:265-287) removes from both lists
 *     WHILE iterating over the first (Python skips the element that slides into the freed index): restated
 *     literally.  counts[12] = TA, FR, FA, TR-correct, TR-wrong, n_canonical, FR and FA under the m_speaker
 *     rule (:270-275 add 1 instead of the list length), 0...
 * ---------------------------------------------------------------------------------------------- */
static int orc_list_find(const int* a, int n, int v) { for (int i = 0; i < n; ++i) if (a[i] == v) return i; return -1; }
static void orc_list_remove(int* a, int* n, int at) { for (int i = at; i + 1 < *n; ++i) a[i] = a[i + 1]; --*n; }
/* d code: 0 '-', 1 'D', 2 + token 'S'+token */
static int orc_dmap(const int32_t* hyp, int Lh, const uint8_t* ops, int n, int* d, int* ilist) {
    int i = 0, j = 0, ni = 0;
    for (int k = 0; k < n; ++k) {
        if (ops[k] == '-') { d[j++] = 0; ++i; }
        else if (ops[k] == 'S') { d[j++] = 2 + (i < Lh ? hyp[i] : -1); ++i; }
        else if (ops[k] == 'D') { d[j++] = 1; }
        else { ilist[ni++] = j; ++i; }
    }
    return ni;
}
ORC_API int orc_corpus_counts(const int32_t* lab, int Ll, const uint8_t* lc_ops, int n_lc,
                              const int32_t* dec, int Ld, const uint8_t* dc_ops, int n_dc, int32_t* counts) {
    int* d1 = (int*)malloc(sizeof(int) * (size_t)(n_lc + n_dc + 2) * 4);
    int* d2 = d1 + (n_lc + n_dc + 2);
    int* i1 = d2 + (n_lc + n_dc + 2);
    int* i2 = i1 + (n_lc + n_dc + 2);
    int n1 = orc_dmap(lab, Ll, lc_ops, n_lc, d1, i1);
    int n2 = orc_dmap(dec, Ld, dc_ops, n_dc, d2, i2);
    int Lc = n_lc - n1;                       /* canonical positions */
    int ta = 0, fr = 0, fa = 0, trc = 0, trw = 0;
    for (int k = 0; k < Lc; ++k) {
        if (d1[k] == 0 && d2[k] == 0) ++ta;
        else if (d1[k] == 0) ++fr;
        else if (d2[k] == 0) ++fa;
        else if (d1[k] == d2[k]) ++trc;
        else ++trw;
    }
    int m_fr = fr, m_fa = fa;
    if (n1 && !n2) { fa += n1; m_fa += 1; }
    else if (!n1 && n2) { fr += n2; m_fr += 1; }
    else if (n1 && n2) {
        for (int idx = 0; idx < n1; ++idx) {                  /* for e in d1['I'] (live list) */
            int e = i1[idx];
            int at2 = orc_list_find(i2, n2, e);
            if (at2 >= 0) {
                orc_list_remove(i1, &n1, orc_list_find(i1, n1, e));
                orc_list_remove(i2, &n2, at2);
                ++trc;
            }
        }
        fa += n1; fr += n2; m_fa += n1; m_fr += n2;
    }
    counts[0] = ta; counts[1] = fr; counts[2] = fa; counts[3] = trc; counts[4] = trw; counts[5] = Lc;
    counts[6] = m_fr; counts[7] = m_fa;
    for (int k = 8; k < 12; ++k) counts[k] = 0;
    free(d1);
    return 0;
}

/* ------------------------------------------------------------------------------------------------
 * 6. CTC prefix scorer (not in the reference).  One scoring step:
 *    x: (T,C) log-probs of one utterance (fp64), r_prev: (T,2) [r_n, r_b] of hypothesis g,
 *    last = last label of g (-1 if empty), glen = |g|.  For every candidate c (0..C-1):
 *    psi[c] = log p_ctc(g.c ...), r_out[c]: (T,2).  psi[blank] = -inf.  psi_eos = r_sum[T_b-1].
 * ---------------------------------------------------------------------------------------------- */
static double orc_lse2d(double a, double b) {
    if (a == -INFINITY) return b;
    if (b == -INFINITY) return a;
    double m = a > b ? a : b;
    return m + log(exp(a - m) + exp(b - m));
}
ORC_API int orc_prefix_init(const double* x, int Tb, int C, int blank, double* r0) {
    double acc = 0;
    for (int t = 0; t < Tb; ++t) { acc += x[(size_t)t * C + blank]; r0[2 * t] = -INFINITY; r0[2 * t + 1] = acc; }
    return 0;
}
ORC_API int orc_prefix_score(const double* x, int Tb, int C, int blank, const double* r_prev, int last, int glen,
                             double* psi, double* r_out, double* psi_eos) {
    int start = glen > 1 ? glen : 1;
    for (int c = 0; c < C; ++c) {
        double* r = r_out ? r_out + (size_t)c * Tb * 2 : NULL;
        if (c == blank) { psi[c] = -INFINITY; if (r) for (int t = 0; t < 2 * Tb; ++t) r[t] = -INFINITY; continue; }
        double rn = -INFINITY, rb = -INFINITY;
        if (r) for (int t = 0; t < 2 * Tb; ++t) r[t] = -INFINITY;
        if (start - 1 < Tb) {
            if (glen == 0) rn = x[c];                      /* r_n[0] = x[0,c] */
            if (r) { r[2 * (start - 1)] = rn; r[2 * (start - 1) + 1] = rb; }
        }
        double ps = rn;
        for (int t = start; t < Tb; ++t) {
            double rsum = orc_lse2d(r_prev[2 * (t - 1)], r_prev[2 * (t - 1) + 1]);
            double phi = (c == last) ? r_prev[2 * (t - 1) + 1] : rsum;
            double nrn = orc_lse2d(rn, phi) + x[(size_t)t * C + c];
            double nrb = orc_lse2d(rn, rb) + x[(size_t)t * C + blank];
            ps = orc_lse2d(ps, phi + x[(size_t)t * C + c]);
            rn = nrn; rb = nrb;
            if (r) { r[2 * t] = rn; r[2 * t + 1] = rb; }
        }
        psi[c] = ps;
    }
    if (psi_eos) *psi_eos = Tb > 0 ? orc_lse2d(r_prev[2 * (Tb - 1)], r_prev[2 * (Tb - 1) + 1]) : -INFINITY;
    return 0;
}

/* ------------------------------------------------------------------------------------------------
 * 7. One step of the label-synchronous joint CTC/attention beam search (not in the
 *    reference -- nearest code: the sort + [:beamWidth] of BeamSearch.py:52-57 around the recursion above).
 *    Per utterance, over the W hypothesis slots and C labels:
 *      value(w,c) = score[w] + lam * (psi(g_w.c) - psi(g_w)) + (1 - lam) * att[w,c]       c != blank
 *    c == eos is scored with psi_eos(g_w); a finished slot (flag 1) is carried unchanged as its (w, eos) candidate;
 *    dead slots (flag 2) and -inf / NaN values yield nothing.  The W best (ties: smaller w*C + c) form the new
 *    beam; a slot without a candidate is dead.  State per slot: r (T,2), last, glen, psi, score, flag,
 *    tokens (max_len, -1 padded).  x: (T,B,C) fp64, lens[B]; all state arrays are (B,W,...) and are updated in place.
 * ---------------------------------------------------------------------------------------------- */
#include <pthread.h>
struct orc_joint_args {
    const double* x; int T, B, C; const int64_t* lens; int blank, eos, W; double lam; int max_len; const float* att;
    double* r; int32_t* last; int32_t* glen; double* psi_st; double* score; int32_t* flag; int32_t* tokens;
    int first, stride, rc;
};
static void* orc_joint_worker(void* vp) {
    struct orc_joint_args* a = (struct orc_joint_args*)vp;
    const double* x = a->x; const int T = a->T, B = a->B, C = a->C, blank = a->blank, eos = a->eos, W = a->W;
    const int max_len = a->max_len; const double lam = a->lam; const float* att = a->att; const int64_t* lens = a->lens;
    double* r = a->r; int32_t* last = a->last; int32_t* glen = a->glen; double* psi_st = a->psi_st;
    double* score = a->score; int32_t* flag = a->flag; int32_t* tokens = a->tokens;
    for (int b = a->first; b < B; b += a->stride) {
        int Tb = lens ? (int)lens[b] : T;
        if (Tb < 0) Tb = 0;
        if (Tb > T) Tb = T;
        double* xb = (double*)malloc(sizeof(double) * (size_t)(Tb > 0 ? Tb : 1) * C);
        double* psi = (double*)malloc(sizeof(double) * (size_t)W * C);
        double* pe = (double*)malloc(sizeof(double) * W);
        double* rall = (double*)malloc(sizeof(double) * (size_t)W * C * (Tb > 0 ? Tb : 1) * 2);
        double* cand = (double*)malloc(sizeof(double) * (size_t)W * C);
        double* r_new = (double*)malloc(sizeof(double) * (size_t)W * T * 2);
        int32_t* tok_new = (int32_t*)malloc(sizeof(int32_t) * (size_t)W * max_len);
        int32_t n_last[64], n_glen[64], n_flag[64];
        double n_psi[64], n_score[64];
        if (!xb || !psi || !pe || !rall || !cand || !r_new || !tok_new || W > 64) { a->rc = -1; goto done; }
        for (int t = 0; t < Tb; ++t) memcpy(xb + (size_t)t * C, x + ((size_t)t * B + b) * C, sizeof(double) * C);
        for (int w = 0; w < W; ++w) {
            const int s = b * W + w;
            for (int c = 0; c < C; ++c) cand[w * C + c] = -INFINITY;
            if (flag[s] == 1) { cand[w * C + (eos >= 0 ? eos : blank)] = score[s]; continue; }
            if (flag[s] != 0 || !(score[s] > -INFINITY)) continue;
            /* the slot's r lives at stride T in the state; orc_prefix_score wants (Tb,2): the first Tb rows */
            orc_prefix_score(xb, Tb, C, blank, r + (size_t)s * T * 2, last[s], glen[s], psi + (size_t)w * C,
                             rall + (size_t)w * C * Tb * 2, &pe[w]);
            for (int c = 0; c < C; ++c) {
                if (c == blank) continue;
                const double pc = (c == eos) ? pe[w] : psi[w * C + c];
                double v = -INFINITY;
                if (pc > -INFINITY && psi_st[s] > -INFINITY)
                    v = score[s] + lam * (pc - psi_st[s]) + (1.0 - lam) * (double)att[(size_t)s * C + c];
                if (v != v) v = -INFINITY;
                cand[w * C + c] = v;
            }
        }
        for (int k = 0; k < W; ++k) {
            int bi = -1;
            double bv = -INFINITY;
            for (int i = 0; i < W * C; ++i)
                if (cand[i] > bv) { bv = cand[i]; bi = i; }
            int32_t* tk = tok_new + (size_t)k * max_len;
            if (bi < 0) {
                n_last[k] = -1; n_glen[k] = 0; n_flag[k] = 2; n_psi[k] = -INFINITY; n_score[k] = -INFINITY;
                for (int i = 0; i < max_len; ++i) tk[i] = -1;
                continue;
            }
            const int w = bi / C, c = bi % C, s = b * W + w;
            const int pf = flag[s] != 0;
            cand[bi] = -INFINITY;
            n_last[k] = c;
            n_glen[k] = glen[s] + (pf ? 0 : 1);
            n_flag[k] = (pf || c == eos) ? 1 : 0;
            n_score[k] = bv;
            n_psi[k] = pf ? psi_st[s] : (c == eos ? pe[w] : psi[w * C + c]);
            memcpy(tk, tokens + (size_t)s * max_len, sizeof(int32_t) * max_len);
            if (!pf && glen[s] < max_len) tk[glen[s]] = c;
            double* rn = r_new + (size_t)k * T * 2;
            for (int t = 0; t < 2 * T; ++t) rn[t] = -INFINITY;
            if (!n_flag[k]) memcpy(rn, rall + ((size_t)w * C + c) * Tb * 2, sizeof(double) * Tb * 2);
        }
        for (int k = 0; k < W; ++k) {
            const int s = b * W + k;
            last[s] = n_last[k]; glen[s] = n_glen[k]; flag[s] = n_flag[k]; psi_st[s] = n_psi[k]; score[s] = n_score[k];
        }
        memcpy(r + (size_t)b * W * T * 2, r_new, sizeof(double) * (size_t)W * T * 2);
        memcpy(tokens + (size_t)b * W * max_len, tok_new, sizeof(int32_t) * (size_t)W * max_len);
    done:
        free(xb); free(psi); free(pe); free(rall); free(cand); free(r_new); free(tok_new);
    }
    return NULL;
}
ORC_API int orc_joint_step(const double* x, int T, int B, int C, const int64_t* lens, int blank, int eos, int W,
                           double lam, int max_len, const float* att, double* r, int32_t* last, int32_t* glen,
                           double* psi_st, double* score, int32_t* flag, int32_t* tokens) {
    enum { NTH = 16 };
    pthread_t th[NTH];
    struct orc_joint_args args[NTH];
    int n = B < NTH ? B : NTH, rc = 0;
    for (int i = 0; i < n; ++i) {
        struct orc_joint_args a = {x, T, B, C, lens, blank, eos, W, lam, max_len, att, r, last, glen, psi_st, score, flag,
                                   tokens, i, n, 0};
        args[i] = a;
        if (pthread_create(&th[i], NULL, orc_joint_worker, &args[i]) != 0) { orc_joint_worker(&args[i]); th[i] = 0; }
    }
    for (int i = 0; i < n; ++i) { if (th[i]) pthread_join(th[i], NULL); if (args[i].rc) rc = args[i].rc; }
    return rc;
}
