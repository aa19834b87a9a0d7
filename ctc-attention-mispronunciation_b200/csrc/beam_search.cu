// beam_search.cu -- CTC prefix beam search with bigram-LM fusion, bit-compatible with the reference decoder.
//
// Replaces ctcBeamSearch.decode (egs/attention_aug/utils/BeamSearch.py:73-153; calcExtPr :52-66, log_add_prob
// :43-50, BeamState.sort/norm :23-33) as called by BeamDecoder.decode (utils/ctcDecoder.py:215-226).
//
// One CTA per utterance.  Facts about the reference loop that make a parallel evaluation exact:
//   * only the top-W prefixes of a frame (BHat) are ever read again, so the state is W entries in sorted order;
//   * a prefix of the next frame receives at most two contributions: its own update (if it is in BHat) and the
//     extension of its parent (if the parent is in BHat); they are applied in BHat-rank order because
//     log_add_prob is symmetric except when an operand is <= LOG_ZERO;
//   * the dict insertion order that breaks ties in sorted() is (rank of the touching BHat prefix, 0 for the
//     prefix itself / 1+k for extension k), minimised over the contributions;
//   * prefix identity (tuple equality) is a node of a per-utterance trie; nodes are canonical through a
//     (parent,label) hash table in the workspace, so a prefix that leaves the beam and is re-created gets the
//     same id.
// Scores are float64 computed from float32 probabilities exactly in the reference's operation order
// ((log p + lm*alpha) + base; x + log(1 + exp(y - x))); thresholds use numpy>=2 float32 semantics.
#include "common.cuh"

namespace mdd {
namespace {

constexpr int NT = 128;
constexpr double LOG_ZERO = -99999999.0;

struct BeamParams {
    const float* probs;     // (B,T,C)
    int B, T, C, W, blank;
    const int64_t* in_len;
    const double* lm;       // (C+1)^2 or null
    double lm_alpha;
    int32_t* tokens;
    int tok_stride;
    int32_t* out_len;
    int32_t* status;
    double* score;
    unsigned char* ws;
    size_t ws_per_utt;
    int node_cap, hash_cap;
};

__device__ __forceinline__ double log_add(double x, double y) {     // BeamSearch.py:43-50
    if (x <= LOG_ZERO) return y;
    if (y <= LOG_ZERO) return x;
    if ((y - x) > 0.0) { double t = x; x = y; y = t; }
    return __dadd_rn(x, log(__dadd_rn(1.0, exp(__dsub_rn(y, x)))));
}

__device__ __forceinline__ unsigned hash_pl(int parent, int label) {
    unsigned long long x = ((unsigned long long)(unsigned)parent << 8) ^ (unsigned)label;
    x *= 0x9E3779B97F4A7C15ull;
    x ^= x >> 29;
    return (unsigned)x;
}

struct Beam {            // W entries, sorted by rank
    int* node; int* label; int* len; int* parent;
    double* pb; double* pnb; double* pt;
};

__global__ void __launch_bounds__(NT) beam_search_kernel(BeamParams p) {
    const int b = blockIdx.x, tid = threadIdx.x;
    const int C = p.C, W = p.W, blank = p.blank, T = p.T;
    const int NS = W * C;
    const float* mat = p.probs + (size_t)b * T * C;
    int Tb = p.in_len ? (int)max(0ll, min((long long)p.in_len[b], (long long)T)) : T;

    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* c_pt = reinterpret_cast<double*>(smem_raw);        // [NS]
    double* lg = c_pt + NS;                                    // [C]
    double* bm_d = lg + C;                                     // 2 beams x 3 x W doubles
    double* red_v = bm_d + 6 * W;                              // [NT/32]
    int* bm_i = reinterpret_cast<int*>(red_v + NT / 32);       // 2 beams x 4 x W ints
    int* merge_with = bm_i + 8 * W;                            // [W]  rank of the parent that extends into me, or -1
    int* self_order = merge_with + W;                          // [W]
    int* sel_slot = self_order + W;                            // [W]
    int* red_o = sel_slot + W;                                 // [NT/32]
    int* red_s = red_o + NT / 32;                              // [NT/32]
    float* rowbuf = reinterpret_cast<float*>(red_s + NT / 32); // [2][C]
    unsigned short* frames = reinterpret_cast<unsigned short*>(rowbuf + 2 * C);   // [T] processed frame list
    unsigned char* c_dead = reinterpret_cast<unsigned char*>(frames + T + (T & 1));   // [NS]
    unsigned char* prevlt = c_dead + NS;                       // [T] mat[t-1,blank] < 0.9 for processed frames

    __shared__ int s_nframes, s_err, s_nb, s_nnodes, s_win_slot;

    Beam last{bm_i, bm_i + W, bm_i + 2 * W, bm_i + 3 * W, bm_d, bm_d + W, bm_d + 2 * W};
    Beam next{bm_i + 4 * W, bm_i + 5 * W, bm_i + 6 * W, bm_i + 7 * W, bm_d + 3 * W, bm_d + 4 * W, bm_d + 5 * W};

    // workspace: trie nodes + hash table
    unsigned char* ws = p.ws + (size_t)b * p.ws_per_utt;
    int2* nodes = reinterpret_cast<int2*>(ws);                                 // (parent, label | len << 8)
    int* htab = reinterpret_cast<int*>(ws + (size_t)p.node_cap * sizeof(int2));
    for (int i = tid; i < p.hash_cap; i += NT) htab[i] = -1;
    if (tid == 0) {
        nodes[0] = make_int2(-1, 0);
        s_nframes = 0; s_err = 0x7fffffff; s_nb = 1; s_nnodes = 1;
        last.node[0] = 0; last.label[0] = -1; last.len[0] = 0; last.parent[0] = -1;
        last.pb[0] = 0.0; last.pt[0] = 0.0; last.pnb[0] = LOG_ZERO;
    }
    __syncthreads();

    // ---- pre-pass: which frames are processed (BeamSearch.py:93) and the repeat-rule flag (BeamSearch.py:63)
    for (int base = 0; base < Tb; base += NT) {
        const int t = base + tid;
        bool proc = false, plt = false;
        if (t < Tb) {
            const float pbk = mat[(size_t)t * C + blank];
            proc = !(__fsub_rn(1.0f, pbk) < 0.1f);
            const int tp = t > 0 ? t - 1 : T - 1;                  // python negative index wraps to the padded end
            plt = mat[(size_t)tp * C + blank] < 0.9f;
        }
        // ordered compaction across the block
        const unsigned m = __ballot_sync(0xffffffffu, proc);
        __shared__ int wcnt[NT / 32];
        if ((tid & 31) == 0) wcnt[tid >> 5] = __popc(m);
        __syncthreads();
        int off = s_nframes;
        for (int w = 0; w < (tid >> 5); ++w) off += wcnt[w];
        if (proc) {
            const int pos = off + __popc(m & ((1u << (tid & 31)) - 1));
            frames[pos] = (unsigned short)t;
            prevlt[pos] = plt;
        }
        __syncthreads();
        if (tid == 0) { int s = 0; for (int w = 0; w < NT / 32; ++w) s += wcnt[w]; s_nframes += s; }
        __syncthreads();
    }
    const int nframes = s_nframes;
    __threadfence_block();

    if (nframes > 0 && tid < C) rowbuf[tid] = mat[(size_t)frames[0] * C + tid];
    __syncthreads();

    for (int f = 0; f < nframes; ++f) {
        const float* row = rowbuf + (f & 1) * C;
        const int nb = s_nb;
        const bool plt = prevlt[f] != 0;
        // prefetch the next processed row
        if (f + 1 < nframes && tid < C) rowbuf[((f + 1) & 1) * C + tid] = mat[(size_t)frames[f + 1] * C + tid];
        // S0: logs, merge map
        if (tid < C) lg[tid] = log((double)row[tid]);
        for (int r = tid; r < nb; r += NT) {
            int mw = -1;
            if (last.len[r] > 0) {
                const int par = last.parent[r];
                for (int q = 0; q < nb; ++q) if (last.node[q] == par) mw = q;
            }
            merge_with[r] = mw;
        }
        __syncthreads();
        // S1: evaluate every candidate slot (r,k); slot (r,blank) is the prefix itself
        for (int s = tid; s < nb * C; s += NT) {
            const int r = s / C, k = s - r * C;
            const int lab = last.label[r], len = last.len[r];
            double pt;
            bool dead = false;
            if (k == blank) {
                // error positions: own-label log (0), blank log (1)
                if (len > 0 && row[lab] <= 0.f) atomicMin(&s_err, ((r * (2 * C + 4) + 0) << 3) | MDD_ST_LOG0);
                if (row[blank] <= 0.f) atomicMin(&s_err, ((r * (2 * C + 4) + 1) << 3) | MDD_ST_LOG0);
                const double pnb_s = len > 0 ? __dadd_rn(last.pnb[r], lg[lab]) : LOG_ZERO;
                const double pb_s = __dadd_rn(last.pt[r], lg[blank]);
                const double tot_s = log_add(pb_s, pnb_s);
                const int rp = merge_with[r];
                int order = r * (C + 1);
                if (rp >= 0) {
                    double bg = 0.0;
                    if (p.lm) {
                        const int c1 = last.len[rp] ? last.label[rp] : C;
                        bg = __dmul_rn(p.lm[(size_t)c1 * (C + 1) + lab], p.lm_alpha);
                    }
                    const bool rep = last.len[rp] && last.label[rp] == lab && plt;
                    const double basev = rep ? last.pb[rp] : last.pt[rp];
                    const double e = __dadd_rn(__dadd_rn(lg[lab], bg), basev);
                    if (rp < r) {       // extension touched the entry first
                        pt = log_add(log_add(LOG_ZERO, e), tot_s);
                        order = rp * (C + 1) + 1 + lab;
                    } else {
                        pt = log_add(log_add(LOG_ZERO, tot_s), e);
                    }
                } else {
                    pt = log_add(LOG_ZERO, tot_s);
                }
                self_order[r] = order;
            } else {
                double bg = 0.0;
                if (p.lm) {
                    const int c1 = len ? lab : C;
                    const double v = p.lm[(size_t)c1 * (C + 1) + k];
                    if (v != v) atomicMin(&s_err, ((r * (2 * C + 4) + 2 + 2 * k) << 3) | MDD_ST_LMKEY);
                    bg = __dmul_rn(v, p.lm_alpha);
                }
                if (row[k] <= 0.f) atomicMin(&s_err, ((r * (2 * C + 4) + 3 + 2 * k) << 3) | MDD_ST_LOG0);
                const bool rep = len && lab == k && plt;
                const double basev = rep ? last.pb[r] : last.pt[r];
                pt = log_add(LOG_ZERO, __dadd_rn(__dadd_rn(lg[k], bg), basev));
                // merged into the self slot of the child prefix if that child is in BHat
                for (int q = 0; q < nb; ++q)
                    if (merge_with[q] == r && last.label[q] == k) dead = true;
            }
            c_pt[s] = pt;
            c_dead[s] = dead;
        }
        __syncthreads();
        if (s_err != 0x7fffffff) break;
        // S2: top-W by (pt desc, insertion order asc)
        const int n_slots = nb * C;
        int n_sel = 0;
        for (int round = 0; round < W; ++round) {
            double bv = 0.0; int bo = 0x7fffffff, bs = -1;
            for (int s = tid; s < n_slots; s += NT) {
                if (c_dead[s]) continue;
                const int r = s / C, k = s - r * C;
                const int o = (k == blank) ? self_order[r] : r * (C + 1) + 1 + k;
                const double v = c_pt[s];
                if (bs < 0 || v > bv || (v == bv && o < bo)) { bv = v; bo = o; bs = s; }
            }
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) {
                const double ov = __shfl_xor_sync(0xffffffffu, bv, d);
                const int oo = __shfl_xor_sync(0xffffffffu, bo, d);
                const int os = __shfl_xor_sync(0xffffffffu, bs, d);
                if (os >= 0 && (bs < 0 || ov > bv || (ov == bv && oo < bo))) { bv = ov; bo = oo; bs = os; }
            }
            if ((tid & 31) == 0) { red_v[tid >> 5] = bv; red_o[tid >> 5] = bo; red_s[tid >> 5] = bs; }
            __syncthreads();
            if (tid == 0) {
                for (int w = 1; w < NT / 32; ++w) {
                    const double ov = red_v[w]; const int oo = red_o[w], os = red_s[w];
                    if (os >= 0 && (bs < 0 || ov > bv || (ov == bv && oo < bo))) { bv = ov; bo = oo; bs = os; }
                }
                s_win_slot = bs;
                if (bs >= 0) { sel_slot[round] = bs; c_dead[bs] = 1; }
            }
            __syncthreads();
            if (s_win_slot < 0) break;
            ++n_sel;
        }
        // S3: materialise the selected entries (recompute pb/pnb in the reference's order) and resolve trie nodes
        for (int i = tid; i < n_sel; i += NT) {
            const int s = sel_slot[i];
            const int r = s / C, k = s - r * C;
            const int lab = last.label[r], len = last.len[r];
            if (k == blank) {
                const double pnb_s = len > 0 ? __dadd_rn(last.pnb[r], lg[lab]) : LOG_ZERO;
                const double pb_s = __dadd_rn(last.pt[r], lg[blank]);
                double pnb = log_add(LOG_ZERO, pnb_s);
                const int rp = merge_with[r];
                if (rp >= 0) {
                    double bg = 0.0;
                    if (p.lm) {
                        const int c1 = last.len[rp] ? last.label[rp] : C;
                        bg = __dmul_rn(p.lm[(size_t)c1 * (C + 1) + lab], p.lm_alpha);
                    }
                    const bool rep = last.len[rp] && last.label[rp] == lab && plt;
                    const double basev = rep ? last.pb[rp] : last.pt[rp];
                    const double e = __dadd_rn(__dadd_rn(lg[lab], bg), basev);
                    pnb = (rp < r) ? log_add(log_add(LOG_ZERO, e), pnb_s) : log_add(pnb, e);
                }
                next.node[i] = last.node[r]; next.label[i] = lab; next.len[i] = len; next.parent[i] = last.parent[r];
                next.pb[i] = log_add(LOG_ZERO, pb_s); next.pnb[i] = pnb; next.pt[i] = c_pt[s];
            } else {
                const int par = last.node[r];
                const int packed = k | ((len + 1) << 8);
                unsigned h = hash_pl(par, k) & (unsigned)(p.hash_cap - 1);
                int id = -1, mine = -1;
                while (true) {
                    int cur = __ldcg(&htab[h]);
                    if (cur == -1) {
                        if (mine < 0) {
                            mine = atomicAdd(&s_nnodes, 1);
                            if (mine < p.node_cap) { nodes[mine] = make_int2(par, packed); __threadfence(); }
                        }
                        if (mine >= p.node_cap) { id = 0; atomicMin(&s_err, MDD_ST_BADLEN); break; }
                        const int old = atomicCAS(&htab[h], -1, mine);
                        if (old == -1) { id = mine; break; }
                        cur = old;
                    }
                    const int2 nd = __ldcg(&nodes[cur]);
                    if (nd.x == par && nd.y == packed) { id = cur; break; }
                    h = (h + 1) & (unsigned)(p.hash_cap - 1);
                }
                next.node[i] = id; next.label[i] = k; next.len[i] = len + 1; next.parent[i] = par;
                next.pb[i] = LOG_ZERO; next.pnb[i] = c_pt[s]; next.pt[i] = c_pt[s];
            }
        }
        __syncthreads();
        for (int i = tid; i < n_sel; i += NT) {
            last.node[i] = next.node[i]; last.label[i] = next.label[i]; last.len[i] = next.len[i];
            last.parent[i] = next.parent[i];
            last.pb[i] = next.pb[i]; last.pnb[i] = next.pnb[i]; last.pt[i] = next.pt[i];
        }
        if (tid == 0) s_nb = n_sel;
        __threadfence();
        __syncthreads();
    }

    // ---- end of sentence LM term, length normalisation, best prefix (BeamSearch.py:127-152)
    if (tid == 0) {
        int st = MDD_ST_OK, n_tok = 0;
        double best = 0.0;
        if (s_err != 0x7fffffff) st = s_err & 7;
        else {
            int best_r = -1;
            const int nb = s_nb;
            for (int r = 0; r < nb; ++r) {
                const int len = last.len[r];
                if (len == 0) { st = MDD_ST_EMPTY; break; }
                double v = 0.0;
                if (p.lm) {
                    v = p.lm[(size_t)last.label[r] * (C + 1) + C];
                    if (v != v) { st = MDD_ST_LMKEY; break; }
                }
                double tot = log_add(LOG_ZERO, __dadd_rn(last.pt[r], __dmul_rn(v, p.lm_alpha)));
                tot = __dmul_rn(tot, __ddiv_rn(1.0, (double)len));
                if (best_r < 0 || tot > best) { best = tot; best_r = r; }
            }
            if (st == MDD_ST_OK && best_r >= 0) {
                int node = last.node[best_r];
                n_tok = last.len[best_r];
                int32_t* tk = p.tokens + (size_t)b * p.tok_stride;
                for (int i = n_tok - 1; i >= 0 && node > 0; --i) {
                    const int2 nd = __ldcg(&nodes[node]);
                    if (i < p.tok_stride) tk[i] = nd.y & 0xff;
                    node = nd.x;
                }
                if (node != 0) st = MDD_E_WORKSPACE;     // trie overflow (cannot happen with the documented workspace size)
            }
        }
        p.out_len[b] = st == MDD_ST_OK ? n_tok : 0;
        p.status[b] = st;
        if (p.score) p.score[b] = best;
    }
}

size_t beam_smem_bytes(int T, int C, int W) {
    const size_t NS = (size_t)W * C;
    size_t bytes = (NS + C + 6 * W + NT / 32) * 8;
    bytes += (8 * W + 3 * W + 2 * (NT / 32)) * 4;
    bytes += 2 * C * 4;
    bytes += (T + (T & 1)) * 2;
    bytes += NS + T;
    return align_up(bytes, 16);
}

void beam_ws_layout(int T, int W, int* node_cap, int* hash_cap, size_t* per_utt) {
    const long long nc = (long long)W * T + 2;
    long long hc = 1024;
    while (hc < 2 * nc) hc <<= 1;
    *node_cap = (int)nc;
    *hash_cap = (int)hc;
    *per_utt = align_up((size_t)nc * sizeof(int2) + (size_t)hc * 4, 256);
}

}  // namespace
}  // namespace mdd

using namespace mdd;

extern "C" int mdd_beam_search_workspace_bytes(int B, int T, int C, int beam_width, size_t* bytes) {
    if (!bytes || B < 0 || T < 0 || C <= 0 || beam_width <= 0) return MDD_E_ARG;
    int nc, hc; size_t per;
    beam_ws_layout(T, beam_width, &nc, &hc, &per);
    *bytes = per * (size_t)B;
    return MDD_OK;
}

extern "C" int mdd_beam_search(const float* probs, int B, int T, int C, const int64_t* input_lengths, int beam_width,
                               int blank, const double* lm, double lm_alpha, int32_t* tokens, int tok_stride,
                               int32_t* out_len, int32_t* status, double* score, void* workspace,
                               size_t workspace_bytes, mdd_stream_t stream) {
    if (B == 0) return MDD_OK;
    if (B < 0 || T < 0 || C <= 0 || C > MDD_MAX_CLASSES || beam_width <= 0 || blank < 0 || blank >= C) return MDD_E_ARG;
    if ((!probs && T > 0) || !tokens || !out_len || !status || !workspace) return MDD_E_ARG;
    if (T > 65535) return MDD_E_UNSUPPORTED;
    BeamParams p{};
    p.probs = probs; p.B = B; p.T = T; p.C = C; p.W = beam_width; p.blank = blank;
    p.in_len = input_lengths; p.lm = lm; p.lm_alpha = lm_alpha;
    p.tokens = tokens; p.tok_stride = tok_stride; p.out_len = out_len; p.status = status; p.score = score;
    beam_ws_layout(T, beam_width, &p.node_cap, &p.hash_cap, &p.ws_per_utt);
    if (workspace_bytes < p.ws_per_utt * (size_t)B) return MDD_E_WORKSPACE;
    p.ws = (unsigned char*)workspace;
    const size_t smem = beam_smem_bytes(T, C, beam_width);
    if (smem > 200 * 1024) return MDD_E_UNSUPPORTED;
    if (smem > 48 * 1024 &&
        cudaFuncSetAttribute(beam_search_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
        cudaGetLastError();
        return MDD_E_CUDA;
    }
    beam_search_kernel<<<B, NT, smem, (cudaStream_t)stream>>>(p);
    MDD_CHECK_LAUNCH();
    return MDD_OK;
}
