// edit_align.cu -- batched Levenshtein distance + op-path traceback + MDD post-processing, one warp per pair.
//
// Replaces Decoder.wer/_edit_distance/printChanges (egs/attention_aug/utils/ctcDecoder.py:118-184), the
// editdistance.eval call of compute_wer (models/model_ctc.py:240) and the per-utterance post-processing of
// infer.py (:310-342 filtering + counts, :162-209 align_canonical_decoded, :405-433 stastics).
//
// Rows i = hypothesis (s1), columns j = reference (s2).  Lane l owns WPL consecutive columns; the DP runs as a
// skewed wavefront (lane l handles row k-l at step k) with the column-boundary values handed over by
// __shfl_up_sync, so there is no block barrier in the recurrence.  Each cell stores only its 2-bit traceback
// decision, chosen with the reference's priority ('-' if equal, else S if d==diag+1, else I if d==up+1, else D);
// decisions are packed per (row, lane) and kept in shared memory (global workspace for very long pairs).
#include "common.cuh"

namespace mdd {
namespace {

template <int WPL> struct DirWord;
template <> struct DirWord<1> { using type = uint8_t; };
template <> struct DirWord<2> { using type = uint8_t; };
template <> struct DirWord<4> { using type = uint8_t; };
template <> struct DirWord<8> { using type = uint16_t; };
template <> struct DirWord<16> { using type = uint32_t; };
template <> struct DirWord<32> { using type = unsigned long long; };

struct EditParams {
    const int32_t *hyp, *hyp_len, *ref, *ref_len;
    int hyp_stride, ref_stride, B, L1max, L2max;
    const int32_t *hyp_map, *ref_map;
    int n_map;
    int32_t* dist;
    uint8_t* ops;
    int ops_stride;
    int32_t* ops_len;
    int32_t* status;
    mdd_post_out post;
    int has_post;
    const uint8_t* ops_in;      // optional caller-supplied op paths (skips the DP): (B, ops_in_stride)
    int ops_in_stride;
    const int32_t* ops_in_len;
    unsigned char* dir_ws;      // global direction storage (null -> shared memory)
    size_t dir_bytes_per_pair;
    size_t smem_per_warp;
};

// load + optional map + compaction of one token row into shared memory; returns the filtered length
__device__ int load_tokens(const int32_t* src, int n, const int32_t* map, int n_map, int32_t* dst, int lane) {
    int out = 0;
    for (int base = 0; base < n; base += 32) {
        int k = base + lane;
        int v = -1;
        bool keep = false;
        if (k < n) {
            v = src[k];
            if (map) v = (v >= 0 && v < n_map) ? map[v] : -1;
            keep = map ? (v >= 0) : true;
        }
        unsigned m = __ballot_sync(0xffffffffu, keep);
        if (keep) dst[out + __popc(m & ((1u << lane) - 1))] = v;
        out += __popc(m);
    }
    __syncwarp();
    return out;
}

template <int WPL>
__global__ void __launch_bounds__(128) edit_align_kernel(EditParams p) {
    using DirT = typename DirWord<WPL>::type;
    const int warps_per_cta = blockDim.x >> 5;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int b = blockIdx.x * warps_per_cta + warp;
    if (b >= p.B) return;

    extern __shared__ __align__(16) unsigned char smem_raw[];
    unsigned char* base = smem_raw + (size_t)warp * p.smem_per_warp;
    int32_t* hs = reinterpret_cast<int32_t*>(base);                  // [L1max]
    int32_t* rs = hs + p.L1max;                                      // [L2max]
    const size_t nops = align_up((size_t)p.L1max + p.L2max, 16);
    uint8_t* rev = base + align_up(((size_t)p.L1max + p.L2max) * 4, 16);   // [L1max+L2max] reversed ops
    uint8_t* fwd = rev + nops;                                             // [L1max+L2max] forward ops
    DirT* dirs = p.dir_ws ? reinterpret_cast<DirT*>(p.dir_ws + (size_t)b * p.dir_bytes_per_pair)
                          : reinterpret_cast<DirT*>(fwd + nops);           // [L1][32]

    int n1 = p.hyp_len[b], n2 = p.ref_len[b];
    n1 = max(0, min(n1, p.L1max));
    n2 = max(0, min(n2, p.L2max));
    const int L1 = load_tokens(p.hyp + (size_t)b * p.hyp_stride, n1, p.hyp_map, p.n_map, hs, lane);
    const int L2 = load_tokens(p.ref + (size_t)b * p.ref_stride, n2, p.ref_map, p.n_map, rs, lane);
    if (p.has_post && lane == 0) { p.post.counts[(size_t)b * 8 + 6] = L1; p.post.counts[(size_t)b * 8 + 7] = L2; }

    if (!p.ops_in && (L1 == 0 || L2 == 0)) {       // ctcDecoder.py:137-138: bare int return -> TypeError in wer()
        if (lane == 0) {
            p.dist[b] = L1 == 0 ? L2 : L1;
            if (p.status) p.status[b] = MDD_ST_EMPTY;
            if (p.ops_len) p.ops_len[b] = 0;
            if (p.has_post) for (int q = 0; q < 6; ++q) p.post.counts[(size_t)b * 8 + q] = 0;
        }
        return;
    }

    int n = 0;
    if (p.ops_in) {             // post-processing of a caller-supplied path (align_canonical_decoded drop-in)
        n = max(0, min(p.ops_in_len[b], p.L1max + p.L2max));
        for (int k = lane; k < n; k += 32) fwd[k] = p.ops_in[(size_t)b * p.ops_in_stride + k];
        __syncwarp();
    } else {
    // ---------------------------------------------------------------- DP wavefront
    int refv[WPL];
    int cur[WPL];          // d[i-1][j] for my columns
#pragma unroll
    for (int q = 0; q < WPL; ++q) {
        const int j = lane * WPL + q + 1;
        refv[q] = j <= L2 ? rs[j - 1] : -0x7fffffff;
        cur[q] = j;
    }
    int prev_left = 0;     // d[i-1][lane*WPL] as received in the previous step
    int my_last = lane * WPL + WPL;   // d[row handled last step][my last column]; row 0 initially
    const int steps = L1 + 31;
    for (int k = 0; k < steps; ++k) {
        int left_in = __shfl_up_sync(0xffffffffu, my_last, 1);
        const int i = k - lane + 1;
        if (lane == 0) left_in = i;
        const bool active = (i >= 1 && i <= L1);
        if (active) {
            const int a = hs[i - 1];
            int diag = (lane == 0) ? i - 1 : prev_left;
            int left = left_in;
            DirT word = 0;
#pragma unroll
            for (int q = 0; q < WPL; ++q) {
                const int up = cur[q];
                const bool eq = (a == refv[q]);
                int v = min(min(left + 1, up + 1), diag + (eq ? 0 : 1));
                int code = eq ? 0 : (v == diag + 1 ? 1 : (v == up + 1 ? 2 : 3));
                word |= (DirT)code << (2 * q);
                diag = up;
                cur[q] = v;
                left = v;
            }
            dirs[(size_t)(i - 1) * 32 + lane] = word;
            my_last = cur[WPL - 1];
        }
        // hand-over for the next step: what I received now is d[i][lane*WPL] -> next step's diagonal
        prev_left = left_in;
    }
    __syncwarp();
    // distance = d[L1][L2] lives in lane (L2-1)/WPL, slot (L2-1)%WPL
    const int own = (L2 - 1) / WPL, slot = (L2 - 1) % WPL;
    int dval = 0;
#pragma unroll
    for (int q = 0; q < WPL; ++q) if (q == slot) dval = cur[q];
    dval = __shfl_sync(0xffffffffu, dval, own);
    if (lane == 0) { p.dist[b] = dval; if (p.status) p.status[b] = MDD_ST_OK; }
    if (!p.ops && !p.has_post) return;

    // ---------------------------------------------------------------- traceback (ctcDecoder.py:154-184)
    if (lane == 0) {
        int i = L1, j = L2;
        while (i > 0 || j > 0) {
            uint8_t op;
            if (i == 0) { op = 'D'; --j; }
            else if (j == 0) { op = 'I'; --i; }
            else {
                const DirT w = dirs[(size_t)(i - 1) * 32 + (j - 1) / WPL];
                const int code = (int)((w >> (2 * ((j - 1) % WPL))) & 3);
                if (code == 0) { op = '-'; --i; --j; }
                else if (code == 1) { op = 'S'; --i; --j; }
                else if (code == 2) { op = 'I'; --i; }
                else { op = 'D'; --j; }
            }
            rev[n++] = op;
        }
    }
    n = __shfl_sync(0xffffffffu, n, 0);
    __syncwarp();
    for (int k = lane; k < n; k += 32) fwd[k] = rev[n - 1 - k];
    __syncwarp();
    }   // !ops_in
    if (p.ops) {
        uint8_t* o = p.ops + (size_t)b * p.ops_stride;
        for (int k = lane; k < n && k < p.ops_stride; k += 32) o[k] = fwd[k];
        if (lane == 0 && p.ops_len) p.ops_len[b] = n;
    }
    if (!p.has_post) return;

    // ---------------------------------------------------------------- MDD post-processing (infer.py:162-209,405-433)
    // leading insertions: d['I'][i] == i holds exactly on the leading run of 'I'
    int lead = n;
    for (int basek = 0; basek < n; basek += 32) {
        const int k = basek + lane;
        unsigned m = __ballot_sync(0xffffffffu, k < n && fwd[k] != 'I');
        if (m) { lead = basek + __ffs(m) - 1; break; }
    }
    int start = lead > 0 ? lead - 1 : 0;
    // decoded-row value at aligned position k: PAD_D for 'D', else hyp[#non-D before k]
    // the duplicate-leading-phone rule needs a_dec[start], a_dec[start+1]
    {
        // ops[0..start] are all 'I' (when lead > 0), so #non-D before start = start
        int d0 = (fwd[start] == 'D') ? MDD_PAD_D : hs[start];
        bool drop = false;
        if (n - start >= 2 && fwd[start] == 'I') {
            int d1 = (fwd[start + 1] == 'D') ? MDD_PAD_D : hs[start + 1];
            drop = (d0 == d1);
        }
        if (drop) ++start;
    }
    const size_t row = (size_t)b * p.ops_stride;
    int ci = 0, cj = 0, n_ins = 0, n_sub = 0, n_del = 0, cc = 0;
    for (int basek = 0; basek < n; basek += 32) {
        const int k = basek + lane;
        const bool in = k < n;
        const uint8_t op = in ? fwd[k] : 0;
        const bool isD = in && op == 'D', isI = in && op == 'I', isS = in && op == 'S', isM = in && op == '-';
        const unsigned lt = (1u << lane) - 1;
        const unsigned mD = __ballot_sync(0xffffffffu, isD), mI = __ballot_sync(0xffffffffu, isI);
        const unsigned mIn = __ballot_sync(0xffffffffu, in);
        const int i_k = ci + __popc(mIn & ~mD & lt);       // hyp index consumed by op k
        const int j_k = cj + __popc(mIn & ~mI & lt);       // ref index consumed by op k
        const int adec = isD ? MDD_PAD_D : (in ? hs[max(0, min(i_k, L1 - 1))] : 0);
        const int acan = isI ? MDD_PAD_I : (in ? rs[max(0, min(j_k, L2 - 1))] : 0);
        const bool live = in && k >= start;
        if (live) {
            p.post.a_dec[row + k - start] = adec;
            p.post.a_can[row + k - start] = acan;
            p.post.a_ops[row + k - start] = op;
        }
        const unsigned mi = __ballot_sync(0xffffffffu, live && isI);
        const unsigned ms = __ballot_sync(0xffffffffu, live && isS);
        const unsigned md = __ballot_sync(0xffffffffu, live && isD);
        const unsigned mm = __ballot_sync(0xffffffffu, live && isM);
        if (live && isI) p.post.ins[row + n_ins + __popc(mi & lt)] = adec;
        if (live && isS) p.post.sub[row + n_sub + __popc(ms & lt)] = acan;
        if (live && isD) p.post.del[row + n_del + __popc(md & lt)] = acan;
        n_ins += __popc(mi); n_sub += __popc(ms); n_del += __popc(md); cc += __popc(mm);
        ci += __popc(mIn & ~mD); cj += __popc(mIn & ~mI);
    }
    if (lane == 0) {
        int32_t* c = p.post.counts + (size_t)b * 8;
        c[0] = n - start; c[1] = n_ins; c[2] = n_sub; c[3] = n_del; c[4] = cc; c[5] = n_sub + n_del;
    }
}

template <int WPL>
int launch_edit(EditParams p, cudaStream_t st) {
    using DirT = typename DirWord<WPL>::type;
    const size_t fixed = align_up(((size_t)p.L1max + p.L2max) * 4, 16) + 2 * align_up((size_t)p.L1max + p.L2max, 16);
    const size_t dir_bytes = (size_t)p.L1max * 32 * sizeof(DirT);
    size_t per_warp = align_up(fixed + dir_bytes, 16);
    int warps = 4;
    bool use_global = false;
    while (warps > 1 && per_warp * warps > 96 * 1024) warps >>= 1;
    if (per_warp * warps > 160 * 1024) {           // very long pairs: decisions go to the global workspace
        if (!p.dir_ws) return MDD_E_WORKSPACE;
        use_global = true;
        per_warp = align_up(fixed, 16);
        warps = 4;
        while (warps > 1 && per_warp * warps > 96 * 1024) warps >>= 1;
    }
    if (!use_global) p.dir_ws = nullptr;
    p.dir_bytes_per_pair = dir_bytes;
    p.smem_per_warp = per_warp;
    const size_t smem = per_warp * warps;
    if (smem > 200 * 1024) return MDD_E_UNSUPPORTED;
    if (smem > 48 * 1024) {
        if (cudaFuncSetAttribute(edit_align_kernel<WPL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
            cudaGetLastError();
            return MDD_E_CUDA;
        }
    }
    const int grid = ceil_div(p.B, warps);
    edit_align_kernel<WPL><<<grid, warps * 32, smem, st>>>(p);
    MDD_CHECK_LAUNCH();
    return MDD_OK;
}

int pick_wpl(int L2max) {
    int w = 1;
    while (w * 32 < L2max) w <<= 1;
    return w;
}

}  // namespace
}  // namespace mdd

using namespace mdd;

extern "C" int mdd_edit_align_workspace_bytes(int B, int max_hyp_len, int max_ref_len, size_t* bytes) {
    if (!bytes || B < 0 || max_hyp_len < 0 || max_ref_len < 0) return MDD_E_ARG;
    const int w = pick_wpl(max_ref_len > 0 ? max_ref_len : 1);
    if (w > 32) return MDD_E_UNSUPPORTED;
    const size_t word = w <= 4 ? 1 : w == 8 ? 2 : w == 16 ? 4 : 8;
    const size_t dir_bytes = (size_t)(max_hyp_len > 0 ? max_hyp_len : 1) * 32 * word;
    // only needed when the decisions do not fit shared memory; always safe to provide
    *bytes = (dir_bytes > 36 * 1024) ? (size_t)B * dir_bytes : 0;
    return MDD_OK;
}

static int edit_align_impl(const int32_t* hyp, int hyp_stride, const int32_t* hyp_len, const int32_t* ref,
                           int ref_stride, const int32_t* ref_len, int B, int max_hyp_len, int max_ref_len,
                           const int32_t* hyp_map, const int32_t* ref_map, int n_map, int32_t* dist, uint8_t* ops,
                           int ops_stride, int32_t* ops_len, int32_t* status, const mdd_post_out* post,
                           const uint8_t* ops_in, int ops_in_stride, const int32_t* ops_in_len, void* workspace,
                           size_t workspace_bytes, mdd_stream_t stream);

extern "C" int mdd_edit_align(const int32_t* hyp, int hyp_stride, const int32_t* hyp_len, const int32_t* ref,
                              int ref_stride, const int32_t* ref_len, int B, int max_hyp_len, int max_ref_len,
                              const int32_t* hyp_map, const int32_t* ref_map, int n_map, int32_t* dist, uint8_t* ops,
                              int ops_stride, int32_t* ops_len, int32_t* status, const mdd_post_out* post,
                              void* workspace, size_t workspace_bytes, mdd_stream_t stream) {
    return edit_align_impl(hyp, hyp_stride, hyp_len, ref, ref_stride, ref_len, B, max_hyp_len, max_ref_len, hyp_map,
                           ref_map, n_map, dist, ops, ops_stride, ops_len, status, post, nullptr, 0, nullptr, workspace,
                           workspace_bytes, stream);
}

extern "C" int mdd_post_align(const int32_t* dec, int dec_stride, const int32_t* dec_len, const int32_t* can,
                              int can_stride, const int32_t* can_len, int B, int max_dec_len, int max_can_len,
                              const uint8_t* ops_in, int ops_in_stride, const int32_t* ops_in_len,
                              const mdd_post_out* post, int out_stride, mdd_stream_t stream) {
    if (!ops_in || !ops_in_len || !post || !post->counts) return MDD_E_ARG;
    // dist is not produced in this mode; counts[...] carries everything
    return edit_align_impl(dec, dec_stride, dec_len, can, can_stride, can_len, B, max_dec_len, max_can_len, nullptr,
                           nullptr, 0, post->counts /*unused scratch*/, nullptr, out_stride, nullptr, nullptr, post,
                           ops_in, ops_in_stride, ops_in_len, nullptr, 0, stream);
}

static int edit_align_impl(const int32_t* hyp, int hyp_stride, const int32_t* hyp_len, const int32_t* ref,
                              int ref_stride, const int32_t* ref_len, int B, int max_hyp_len, int max_ref_len,
                              const int32_t* hyp_map, const int32_t* ref_map, int n_map, int32_t* dist, uint8_t* ops,
                           int ops_stride, int32_t* ops_len, int32_t* status, const mdd_post_out* post,
                           const uint8_t* ops_in, int ops_in_stride, const int32_t* ops_in_len, void* workspace,
                           size_t workspace_bytes, mdd_stream_t stream) {
    if (B == 0) return MDD_OK;
    if (!hyp || !hyp_len || !ref || !ref_len || !dist || B < 0 || max_hyp_len < 0 || max_ref_len < 0) return MDD_E_ARG;
    if ((ops || post) && ops_stride < max_hyp_len + max_ref_len) return MDD_E_ARG;
    if ((hyp_map || ref_map) && n_map <= 0) return MDD_E_ARG;
    EditParams p{};
    p.hyp = hyp; p.hyp_len = hyp_len; p.ref = ref; p.ref_len = ref_len;
    p.hyp_stride = hyp_stride; p.ref_stride = ref_stride; p.B = B;
    p.L1max = max_hyp_len > 0 ? max_hyp_len : 1; p.L2max = max_ref_len > 0 ? max_ref_len : 1;
    p.hyp_map = hyp_map; p.ref_map = ref_map; p.n_map = n_map;
    p.dist = dist; p.ops = ops; p.ops_stride = ops_stride; p.ops_len = ops_len; p.status = status;
    p.has_post = post != nullptr;
    if (post) p.post = *post;
    p.ops_in = ops_in; p.ops_in_stride = ops_in_stride; p.ops_in_len = ops_in_len;
    size_t need = 0;
    int rc = mdd_edit_align_workspace_bytes(B, max_hyp_len, max_ref_len, &need);
    if (rc) return rc;
    p.dir_ws = (need && workspace && workspace_bytes >= need) ? (unsigned char*)workspace : nullptr;
    cudaStream_t st = (cudaStream_t)stream;
    switch (pick_wpl(p.L2max)) {
        case 1: return launch_edit<1>(p, st);
        case 2: return launch_edit<2>(p, st);
        case 4: return launch_edit<4>(p, st);
        case 8: return launch_edit<8>(p, st);
        case 16: return launch_edit<16>(p, st);
        case 32: return launch_edit<32>(p, st);
        default: return MDD_E_UNSUPPORTED;
    }
}
