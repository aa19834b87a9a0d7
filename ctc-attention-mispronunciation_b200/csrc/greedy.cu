// greedy.cu -- frame argmax over (T,B,C) log-probs + CTC collapse.
//
// Replaces GreedyDecoder.decode (egs/attention_aug/utils/ctcDecoder.py:186-200, 80-116) and the
// torch.max + collapse half of compute_wer (steps/train_ctc.py:77, models/model_ctc.py:227-244).
//
// argmax: HBM-bound stream (4*C B read, 1 B written per frame).  (T,B,C) contiguous is one flat array of T*B rows, so
// persistent CTAs walk 128-row tiles; each tile is one contiguous run staged into shared memory by the TMA engine
// (cp.async.bulk + mbarrier, 3 stages), then one thread scans one row.  First maximum wins (torch.max); NaN wins over
// everything (torch propagates NaN).
// collapse: one warp per utterance, 32 frames per step, ballot/popc compaction.
#include "common.cuh"

namespace mdd {
namespace {

constexpr int ROWS = 128;
constexpr int STAGES = 3;

// exact rule of torch.max: strictly greater wins, the first NaN wins over everything (rows holding a NaN only)
__device__ __noinline__ int row_argmax_exact(const float* r, int C) {
    float best = r[0];
    int bi = 0;
    for (int c = 1; c < C; ++c) {
        const float v = r[c];
        if (v > best || (v != v && best == best)) { best = v; bi = c; }
    }
    return bi;
}

// One thread, one row: the maximum by FMNMX (which skips NaNs), then the first index that holds it.  The stream is
// bound by instruction issue next to HBM, so the scan costs 2-3 instructions per class instead of 4-5; a row with a
// NaN (the sum of the row is NaN) takes the exact rule above.
__device__ __forceinline__ int row_argmax(const float* r, int C) {
    float m = -CUDART_INF_F, acc = 0.f;
    int bi = -1;
    if ((C & 1) == 0 && (reinterpret_cast<uintptr_t>(r) & 7) == 0) {
        const float2* r2 = reinterpret_cast<const float2*>(r);
        const int n2 = C >> 1;
        for (int i = 0; i < n2; ++i) {
            const float2 v = r2[i];
            m = fmaxf(m, fmaxf(v.x, v.y));
            acc += v.x + v.y;
        }
        for (int i = n2 - 1; i >= 0; --i) {
            const float2 v = r2[i];
            if (v.y == m) bi = 2 * i + 1;
            if (v.x == m) bi = 2 * i;
        }
    } else {
        for (int c = 0; c < C; ++c) { m = fmaxf(m, r[c]); acc += r[c]; }
        for (int c = C - 1; c >= 0; --c)
            if (r[c] == m) bi = c;
    }
    if (acc != acc) {                       // NaN, or +inf and -inf in one row: settle it exactly
        bool nan = false;
        for (int c = 0; c < C; ++c) nan |= r[c] != r[c];
        if (nan) return row_argmax_exact(r, C);
    }
    return bi < 0 ? 0 : bi;
}

template <bool BULK>
__global__ void __launch_bounds__(ROWS) argmax_rows_kernel(const float* __restrict__ lp, long long n_rows, int C,
                                                           uint8_t* __restrict__ out) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float* tiles = reinterpret_cast<float*>(smem_raw);                   // [STAGES][ROWS*C]
    const int tile_elems = ROWS * C;
    __shared__ __align__(8) uint64_t bars[STAGES];
    const int tid = threadIdx.x;
    const long long n_tiles = (n_rows + ROWS - 1) / ROWS;

    if (!BULK) {
        for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
            const long long r0 = tile * ROWS;
            const int rows = (int)min((long long)ROWS, n_rows - r0);
            const float* src = lp + r0 * C;
            __syncthreads();
            for (int i = tid; i < rows * C; i += ROWS) tiles[i] = src[i];
            __syncthreads();
            if (tid < rows) out[r0 + tid] = (uint8_t)row_argmax(tiles + tid * C, C);
        }
        return;
    }

    if (tid == 0) {
        for (int s = 0; s < STAGES; ++s) mbar_init(&bars[s], 1);
        mbar_fence_init();
    }
    __syncthreads();

    auto issue = [&](long long tile, int stage) {       // thread 0 only
        const long long r0 = tile * ROWS;
        const int rows = (int)min((long long)ROWS, n_rows - r0);
        const uint32_t bytes = (uint32_t)rows * C * 4u;
        const uint32_t b16 = bytes & ~15u;
        float* dst = tiles + (size_t)stage * tile_elems;
        const float* src = lp + r0 * C;
        if (b16) {
            mbar_expect_tx(&bars[stage], b16);
            bulk_g2s(dst, src, b16, &bars[stage]);
        } else {
            mbar_arrive(&bars[stage]);
        }
        for (uint32_t i = b16 / 4; i < bytes / 4; ++i) dst[i] = src[i];   // <= 3 trailing floats of the last tile
    };

    long long it = 0;
    if (tid == 0)
        for (int s = 0; s < STAGES; ++s) {
            const long long tile = blockIdx.x + (long long)s * gridDim.x;
            if (tile < n_tiles) issue(tile, s);
        }
    for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
        const int stage = (int)(it % STAGES);
        const uint32_t parity = (uint32_t)((it / STAGES) & 1);
        const long long r0 = tile * ROWS;
        const int rows = (int)min((long long)ROWS, n_rows - r0);
        mbar_wait(&bars[stage], parity);
        if (rows < ROWS) __syncthreads();       // trailing floats were written with plain stores by thread 0
        if (tid < rows) out[r0 + tid] = (uint8_t)row_argmax(tiles + (size_t)stage * tile_elems + tid * C, C);
        __syncthreads();                        // everyone is done with this stage
        const long long nxt = tile + (long long)STAGES * gridDim.x;
        if (tid == 0 && nxt < n_tiles) issue(nxt, stage);
    }
}

__global__ void __launch_bounds__(128) collapse_kernel(const uint8_t* __restrict__ amax, int T, int B,
                                                       const int64_t* __restrict__ in_len, int blank,
                                                       const int32_t* __restrict__ word_id, int32_t* __restrict__ tokens,
                                                       int tok_stride, int32_t* __restrict__ out_len) {
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= B) return;
    const int b = warp;
    long long len64 = in_len[b];
    const int len = (int)max(0ll, min(len64, (long long)T));
    const int wblank = word_id ? word_id[blank] : blank;
    int n = 0, carry = -1;
    int32_t* tk = tokens + (size_t)b * tok_stride;
    for (int base = 0; base < len; base += 32) {
        const int t = base + lane;
        int idx = blank;
        if (t < len) idx = amax[(size_t)t * B + b];
        const int w = word_id ? word_id[idx] : idx;
        int prev = __shfl_up_sync(0xffffffffu, w, 1);
        if (lane == 0) prev = carry;
        const bool keep = (t < len) && (w != wblank) && !(t != 0 && w == prev);
        const unsigned m = __ballot_sync(0xffffffffu, keep);
        if (keep) {
            const int pos = n + __popc(m & ((1u << lane) - 1));
            if (pos < tok_stride) tk[pos] = idx;
        }
        n += __popc(m);
        carry = __shfl_sync(0xffffffffu, w, 31);
    }
    if (lane == 0) out_len[b] = n;
}

__global__ void exp_transpose_kernel(const float* __restrict__ lp, int T, int B, int C, float* __restrict__ probs) {
    // probs[b,t,c] = expf(lp[t,b,c]); one thread per element of the output, reads are C-contiguous
    const long long n = (long long)T * B * C;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % C);
        const long long bt = i / C;
        const int t = (int)(bt % T);
        const int b = (int)(bt / T);
        probs[i] = expf(lp[((size_t)t * B + b) * C + c]);
    }
}

}  // namespace
}  // namespace mdd

using namespace mdd;

extern "C" int mdd_greedy_decode(const float* log_probs, int T, int B, int C, const int64_t* input_lengths, int blank,
                                 const int32_t* word_id, uint8_t* argmax, int32_t* tokens, int tok_stride,
                                 int32_t* out_len, mdd_stream_t stream) {
    if (T < 0 || B < 0 || C <= 0 || C > MDD_MAX_CLASSES || blank < 0 || blank >= C) return MDD_E_ARG;
    if (B == 0) return MDD_OK;
    if (!log_probs && T > 0) return MDD_E_ARG;
    if (!argmax || (tokens && (!input_lengths || !out_len))) return MDD_E_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    const long long n_rows = (long long)T * B;
    if (n_rows > 0) {
        const size_t smem = (size_t)STAGES * ROWS * C * 4;
        const long long n_tiles = (n_rows + ROWS - 1) / ROWS;
        const bool bulk = ((uintptr_t)log_probs % 16 == 0) && ((ROWS * C * 4) % 16 == 0);
        auto kern = bulk ? argmax_rows_kernel<true> : argmax_rows_kernel<false>;
        if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
            cudaGetLastError();
            return MDD_E_CUDA;
        }
        int per_sm = (int)((200 * 1024) / (smem + 1024));
        if (per_sm < 1) per_sm = 1;
        if (per_sm > 8) per_sm = 8;
        const int grid = (int)min(n_tiles, (long long)kNumSM * per_sm);
        kern<<<grid, ROWS, smem, st>>>(log_probs, n_rows, C, argmax);
        MDD_CHECK_LAUNCH();
    }
    if (tokens) {
        const int warps_per_cta = 4;
        collapse_kernel<<<ceil_div(B, warps_per_cta), warps_per_cta * 32, 0, st>>>(argmax, T, B, input_lengths, blank,
                                                                                 word_id, tokens, tok_stride, out_len);
        MDD_CHECK_LAUNCH();
    }
    return MDD_OK;
}

extern "C" int mdd_collapse(const uint8_t* argmax, int T, int B, const int64_t* input_lengths, int blank,
                            const int32_t* word_id, int32_t* tokens, int tok_stride, int32_t* out_len,
                            mdd_stream_t stream) {
    if (T < 0 || B < 0 || blank < 0 || (!argmax && T > 0) || !input_lengths || !tokens || !out_len) return MDD_E_ARG;
    if (B == 0) return MDD_OK;
    const int warps_per_cta = 4;
    collapse_kernel<<<ceil_div(B, warps_per_cta), warps_per_cta * 32, 0, (cudaStream_t)stream>>>(
        argmax, T, B, input_lengths, blank, word_id, tokens, tok_stride, out_len);
    MDD_CHECK_LAUNCH();
    return MDD_OK;
}

extern "C" int mdd_exp_transpose(const float* log_probs, int T, int B, int C, float* probs, mdd_stream_t stream) {
    if (!log_probs || !probs || T < 0 || B < 0 || C <= 0) return MDD_E_ARG;
    const long long n = (long long)T * B * C;
    if (n == 0) return MDD_OK;
    const int grid = (int)min((n + 255) / 256, (long long)kNumSM * 16);
    exp_transpose_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(log_probs, T, B, C, probs);
    MDD_CHECK_LAUNCH();
    return MDD_OK;
}
