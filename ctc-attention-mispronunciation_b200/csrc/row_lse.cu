// row_lse.cu -- per-frame log-sum-exp (+ argmax) of the output layer's logits: the pre-pass of mdd_ctc_loss_logits.
//
// Replaces the tail of the reference models' forward, `out = F.log_softmax(out, dim=-1)`
// (egs/attention_aug/models/model_ctc.py:214-218, egs/cnn-rnn-ctc/models/cnn_rnn.py:168-172), and the train-time
// `torch.max(out, dim=-1)` (steps/train_ctc.py:77): one stream over the (T,B,C) logits writes 4 B (lse) + 1 B (argmax)
// per frame instead of a second (T,B,C) array.  The CTC kernels then exponentiate `logit - lse` on the fly, and what
// they emit, (softmax - occupancy) * g, already IS the gradient with respect to the logits.
//
// HBM-bound stream: 4*C B read per frame.  The contiguous case walks 128-row tiles staged by the TMA engine
// (cp.async.bulk + mbarrier, 3 stages, persistent CTAs); one thread scans one row with 8-byte shared-memory loads
// (row stride 4*C B: with C even and C/2 odd the 16 threads of a half-warp hit 16 different 8-byte banks).
// lse = m + log(sum exp(x - m)) with m the row maximum (as torch's log_softmax; exponentials by MUFU.EX2, 2 ulp);
// argmax: first maximum wins, NaN wins (torch.max).
#include "common.cuh"

namespace mdd {
namespace {

constexpr int ROWS = 128;
constexpr int STAGES = 3;

__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// exact rule of torch.max for rows that hold a NaN (first NaN wins); the fast scan below ignores NaNs
__device__ __noinline__ int row_argmax_nan(const float* r, int C) {
    for (int c = 0; c < C; ++c)
        if (r[c] != r[c]) return c;
    return 0;
}

// One thread, one row.  Pass 1: the maximum (FMNMX).  Pass 2: sum of 2^((x - m) * log2 e) (FFMA + MUFU.EX2, 2 ulp) and
// the first index that holds the maximum.  The kernel is bound by instruction issue next to HBM, hence the approximate
// exponential: the log-sum-exp stays within 3e-7 relative of torch.logsumexp.
template <bool VEC2>
__device__ __forceinline__ void row_scan(const float* r, int C, float& lse, int& amax) {
    constexpr float L2E = 1.4426950408889634f;
    float m = -CUDART_INF_F, s = 0.f;
    int bi = -1;
    if (VEC2) {
        const float2* r2 = reinterpret_cast<const float2*>(r);
        const int n2 = C >> 1;
#pragma unroll 7
        for (int i = 0; i < n2; ++i) {
            const float2 v = r2[i];
            m = fmaxf(m, fmaxf(v.x, v.y));
        }
        const float nm = -m * L2E;
#pragma unroll 7
        for (int i = 0; i < n2; ++i) {
            const float2 v = r2[i];
            s += ex2_approx(fmaf(v.x, L2E, nm));
            s += ex2_approx(fmaf(v.y, L2E, nm));
            if (v.x == m && bi < 0) bi = 2 * i;
            if (v.y == m && bi < 0) bi = 2 * i + 1;
        }
    } else {
        for (int c = 0; c < C; ++c) m = fmaxf(m, r[c]);
        const float nm = -m * L2E;
        for (int c = 0; c < C; ++c) {
            const float v = r[c];
            s += ex2_approx(fmaf(v, L2E, nm));
            if (v == m && bi < 0) bi = c;
        }
    }
    if (s != s || m == CUDART_INF_F) {              // NaN in the row, or +inf (inf - inf): the slow exact path
        float q = 0.f;
        bool nan = false;
        for (int c = 0; c < C; ++c) { nan |= r[c] != r[c]; q += expf(r[c] - m); }
        s = q;
        if (nan) bi = row_argmax_nan(r, C);
    }
    lse = (m == -CUDART_INF_F) ? -CUDART_INF_F : m + logf(s);
    amax = bi < 0 ? 0 : bi;
}

// contiguous (T*B, C) rows
template <bool BULK, bool VEC2>
__global__ void __launch_bounds__(ROWS) row_lse_kernel(const float* __restrict__ x, long long n_rows, int C,
                                                       float* __restrict__ lse_out, uint8_t* __restrict__ amax_out) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float* tiles = reinterpret_cast<float*>(smem_raw);                   // [STAGES][ROWS*C]
    const int tile_elems = ROWS * C;
    __shared__ __align__(8) uint64_t bars[STAGES];
    const int tid = threadIdx.x;
    const long long n_tiles = (n_rows + ROWS - 1) / ROWS;

    auto emit = [&](const float* tile, long long r0, int rows) {
        if (tid < rows) {
            float l;
            int a;
            row_scan<VEC2>(tile + tid * C, C, l, a);
            lse_out[r0 + tid] = l;
            if (amax_out) amax_out[r0 + tid] = (uint8_t)a;
        }
    };

    if (!BULK) {
        for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
            const long long r0 = tile * ROWS;
            const int rows = (int)min((long long)ROWS, n_rows - r0);
            const float* src = x + r0 * C;
            __syncthreads();
            for (int i = tid; i < rows * C; i += ROWS) tiles[i] = src[i];
            __syncthreads();
            emit(tiles, r0, rows);
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
        const float* src = x + r0 * C;
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
        emit(tiles + (size_t)stage * tile_elems, r0, rows);
        __syncthreads();                        // everyone is done with this stage
        const long long nxt = tile + (long long)STAGES * gridDim.x;
        if (tid == 0 && nxt < n_tiles) issue(nxt, stage);
    }
}

// strided (T,B,C) views (stride_c == 1): one warp per row, lanes over classes
__global__ void __launch_bounds__(256) row_lse_strided_kernel(const float* __restrict__ x, int T, int B, int C,
                                                              int64_t stride_t, int64_t stride_b,
                                                              float* __restrict__ lse_out, uint8_t* __restrict__ amax_out) {
    const int lane = threadIdx.x & 31;
    const long long n_rows = (long long)T * B;
    const long long wpg = (long long)gridDim.x * (blockDim.x >> 5);
    for (long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); row < n_rows; row += wpg) {
        const int t = (int)(row / B), b = (int)(row % B);
        const float* r = x + (size_t)t * stride_t + (size_t)b * stride_b;
        float best = -CUDART_INF_F;
        int bi = 0x7fffffff;
        bool seen = false;
        for (int c = lane; c < C; c += 32) {
            const float v = r[c];
            if (!seen || v > best || (v != v && best == best)) { best = v; bi = c; seen = true; }
        }
        // (value, index) arg-max across lanes: NaN beats numbers, the smaller index wins ties
        for (int o = 16; o > 0; o >>= 1) {
            const float ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            const bool mine_nan = best != best, other_nan = ob != ob;
            bool take;
            if (oi == 0x7fffffff) take = false;
            else if (bi == 0x7fffffff) take = true;
            else if (mine_nan || other_nan) take = other_nan && (!mine_nan || oi < bi);
            else take = ob > best || (ob == best && oi < bi);
            if (take) { best = ob; bi = oi; }
        }
        float s = 0.f;
        for (int c = lane; c < C; c += 32) s += expf(r[c] - best);
        s = warp_sum(s);
        if (lane == 0) {
            lse_out[row] = (best == -CUDART_INF_F) ? -CUDART_INF_F : best + logf(s);
            if (amax_out) amax_out[row] = (uint8_t)bi;
        }
    }
}

}  // namespace

int launch_row_lse(const float* x, int T, int B, int C, int64_t stride_t, int64_t stride_b, float* lse, uint8_t* amax,
                   cudaStream_t st) {
    const long long n_rows = (long long)T * B;
    if (n_rows == 0) return MDD_OK;
    const bool contiguous = stride_b == C && (stride_t == (int64_t)B * C || T == 1);
    if (!contiguous) {
        const int grid = (int)min((n_rows + 7) / 8, (long long)kNumSM * 8);
        row_lse_strided_kernel<<<grid, 256, 0, st>>>(x, T, B, C, stride_t, stride_b, lse, amax);
        MDD_CHECK_LAUNCH();
        return MDD_OK;
    }
    const size_t smem = (size_t)STAGES * ROWS * C * 4;
    const long long n_tiles = (n_rows + ROWS - 1) / ROWS;
    const bool bulk = ((uintptr_t)x % 16 == 0) && ((ROWS * C * 4) % 16 == 0);
    const bool vec2 = (C % 2 == 0) && ((uintptr_t)x % 8 == 0);
    auto kern = bulk ? (vec2 ? row_lse_kernel<true, true> : row_lse_kernel<true, false>)
                     : (vec2 ? row_lse_kernel<false, true> : row_lse_kernel<false, false>);
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
        cudaGetLastError();
        return MDD_E_CUDA;
    }
    int per_sm = (int)((200 * 1024) / (smem + 1024));
    if (per_sm < 1) per_sm = 1;
    if (per_sm > 8) per_sm = 8;
    const int grid = (int)min(n_tiles, (long long)kNumSM * per_sm);
    kern<<<grid, ROWS, smem, st>>>(x, n_rows, C, lse, amax);
    MDD_CHECK_LAUNCH();
    return MDD_OK;
}

}  // namespace mdd

using namespace mdd;

extern "C" int mdd_row_lse(const float* logits, int T, int B, int C, int64_t stride_t, int64_t stride_b, float* row_lse,
                           uint8_t* argmax, mdd_stream_t stream) {
    if (T < 0 || B < 0 || C <= 0 || C > MDD_MAX_CLASSES || !row_lse) return MDD_E_ARG;
    if ((long long)T * B == 0) return MDD_OK;
    if (!logits) return MDD_E_ARG;
    return launch_row_lse(logits, T, B, C, stride_t, stride_b, row_lse, argmax, (cudaStream_t)stream);
}
