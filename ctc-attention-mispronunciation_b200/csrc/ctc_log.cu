// ctc_log.cu -- log-space CTC alpha/beta forward-backward (robust path and per-utterance fallback).
//
// Semantics: torch aten ctc_loss / _ctc_loss_backward as reached from nn.CTCLoss at
// egs/attention_aug/steps/train_ctc.py:186,72,96.
//
// One CTA per utterance.  The extended-label lattice never goes to HBM as a whole: the forward sweep keeps two
// alpha rows in shared memory and writes one checkpoint row every K frames (K*S*4 B fits shared memory); the backward
// sweep walks chunks from the end, recomputes alpha for the chunk into shared memory, runs beta and emits
//     grad[t,c] = (exp(lp[t,c]) - exp(LSE_{s: l'_s = c}(alpha_t(s) + beta_t(s)) + nll - lp[t,c])) * g_b
// with a deterministic per-class reduction (CSR class -> states built once per utterance; no atomics).
//
// Precision: alpha/beta are kept *normalised* -- each row is shifted by the maximum of the previous row (a lagged
// max costs no extra barrier) and the shifts are accumulated in fp64 -- so stored magnitudes stay O(10) instead of
// O(nll) and fp32 rounding does not grow with T.  The per-frame constant that ties alpha~ + beta~ back to the
// occupancy is not formed as a difference of large numbers either: it is the frame's own normaliser
//     Z_t = LSE_s(alpha~_t(s) + beta~_t(s) - lp[t, l'_s])   (sum_s alpha_t(s) beta_t(s) / p_t(l'_s) = P for every t).
#include "common.cuh"
#include "ctc_params.cuh"

namespace mdd {

namespace {

constexpr int NT = 256;

__device__ __forceinline__ float block_lse_blank(const float* ab, int S, float* red, int tid) {
    // LSE over even states (the blank class), computed by warp 0
    float m = -CUDART_INF_F;
    for (int s = 2 * tid; s < S; s += 64) m = fmaxf(m, ab[s]);
    m = warp_max(m);
    float acc = 0.f;
    if (m != -CUDART_INF_F)
        for (int s = 2 * tid; s < S; s += 64) acc += expf(ab[s] - m);
    acc = warp_sum(acc);
    (void)red;
    return (m == -CUDART_INF_F) ? -CUDART_INF_F : m + logf(acc);
}

__global__ void __launch_bounds__(NT) ctc_log_kernel(CtcParams p) {
    const int b = blockIdx.x;
    const int tid = threadIdx.x;
    if (p.only_flagged && p.flags[b] == 0) return;

    const int Tb64 = (int)p.in_len[b];
    const int L = (int)p.tgt_len[b];
    const bool bad = Tb64 < 0 || Tb64 > p.T || L < 0 || L > p.Lmax;
    float* grad_b = p.grad ? p.grad + (size_t)b * p.C : nullptr;   // + t*B*C
    const size_t gstride = (size_t)p.B * p.C;
    if (bad) {
        if (tid == 0) { p.nll[b] = CUDART_NAN_F; if (p.status) p.status[b] = MDD_ST_BADLEN; }
        if (grad_b) for (int i = tid; i < p.T * p.C; i += NT) grad_b[(size_t)(i / p.C) * gstride + i % p.C] = 0.f;
        return;
    }
    const int Tb = Tb64;
    const int S = 2 * L + 1;
    const int Smax = 2 * p.Lmax + 1;
    const int K = p.K;
    const int C = p.C;
    const int blank = p.blank;

    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* chunk = reinterpret_cast<float*>(smem_raw);          // [K][Smax]
    float* bbuf = chunk + (size_t)K * Smax;                     // [2][Smax]  (alpha ping-pong in forward, beta in backward)
    float* ab = bbuf + 2 * Smax;                                // [Smax]
    float* srow = ab + Smax;                                    // [2][C]
    int* lab = reinterpret_cast<int*>(srow + 2 * C);            // [Lmax]
    int* cls_start = lab + p.Lmax;                              // [C+1]
    int* cls_list = cls_start + C + 1;                          // [Lmax] label index i (state 2i+1) grouped by class
    __shared__ float s_nll;
    __shared__ int s_badlabel;
    __shared__ float s_pmax[2][NT / 32];     // per-warp maxima of the row written in the previous step
    __shared__ float s_lc[MDD_MAX_CLASSES];  // per-class LSE(alpha~+beta~) - lp
    const int warp = tid >> 5, lane = tid & 31;
    auto row_shift = [&](int parity) {       // max of the previous row (0 if the row is all -inf)
        float m = s_pmax[parity][0];
#pragma unroll
        for (int w = 1; w < NT / 32; ++w) m = fmaxf(m, s_pmax[parity][w]);
        return m == -CUDART_INF_F ? 0.f : m;
    };

    const int64_t* tg = p.tgt_stride ? p.targets + (int64_t)b * p.tgt_stride : p.targets + p.tgt_offsets[b];
    if (tid == 0) s_badlabel = 0;
    __syncthreads();
    for (int i = tid; i < L; i += NT) {
        int64_t v = tg[i];
        if (v < 0 || v >= C) { s_badlabel = 1; v = 0; }
        lab[i] = (int)v;
    }
    for (int c = tid; c <= C; c += NT) cls_start[c] = 0;
    __syncthreads();
    if (s_badlabel) {
        if (tid == 0) { p.nll[b] = CUDART_NAN_F; if (p.status) p.status[b] = MDD_ST_BADLEN; }
        if (grad_b) for (int i = tid; i < p.T * C; i += NT) grad_b[(size_t)(i / C) * gstride + i % C] = 0.f;
        return;
    }
    // CSR class -> label positions (deterministic: positions ascending inside a class)
    if (tid == 0) {
        for (int i = 0; i < L; ++i) cls_start[lab[i] + 1]++;
        for (int c = 0; c < C; ++c) cls_start[c + 1] += cls_start[c];
    }
    __syncthreads();
    if (tid < C) {
        int c = tid, n = cls_start[c];
        for (int i = 0; i < L; ++i) if (lab[i] == c) cls_list[n++] = i;
    }

    const float* lp_b = p.lp + (size_t)b * p.stride_b;
    const float g = (p.grad_out ? p.grad_out[b] : 1.f) * p.gscalar;

    // rows t >= Tb of the gradient are zero (torch semantics)
    if (grad_b)
        for (int i = tid; i < (p.T - Tb) * C; i += NT) grad_b[(size_t)(Tb + i / C) * gstride + i % C] = 0.f;

    if (Tb == 0) {
        if (tid == 0) {
            float v = (L == 0) ? 0.f : CUDART_INF_F;
            if (p.zero_inf && v == CUDART_INF_F) v = 0.f;
            p.nll[b] = v;
            if (p.status) p.status[b] = MDD_ST_OK;
        }
        return;
    }

    float* ck = p.ckpt + (size_t)b * p.n_ckpt * Smax;

    // ------------------------------------------------------------------ forward: alpha, checkpoints, nll
    if (tid < C) srow[tid] = lp_b[tid] - (p.lse ? p.lse[b] : 0.f);
    __syncthreads();
    {
        float pm = -CUDART_INF_F;
        for (int s = tid; s < S; s += NT) {
            float v = -CUDART_INF_F;
            if (s == 0) v = srow[blank];
            else if (s == 1) v = srow[lab[0]];
            bbuf[s] = v;
            ck[s] = v;
            pm = fmaxf(pm, v);
        }
        pm = warp_max(pm);
        if (lane == 0) s_pmax[0][warp] = pm;
    }
    double shift_sum = 0.0;                  // identical in every thread
    for (int t = 1; t < Tb; ++t) {
        float* rw = srow + (t & 1) * C;
        if (tid < C) rw[tid] = lp_b[(size_t)t * p.stride_t + tid] - (p.lse ? p.lse[(size_t)t * p.B + b] : 0.f);
        __syncthreads();
        const float shift = row_shift((t - 1) & 1);
        shift_sum += (double)shift;
        const float* prev = bbuf + ((t - 1) & 1) * Smax;
        float* cur = bbuf + (t & 1) * Smax;
        const bool store = (t % K) == 0;
        float* ckt = ck + (size_t)(t / K) * Smax;
        float pm = -CUDART_INF_F;
        for (int s = tid; s < S; s += NT) {
            float a1 = prev[s];
            float a2 = s > 0 ? prev[s - 1] : -CUDART_INF_F;
            float a3 = -CUDART_INF_F;
            int cls = blank;
            if (s & 1) {
                int i = s >> 1;
                cls = lab[i];
                if (i > 0 && lab[i - 1] != cls) a3 = prev[s - 2];
            }
            float v = lse3_acc(a1, a2, a3) + (rw[cls] - shift);
            cur[s] = v;
            if (store) ckt[s] = v;
            pm = fmaxf(pm, v);
        }
        pm = warp_max(pm);
        if (lane == 0) s_pmax[t & 1][warp] = pm;
    }
    __syncthreads();
    if (tid == 0) {
        const float* lastrow = bbuf + ((Tb - 1) & 1) * Smax;
        float ll = S > 1 ? lse2_acc(lastrow[S - 1], lastrow[S - 2]) : lastrow[S - 1];
        s_nll = (float)(-((double)ll + shift_sum));
    }
    __syncthreads();
    const float nll = s_nll;
    const bool is_inf = (nll == CUDART_INF_F);
    if (tid == 0) {
        p.nll[b] = (p.zero_inf && is_inf) ? 0.f : nll;
        if (p.status) p.status[b] = MDD_ST_OK;
    }
    if (!grad_b) return;
    if (p.zero_inf && is_inf) {
        for (int i = tid; i < Tb * C; i += NT) grad_b[(size_t)(i / C) * gstride + i % C] = 0.f;
        return;
    }

    // ------------------------------------------------------------------ backward: chunks from the end
    const int n_chunks = (Tb + K - 1) / K;
    for (int j = n_chunks - 1; j >= 0; --j) {
        const int t0 = j * K, t1 = min(Tb, t0 + K);
        __syncthreads();   // previous chunk's readers of `chunk` are done
        {
            float pm = -CUDART_INF_F;
            for (int s = tid; s < S; s += NT) { const float v = ck[(size_t)j * Smax + s]; chunk[s] = v; pm = fmaxf(pm, v); }
            pm = warp_max(pm);
            if (lane == 0) s_pmax[t0 & 1][warp] = pm;
        }
        for (int t = t0 + 1; t < t1; ++t) {
            float* rw = srow + (t & 1) * C;
            if (tid < C) rw[tid] = lp_b[(size_t)t * p.stride_t + tid] - (p.lse ? p.lse[(size_t)t * p.B + b] : 0.f);
            __syncthreads();
            const float shift = row_shift((t - 1) & 1);
            const float* prev = chunk + (size_t)(t - 1 - t0) * Smax;
            float* cur = chunk + (size_t)(t - t0) * Smax;
            float pm = -CUDART_INF_F;
            for (int s = tid; s < S; s += NT) {
                float a1 = prev[s];
                float a2 = s > 0 ? prev[s - 1] : -CUDART_INF_F;
                float a3 = -CUDART_INF_F;
                int cls = blank;
                if (s & 1) {
                    int i = s >> 1;
                    cls = lab[i];
                    if (i > 0 && lab[i - 1] != cls) a3 = prev[s - 2];
                }
                const float v = lse3_acc(a1, a2, a3) + (rw[cls] - shift);
                cur[s] = v;
                pm = fmaxf(pm, v);
            }
            pm = warp_max(pm);
            if (lane == 0) s_pmax[t & 1][warp] = pm;
        }
        __syncthreads();   // s_pmax is reused (with its own parity) by the beta sweep
        for (int t = t1 - 1; t >= t0; --t) {
            float* rw = srow + (t & 1) * C;
            if (tid < C) rw[tid] = lp_b[(size_t)t * p.stride_t + tid] - (p.lse ? p.lse[(size_t)t * p.B + b] : 0.f);
            __syncthreads();
            const float* al = chunk + (size_t)(t - t0) * Smax;
            float* bn = bbuf + (t & 1) * Smax;
            const float* bo = bbuf + ((t + 1) & 1) * Smax;
            float pm = -CUDART_INF_F;
            if (t == Tb - 1) {
                for (int s = tid; s < S; s += NT) {
                    float v = -CUDART_INF_F;
                    if (s == S - 1) v = rw[blank];
                    else if (s == S - 2) v = rw[lab[L - 1]];
                    bn[s] = v;
                    ab[s] = al[s] + v;
                    pm = fmaxf(pm, v);
                }
            } else {
                const float shift = row_shift((t + 1) & 1);
                for (int s = tid; s < S; s += NT) {
                    float b1 = bo[s];
                    float b2 = s + 1 < S ? bo[s + 1] : -CUDART_INF_F;
                    float b3 = -CUDART_INF_F;
                    int cls = blank;
                    if (s & 1) {
                        int i = s >> 1;
                        cls = lab[i];
                        if (i + 1 < L && lab[i + 1] != cls) b3 = bo[s + 2];
                    }
                    float v = lse3_acc(b1, b2, b3) + (rw[cls] - shift);
                    bn[s] = v;
                    ab[s] = al[s] + v;
                    pm = fmaxf(pm, v);
                }
            }
            pm = warp_max(pm);
            if (lane == 0) s_pmax[t & 1][warp] = pm;
            __syncthreads();
            // per-class LSE_{s in c}(alpha~ + beta~) - lp[c]   (alpha and beta both carry the emission at t)
            if (tid < 32) {
                float lcab = block_lse_blank(ab, S, nullptr, tid);
                if (tid == 0) {
                    // targets equal to the blank index (legal for torch): their label states also count for class blank
                    for (int n = cls_start[blank]; n < cls_start[blank + 1]; ++n) lcab = lse2_acc(lcab, ab[2 * cls_list[n] + 1]);
                    s_lc[blank] = lcab - rw[blank];
                }
            } else if (tid < 32 + C) {
                const int c = tid - 32;
                if (c != blank) {
                    float m = -CUDART_INF_F;
                    const int n0 = cls_start[c], n1 = cls_start[c + 1];
                    for (int n = n0; n < n1; ++n) m = fmaxf(m, ab[2 * cls_list[n] + 1]);
                    float lcab = -CUDART_INF_F;
                    if (m != -CUDART_INF_F) {
                        float acc = 0.f;
                        for (int n = n0; n < n1; ++n) acc += expf(ab[2 * cls_list[n] + 1] - m);
                        lcab = m + logf(acc);
                    }
                    s_lc[c] = lcab - rw[c];
                }
            }
            __syncthreads();
            // frame normaliser Z_t = LSE_c(s_lc[c]) (= log P up to the shifts), then grad = (softmax prob - occupancy) * g
            if (tid < 32) {
                float m = -CUDART_INF_F;
                for (int c = tid; c < C; c += 32) m = fmaxf(m, s_lc[c]);
                m = warp_max(m);
                float acc = 0.f;
                if (m != -CUDART_INF_F)
                    for (int c = tid; c < C; c += 32) acc += expf(s_lc[c] - m);
                acc = warp_sum(acc);
                const float Z = (m == -CUDART_INF_F) ? -CUDART_INF_F : m + logf(acc);
                float* gr = grad_b + (size_t)t * gstride;
                for (int c = tid; c < C; c += 32) gr[c] = (expf(rw[c]) - expf(s_lc[c] - Z)) * g;   // -inf - -inf = NaN: infeasible
            }
        }
    }
}

// exclusive prefix sum of target_lengths (1-D concatenated targets), one block
__global__ void __launch_bounds__(1024) tgt_offsets_kernel(const int64_t* tgt_len, int B, int64_t* off) {
    __shared__ int64_t part[1024];
    const int tid = threadIdx.x;
    const int per = (B + 1023) / 1024;
    const int lo = min(B, tid * per), hi = min(B, lo + per);
    int64_t s = 0;
    for (int i = lo; i < hi; ++i) s += tgt_len[i] > 0 ? tgt_len[i] : 0;
    part[tid] = s;
    __syncthreads();
    for (int d = 1; d < 1024; d <<= 1) {
        int64_t v = tid >= d ? part[tid - d] : 0;
        __syncthreads();
        part[tid] += v;
        __syncthreads();
    }
    int64_t run = tid ? part[tid - 1] : 0;
    for (int i = lo; i < hi; ++i) { off[i] = run; run += tgt_len[i] > 0 ? tgt_len[i] : 0; }
}

}  // namespace

int ctc_log_chunk(int Lmax) {
    const int Smax = 2 * Lmax + 1;
    int K = (64 * 1024) / (4 * Smax);
    if (K > 64) K = 64;
    if (K < 4) K = 4;
    return K;
}

size_t ctc_log_smem_bytes(int C, int Lmax, int K) {
    const size_t Smax = 2 * (size_t)Lmax + 1;
    return (K * Smax + 3 * Smax + 2 * (size_t)C) * 4 + ((size_t)Lmax * 2 + C + 1) * 4 + 16;
}

size_t ctc_log_ckpt_bytes(int T, int B, int Lmax) {
    const int K = ctc_log_chunk(Lmax);
    const size_t n_ckpt = (size_t)(T + K - 1) / K + 1;
    return (size_t)B * n_ckpt * (2 * (size_t)Lmax + 1) * 4;
}

int launch_tgt_offsets(const int64_t* tgt_len, int B, int64_t* off, cudaStream_t st) {
    tgt_offsets_kernel<<<1, 1024, 0, st>>>(tgt_len, B, off);
    MDD_CHECK_LAUNCH();
    return MDD_OK;
}

int launch_ctc_log(CtcParams p, cudaStream_t st) {
    p.K = ctc_log_chunk(p.Lmax);
    p.n_ckpt = (p.T + p.K - 1) / p.K + 1;
    const size_t smem = ctc_log_smem_bytes(p.C, p.Lmax, p.K);
    if (smem > 200 * 1024) return MDD_E_UNSUPPORTED;
    static size_t configured = 0;
    if (smem > configured) {
        if (cudaFuncSetAttribute(ctc_log_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
            cudaGetLastError();
            return MDD_E_CUDA;
        }
        configured = smem;
    }
    ctc_log_kernel<<<p.B, NT, smem, st>>>(p);
    MDD_CHECK_LAUNCH();
    return MDD_OK;
}

}  // namespace mdd
