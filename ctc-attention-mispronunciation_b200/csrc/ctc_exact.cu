// ctc_exact.cu -- CTC alpha/beta forward-backward in the linear domain, fp64 (robust path `algo = 1` and the
// per-utterance fallback behind the packed fp32 kernel of ctc_pair.cu).
//
// Semantics: torch aten ctc_loss / _ctc_loss_backward as reached from nn.CTCLoss at
// egs/attention_aug/steps/train_ctc.py:186,72,96, including inf losses / NaN gradient rows
// of infeasible alignments and zero_infinity.
//
// Why not log space: a log-domain recursion in fp32 rounds every stored value to ulp(|value|), i.e. to an ABSOLUTE
// error that grows with the magnitude of the log-likelihoods involved, and those absolute errors are relative errors
// of the occupancies; on high-loss inputs (> 4 nats per frame) the round-1 log-space kernel reached 2.5e-4 against
// torch fp64.  Here alpha and beta are plain probabilities in fp64: every operation has a RELATIVE error of 1.1e-16
// whatever the magnitude, there is no exp / log per lattice state at all (three DP operations per state and frame),
// and the gradient agrees with torch's fp64 path to the rounding of the fp32 output.
//
// Range: each row is rescaled by an exact power of two taken from the largest exponent of the previous row (lagged by
// one frame, so it costs no extra barrier; the exponents are summed as integers for the loss).  A row therefore keeps
// its maximum within one frame's emission of 1, and states more than 2^-1022 below it are dropped (relative mass
// < 1e-307).  Emissions are exp(lp) in fp64: exact zeros for lp = -inf, representable down to lp = -708.
//
// One CTA per utterance (persistent CTAs stride over the batch, so a sweep that has nothing to do costs a few flag
// reads).  The lattice never goes to HBM as a whole: the forward sweep writes one checkpoint row every K frames; the
// backward sweep walks chunks from the end, recomputes alpha for the chunk into shared memory, runs beta and emits
//     grad[t,c] = (exp(lp[t,c]) - (sum_{s: l'_s = c} alpha_t(s) u_t(s)) / Z_t) * g_b,   Z_t = sum_s alpha_t(s) u_t(s)
// (u_t = beta_t without the emission at t; Z_t is the frame's own normaliser, so no scale has to be carried between
// the two sweeps) with a deterministic per-class reduction (CSR class -> states, fixed order, no atomics).
#include "common.cuh"
#include "ctc_params.cuh"

namespace mdd {

namespace {

constexpr int NT = 256;
constexpr int NW = NT / 32;
constexpr int STG0 = NT - 128;             // threads STG0 + c stage the emission of class c (C <= 128)

__device__ __forceinline__ int biased_exp(double v) { return (__double2hiint(v) >> 20) & 0x7ff; }

// 2^(1023 - e) for the biased exponent e of a row maximum; 1 when the row is all zero / not finite
__device__ __forceinline__ double row_scale(int e, long long& shift) {
    if (e <= 0 || e >= 0x7ff) return 1.0;
    const int f = max(2046 - e, 1);
    shift += (long long)(1023 - f);        // = e - 1023 unless clamped
    return __hiloint2double(f << 20, 0);
}

__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__global__ void __launch_bounds__(NT) ctc_exact_kernel(CtcParams p) {
    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int Smax = 2 * p.Lmax + 1;
    const int K = p.K;
    const int C = p.C;
    const int blank = p.blank;

    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* chunk = reinterpret_cast<double*>(smem_raw);        // [K][Smax]  alpha of the current chunk
    double* bbuf = chunk + (size_t)K * Smax;                    // [2][Smax]  alpha ping-pong (forward), beta (backward)
    double* ab = bbuf + 2 * Smax;                               // [Smax]     alpha_t(s) * u_t(s)
    double* pe = ab + Smax;                                     // [2][C]     exp(lp[t, c])
    double* s_cs = pe + 2 * C;                                  // [C]        class sums of ab
    int* lab = reinterpret_cast<int*>(s_cs + C);                // [Lmax]
    int* cls_start = lab + p.Lmax;                              // [C+1]
    int* cls_list = cls_start + C + 1;                          // [Lmax] label index i (state 2i+1) grouped by class
    __shared__ int s_amax[2][NW];          // per-warp largest biased exponent of the alpha row written last / before
    __shared__ int s_bmax[2][NW];          // the same for beta (its own slots: the two sweeps alternate inside a chunk)
    __shared__ double s_ll;
    __shared__ int s_badlabel;

    auto row_max = [&](const int (*m)[NW], int parity) {
        int e = m[parity][0];
#pragma unroll
        for (int w = 1; w < NW; ++w) e = max(e, m[parity][w]);
        return e;
    };

    if (p.only_flagged) pdl_wait();        // launched behind the packed tiers of the same call
    for (int b = blockIdx.x; b < p.B; b += gridDim.x) {
        if (p.only_flagged && p.flags[b] == 0) continue;
        __syncthreads();                   // the previous utterance of this CTA is done with shared memory

        const long long Tb64 = p.in_len[b], L64 = p.tgt_len[b];
        const bool bad = Tb64 < 0 || Tb64 > p.T || L64 < 0 || L64 > p.Lmax;
        float* grad_b = p.grad ? p.grad + (size_t)b * C : nullptr;   // + t*B*C
        const size_t gstride = (size_t)p.B * C;
        if (bad) {
            if (tid == 0) { p.nll[b] = CUDART_NAN_F; if (p.status) p.status[b] = MDD_ST_BADLEN; }
            if (grad_b) for (int i = tid; i < p.T * C; i += NT) grad_b[(size_t)(i / C) * gstride + i % C] = 0.f;
            continue;
        }
        const int Tb = (int)Tb64, L = (int)L64;
        const int S = 2 * L + 1;

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
            continue;
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
        const double g = (double)((p.grad_out ? p.grad_out[b] : 1.f) * p.gscalar);

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
            continue;
        }

        // emission rows: thread STG0 + c owns class c; the row of frame t lands in pe[t & 1] one step ahead of its
        // use, its log-prob is fetched one step before that
        const int sc = tid - STG0;
        const bool stager = sc >= 0 && sc < C;
        float lp_next = 0.f;
        auto fetch = [&](int t) {
            if (stager && t >= 0 && t < Tb)
                lp_next = lp_b[(size_t)t * p.stride_t + sc] - (p.lse ? p.lse[(size_t)t * p.B + b] : 0.f);
        };
        auto stage = [&](int t) {          // pe[t & 1][c] = exp(lp[t, c]) from the value fetched earlier
            if (stager && t >= 0 && t < Tb) pe[(t & 1) * C + sc] = exp((double)lp_next);
        };
        // my states: class, skip mask (computed per use from `lab`, which is shared memory)
        double* ck = reinterpret_cast<double*>(p.ckpt) + (size_t)b * p.n_ckpt * Smax;

        // ------------------------------------------------------------------ forward: alpha, checkpoints, nll
        fetch(0);
        stage(0);
        fetch(1);
        __syncthreads();
        {
            int em = 0;
            for (int s = tid; s < S; s += NT) {
                double v = 0.0;
                if (s == 0) v = pe[blank];
                else if (s == 1) v = pe[lab[0]];
                bbuf[s] = v;
                ck[s] = v;
                em = max(em, biased_exp(v));
            }
            em = __reduce_max_sync(0xffffffffu, em);
            if (lane == 0) s_amax[0][warp] = em;
        }
        stage(1);
        fetch(2);
        long long shift = 0;               // identical in every thread
        for (int t = 1; t < Tb; ++t) {
            __syncthreads();               // row t-1, its maxima and pe[t & 1] are complete
            const double scale = row_scale(row_max(s_amax, (t - 1) & 1), shift);
            const double* prev = bbuf + ((t - 1) & 1) * Smax;
            double* cur = bbuf + (t & 1) * Smax;
            const double* rw = pe + (t & 1) * C;
            const bool store = (t % K) == 0;
            double* ckt = ck + (size_t)(t / K) * Smax;
            int em = 0;
            for (int s = tid; s < S; s += NT) {
                double a = prev[s];
                if (s > 0) a += prev[s - 1];
                int cls = blank;
                if (s & 1) {
                    const int i = s >> 1;
                    cls = lab[i];
                    if (i > 0 && lab[i - 1] != cls) a += prev[s - 2];
                }
                const double v = (a * scale) * rw[cls];
                cur[s] = v;
                if (store) ckt[s] = v;
                em = max(em, biased_exp(v));
            }
            em = __reduce_max_sync(0xffffffffu, em);
            if (lane == 0) s_amax[t & 1][warp] = em;
            stage(t + 1);                  // into the other row buffer; read after the next barrier
            fetch(t + 2);
        }
        __syncthreads();
        if (tid == 0) {
            const double* lastrow = bbuf + ((Tb - 1) & 1) * Smax;
            const double tot = S > 1 ? lastrow[S - 1] + lastrow[S - 2] : lastrow[S - 1];
            s_ll = log(tot) + (double)shift * 0.69314718055994530942;     // log(0) = -inf: infeasible
        }
        __syncthreads();
        const float nll = (float)(-s_ll);
        const bool is_inf = (nll == CUDART_INF_F);
        if (tid == 0) {
            p.nll[b] = (p.zero_inf && is_inf) ? 0.f : nll;
            if (p.status) p.status[b] = MDD_ST_OK;
        }
        if (!grad_b) continue;
        if (p.zero_inf && is_inf) {
            for (int i = tid; i < Tb * C; i += NT) grad_b[(size_t)(i / C) * gstride + i % C] = 0.f;
            continue;
        }

        // ------------------------------------------------------------------ backward: chunks from the end
        const int n_chunks = (Tb + K - 1) / K;
        long long bshift = 0;              // bookkeeping only (Z_t normalises each frame)
        for (int j = n_chunks - 1; j >= 0; --j) {
            const int t0 = j * K, t1 = min(Tb, t0 + K);
            __syncthreads();               // the previous chunk's readers of `chunk` and `pe` are done
            {
                int em = 0;
                for (int s = tid; s < S; s += NT) {
                    const double v = ck[(size_t)j * Smax + s];
                    chunk[s] = v;
                    em = max(em, biased_exp(v));
                }
                em = __reduce_max_sync(0xffffffffu, em);
                if (lane == 0) s_amax[t0 & 1][warp] = em;
            }
            fetch(t0 + 1);
            stage(t0 + 1);
            fetch(t0 + 2);
            for (int t = t0 + 1; t < t1; ++t) {
                __syncthreads();
                long long dummy = 0;
                const double scale = row_scale(row_max(s_amax, (t - 1) & 1), dummy);
                const double* prev = chunk + (size_t)(t - 1 - t0) * Smax;
                double* cur = chunk + (size_t)(t - t0) * Smax;
                const double* rw = pe + (t & 1) * C;
                int em = 0;
                for (int s = tid; s < S; s += NT) {
                    double a = prev[s];
                    if (s > 0) a += prev[s - 1];
                    int cls = blank;
                    if (s & 1) {
                        const int i = s >> 1;
                        cls = lab[i];
                        if (i > 0 && lab[i - 1] != cls) a += prev[s - 2];
                    }
                    const double v = (a * scale) * rw[cls];
                    cur[s] = v;
                    em = max(em, biased_exp(v));
                }
                em = __reduce_max_sync(0xffffffffu, em);
                if (lane == 0) s_amax[t & 1][warp] = em;
                if (t + 1 < t1) { stage(t + 1); fetch(t + 2); }
            }
            __syncthreads();               // the chunk's alpha rows are complete; pe is free again
            fetch(t1 - 1);
            stage(t1 - 1);
            fetch(t1 - 2);
            for (int t = t1 - 1; t >= t0; --t) {
                __syncthreads();           // (1) pe[t & 1] and the beta row t+1 (with its maxima) are complete
                const double* rw = pe + (t & 1) * C;
                const double* al = chunk + (size_t)(t - t0) * Smax;
                double* bn = bbuf + (t & 1) * Smax;
                const double* bo = bbuf + ((t + 1) & 1) * Smax;
                int em = 0;
                if (t == Tb - 1) {
                    for (int s = tid; s < S; s += NT) {
                        const double u = (s == S - 1 || s == S - 2) ? 1.0 : 0.0;
                        int cls = blank;
                        if (s & 1) cls = lab[s >> 1];
                        const double v = u * rw[cls];
                        bn[s] = v;
                        ab[s] = al[s] * u;
                        em = max(em, biased_exp(v));
                    }
                } else {
                    const double scale = row_scale(row_max(s_bmax, (t + 1) & 1), bshift);
                    for (int s = tid; s < S; s += NT) {
                        double u = bo[s];
                        if (s + 1 < S) u += bo[s + 1];
                        int cls = blank;
                        if (s & 1) {
                            const int i = s >> 1;
                            cls = lab[i];
                            if (i + 1 < L && lab[i + 1] != cls) u += bo[s + 2];
                        }
                        u *= scale;
                        const double v = u * rw[cls];
                        bn[s] = v;
                        ab[s] = al[s] * u;
                        em = max(em, biased_exp(v));
                    }
                }
                em = __reduce_max_sync(0xffffffffu, em);
                if (lane == 0) s_bmax[t & 1][warp] = em;
                __syncthreads();           // (2) ab complete
                // per-class sums of alpha_t(s) u_t(s), fixed order
                if (tid < 32) {
                    double acc = 0.0;
                    for (int s = 2 * tid; s < S; s += 64) acc += ab[s];
                    acc = warp_sum_d(acc);
                    if (tid == 0) {
                        // targets equal to the blank index (legal for torch): their label states also count as blank
                        for (int n = cls_start[blank]; n < cls_start[blank + 1]; ++n) acc += ab[2 * cls_list[n] + 1];
                        s_cs[blank] = acc;
                    }
                } else if (tid < 32 + C) {
                    const int c = tid - 32;
                    if (c != blank) {
                        double acc = 0.0;
                        for (int n = cls_start[c]; n < cls_start[c + 1]; ++n) acc += ab[2 * cls_list[n] + 1];
                        s_cs[c] = acc;
                    }
                }
                if (t - 1 >= t0) { stage(t - 1); fetch(t - 2); }     // the stagers sit in the upper warps
                __syncthreads();           // (3) class sums complete
                if (tid < 32) {
                    double z = 0.0;
                    for (int c = tid; c < C; c += 32) z += s_cs[c];
                    z = warp_sum_d(z);
                    const double zi = 1.0 / z;                        // 1/0 = inf, 0 * inf = NaN: infeasible rows are NaN
                    float* gr = grad_b + (size_t)t * gstride;
                    for (int c = tid; c < C; c += 32) gr[c] = (float)((rw[c] - s_cs[c] * zi) * g);
                }
            }
        }
    }
    // Loss sum in the same launch: the last CTA to get here adds up nll[B] in a fixed order (thread i takes i, i + NT,
    // ..., then a fixed tree) -- bitwise reproducible, and one launch fewer on the path where nothing was flagged.
    if (p.ticket && p.loss_sum) {
        __shared__ int s_last;
        __shared__ float s_part[NT];
        __syncthreads();
        if (tid == 0) {
            __threadfence();
            s_last = atomicAdd(p.ticket, 1u) == gridDim.x - 1;
        }
        __syncthreads();
        if (s_last) {
            __threadfence();
            float s = 0.f;
            for (int i = tid; i < p.B; i += NT) s += __ldcg(p.nll + i);
            s_part[tid] = s;
            __syncthreads();
            for (int d = NT / 2; d > 0; d >>= 1) {
                if (tid < d) s_part[tid] += s_part[tid + d];
                __syncthreads();
            }
            if (tid == 0) *p.loss_sum = s_part[0];
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

size_t exact_smem_bytes(int C, int Lmax, int K) {
    const size_t Smax = 2 * (size_t)Lmax + 1;
    return (K * Smax + 3 * Smax + 3 * (size_t)C) * 8 + ((size_t)Lmax * 2 + C + 1) * 4 + 16;
}

}  // namespace

int ctc_exact_chunk(int Lmax) {
    const int Smax = 2 * Lmax + 1;
    int K = (64 * 1024) / (8 * Smax);
    if (K > 32) K = 32;
    if (K < 2) K = 2;
    return K;
}

size_t ctc_exact_ckpt_bytes(int T, int B, int Lmax) {
    const int K = ctc_exact_chunk(Lmax);
    const size_t n_ckpt = (size_t)(T + K - 1) / K + 1;
    return (size_t)B * n_ckpt * (2 * (size_t)Lmax + 1) * 8;
}

int launch_tgt_offsets(const int64_t* tgt_len, int B, int64_t* off, cudaStream_t st) {
    tgt_offsets_kernel<<<1, 1024, 0, st>>>(tgt_len, B, off);
    MDD_CHECK_LAUNCH();
    return MDD_OK;
}

int launch_ctc_exact(CtcParams p, cudaStream_t st) {
    p.K = ctc_exact_chunk(p.Lmax);
    p.n_ckpt = (p.T + p.K - 1) / p.K + 1;
    const size_t smem = exact_smem_bytes(p.C, p.Lmax, p.K);
    if (smem > 225 * 1024) return MDD_E_UNSUPPORTED;
    if (cudaFuncSetAttribute(ctc_exact_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
        cudaGetLastError();
        return MDD_E_CUDA;
    }
    // persistent CTAs: a fallback sweep with nothing flagged reads B flags and leaves
    const int per_sm = (int)std::max<size_t>(1, std::min<size_t>(8, (220 * 1024) / (smem + 1024)));
    const int grid = std::min(p.B, sm_count() * per_sm);
    if (p.only_flagged) {
        if (launch_dependent(ctc_exact_kernel, dim3(grid), dim3(NT), smem, st, p) != cudaSuccess) {
            cudaGetLastError();
            return MDD_E_CUDA;
        }
    } else {
        ctc_exact_kernel<<<grid, NT, smem, st>>>(p);
    }
    MDD_CHECK_LAUNCH();
    return MDD_OK;
}

}  // namespace mdd
