// ctc_params.cuh -- launch parameters shared by the CTC loss kernels
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace mdd {

struct CtcParams {
    const float* lp;          // (T,B,C), strides (stride_t, stride_b, 1)
    const float* lse;         // null: lp holds log-probs; else lp holds LOGITS and lse[t*B+b] their row log-sum-exp
    int T, B, C;
    int64_t stride_t, stride_b;
    const int64_t* targets;
    int64_t tgt_stride;       // 0 -> concatenated, use tgt_offsets
    const int64_t* tgt_offsets;
    const int64_t* in_len;
    const int64_t* tgt_len;
    int Lmax;
    int blank;
    int zero_inf;
    const float* grad_out;    // [B] or null
    float gscalar;
    float* nll;               // [B]
    float* grad;              // (T,B,C) contiguous or null
    int32_t* status;          // [B] or null
    // fp64 kernel (ctc_exact.cu)
    float* ckpt;              // [B][n_ckpt][2*Lmax+1] doubles
    int K, n_ckpt;
    int only_flagged;         // process only utterances with flags[b] != 0
    int32_t* flags;           // [B] fast path -> fallback requests
    // pair kernel: forward/backward likelihood agreement demanded of an utterance (relative), and the earlier a-priori
    // squash criterion (see renorm_alpha)
    float pbeta_tol;
    int strict_squash;
    // second tier of the pair kernel: compacted list of the utterances the first pass flagged (null: no second tier)
    int32_t* list;
    int32_t* list_count;
    int list_cap_pairs;
    int list_team_max;        // second tier: up to this many flagged utterances go to the team kernel, more to the pair kernel
    int decay_max;            // largest per-chunk decay (bits) of the lanes holding the bulk of alpha
    float4* rec;              // [B] forward kernel -> backward kernel: 1 / mantissa(P), exponent, flag
    // few utterances (ctc_dual.cu + the SEG instantiation of ctc_pair.cu): beta checkpoints, the emission-row layout of
    // every pair ([npairs][2][72] positions), segments per pair of the backward kernel
    float* ckb;
    int* cpos_ws;
    int nseg;
    float* erow;              // [npairs][T][128] exponentiated emission rows as the sweep kernel committed them
    // loss sum folded into the fp64 kernel's launch: the first kernel of a call zeroes the ticket, the last CTA of the
    // fp64 sweep to finish adds up nll[B] in a fixed order (null: the caller launches sum_nll_kernel instead)
    unsigned* ticket;
    float* loss_sum;
};

// fp64 linear-domain kernel, one CTA per utterance (ctc_exact.cu): robust path and per-utterance fallback
int ctc_exact_chunk(int Lmax);
size_t ctc_exact_ckpt_bytes(int T, int B, int Lmax);
int launch_tgt_offsets(const int64_t* tgt_len, int B, int64_t* off, cudaStream_t st);
int launch_ctc_exact(CtcParams p, cudaStream_t st);

// two utterances per warp, packed FP32 (ctc_pair.cu): the default fast path
bool ctc_pair_supported(int C, int Lmax);
size_t ctc_pair_ws_bytes(int T, int B, int Lmax);
int launch_ctc_pair(CtcParams p, void* ws, cudaStream_t st);

// one CTA per pair, lattice warp + helper warps (ctc_team.cu): few utterances per SM, and the second tier
bool ctc_team_supported(int C, int Lmax);
int launch_ctc_team(const CtcParams& p, void* ws, cudaStream_t st, bool list, int nblocks);

// few utterances: alpha and beta swept concurrently, checkpoints of both (ctc_dual.cu); the gradient is then the backward
// kernel of ctc_pair.cu on many segments of every pair at once
int launch_ctc_dual(const CtcParams& p, float* cka, float* ckb, cudaStream_t st, int npairs);

// per-frame log-sum-exp (+ argmax) of logits (row_lse.cu): the pre-pass of mdd_ctc_loss_logits
int launch_row_lse(const float* x, int T, int B, int C, int64_t stride_t, int64_t stride_b, float* lse, uint8_t* amax,
                   cudaStream_t st);

// deterministic sum of nll[B] -> loss_sum (reduce.cu)
int launch_sum_nll(const float* nll, int B, float* out, cudaStream_t st);
int launch_scale_grad(float* grad, int T, int B, int C, const float* scale, float scalar, cudaStream_t st);

}  // namespace mdd
