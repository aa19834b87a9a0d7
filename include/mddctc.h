/*
 * mddctc.h -- C ABI of the B200-native CTC hot path (libmddctc.so, sm_100a).
 *
 * Drop-in boundary for the CTC path of dyustc/CTC-Attention-Mispronunciation.  The reference is pure Python on
 * top of torch; each entry point below names the reference interface it replaces (paths relative to the reference
 * root).  The Python host side (ctc-attention-mispronunciation_b200/*.py) binds these with ctypes, passing
 * tensor.data_ptr() values and the current cudaStream_t; INTEGRATION.md shows the binding.
 *
 * Conventions
 *   - extern "C", plain pointers and sizes, no torch / C++ types.  All data pointers are DEVICE pointers unless the
 *     parameter name ends in _host.  mdd_stream_t is a cudaStream_t.
 *   - Every call is asynchronous on the given stream, never allocates and never takes ownership: scratch space is
 *     supplied by the caller, sized by the matching *_workspace_bytes query.
 *   - Return value: 0 = ok, <0 = argument error (MDD_E_*); no C++ exception crosses the ABI.  Per-utterance
 *     conditions the reference turns into Python exceptions are reported in an int32 status[B] output
 *     (MDD_ST_*), which the Python wrappers map back to the same exception types.
 *   - There is no CPU fallback: on a machine without a CUDA device the compute entry points return MDD_E_CUDA.
 */
#ifndef MDDCTC_H_
#define MDDCTC_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* mdd_stream_t; /* cudaStream_t */

enum {
    MDD_OK = 0,
    MDD_E_ARG = -1,       /* null pointer / negative size / C out of range            */
    MDD_E_WORKSPACE = -2, /* workspace too small                                       */
    MDD_E_CUDA = -3,      /* CUDA runtime error (launch failure, no device)            */
    MDD_E_UNSUPPORTED = -4,
    MDD_E_STATIC_LIMIT = -5 /* mdd_ctc_loss_static: the shape needs more than MDD_CTC_STATIC_WS_BYTES of scratch */
};

enum {                    /* per-utterance status codes                                 */
    MDD_ST_OK = 0,
    MDD_ST_EMPTY = 1,     /* empty side in wer() -> TypeError (utils/ctcDecoder.py:125,137-138);
                             empty prefix in the final beam -> IndexError (utils/BeamSearch.py:135) */
    MDD_ST_LOG0 = 2,      /* math.log(0) -> ValueError (utils/BeamSearch.py:64,66,103,106)        */
    MDD_ST_LMKEY = 3,     /* LM vocabulary miss -> KeyError (utils/NgramLM.py:76)                  */
    MDD_ST_BADLEN = 4     /* length out of range (torch raises RuntimeError at the call site)      */
};

#define MDD_MAX_CLASSES 128
#define MDD_PAD_D (-2)    /* decoded-row pad where the path says 'D' (infer.py:179-182) */
#define MDD_PAD_I (-3)    /* canonical-row pad where the path says 'I' (infer.py:183-185) */

int mdd_version(void);
const char* mdd_error_string(int code);
/* Number of kernels this library has launched since load / since the last reset (bench.py "gpu_launches"). */
int64_t mdd_launch_count(void);
void mdd_launch_count_reset(void);

/* ------------------------------------------------------------------------------------------------------------------
 * CTC loss + gradient.
 * Replaces: nn.CTCLoss(reduction='sum')(out, targets, input_sizes, target_sizes) and loss.backward()
 *           egs/attention_aug/steps/train_ctc.py:186,72,96 ; egs/cnn-rnn-ctc/steps/train_ctc.py:148,51,67
 *           (torch aten ctc_loss / _ctc_loss_backward semantics).
 *
 * log_probs : (T,B,C) fp32, element strides (stride_t, stride_b, 1).
 * targets   : int64; row b starts at targets + b*tgt_stride; tgt_stride == 0 means 1-D concatenated targets.
 * input_lengths, target_lengths : int64[B] on the DEVICE (train_ctc.py:45-47 moves them there).
 * max_target_len : upper bound on target_lengths (host-known padding width); rows violating it get MDD_ST_BADLEN.
 * grad_out  : fp32[B] upstream gradient of nll[b] (NULL = 1).  grad_out_scalar multiplies it (use 1.0f).
 * nll       : fp32[B] out.  With zero_infinity, infinite losses are written as 0 and their gradient rows as 0;
 *             without it they are +inf and the gradient rows are NaN, as in torch.
 * grad      : (T,B,C) contiguous fp32 out, or NULL for loss only.  Rows t >= input_lengths[b] are written as 0.
 * loss_sum  : optional fp32 scalar out = sum_b nll[b] (deterministic fixed-order tree), e.g. an NCCL send buffer.
 * status    : optional int32[B].
 * algo      : 0 = auto: packed two-utterances-per-warp fp32 linear-domain kernel at 8 frames per chunk (for up to
 *                 444 pairs: alpha and beta swept concurrently, one CTA per pair, then the gradient on independent
 *                 segments of every pair -- same arithmetic, a third of the dependent chain); the utterances
 *                 it flags (dynamic range, malformed rows) again at 2 frames per chunk; what is still flagged by the
 *                 fp64 linear-domain kernel (one CTA per utterance, exact to the rounding of the fp32 outputs),
 *             1 = fp64 kernel only, 2 = first packed pass only (no second tier, no fallback; for profiling).
 * ------------------------------------------------------------------------------------------------------------------ */
int mdd_ctc_loss_workspace_bytes(int T, int B, int C, int max_target_len, int algo, size_t* bytes);
int mdd_ctc_loss(const float* log_probs, int T, int B, int C, int64_t stride_t, int64_t stride_b,
                 const int64_t* targets, int64_t tgt_stride, const int64_t* input_lengths,
                 const int64_t* target_lengths, int max_target_len, int blank, int zero_infinity,
                 const float* grad_out, float grad_out_scalar, float* nll, float* grad, float* loss_sum,
                 int32_t* status, int algo, void* workspace, size_t workspace_bytes, mdd_stream_t stream);

/* ------------------------------------------------------------------------------------------------------------------
 * CTC loss + gradient straight from the output layer's LOGITS.
 * Replaces: out = F.log_softmax(out, dim=-1)      egs/attention_aug/models/model_ctc.py:214-218,
 *                                                 egs/cnn-rnn-ctc/models/cnn_rnn.py:168-172
 *           followed by nn.CTCLoss / loss.backward() (call sites above) and, with `argmax`, the train-time
 *           torch.max(out, dim=-1)                steps/train_ctc.py:77
 *
 * Same arguments as mdd_ctc_loss except:
 * logits      : (T,B,C) fp32, element strides (stride_t, stride_b, 1): the input of log_softmax.
 * grad_logits : (T,B,C) contiguous fp32 out or NULL: d(sum_b grad_out[b]*nll[b]) / d logits
 *               = (softmax - occupancy) * grad_out[b], i.e. what autograd obtains by back-propagating torch's CTC
 *               gradient through log_softmax (its rows sum to zero).  Rows t >= input_lengths[b] are 0.
 * row_lse     : fp32 (T,B) out, required: log sum_c exp(logits[t,b,c]); log_probs[t,b,c] = logits[t,b,c] - row_lse[t,b].
 * argmax      : uint8 (T,B) out or NULL: frame arg-max (first maximum wins), ready for mdd_collapse.
 * algo        : 0, 1 or 2.  Workspace: mdd_ctc_loss_workspace_bytes.
 * One pre-pass kernel streams the logits once (4*C B/frame) and writes 5 B/frame; no (T,B,C) log-prob array and no
 * log_softmax backward pass exist.
 * ------------------------------------------------------------------------------------------------------------------ */
int mdd_ctc_loss_logits(const float* logits, int T, int B, int C, int64_t stride_t, int64_t stride_b,
                        const int64_t* targets, int64_t tgt_stride, const int64_t* input_lengths,
                        const int64_t* target_lengths, int max_target_len, int blank, int zero_infinity,
                        const float* grad_out, float grad_out_scalar, float* nll, float* grad_logits, float* loss_sum,
                        int32_t* status, float* row_lse, uint8_t* argmax, int algo, void* workspace,
                        size_t workspace_bytes, mdd_stream_t stream);
/* mdd_ctc_loss without a workspace argument: the scratch is a static device array inside the library.
 * This entry point has no counterpart in the reference; it was added for callers
 * that cannot allocate device memory (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  It calls no cudaMalloc / cudaFree and no allocating library routine; arguments, results and kernels are
 * those of mdd_ctc_loss.
 *   - Limit: shapes whose mdd_ctc_loss_workspace_bytes(T, B, C, max_target_len, algo) exceeds MDD_CTC_STATIC_WS_BYTES
 *     are rejected with MDD_E_STATIC_LIMIT.  16 MiB holds e.g. B = 64 at T = 500 with targets of up to 63 labels
 *     (14.9 MiB with algo 0) or B = 16 at T = 2000 (14.0 MiB); B = 64 at T = 500 with 120 labels needs 20.8 MiB.
 *   - Every GPU that loads the library holds the 16 MiB until the process exits.
 *   - All calls on a device share the one array, which keeps the previous call's contents (the entry point clears the
 *     part that must start at zero on `stream`): calls that overlap on the GPU -- different streams, or threads --
 *     race; serialise them or use mdd_ctc_loss with separate workspaces.
 *   - The first call on each device looks up the array's device address (cudaGetSymbolAddress); make that call before
 *     starting a graph capture. */
#define MDD_CTC_STATIC_WS_BYTES (16u << 20)
int mdd_ctc_loss_static(const float* log_probs, int T, int B, int C, int64_t stride_t, int64_t stride_b,
                        const int64_t* targets, int64_t tgt_stride, const int64_t* input_lengths,
                        const int64_t* target_lengths, int max_target_len, int blank, int zero_infinity,
                        const float* grad_out, float grad_out_scalar, float* nll, float* grad, float* loss_sum,
                        int32_t* status, int algo, mdd_stream_t stream);

/* The pre-pass alone: row_lse (T,B) and optional argmax (T,B) of (T,B,C) logits with strides (stride_t, stride_b, 1). */
int mdd_row_lse(const float* logits, int T, int B, int C, int64_t stride_t, int64_t stride_b, float* row_lse,
                uint8_t* argmax, mdd_stream_t stream);

/* Host -> device copy of a batch slice of a (T,B,C) array: `height` rows of `width_bytes`, row pitches in bytes
 * (cudaMemcpy2DAsync on `stream`; `src_host` should be page-locked).  Lets a caller whose log-probs live in host memory
 * (the reference keeps them there for decoding, infer.py:300) overlap the transfer of one batch slice with the kernels of
 * the previous one: mddctc.loss.ctc_loss_streamed. */
int mdd_copy_h2d_2d(void* dst, size_t dst_pitch, const void* src_host, size_t src_pitch, size_t width_bytes,
                    size_t height, mdd_stream_t stream);

/* grad[t,b,c] *= scale[b] * scalar   (autograd backward when the upstream gradient was not known in forward). */
int mdd_ctc_scale_grad(float* grad, int T, int B, int C, const float* scale, float scalar, mdd_stream_t stream);

/* ------------------------------------------------------------------------------------------------------------------
 * Greedy decode: frame argmax + collapse.
 * Replaces: GreedyDecoder.decode          egs/attention_aug/utils/ctcDecoder.py:186-200 (+ :80-116)
 *           torch.max(out,-1) + compute_wer's collapse   steps/train_ctc.py:77, models/model_ctc.py:227-244
 *
 * log_probs : (T,B,C) fp32 contiguous.   input_lengths: int64[B] device.
 * word_id   : optional int32[C]; classes with equal word_id compare equal in the repeat rule (the reference compares
 *             word strings, ctcDecoder.py:83-84).  NULL = identity.
 * argmax    : uint8 (T,B) out, required (it is also the scratch the collapse reads); first maximum wins, like torch.max.
 * tokens    : int32 (B, tok_stride) out, out_len int32[B] out (both optional: NULL = argmax only).
 * ------------------------------------------------------------------------------------------------------------------ */
int mdd_greedy_decode(const float* log_probs, int T, int B, int C, const int64_t* input_lengths, int blank,
                      const int32_t* word_id, uint8_t* argmax, int32_t* tokens, int tok_stride, int32_t* out_len,
                      mdd_stream_t stream);

/* Collapse only, on caller-supplied frame argmax ids (uint8, (T,B) layout): the compute_wer drop-in receives
 * torch.max indices from its caller (steps/train_ctc.py:77-81). */
int mdd_collapse(const uint8_t* argmax, int T, int B, const int64_t* input_lengths, int blank,
                 const int32_t* word_id, int32_t* tokens, int tok_stride, int32_t* out_len, mdd_stream_t stream);

/* ------------------------------------------------------------------------------------------------------------------
 * Levenshtein distance + op-path traceback + MDD post-processing.
 * Replaces: Decoder.wer/_edit_distance/printChanges    egs/attention_aug/utils/ctcDecoder.py:118-184
 *           editdistance.eval in compute_wer           egs/attention_aug/models/model_ctc.py:240
 *           sil/err filtering, align_canonical_decoded, stastics, counts   egs/attention_aug/infer.py:310-342,162-209,405-433
 *
 * hyp (B,hyp_stride), ref (B,ref_stride) int32 token rows with lengths hyp_len/ref_len int32[B].
 * hyp_map/ref_map : optional int32[n_map] token -> id tables applied first; id < 0 drops the token (sil / err).
 * dist   : int32[B].      status: int32[B] (MDD_ST_EMPTY when a side is empty after filtering; dist is still set).
 * ops    : optional uint8 (B, ops_stride) '-','S','I','D' in forward order, ops_len int32[B]; ops_stride >= Lh+Lr.
 * post   : optional mdd_post_out; all rows have stride ops_stride.  counts (B,8) int32 =
 *          [n_aligned, n_ins, n_sub, n_del, correct_cnt, del_sub_cnt, hyp_len_filtered, ref_len_filtered].
 * ------------------------------------------------------------------------------------------------------------------ */
typedef struct {
    int32_t* a_dec;  /* (B, ops_stride) decoded row with MDD_PAD_D pads, after the leading-insertion trims */
    int32_t* a_can;  /* (B, ops_stride) canonical row with MDD_PAD_I pads                                 */
    uint8_t* a_ops;  /* (B, ops_stride) trimmed op path                                                    */
    int32_t* ins;    /* (B, ops_stride) inserted (decoded) ids                                             */
    int32_t* sub;    /* (B, ops_stride) substituted canonical ids                                          */
    int32_t* del;    /* (B, ops_stride) deleted canonical ids                                              */
    int32_t* counts; /* (B, 8)                                                                             */
} mdd_post_out;

int mdd_edit_align_workspace_bytes(int B, int max_hyp_len, int max_ref_len, size_t* bytes);
int mdd_edit_align(const int32_t* hyp, int hyp_stride, const int32_t* hyp_len, const int32_t* ref, int ref_stride,
                   const int32_t* ref_len, int B, int max_hyp_len, int max_ref_len, const int32_t* hyp_map,
                   const int32_t* ref_map, int n_map, int32_t* dist, uint8_t* ops, int ops_stride, int32_t* ops_len,
                   int32_t* status, const mdd_post_out* post, void* workspace, size_t workspace_bytes,
                   mdd_stream_t stream);

/* Post-processing only, on a caller-supplied op path -- the literal align_canonical_decoded(s1, s2, l) / stastics
 * call surface of egs/attention_aug/infer.py:162-209,405-433 (the path normally comes from mdd_edit_align).
 * dec/can are the token rows, ops_in (B, ops_in_stride) the paths; outputs as in mdd_post_out with row stride
 * out_stride >= max_dec_len + max_can_len. */
int mdd_post_align(const int32_t* dec, int dec_stride, const int32_t* dec_len, const int32_t* can, int can_stride,
                   const int32_t* can_len, int B, int max_dec_len, int max_can_len, const uint8_t* ops_in,
                   int ops_in_stride, const int32_t* ops_in_len, const mdd_post_out* post, int out_stride,
                   mdd_stream_t stream);

/* ------------------------------------------------------------------------------------------------------------------
 * Corpus MDD evaluation counters.
 * Replaces: the per-utterance body of test()    egs/attention_aug/steps/test_ctc_nosil.py:218-297
 *           print_align_space_canonical_origin  egs/attention_aug/steps/test_ctc_nosil.py:33-82 (its dict d)
 *
 * lab / dec : human-transcription and recognised token rows (int32, row strides lab_stride / dec_stride) with
 *             lengths lab_len / dec_len int32[B]; lab_map / dec_map as in mdd_edit_align (id < 0 drops, e.g. 'sil',
 *             :198-205) -- pass the maps the paths were computed with.
 * lc_ops    : op paths of wer(labels, canonical) (:219), dc_ops: of wer(decoded, canonical) (:221), both from
 *             mdd_edit_align with the canonical row as `ref`; uint8 (B, *_stride) with lengths int32[B].
 * dist_dl   : optional int32[B] = distance of wer(decoded, labels) (:220), copied into counts[.,9] / the totals.
 * group     : optional uint8[B]; non-zero marks the utterances of the reference's m_speaker list (:152,236).
 * counts    : int32 (B,12) out = [TA, FR, FA, TR_correct, TR_wrong, n_canonical, FR_m, FA_m, n_labels, dist_dl, 0, 0];
 *             FR_m / FA_m follow the m_speaker branch, which adds 1 where the general branch adds the list length
 *             (:270-275).  Rows with status != 0 are zero.
 * totals    : optional int64[2][12], ACCUMULATED (zero it first): row 0 = sum of counts over all utterances, row 1 =
 *             over the group with FR/FA taken from FR_m/FA_m (the "mandarin" block, :318-343).
 * status    : optional int32[B]: MDD_ST_EMPTY where decoder.wer would raise TypeError (an empty side),
 *             MDD_ST_BADLEN where the two paths disagree about the canonical length.
 * ------------------------------------------------------------------------------------------------------------------ */
int mdd_corpus_counts(const int32_t* lab, int lab_stride, const int32_t* lab_len, const int32_t* dec, int dec_stride,
                      const int32_t* dec_len, int B, int max_lab_len, int max_dec_len, int max_can_len,
                      const int32_t* lab_map, const int32_t* dec_map, int n_map, const uint8_t* lc_ops, int lc_stride,
                      const int32_t* lc_len, const uint8_t* dc_ops, int dc_stride, const int32_t* dc_len,
                      const int32_t* dist_dl, const uint8_t* group, int32_t* counts, int64_t* totals, int32_t* status,
                      mdd_stream_t stream);

/* ------------------------------------------------------------------------------------------------------------------
 * CTC prefix beam search with bigram-LM fusion (the reference's joint decoder).
 * Replaces: BeamDecoder.decode -> ctcBeamSearch.decode   utils/ctcDecoder.py:215-226, utils/BeamSearch.py:73-153
 *
 * probs : (B,T,C) fp32 PROBABILITIES, contiguous (ctcDecoder.py:223-225 applies transpose + exp first; use
 *         mdd_exp_transpose for that step).  input_lengths: int64[B] device or NULL (= T).
 * lm    : fp64 (C+1)x(C+1) table of LanguageModel.get_bi_prob (utils/NgramLM.py:65-78), index C = '' (<s> / </s>),
 *         NaN where the lookup raises KeyError.  lm_alpha multiplies each lookup at use (BeamSearch.py:60).
 * tokens (B,tok_stride) int32, out_len int32[B], status int32[B], score fp64[B] (optional, length-normalised).
 * ------------------------------------------------------------------------------------------------------------------ */
int mdd_beam_search_workspace_bytes(int B, int T, int C, int beam_width, size_t* bytes);
int mdd_beam_search(const float* probs, int B, int T, int C, const int64_t* input_lengths, int beam_width,
                    int blank, const double* lm, double lm_alpha, int32_t* tokens, int tok_stride, int32_t* out_len,
                    int32_t* status, double* score, void* workspace, size_t workspace_bytes, mdd_stream_t stream);

/* probs[b,t,c] = expf(log_probs[t,b,c])  -- torch.exp(prob_tensor.transpose(0,1)), ctcDecoder.py:223-225.
 * Device expf is not bit-identical to torch's CPU exp; bit-exactness of the beam search is defined at its
 * probability input (the reference boundary ctcBeamSearch.decode(probs, lens)) -- see DESIGN.md. */
int mdd_exp_transpose(const float* log_probs, int T, int B, int C, float* probs, mdd_stream_t stream);

/* ------------------------------------------------------------------------------------------------------------------
 * CTC prefix scorer for CTC/attention joint decoding (named by BASELINE.json's north_star,
 * absent from the reference as a function -- nearest reference code is BeamSearch.py:52-66,99-125).
 *
 * x      : (T,B,C) fp32 log-probs contiguous; input_lengths int64[B] or NULL.
 * p_lin  : (B,T,C) fp64 emission probabilities exp(x), utterance-major, made once per utterance batch by
 *          mdd_prefix_probs (caller-allocated, T*B*C doubles); score and advance read it instead of x.
 * Arithmetic and state are fp64 (log-likelihoods of magnitude |nll|; the beam is chosen by their differences).  The
 * scans over t run in the linear domain relative to each hypothesis' own maximum; state entries more than e^-708
 * below that maximum are stored as -inf (csrc/prefix_score.cu).
 * r_prev : (B,W,T,2) fp64 [log r_n, log r_b] per hypothesis (mdd_prefix_init fills the empty-prefix state, w = 0..W-1).
 * last   : int32 (B,W) last label of each hypothesis (-1 = empty), glen int32 (B,W) hypothesis lengths.
 * skip   : optional int32 (B,W): non-zero = hypothesis is finished (1) or dead (2); it is not extended (psi = -inf).
 * psi    : (B,W,C) fp64 out: log p_ctc(g.c, ...); psi[...,blank] = -inf.   psi_eos (B,W) = log p_ctc(g) (optional).
 * mdd_prefix_advance computes and stores the state of the chosen (parent hypothesis, label) pairs:
 * sel_hyp, sel_label int32 (B,W) -> r_new (B,W,T,2); slots with skip_new != 0 or sel_hyp < 0 are left untouched.
 * T is limited by shared memory (one hypothesis' linearised state per CTA): T <= 12000, else MDD_E_UNSUPPORTED.
 * mdd_prefix_select is the beam step between the two (nearest reference code: the sort + [:beamWidth] of
 * BeamSearch.py:52-57): per utterance, over all W*C candidates,
 *     value(w,c) = score[w] + ctc_weight * (psi[w,c] - psi_prev[w]) + (1 - ctc_weight) * att[w,c]     (c != blank)
 * with c == eos scored by psi_eos[w], finished hypotheses carried unchanged, NaN and -inf excluded; the W best
 * (ties: smaller w*C + c) become the new beam: sel_hyp / sel_label (-1 for a dead slot), score_new, psi_new, glen_new,
 * finished_new (0 live, 1 finished, 2 dead) and the token histories tokens_new (B,W,max_len) int32 (-1 padded).
 * eos < 0: no end-of-sequence candidate.  att: (B,W,C) fp32 attention-decoder log-probs.
 * ------------------------------------------------------------------------------------------------------------------ */
int mdd_prefix_init(const float* x, int T, int B, int C, const int64_t* input_lengths, int blank, int W,
                    double* r0, mdd_stream_t stream);
int mdd_prefix_probs(const float* x, int T, int B, int C, double* p_lin, mdd_stream_t stream);
int mdd_prefix_score(const double* p_lin, int T, int B, int C, const int64_t* input_lengths, int blank, int W,
                     const double* r_prev, const int32_t* last, const int32_t* glen, const int32_t* skip, double* psi,
                     double* psi_eos, mdd_stream_t stream);
int mdd_prefix_advance(const double* p_lin, int T, int B, int C, const int64_t* input_lengths, int blank, int W,
                       const double* r_prev, const int32_t* last, const int32_t* glen, const int32_t* sel_hyp,
                       const int32_t* sel_label, const int32_t* skip_new, double* r_new, mdd_stream_t stream);
int mdd_prefix_select(int B, int W, int C, int blank, int eos, double ctc_weight, int max_len, const double* psi,
                      const double* psi_eos, const float* att, const double* score, const double* psi_prev,
                      const int32_t* glen, const int32_t* finished, const int32_t* tokens, int32_t* sel_hyp,
                      int32_t* sel_label, double* score_new, double* psi_new, int32_t* glen_new, int32_t* finished_new,
                      int32_t* tokens_new, mdd_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* MDDCTC_H_ */
