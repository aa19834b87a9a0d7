// capi.cu -- C ABI entry points for the CTC loss (include/mddctc.h) + library bookkeeping.
#include <cstdlib>

#include "common.cuh"
#include "ctc_params.cuh"

namespace mdd {
int64_t g_launch_count = 0;

bool pdl_enabled() {
    static int on = [] { const char* e = getenv("MDD_PDL"); return e ? atoi(e) : 1; }();
    return on != 0;
}

int sm_count() {
    static int cached[64];                                       // per device: one process may drive several GPUs
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return kNumSM;
    if (!cached[dev]) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = kNumSM;
        cached[dev] = n;
    }
    return cached[dev];
}
}
using namespace mdd;

extern "C" int mdd_version(void) { return 100; }

extern "C" const char* mdd_error_string(int code) {
    switch (code) {
        case MDD_OK: return "ok";
        case MDD_E_ARG: return "invalid argument";
        case MDD_E_WORKSPACE: return "workspace too small";
        case MDD_E_CUDA: return "CUDA runtime error (no device, or launch failure)";
        case MDD_E_UNSUPPORTED: return "shape not supported by this build";
        case MDD_E_STATIC_LIMIT:
            return "shape needs more scratch than the library's static array holds (MDD_CTC_STATIC_WS_BYTES); "
                   "call mdd_ctc_loss with a caller-supplied workspace";
        default: return "unknown error";
    }
}

extern "C" int64_t mdd_launch_count(void) { return g_launch_count; }
extern "C" void mdd_launch_count_reset(void) { g_launch_count = 0; }

namespace {
// which kernel serves this shape: 1 = the packed fp32 pair kernel where it applies, 0 = the fp64 kernel alone
int ctc_kernel_choice(int C, int Lmax, int algo) {
    if (algo == 1) return 0;
    return ctc_pair_supported(C, Lmax) ? 1 : 0;
}
struct CtcWs {
    size_t off_offsets, off_flags, off_list, off_ckpt, off_fast, total;
};
CtcWs ctc_ws_layout(int T, int B, int C, int Lmax, int algo) {
    CtcWs w{};
    size_t o = 0;
    w.off_offsets = o; o = align_up(o + (size_t)B * 8, 256);
    w.off_flags = o;   o = align_up(o + (size_t)B * 4, 256);
    w.off_list = o;    o = align_up(o + (size_t)(B + 2) * 4, 256);     // second tier: flagged utterances + their count; ticket
    w.off_ckpt = o;
    const int kern = ctc_kernel_choice(C, Lmax, algo);           // 0 fp64 kernel only, 1 pair kernel
    if (algo != 2) o = align_up(o + ctc_exact_ckpt_bytes(T, B, Lmax), 256);
    w.off_fast = o;
    if (kern == 1) o = align_up(o + ctc_pair_ws_bytes(T, B, Lmax), 256);
    w.total = o;
    return w;
}
}  // namespace

extern "C" int mdd_ctc_loss_workspace_bytes(int T, int B, int C, int max_target_len, int algo, size_t* bytes) {
    if (!bytes || T < 0 || B < 0 || C <= 0 || C > MDD_MAX_CLASSES || max_target_len < 0 || algo < 0 || algo > 2)
        return MDD_E_ARG;
    *bytes = ctc_ws_layout(T, B, C, max_target_len, algo).total;
    return MDD_OK;
}

namespace {
// row_lse == null: `log_probs` holds log-probs; else it holds logits and row_lse (T*B) receives their row log-sum-exp
int ctc_loss_impl(const float* log_probs, int T, int B, int C, int64_t stride_t, int64_t stride_b,
                  const int64_t* targets, int64_t tgt_stride, const int64_t* input_lengths,
                  const int64_t* target_lengths, int max_target_len, int blank, int zero_infinity,
                  const float* grad_out, float grad_out_scalar, float* nll, float* grad, float* loss_sum,
                  int32_t* status, float* row_lse, uint8_t* argmax, int algo, void* workspace,
                  size_t workspace_bytes, mdd_stream_t stream) {
    if (T < 0 || B < 0 || C <= 0 || C > MDD_MAX_CLASSES || max_target_len < 0 || blank < 0 || blank >= C || algo < 0 ||
        algo > 2)
        return MDD_E_ARG;
    if (B == 0) return MDD_OK;
    if ((!log_probs && T > 0) || !input_lengths || !target_lengths || !nll || (!targets && max_target_len > 0))
        return MDD_E_ARG;
    if (max_target_len > 2047) return MDD_E_UNSUPPORTED;
    const CtcWs w = ctc_ws_layout(T, B, C, max_target_len, algo);
    if (!workspace || workspace_bytes < w.total) return MDD_E_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    unsigned char* ws = (unsigned char*)workspace;

    CtcParams p{};
    p.lp = log_probs; p.lse = row_lse; p.T = T; p.B = B; p.C = C; p.stride_t = stride_t; p.stride_b = stride_b;
    p.targets = targets; p.tgt_stride = tgt_stride; p.in_len = input_lengths; p.tgt_len = target_lengths;
    p.Lmax = max_target_len; p.blank = blank; p.zero_inf = zero_infinity;
    p.grad_out = grad_out; p.gscalar = grad_out_scalar; p.nll = nll; p.grad = grad; p.status = status;
    p.tgt_offsets = reinterpret_cast<const int64_t*>(ws + w.off_offsets);
    p.flags = reinterpret_cast<int32_t*>(ws + w.off_flags);
    p.ckpt = reinterpret_cast<float*>(ws + w.off_ckpt);
    if (algo == 0) {                                             // the packed kernel's second tier (ctc_pair.cu)
        p.list = reinterpret_cast<int32_t*>(ws + w.off_list);
        p.list_count = p.list + B;
        if (loss_sum) p.ticket = reinterpret_cast<unsigned*>(p.list + B + 1);
    }
    int rc;
    if (tgt_stride == 0) {
        rc = launch_tgt_offsets(target_lengths, B, reinterpret_cast<int64_t*>(ws + w.off_offsets), st);
        if (rc) return rc;
    }
    int kern = T > 0 ? ctc_kernel_choice(C, max_target_len, algo) : 0;
    // the packed kernel addresses rows with 32-bit byte pitches
    if (kern && (stride_t * 4 >= (1ll << 32) || (int64_t)B * C * 4 >= (1ll << 32) || stride_t < 0 || stride_b < 0)) kern = 0;
    if (algo == 2 && kern != 1) return MDD_E_UNSUPPORTED;
    if (row_lse) {
        rc = launch_row_lse(log_probs, T, B, C, stride_t, stride_b, row_lse, argmax, st);
        if (rc) return rc;
    }
    if (kern) {
        // flags[b] != 0 on return -> the utterance is redone by the fp64 kernel
        rc = launch_ctc_pair(p, ws + w.off_fast, st);
        if (rc) return rc;
        if (algo == 0) {
            p.only_flagged = 1;
            p.loss_sum = loss_sum;                               // summed by the last CTA of this sweep (ticket zeroed above)
            rc = launch_ctc_exact(p, st);
            if (rc) return rc;
            if (loss_sum) return MDD_OK;
        }
    } else {
        p.only_flagged = 0;
        rc = launch_ctc_exact(p, st);
        if (rc) return rc;
    }
    if (loss_sum) {
        rc = launch_sum_nll(nll, B, loss_sum, st);
        if (rc) return rc;
    }
    return MDD_OK;
}
}  // namespace

extern "C" int mdd_ctc_loss(const float* log_probs, int T, int B, int C, int64_t stride_t, int64_t stride_b,
                            const int64_t* targets, int64_t tgt_stride, const int64_t* input_lengths,
                            const int64_t* target_lengths, int max_target_len, int blank, int zero_infinity,
                            const float* grad_out, float grad_out_scalar, float* nll, float* grad, float* loss_sum,
                            int32_t* status, int algo, void* workspace, size_t workspace_bytes, mdd_stream_t stream) {
    return ctc_loss_impl(log_probs, T, B, C, stride_t, stride_b, targets, tgt_stride, input_lengths, target_lengths,
                         max_target_len, blank, zero_infinity, grad_out, grad_out_scalar, nll, grad, loss_sum, status,
                         nullptr, nullptr, algo, workspace, workspace_bytes, stream);
}

extern "C" int mdd_ctc_loss_logits(const float* logits, int T, int B, int C, int64_t stride_t, int64_t stride_b,
                                   const int64_t* targets, int64_t tgt_stride, const int64_t* input_lengths,
                                   const int64_t* target_lengths, int max_target_len, int blank, int zero_infinity,
                                   const float* grad_out, float grad_out_scalar, float* nll, float* grad_logits,
                                   float* loss_sum, int32_t* status, float* row_lse, uint8_t* argmax, int algo,
                                   void* workspace, size_t workspace_bytes, mdd_stream_t stream) {
    if (!row_lse && (long long)T * B > 0) return MDD_E_ARG;
    return ctc_loss_impl(logits, T, B, C, stride_t, stride_b, targets, tgt_stride, input_lengths, target_lengths,
                         max_target_len, blank, zero_infinity, grad_out, grad_out_scalar, nll, grad_logits, loss_sum,
                         status, row_lse, argmax, algo, workspace, workspace_bytes, stream);
}

// ---------------------------------------------------------------------------------------------------------------------
// mdd_ctc_loss_static: mdd_ctc_loss with its scratch in a file-scope device array, for callers that cannot allocate
// (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// In host code the array's NAME is only a handle for the runtime; its device address comes from
// cudaGetSymbolAddress, taken once per device with that device current.
namespace {
__device__ __align__(256) unsigned char g_ctc_static_ws[MDD_CTC_STATIC_WS_BYTES];

int ctc_static_ws_address(void** out) {
    static void* addr[64];                                       // per device; every writer stores the same value
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { cudaGetLastError(); return MDD_E_CUDA; }
    void* p = addr[dev];
    if (!p) {
        if (cudaGetSymbolAddress(&p, g_ctc_static_ws) != cudaSuccess || !p) { cudaGetLastError(); return MDD_E_CUDA; }
        addr[dev] = p;
    }
    *out = p;
    return MDD_OK;
}
}  // namespace

extern "C" int mdd_ctc_loss_static(const float* log_probs, int T, int B, int C, int64_t stride_t, int64_t stride_b,
                                   const int64_t* targets, int64_t tgt_stride, const int64_t* input_lengths,
                                   const int64_t* target_lengths, int max_target_len, int blank, int zero_infinity,
                                   const float* grad_out, float grad_out_scalar, float* nll, float* grad,
                                   float* loss_sum, int32_t* status, int algo, mdd_stream_t stream) {
    if (T < 0 || B < 0 || C <= 0 || C > MDD_MAX_CLASSES || max_target_len < 0 || blank < 0 || blank >= C || algo < 0 ||
        algo > 2)
        return MDD_E_ARG;
    if (B == 0) return MDD_OK;
    const CtcWs w = ctc_ws_layout(T, B, C, max_target_len, algo);
    if (w.total > (size_t)MDD_CTC_STATIC_WS_BYTES) return MDD_E_STATIC_LIMIT;
    void* ws = nullptr;
    const int rc = ctc_static_ws_address(&ws);
    if (rc) return rc;
    // the array keeps the previous call's contents: target offsets, flags and the flagged list start from zero
    if (cudaMemsetAsync(ws, 0, w.off_ckpt, (cudaStream_t)stream) != cudaSuccess) { cudaGetLastError(); return MDD_E_CUDA; }
    return ctc_loss_impl(log_probs, T, B, C, stride_t, stride_b, targets, tgt_stride, input_lengths, target_lengths,
                         max_target_len, blank, zero_infinity, grad_out, grad_out_scalar, nll, grad, loss_sum, status,
                         nullptr, nullptr, algo, ws, (size_t)MDD_CTC_STATIC_WS_BYTES, stream);
}

extern "C" int mdd_copy_h2d_2d(void* dst, size_t dst_pitch, const void* src_host, size_t src_pitch, size_t width_bytes,
                               size_t height, mdd_stream_t stream) {
    if (width_bytes == 0 || height == 0) return MDD_OK;
    if (!dst || !src_host || dst_pitch < width_bytes || src_pitch < width_bytes) return MDD_E_ARG;
    if (cudaMemcpy2DAsync(dst, dst_pitch, src_host, src_pitch, width_bytes, height, cudaMemcpyHostToDevice,
                          (cudaStream_t)stream) != cudaSuccess) {
        cudaGetLastError();
        return MDD_E_CUDA;
    }
    return MDD_OK;
}

extern "C" int mdd_ctc_scale_grad(float* grad, int T, int B, int C, const float* scale, float scalar,
                                  mdd_stream_t stream) {
    if (!grad || T < 0 || B < 0 || C <= 0) return MDD_E_ARG;
    return launch_scale_grad(grad, T, B, C, scale, scalar, (cudaStream_t)stream);
}
