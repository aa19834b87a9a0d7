// ctc_fast.cu -- rescaled linear-domain CTC forward-backward (fast path).  Placeholder until the kernel lands:
// reports "unsupported" so mdd_ctc_loss uses the log-space kernel for every utterance.
#include "common.cuh"
#include "ctc_params.cuh"

namespace mdd {
bool ctc_fast_supported(int, int) { return false; }
size_t ctc_fast_ws_bytes(int, int, int) { return 0; }
int launch_ctc_fast(CtcParams, void*, cudaStream_t) { return MDD_E_UNSUPPORTED; }
}  // namespace mdd
