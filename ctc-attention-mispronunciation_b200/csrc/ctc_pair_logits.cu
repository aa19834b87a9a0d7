// ctc_pair_logits.cu -- the LOGITS instantiations of the CTC pair kernel (ctc_pair.cu), a translation unit of their own so
// that the two halves build in parallel.  See ctc_pair.cu for everything else.
#define MDD_PAIR_TU_LOGITS 1
#include "ctc_pair.cu"
