// SwinIR forward (network_swinir.py:805-840) — graph runner and the shifted-window attention kernel.
#include "model.h"

namespace irb {

int build_swinir(const irb_model_desc& d, const Blob& blob, int device, Model** out) {
    (void)d; (void)blob; (void)device;
    *out = nullptr;
    set_error("SwinIR runner is not linked into this build yet");
    return 1;
}

}  // namespace irb
