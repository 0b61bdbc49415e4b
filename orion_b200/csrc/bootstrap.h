// bootstrap.h -- CKKS bootstrapping (ModRaise, CoeffsToSlots, EvalMod, SlotsToCoeffs); LG/bootstrapper.go:19-93.
#pragma once
#include "scheme.h"

namespace ob {

void boot_new(Scheme& s, const std::vector<int>& logP, int num_slots);
std::unique_ptr<Ciphertext> boot_apply(Scheme& s, const Ciphertext& in, int num_slots);
void boot_reset();
// Bootstrap through a captured CUDA graph (ORION_B200_GRAPHS, bootstrap.cu): same result as boot_apply, bit for bit.
std::unique_ptr<Ciphertext> boot_apply_graphed(Scheme& s, const Ciphertext& in, int num_slots);

}  // namespace ob
