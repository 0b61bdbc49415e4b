// bootstrap.cu -- placeholder until the bootstrapping circuit lands.
#include "bootstrap.h"

namespace ob {
void boot_new(Scheme&, const std::vector<int>&, int) { OB_FATAL("NewBootstrapper: bootstrapping is not implemented yet"); }
std::unique_ptr<Ciphertext> boot_apply(Scheme&, const Ciphertext&, int) { OB_FATAL("Bootstrap: bootstrapping is not implemented yet"); }
void boot_reset() {}
}  // namespace ob
