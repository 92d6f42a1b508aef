// Test-fixture generator (NOT product code): writes a small `.ot` archive through libtorch's own
// torch::serialize::OutputArchive -- the API tch's `VarStore::save` sits on -- so the C++ reader in
// the product is checked against a genuinely libtorch-written file (tests/golden/tiny_libtorch.ot).
// Built by oracle/Makefile into oracle/_ref/ (git-ignored).
#include <torch/serialize/archive.h>
#include <torch/torch.h>

int main(int argc, char** argv) {
  if (argc < 2) return 2;
  torch::manual_seed(0);
  torch::serialize::OutputArchive ar;
  ar.write("conv1|weight", torch::randn({4, 3, 2, 2}), false);
  ar.write("bn1|running_mean", torch::arange(4, torch::kFloat32), false);
  ar.write("layer1|0|conv1|weight", torch::randn({300}), false);          // numel > 255  -> BININT2
  ar.write("fc|bias", torch::arange(66000, torch::kFloat32) * 0.5f, false); // numel > 65535 -> BININT
  auto base = torch::arange(24, torch::kFloat32).reshape({4, 6});
  ar.write("strided|view", base.t(), false);                              // non-contiguous strides
  ar.save_to(argv[1]);
  return 0;
}
