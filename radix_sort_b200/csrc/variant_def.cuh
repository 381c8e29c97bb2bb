// variant_def.cuh -- definition of VariantEntry<FAM, W, V>: one K3 instantiation.
// Included only by inst_variant.cu.
#pragma once
#include "sort_impl.cuh"

namespace b200 {

template <int FAM, int W, int V>
cudaError_t VariantEntry<FAM, W, V>::launch(const PassArgs<typename FamilyTraits<FAM, W>::type>& a, cudaStream_t st) {
    constexpr VariantDesc d = Variants<FAM, W>::v[V];
    return PassLauncher<typename FamilyTraits<FAM, W>::type, d.kernel, d.threads, d.items, d.min_blocks, d.lbb>::launch(a, st);
}

}  // namespace b200
