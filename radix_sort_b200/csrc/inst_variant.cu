// One K3 kernel instantiation per translation unit:
//   nvcc -DB200_FAM=<0 keys|1 records|2 partition> -DB200_W=<element bytes> -DB200_V=<variant index>
#include "variant_def.cuh"
namespace b200 {
template struct VariantEntry<B200_FAM, B200_W, B200_V>;
}
