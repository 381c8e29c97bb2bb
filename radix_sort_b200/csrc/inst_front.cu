// K1 / K2 / copy-back and the launch sequence for one (family, element width):
//   nvcc -DB200_FAM=<0 keys|1 records|2 partition> -DB200_W=<element bytes>
#include "front_def.cuh"
namespace b200 {
template int sort_front<B200_FAM, B200_W>(const SortCall&);
template size_t temp_bytes_front<B200_FAM, B200_W>(size_t);
}
