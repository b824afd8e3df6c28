// fc_plate_v8.cu -- instantiations of the specialised plate kernel: GRAD = false, TY = 8
// (split over translation units so the build parallelises; see fc_plate.cuh).
#include "fc_plate.cuh"
namespace fc {
FC_PLATE_FAST_INSTANTIATE(false, 8)
}  // namespace fc
