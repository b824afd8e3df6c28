// fc_plate_g1.cu -- instantiations of the specialised plate kernel: GRAD = true, TY = 1
// (split over translation units so the build parallelises; see fc_plate.cuh).
#include "fc_plate.cuh"
namespace fc {
FC_PLATE_FAST_INSTANTIATE(true, 1)
}  // namespace fc
