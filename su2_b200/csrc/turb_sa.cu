// a25-a27. Spalart-Allmaras transport on the device (placeholder until the SA kernels land in this file).
#include "internal.hpp"
#include "kernels.hpp"
namespace su2b200 {
void turb_preprocessing(su2b200_ctx*) { throw Error("SA turbulence kernels not built yet"); }
void turb_set_time_step(su2b200_ctx*) { throw Error("SA turbulence kernels not built yet"); }
void turb_space_integration(su2b200_ctx*) { throw Error("SA turbulence kernels not built yet"); }
void turb_complete_implicit(su2b200_ctx*) { throw Error("SA turbulence kernels not built yet"); }
void turb_postprocessing(su2b200_ctx*) { throw Error("SA turbulence kernels not built yet"); }
}  // namespace su2b200
