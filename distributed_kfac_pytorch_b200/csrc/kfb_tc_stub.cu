// Temporary: tensor-core SYRK not built yet -> everything goes to the SIMT SYRK.
#include "kfb_factor.cuh"
namespace kfb {
int launch_syrk(const FactorPlan& plan, void* workspace, Context*, cudaStream_t stream) { return launch_syrk_simt(plan, workspace, stream); }
}
