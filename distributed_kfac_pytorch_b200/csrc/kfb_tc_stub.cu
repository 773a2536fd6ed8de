// Temporary: tensor-core back ends not built yet.
#include "kfb_factor.cuh"
#include "kfb_gemm.cuh"
namespace kfb {
bool gemm_tc_eligible(const GemmProb&) { return false; }
int gemm_grouped_tc(Context*, const GemmProb*, int, float*, cudaStream_t) { KFB_FAIL(KFB_ERR_UNSUPPORTED, "tcgen05 GEMM not built"); }
int launch_syrk(const FactorPlan& plan, void* workspace, Context*, cudaStream_t stream) { return launch_syrk_simt(plan, workspace, stream); }
}
