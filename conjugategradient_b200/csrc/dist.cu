// dist.cu — one-process-per-GPU sharded solve (placeholder until the NCCL path lands in this round).
#include "common.cuh"
using namespace mgcg;
extern "C" {
int mgcg_comm_unique_id(void *) { set_error("mgcg_comm_unique_id: distributed path not built yet"); return MGCG_ERR_UNSUPPORTED; }
int mgcg_comm_init(mgcg_ctx *, const void *, int, int) { set_error("mgcg_comm_init: distributed path not built yet"); return MGCG_ERR_UNSUPPORTED; }
int mgcg_setup_dist(mgcg_ctx *, int, int, int, int, const double *, const int *, const int *, const mgcg_options *) { set_error("mgcg_setup_dist: distributed path not built yet"); return MGCG_ERR_UNSUPPORTED; }
int mgcg_solve_dist(mgcg_ctx *, const double *, double *, mgcg_result *) { set_error("mgcg_solve_dist: distributed path not built yet"); return MGCG_ERR_UNSUPPORTED; }
}
