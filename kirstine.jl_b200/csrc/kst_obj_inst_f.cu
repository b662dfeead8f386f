// explicit instantiations of the objective evaluation kernels (kst_objective_kernel.cuh) and of the persistent
// swarm kernel built on them (kst_pso_fused.cuh)
#include "kst_pso_fused.cuh"

template int kst_obj_launch<KstVuAr>(kst_problem *, const KstObjParams &, dim3, int);
template int kst_fused_launch<KstVuAr>(kst_problem *, KstFusedParams &, dim3);
