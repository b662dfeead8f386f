// explicit instantiations of the objective evaluation kernels (see kst_objective_kernel.cuh)
#include "kst_objective_kernel.cuh"

template int kst_obj_launch<KstLinLogReg<3>>(kst_problem *, const KstObjParams &, dim3, int);
template int kst_obj_launch<KstSinLogReg>(kst_problem *, const KstObjParams &, dim3, int);
