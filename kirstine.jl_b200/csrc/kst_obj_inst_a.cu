// explicit instantiations of the objective evaluation kernels (see kst_objective_kernel.cuh)
#include "kst_objective_kernel.cuh"

template int kst_obj_launch<KstSigEmax>(kst_problem *, const KstObjParams &, dim3, int);
