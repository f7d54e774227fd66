// Curve_BLS12381: all kernels of this parameter set are instantiated in this translation unit.
#include "ctx.cuh"

KZG_GPU_DEFINE_FACTORY(kzg_gpu_make_ctx_bls12381, Curve_BLS12381)
