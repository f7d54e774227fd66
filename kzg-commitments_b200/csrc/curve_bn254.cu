// Curve_BN254: all kernels of this parameter set are instantiated in this translation unit.
#include "ctx.cuh"

KZG_GPU_DEFINE_FACTORY(kzg_gpu_make_ctx_bn254, Curve_BN254)
