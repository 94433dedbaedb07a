// fused_k0.cu -- the fused collision step (fused_kernel.cuh) instantiated for the collision
// kernel KERNEL_GOLOVIN; one translation unit per kernel so that they compile in parallel.
#include "fused_kernel.cuh"

PYSDM_FUSED_DEFINE_DISPATCH(KERNEL_GOLOVIN)
