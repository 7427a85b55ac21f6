// Instantiation of the row-resident quantisation kernels for __nv_bfloat16 activations.
#include "quant_row_impl.cuh"
template int vxk::row_launch<__nv_bfloat16>(int, int, const vxk::RowParams&, int64_t, bool, cudaStream_t);
