// Instantiation of the row-resident quantisation kernels for __half activations.
#include "quant_row_impl.cuh"
template int vxk::row_launch<__half>(int, int, const vxk::RowParams&, int64_t, bool, cudaStream_t);
