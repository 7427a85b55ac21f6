// Instantiation of the row-resident quantisation kernels for float activations.
#include "quant_row_impl.cuh"
template int vxk::row_launch<float>(int, int, const vxk::RowParams&, int64_t, bool, cudaStream_t);
