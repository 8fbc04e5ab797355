/*
 * sw_fill_inst.cu -- one explicit instantiation of the fill kernel.  The
 * Makefile compiles this file once per (INST_T, INST_G, INST_K) of sw_config.h so
 * that the 66 kernels build in parallel.
 */
#include "sw_fill_kernel.cuh"
#if !defined(INST_T) || !defined(INST_G) || !defined(INST_K)
#error "compile with -DINST_T=<S16x2|S32> -DINST_G=<8|32> -DINST_K=<1..24>"
#endif
namespace mglsw {
template cudaError_t launch_fill<INST_T, INST_G, INST_K>(const FillArgs &, u32, cudaStream_t);
}
