/*
 * sw_fill_pipe_inst.cu -- one explicit instantiation of the pipelined long-pair fill kernel
 * (sw_fill_pipe.cuh); compiled once per INST_K by the Makefile.
 */
#include "sw_fill_pipe.cuh"
#if !defined(INST_K)
#error "compile with -DINST_K=<7..24>"
#endif
namespace mglsw {
template cudaError_t launch_fill_pipe<S32, INST_K>(const FillArgs &, u32, bool, cudaStream_t);
}
