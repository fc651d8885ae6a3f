// Batched square FP64 GEMM on the FP64 tensor pipe (DMMA) with fused diagonal scalings.
//
//   C[b] = diag(row) * opA(A[b]) * diag(mid) * opB(B[b]) * diag(col) + diag(add)
//
// This one kernel replaces the reference's hand-written loops
//   vmul!(C, A, B), vmul!(C, A, B'), vmul!(C, A', B)        src/linalg/general.jl:7-45
//   vmul!(C, A, Diagonal), vmul!(C, Diagonal, B), rvadd!     src/linalg/general.jl:16-25, 67-71
// and the materialised slice matrices of slice_matrix!      src/flavors/DQMC/slice_matrices.jl:23-76
// (the exp(+-lambda s) factor is the `mid`/`row`/`col` Scale, never written to memory).
// A stride of 0 broadcasts an operand (the hopping exponentials are shared by all chains).
#pragma once
#include "common.cuh"

namespace mcb {

struct GemmArgs {
    int n = 0, batch = 0;
    const double* A = nullptr; long long strideA = 0; bool transA = false;
    const double* B = nullptr; long long strideB = 0; bool transB = false;
    double* C = nullptr; long long strideC = 0;
    Scale row, mid, col;
    const double* add = nullptr; long long stride_add = 0;
};

int launch_gemm(const GemmArgs& g, cudaStream_t st);
int init_gemm_kernels();
// the same product with TMA-staged tiles (gemm_tma.cu) for sizes a 48 / 64 / 80 tile divides
bool gemm_tma_supported(const GemmArgs& g);
int launch_gemm_tma(const GemmArgs& g, cudaStream_t st);
int init_gemm_tma_kernels();
// true if launch_gemm sends A B' products of this size and batch to the TMA-staged kernel
bool gemm_tma_routes_abt(int n, int batch);

} // namespace mcb
