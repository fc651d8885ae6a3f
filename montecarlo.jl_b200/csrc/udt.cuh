// Batched UDT (column-pivoted Householder QR) and pivoted right triangular solve.
#pragma once
#include "common.cuh"

namespace mcb {

// udt_AVX_pivot! (src/linalg/UDT.jl:192-306): per chain, X -> U (explicit Q), D = |diag R|,
// T = D^-1 R P' (apply_pivot) or the "dirty" D^-1 R (upper triangle valid) + pivot.
// X is clobbered (workspace). pivot is 0-based int32.
int launch_udt(int n, int batch, double* X, long long sX, double* U, long long sU, double* D,
               long long sD, double* T, long long sT, int* pivot, long long sP, int apply_pivot,
               cudaStream_t st);

// rdivp! (src/linalg/general.jl:138-166): out = in[:, pivot] * triu(T)^-1  (out != in)
int launch_rdivp(int n, int batch, const double* in, long long sIn, double* out, long long sOut,
                 const double* T, long long sT, const int* pivot, long long sP, cudaStream_t st);

int init_udt_kernels();
// cluster-per-chain factorisation (udt_cluster.cu): 8 CTAs share one matrix through distributed shared memory
bool udt_cluster_supported(int n);
int init_udt_cluster_kernels();
int launch_udt_cluster(int n, int batch, double* X, double* U, double* D, double* T, int* pivot, int apply_pivot,
                       cudaStream_t st);
// blocked panel-pivoted QR on the tensor pipe (udt_blocked.cu): n % 16 == 0, 128 <= n <= 400
// Stream-ordered scratch of the blocked / resident factorisations: cudaMallocAsync from a pool of the library's own, one
// per device, that keeps what it has been given (release threshold = max). The device's default pool hands its memory
// back to the driver at every synchronisation, so a host that synchronises after each call paid a fresh allocation in
// the next one. MCB_SCRATCH_POOL=0: the default pool (comparisons).
int scratch_alloc(void** p, size_t bytes, cudaStream_t st);
bool udt_blocked_supported(int n);
int init_udt_blocked_kernels();
int launch_udt_blocked(int n, int batch, double* X, double* U, double* D, long long sD, double* T, int* pivot,
                       long long sP, int apply_pivot, cudaStream_t st);
// the same factorisation as one shared-memory-resident kernel per chain (udt_resident.cu): n = 128, 144
bool udt_resident_supported(int n);
int init_udt_resident_kernels();
int launch_udt_resident(int n, int batch, const double* X, double* U, double* D, long long sD, double* T, int* pivot,
                        long long sP, int apply_pivot, cudaStream_t st);
// rdivp! as a blocked forward substitution on the tensor pipe (rdivp_blocked.cu): n % 16 == 0, 256 <= n <= 400
bool rdivp_blocked_supported(int n);
int init_rdivp_blocked_kernels();
int launch_rdivp_blocked(int n, int batch, const double* in, long long sIn, double* out, long long sOut, const double* T,
                         long long sT, const int* pivot, long long sP, cudaStream_t st);
// testing aid: route every size through udt_kernel / rdivp_kernel
void set_force_general_udt(bool on);
bool force_general_udt();

} // namespace mcb
