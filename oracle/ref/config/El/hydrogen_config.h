/* Hand-written stand-in for the header CMake generates from cmake/configure_files/hydrogen_config.h.in
 * (TEST INFRASTRUCTURE ONLY).  Only the switches the reference's GPU BLAS-1 sources
 * (src/hydrogen/blas/gpu/{Copy,Axpy,Transpose,Fill,Hadamard,Scale}.cu) look at: CUDA on, no half types,
 * no Aluminum / MPI / NVSHMEM. */
#ifndef HYDROGEN_CONFIG_H
#define HYDROGEN_CONFIG_H
#define HYDROGEN_VERSION_MAJOR "1"
#define HYDROGEN_VERSION_MINOR "3"
#define HYDROGEN_VERSION_PATCH "4"
#define HYDROGEN_RELEASE_BUILD
#define HYDROGEN_HAVE_CUDA
#define HYDROGEN_HAVE_GPU
#endif
