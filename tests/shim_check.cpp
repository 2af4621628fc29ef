// Compile-only check (tests/test_abi.py): every wrapper of include/El.hpp instantiates for double and float against
// the C ABI.  Nothing here runs.
#include <El.hpp>
template <typename T> void inst(El::Grid& g) {
    El::DistMatrix<T, El::MC, El::MR> A(10, 10, g), C(10, 10, g);
    El::DistMatrix<T, El::MC, El::STAR> As(10, 10, g);
    El::DistMatrix<T, El::STAR, El::MR> Bs(10, 10, g);
    El::DistMatrix<T, El::MR, El::STAR> Bt(10, 10, g);
    El::Gemm(El::NORMAL, El::TRANSPOSE, T(1), A, A, T(0), C, El::GEMM_SUMMA_A);
    El::Gemm(El::NORMAL, El::NORMAL, T(1), A, A, C);
    El::LocalGemm(El::NORMAL, El::NORMAL, T(1), As, Bs, T(0), C);
    El::LocalGemm(El::NORMAL, El::TRANSPOSE, T(1), As, Bt, C);
    El::Contract(As, A);
    El::AxpyContract(T(1), As, A);
    El::TransposeAxpyContract(T(1), Bt, A);
    El::Transpose(A, C);
    El::Copy(A, As);
    As = A;
    El::Scale(T(2), A);
    El::Zero(A);
    El::Fill(A, T(3));
    El::Axpy(T(2), As, A);
    El::TransposeAxpy(T(2), Bt, A);
    El::AdjointAxpy(T(2), A, C);
    El::Hadamard(A, C, C);
    El::SyncInfoGPU si;
    T* p = A.Buffer();
    El::gpu_blas::Gemm(El::NORMAL, El::NORMAL, 2, 2, 2, T(1), p, 2, p, 2, T(0), p, 2, si);
    El::gpu_blas::Copy(El::TRANSPOSE, 2, 2, p, 2, p, 2, si);
    El::gpu_blas::Copy(El::NORMAL, 2, 2, p, 2, p, 2, si);
    El::InterDeviceCopy<El::Device::CPU, El::Device::GPU>::MemCopy2DAsync(p, 2, p, 2, 2, 2, si.stream_);
    El::InterDeviceCopy<El::Device::GPU, El::Device::CPU>::MemCopy1DAsync(p, p, 4, si.stream_);
    El::gpu_blas::Axpy(2, 2, T(1), p, 2, p, 2, si);
    El::gpu_blas::Scale(2, 2, T(1), p, 2, si);
    El::gpu_blas::Axpy(El::TRANSPOSE, 2, 2, T(1), p, 2, p, 2, si);
    El::gpu_blas::Copy(2, 2, p, 1, 2, p, 1, 2, si);
    El::gpu_blas::Axpy(2, 2, T(1), p, 1, 2, p, 1, 2, si);
    El::gpu_blas::Copy(4, p, 1, p, 1, si);
    El::gpu_blas::Axpy(4, T(1), p, 1, p, 1, si);
    El::gpu_blas::Scale(4, T(1), p, 1, si);
    El::hydrogen_gpu::Fill_GPU_impl<T>(2, 2, T(1), p, 2, si.stream_);
    El::hydrogen_gpu::Hadamard_GPU_impl<T>(2, 2, p, 1, 2, p, 1, 2, p, 1, 2, si.stream_);
    El::copy::util::InterleaveMatrix(2, 2, p, 1, 2, p, 1, 2, si);
    El::copy::util::RowStridedPack(2, 2, 0, 1, p, 2, p, 4, si);
    El::copy::util::RowStridedUnpack(2, 2, 0, 1, p, 4, p, 2, si);
    El::copy::util::ColStridedPack(2, 2, 0, 1, p, 2, p, 4, si);
    El::copy::util::ColStridedUnpack(2, 2, 0, 1, p, 4, p, 2, si);
    El::copy::util::PartialRowStridedPack(2, 2, 0, 1, 1, 1, 0, 0, p, 2, p, 4, si);
    El::copy::util::PartialRowStridedUnpack(2, 2, 0, 1, 1, 1, 0, 0, p, 4, p, 2, si);
    El::copy::util::PartialColStridedPack(2, 2, 0, 1, 1, 1, 0, 0, p, 2, p, 4, si);
    El::copy::util::PartialColStridedUnpack(2, 2, 0, 1, 1, 1, 0, 0, p, 4, p, 2, si);
    El::copy::util::StridedPack(2, 2, 0, 1, 0, 1, p, 2, p, 4, si);
    El::copy::util::StridedUnpack(2, 2, 0, 1, 0, 1, p, 4, p, 2, si);
    El::axpy::util::InterleaveMatrixUpdate(T(1), 2, 2, p, 1, 2, p, 1, 2, si);
    El::mpi::AllGather(p, 1, p, 1, g.MRComm(), si);
    El::mpi::AllToAll(p, 1, p, 1, g.MCComm(), si);
    El::mpi::SendRecv(p, 1, 0, p, 1, 0, g.VCComm(), si);
    El::mpi::Broadcast(p, 1, 0, g.VRComm(), si);
    El::mpi::ReduceScatter(p, p, 1, g.RowComm(), si);
    (void)g.ColComm().Size();
    (void)g.ColComm().Rank();
}
template void inst<double>(El::Grid&);
template void inst<float>(El::Grid&);
int main() { return 0; }
