// One translation unit per build of the pairwise instance kernel (compiled with -DCGNN_PW_NT=<threads per CTA>
// -DCGNN_PW_CTAS=<CTAs per SM> -DCGNN_PW_NS=<namespace> -DCGNN_PW_TAG=<suffix>): instantiates it for the four exp
// policies and exports launcher, occupancy and shared-memory size.
#include "pairwise.cuh"

#if !defined(CGNN_PW_TAG)
#error "compile with -DCGNN_PW_NT=... -DCGNN_PW_CTAS=... -DCGNN_PW_NS=... -DCGNN_PW_TAG=..."
#endif
#define CAT2(a, b) a##b
#define CAT(a, b) CAT2(a, b)

using namespace cgnn;
using namespace cgnn::CGNN_PW_NS;

template <int POL>
static cudaError_t launch(const PairwiseArgs& a, int grid, size_t smem, cudaStream_t st) {
    cudaError_t e = cudaFuncSetAttribute(pairwise_kernel<POL>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
    if (e != cudaSuccess) return e;
    pairwise_kernel<POL><<<grid, PW_NT, smem, st>>>(a);
    return cudaGetLastError();
}

template <int POL>
static int occupancy(size_t smem) {
    int n = 0;
    if (cudaFuncSetAttribute(pairwise_kernel<POL>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)) != cudaSuccess ||
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, pairwise_kernel<POL>, PW_NT, smem) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

cudaError_t CAT(cgnn_pw_launch_, CGNN_PW_TAG)(int pol, const PairwiseArgs& a, int grid, size_t smem, cudaStream_t st) {
    switch (pol) {
        case 0: return launch<0>(a, grid, smem, st);
        case 1: return launch<1>(a, grid, smem, st);
        case 2: return launch<2>(a, grid, smem, st);
        case 3: return launch<3>(a, grid, smem, st);
    }
    return cudaErrorInvalidValue;
}

int CAT(cgnn_pw_occupancy_, CGNN_PW_TAG)(int pol, size_t smem) {
    switch (pol) {
        case 0: return occupancy<0>(smem);
        case 1: return occupancy<1>(smem);
        case 2: return occupancy<2>(smem);
        case 3: return occupancy<3>(smem);
    }
    return 0;
}

size_t CAT(cgnn_pw_smem_, CGNN_PW_TAG)(int np_max, int h, int fast_k) { return pw_smem_bytes(np_max, h, fast_k); }
