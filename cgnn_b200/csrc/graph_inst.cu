// One translation unit per padded column count (compiled with -DCGNN_DP=8|16|20|24|32): instantiates the
// graph instance kernel for the four exp policies and exports a launcher + occupancy query.
#include "graph.cuh"

#if !defined(CGNN_DP) || !defined(CGNN_HP)
#error "compile with -DCGNN_DP=<padded number of columns> -DCGNN_HP=<padded hidden width: 20 or 32>"
#endif
#define CAT2(a, b) a##b
#define CAT(a, b) CAT2(a, b)

using namespace cgnn;

template <int POL>
static cudaError_t launch(const GraphArgs& a, int grid, size_t smem, cudaStream_t st) {
    cudaError_t e = cudaFuncSetAttribute(graph_kernel<CGNN_DP, CGNN_HP, POL>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
    if (e != cudaSuccess) return e;
    graph_kernel<CGNN_DP, CGNN_HP, POL><<<grid, GR_NT, smem, st>>>(a);
    return cudaGetLastError();
}

cudaError_t CAT(CAT(cgnn_launch_graph_dp, CGNN_DP), CAT(_hp, CGNN_HP))(int pol, const GraphArgs& a, int grid, size_t smem, cudaStream_t st) {
    switch (pol) {
        case 0: return launch<0>(a, grid, smem, st);
        case 1: return launch<1>(a, grid, smem, st);
        case 2: return launch<2>(a, grid, smem, st);
        case 3: return launch<3>(a, grid, smem, st);
    }
    return cudaErrorInvalidValue;
}
