// One translation unit per (padded column count, padded hidden width, noise mode): instantiates the
// graph instance kernel for the four exp policies and exports a launcher + occupancy query.
#include <algorithm>
#include "graph.cuh"

#if !defined(CGNN_DP) || !defined(CGNN_HP) || !defined(CGNN_REGEN)
#error "compile with -DCGNN_DP=<padded number of columns> -DCGNN_HP=<padded hidden width: 20 or 32> -DCGNN_REGEN=<0|1: noise re-derived at every use>"
#endif
#define CAT2(a, b) a##b
#define CAT(a, b) CAT2(a, b)

using namespace cgnn;

// n_clusters < 0: only report, in *max_clusters, how many clusters of `csz` CTAs can be resident at once.
template <int POL>
static cudaError_t launch(const GraphArgs& a, int n_clusters, int csz, size_t smem, cudaStream_t st, int* max_clusters) {
    cudaError_t e = cudaFuncSetAttribute(graph_kernel<CGNN_DP, CGNN_HP, POL, CGNN_REGEN != 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
    if (e != cudaSuccess) return e;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(unsigned(std::max(n_clusters, 1) * csz));
    cfg.blockDim = dim3(GR_NT);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = unsigned(csz); at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    if (n_clusters < 0) return cudaOccupancyMaxActiveClusters(max_clusters, graph_kernel<CGNN_DP, CGNN_HP, POL, CGNN_REGEN != 0>, &cfg);
    return cudaLaunchKernelEx(&cfg, graph_kernel<CGNN_DP, CGNN_HP, POL, CGNN_REGEN != 0>, a);
}

#if CGNN_REGEN
#define CGNN_SUFFIX _regen
#else
#define CGNN_SUFFIX
#endif
cudaError_t CAT(CAT(CAT(cgnn_launch_graph_dp, CGNN_DP), CAT(_hp, CGNN_HP)), CGNN_SUFFIX)(int pol, const GraphArgs& a, int n_clusters, int csz, size_t smem, cudaStream_t st,
                                                                   int* max_clusters) {
    switch (pol) {
        case 0: return launch<0>(a, n_clusters, csz, smem, st, max_clusters);
        case 1: return launch<1>(a, n_clusters, csz, smem, st, max_clusters);
        case 2: return launch<1>(a, n_clusters, csz, smem, st, max_clusters);   // not built for this kernel: 1 is the nearest more accurate policy
        case 3: return launch<3>(a, n_clusters, csz, smem, st, max_clusters);
    }
    return cudaErrorInvalidValue;
}
