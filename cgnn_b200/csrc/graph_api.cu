// C-ABI of the graph path: cgnn_b200_run_graph, cgnn_b200_graph_step_debug.  See include/cgnn_b200.h.
#include <cmath>
#include <cstdlib>
#include <vector>

#include "ctx.h"
#include "graph.cuh"

using namespace cgnn;

// Gather rows (optionally a subset) of the row-major data matrix into the padded layout T[NP][DP].
// Row indices of a subset are the caller's: one outside [0, n_rows_total) is never dereferenced - the element becomes 0
// and *bad is raised, which the host turns into CGNN_ERR_ARG after the batch.  The sets are walked by a grid-stride loop
// (gridDim.y is capped at 65535, a batch may hold more instances than that).
static __global__ void gather_rows_kernel(const float* __restrict__ data, int n_var, int n_rows_total, const int* __restrict__ subset,
                                   int n_sets, int N, int NP, int DP, float* __restrict__ T, int* __restrict__ bad) {
    for (size_t set = blockIdx.y; set < size_t(n_sets); set += gridDim.y) {
        const int* sub = subset ? subset + set * N : nullptr;
        float* dst = T + set * size_t(NP) * DP;
        for (int q = blockIdx.x * blockDim.x + threadIdx.x; q < NP * DP; q += gridDim.x * blockDim.x) {
            const int r = q / DP, c = q - r * DP;
            float v = 0.f;
            if (r < N && c < n_var) {
                const int src = sub ? sub[r] : r;
                if (src >= 0 && src < n_rows_total) v = data[size_t(src) * n_var + c];
                else *bad = 1;
            }
            dst[q] = v;
        }
    }
}


#define DECL(DP_)                                                                                  \
    cudaError_t cgnn_launch_graph_dp##DP_##_hp20(int, const GraphArgs&, int, int, size_t, cudaStream_t, int*); \
    cudaError_t cgnn_launch_graph_dp##DP_##_hp32(int, const GraphArgs&, int, int, size_t, cudaStream_t, int*); \
    cudaError_t cgnn_launch_graph_dp##DP_##_hp20_regen(int, const GraphArgs&, int, int, size_t, cudaStream_t, int*); \
    cudaError_t cgnn_launch_graph_dp##DP_##_hp32_regen(int, const GraphArgs&, int, int, size_t, cudaStream_t, int*);
DECL(8) DECL(16) DECL(20) DECL(22) DECL(24) DECL(32)
#undef DECL

// padded column count of the generated / observed matrices.  22 (the bundled example's width) has rows of 88 bytes:
// 8-byte vector loads instead of 16-byte ones, but 11 instead of 12 coordinate pairs in every FFMA2 loop; the
// random-feature mode keeps 16-byte rows (its frequency table is read with 128-bit loads).
static int pick_dp(int n_var, bool fast) {
    const int choices[6] = {8, 16, 20, 22, 24, 32};
    for (int c : choices)
        if (n_var <= c && !(fast && c == 22)) return c;
    return 0;
}

static cudaError_t launch_graph(int dp, int hp, int pol, const GraphArgs& a, int n_clusters, int csz, size_t smem,
                                cudaStream_t st, int* max_clusters) {
#define CASE(DP_)                                                                                              \
    case DP_:                                                                                                  \
        if (a.noise_regen)                                                                                     \
            return hp == 20 ? cgnn_launch_graph_dp##DP_##_hp20_regen(pol, a, n_clusters, csz, smem, st, max_clusters)    \
                            : cgnn_launch_graph_dp##DP_##_hp32_regen(pol, a, n_clusters, csz, smem, st, max_clusters);   \
        return hp == 20 ? cgnn_launch_graph_dp##DP_##_hp20(pol, a, n_clusters, csz, smem, st, max_clusters)    \
                        : cgnn_launch_graph_dp##DP_##_hp32(pol, a, n_clusters, csz, smem, st, max_clusters);
    switch (dp) { CASE(8) CASE(16) CASE(20) CASE(22) CASE(24) CASE(32) }
#undef CASE
    return cudaErrorInvalidValue;
}

static int run_graph_impl(cgnn_ctx* ctx, const cgnn_hparams* hp, const float* data, int n_rows_total,
                          const cgnn_graph_batch* gb, int n_runs, const int* row_subsets, int n_rows,
                          unsigned long long uid_base, const unsigned long long* graph_uids, const float* noise, const float* init_params, int ld_params,
                          float* scores, float* loss_trace, float* params_out, float* const dbg_host[5], int dbg_steps,
                          const float* fourier_wz = nullptr) {
    if (!ctx) return CGNN_ERR_ARG;
    ctx->err.clear();
    int rc = check_hp(ctx, hp);
    if (rc) return rc;
    if (hp->h_layer_dim > CGNN_MAX_H_GRAPH)
        return fail(ctx, CGNN_ERR_LIMIT, "graph path: h_layer_dim=%d exceeds CGNN_MAX_H_GRAPH=%d", hp->h_layer_dim, CGNN_MAX_H_GRAPH);
    if (!data || !gb) return fail(ctx, CGNN_ERR_ARG, "data / graph batch is NULL");
    if (!gb->node_col || !gb->n_parents || !gb->n_inputs || !gb->in_ptr || !gb->in_idx)
        return fail(ctx, CGNN_ERR_ARG, "graph batch has a NULL array");
    if (gb->n_graphs < 1 || n_runs < 1) return fail(ctx, CGNN_ERR_ARG, "n_graphs and n_runs must be >= 1");
    const int n_var = gb->n_var, n_noise = gb->n_noise_cols;
    if (n_var < 1 || n_var > CGNN_MAX_VARS) return fail(ctx, CGNN_ERR_LIMIT, "n_var=%d outside [1,%d]", n_var, CGNN_MAX_VARS);
    if (n_noise < n_var) return fail(ctx, CGNN_ERR_ARG, "n_noise_cols=%d < n_var=%d (every node has its own noise column)", n_noise, n_var);
    if (n_noise > 4096) return fail(ctx, CGNN_ERR_LIMIT, "too many noise columns");
    if (n_rows < 1 || n_rows > CGNN_MAX_ROWS_GRAPH)
        return fail(ctx, CGNN_ERR_LIMIT, "n_rows=%d outside [1,%d] (subsample first, CGNN.py:183-185)", n_rows, CGNN_MAX_ROWS_GRAPH);
    if (row_subsets ? (n_rows > n_rows_total) : (n_rows != n_rows_total))
        return fail(ctx, CGNN_ERR_ARG, "n_rows=%d inconsistent with n_rows_total=%d", n_rows, n_rows_total);
    if (!scores && !dbg_host) return fail(ctx, CGNN_ERR_ARG, "scores is NULL");
    CU(cudaSetDevice(ctx->device));

    const int h = hp->h_layer_dim;
    const int HP = h <= 20 ? 20 : 32;                    // padded hidden width of the shared-memory layout
    const int n_graphs = gb->n_graphs;
    const int n_inst = dbg_host ? 1 : n_graphs * n_runs;
    const int epochs = hp->train_epochs + hp->test_epochs;

    // ---- descriptors ---------------------------------------------------------------------------
    std::vector<GraphDesc> descs(n_graphs);
    int p_max = 0, p_pad_max = 0;
    for (int g = 0; g < n_graphs; ++g) {
        GraphDesc& d = descs[g];
        std::memset(&d, 0, sizeof d);
        d.n_var = n_var; d.n_noise = n_noise;
        const int base = gb->in_ptr[g], end = gb->in_ptr[g + 1];
        if (end - base > GR_MAX_IN) return fail(ctx, CGNN_ERR_LIMIT, "graph %d has %d inputs in total (limit %d)", g, end - base, GR_MAX_IN);
        int io = 0, po = 0, ppad = 0;
        unsigned int seen = 0;
        for (int k = 0; k < n_var; ++k) {
            const int col = gb->node_col[g * n_var + k], npar = gb->n_parents[g * n_var + k], nin = gb->n_inputs[g * n_var + k];
            if (col < 0 || col >= n_var || (seen >> col & 1u)) return fail(ctx, CGNN_ERR_ARG, "graph %d: node_col is not a permutation", g);
            if (npar < 0 || nin <= npar || base + io + nin > end) return fail(ctx, CGNN_ERR_ARG, "graph %d node %d: bad input counts", g, k);
            for (int p = 0; p < nin; ++p) {
                const int id = gb->in_idx[base + io + p];
                if (p < npar) {
                    if (id < 0 || id >= n_var || !(seen >> id & 1u))
                        return fail(ctx, CGNN_ERR_ARG, "graph %d node %d: parent column %d is not generated yet (not a build order)", g, k, id);
                } else if (id < 0 || id >= n_noise) {
                    return fail(ctx, CGNN_ERR_ARG, "graph %d node %d: noise column %d out of range", g, k, id);
                }
                d.in_idx[io + p] = short(id);
            }
            seen |= 1u << col;
            d.col[k] = short(col); d.n_par[k] = short(npar); d.n_in[k] = short(nin);
            d.in_off[k] = io; d.par_off[k] = po; d.pad_off[k] = ppad;
            io += nin;
            po += nin * h + 2 * h + 1;
            ppad += (nin + 2) * HP + 4;
        }
        if (base + io != end) return fail(ctx, CGNN_ERR_ARG, "graph %d: in_ptr does not match the sum of n_inputs", g);
        d.n_in_total = io; d.n_params = po; d.n_params_pad = ppad;
        p_max = std::max(p_max, po);
        p_pad_max = std::max(p_pad_max, ppad);
    }
    if ((init_params || params_out) && ld_params < p_max)
        return fail(ctx, CGNN_ERR_ARG, "ld_params=%d smaller than the largest parameter count %d", ld_params, p_max);

    const int DP = pick_dp(n_var, hp->fast_mmd_vectors > 0);
    const int NP = 32 * ((n_rows + 31) / 32);
    GraphArgs a{};
    const void* dp = nullptr;
    void* op = nullptr;

    // ---- observed matrix in the padded layout, one copy per row subset ------------------------------
    if ((rc = stage_in(ctx, ctx->gdata, data, sizeof(float) * size_t(n_rows_total) * n_var, &dp))) return rc;
    const float* d_data = static_cast<const float*>(dp);
    const int n_T = row_subsets ? n_inst : 1;
    const int* d_sub = nullptr;
    if (row_subsets) {
        if ((rc = stage_in(ctx, ctx->gsub, row_subsets, sizeof(int) * size_t(n_inst) * n_rows, &dp))) return rc;
        d_sub = static_cast<const int*>(dp);
    }
    if ((rc = ensure(ctx, ctx->gT, sizeof(float) * size_t(n_T) * NP * DP))) return rc;
    if ((rc = ensure(ctx, ctx->counter, 256))) return rc;
    int* d_bad = static_cast<int*>(ctx->counter.p) + 48;          // words 0 and 32 are the two queue heads
    CU(cudaMemsetAsync(ctx->counter.p, 0, 256, ctx->stream));
    {
        dim3 grid((NP * DP + 255) / 256, std::min(n_T, 65535));
        gather_rows_kernel<<<grid, 256, 0, ctx->stream>>>(d_data, n_var, n_rows_total, d_sub, n_T, n_rows, NP, DP,
                                                           static_cast<float*>(ctx->gT.p), d_bad);
        CU(cudaGetLastError());
    }
    if (row_subsets) {       // refuse the batch before the long kernel runs on zero-filled rows
        int bad = 0;
        CU(cudaMemcpyAsync(&bad, d_bad, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
        CU(cudaStreamSynchronize(ctx->stream));
        if (bad) return fail(ctx, CGNN_ERR_ARG, "row_subsets holds an index outside [0, n_rows_total=%d)", n_rows_total);
    }
    a.T = static_cast<const float*>(ctx->gT.p);
    if (row_subsets) {
        std::vector<int> ti(n_inst);
        for (int k = 0; k < n_inst; ++k) ti[k] = k;
        if ((rc = ensure(ctx, ctx->gtidx, sizeof(int) * n_inst))) return rc;
        CU(cudaMemcpyAsync(ctx->gtidx.p, ti.data(), sizeof(int) * n_inst, cudaMemcpyHostToDevice, ctx->stream));
        CU(cudaStreamSynchronize(ctx->stream));
        a.t_index = static_cast<const int*>(ctx->gtidx.p);
    }
    if ((rc = ensure(ctx, ctx->gdesc, sizeof(GraphDesc) * n_graphs))) return rc;
    CU(cudaMemcpyAsync(ctx->gdesc.p, descs.data(), sizeof(GraphDesc) * n_graphs, cudaMemcpyHostToDevice, ctx->stream));
    a.desc = static_cast<const GraphDesc*>(ctx->gdesc.p);

    // instance table: larger parameter count first
    std::vector<int2> tab;
    tab.reserve(size_t(n_graphs) * n_runs);
    std::vector<int> order(n_graphs);
    for (int g = 0; g < n_graphs; ++g) order[g] = g;
    std::stable_sort(order.begin(), order.end(), [&](int x, int y) { return descs[x].n_params > descs[y].n_params; });
    std::vector<unsigned long long> uids;
    uids.reserve(size_t(n_graphs) * n_runs);
    if (dbg_host) { tab.push_back(make_int2(0, 0)); uids.push_back(uid_base); }
    else
        for (int g : order)
            for (int r = 0; r < n_runs; ++r) {
                tab.push_back(make_int2(g, g * n_runs + r));
                uids.push_back(graph_uids ? graph_uids[g] + (unsigned long long)r : uid_base + (unsigned long long)(g * n_runs + r));
            }
    if ((rc = ensure(ctx, ctx->inst, sizeof(int2) * n_inst))) return rc;
    CU(cudaMemcpyAsync(ctx->inst.p, tab.data(), sizeof(int2) * n_inst, cudaMemcpyHostToDevice, ctx->stream));
    a.inst = static_cast<const int2*>(ctx->inst.p);
    if ((rc = ensure(ctx, ctx->guid, sizeof(unsigned long long) * n_inst))) return rc;
    CU(cudaMemcpyAsync(ctx->guid.p, uids.data(), sizeof(unsigned long long) * n_inst, cudaMemcpyHostToDevice, ctx->stream));
    a.inst_uid = static_cast<const unsigned long long*>(ctx->guid.p);

    if ((rc = stage_in(ctx, ctx->noise, noise, sizeof(float) * size_t(n_inst) * epochs * n_noise * n_rows, &dp))) return rc;
    a.noise = static_cast<const float*>(dp);
    if ((rc = stage_in(ctx, ctx->init, init_params, sizeof(float) * size_t(n_inst) * ld_params, &dp))) return rc;
    a.init_params = static_cast<const float*>(dp);
    if ((rc = stage_out(ctx, ctx->scores, scores, sizeof(float) * n_inst, &op))) return rc;
    a.scores = static_cast<float*>(op);
    if ((rc = stage_out(ctx, ctx->trace, loss_trace, sizeof(float) * size_t(n_inst) * epochs, &op))) return rc;
    a.trace = static_cast<float*>(op);
    if ((rc = stage_out(ctx, ctx->pout, params_out, sizeof(float) * size_t(n_inst) * ld_params, &op))) return rc;
    a.params_out = static_cast<float*>(op);
    a.counter = static_cast<int*>(ctx->counter.p);             // zeroed above, before the gather

    size_t dbg_bytes[5] = {0, 0, 0, 0, 0};
    float* dbg_dev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    if (dbg_host) {
        const size_t P0 = size_t(descs[0].n_params);
        dbg_bytes[0] = sizeof(float) * dbg_steps;
        dbg_bytes[1] = dbg_bytes[2] = sizeof(float) * dbg_steps * size_t(n_rows) * n_var;
        dbg_bytes[3] = dbg_bytes[4] = sizeof(float) * dbg_steps * P0;
        for (int k = 0; k < 5; ++k) {
            if ((rc = stage_out(ctx, ctx->dbg[k], dbg_host[k], dbg_bytes[k], &op))) return rc;
            dbg_dev[k] = static_cast<float*>(op);
        }
        a.dbg_loss = dbg_dev[0]; a.dbg_pred = dbg_dev[1]; a.dbg_dldp = dbg_dev[2]; a.dbg_grad = dbg_dev[3]; a.dbg_params = dbg_dev[4];
    }

    a.fast_k = hp->fast_mmd_vectors;
    if (fourier_wz) {
        if (a.fast_k < 1) return fail(ctx, CGNN_ERR_ARG, "fourier_wz given but hparams.fast_mmd_vectors is 0");
        if ((rc = stage_in(ctx, ctx->wz, fourier_wz, sizeof(float) * size_t(n_inst) * epochs * (n_var + 1) * 7 * a.fast_k, &dp))) return rc;
        a.wz_ext = static_cast<const float*>(dp);
    }
    a.n_inst = n_inst; a.N = n_rows; a.NP = NP; a.n_noise = n_noise; a.p_max = p_pad_max;
    a.h = h; a.train_epochs = hp->train_epochs; a.test_epochs = hp->test_epochs; a.trace_every = hp->trace_every;
    a.lr = hp->learning_rate; a.init_std = hp->init_std; a.ld_params = ld_params;
    a.key = make_uint2((unsigned int)(hp->seed & 0xffffffffu), (unsigned int)(hp->seed >> 32));

    // Shared-memory plan of an instance, fastest first (the arithmetic is the same in all of them; it is decided by
    // the instance shape only):
    //   noise buffered per row block + observed matrix in shared memory  >  ... streamed through the warps' noise
    //   buffers (ring)  >  ... read through L1/L2  >  the same three with the noise re-derived at every use instead of
    //   buffered (wide hidden-variable models at 1500 rows: 16 warps x all noise columns would not fit).
    const size_t smem_max = size_t(ctx->prop.sharedMemPerBlockOptin);
    const int nn_pad = (n_noise + 3) & ~3;
    size_t smem = 0;
    bool placed = false;
    for (int regen = hp->graph_noise_regen ? 1 : 0; regen < 2 && !placed; ++regen)
        for (int tm = 0; tm < 3 && !placed; ++tm) {              // 0: observed matrix in smem, 1: ring, 2: global
            if (a.fast_k > 0 && tm != 2) continue;               // random-feature mode reads the observed rows from global
            const int cols = regen ? (tm == 1 ? DP : 0) : nn_pad;
            if (tm == 1 && cols < DP) continue;                  // the ring needs 2 x 16 rows of DP floats per warp
            const size_t need = gr_smem_bytes(NP, DP, cols, p_pad_max, a.fast_k, tm == 0);
            if (need > smem_max) continue;
            a.noise_regen = regen; a.nbuf_cols = cols; a.t_smem = tm == 0; a.t_ring = tm == 1;
            smem = need;
            placed = true;
        }
    if (!placed)
        return fail(ctx, CGNN_ERR_LIMIT, "graph instance does not fit shared memory (rows %d x cols %d, %d noise columns, %d parameters): "
                    "even without noise and observed-matrix buffers it needs %zu B, device allows %zu", n_rows, DP, n_noise, p_max,
                    gr_smem_bytes(NP, DP, 0, p_pad_max, a.fast_k, 0), smem_max);
    ctx->last_graph_mode = a.noise_regen * 4 + (a.t_smem ? 0 : a.t_ring ? 1 : 2);
    const int pol = hp->exp_policy < 0 ? CGNN_EXP_POLICY_GRAPH : hp->exp_policy;
    // ---- CTAs per instance: a cluster when the batch cannot fill the GPU with one CTA per instance ----------
    const int NB = NP / 32;
    const int min_csz = a.fast_k > 0 ? (NB + GR_NW - 1) / GR_NW : 1;   // random-feature mode: one row block per warp
    if (min_csz > GR_MAX_CLUSTER || (dbg_host && min_csz > 1))
        return fail(ctx, CGNN_ERR_LIMIT, "random-feature loss on the graph path: n_rows=%d needs a cluster of %d CTAs (limit %d)",
                    n_rows, min_csz, dbg_host ? 1 : GR_MAX_CLUSTER);
    int slots_of[GR_MAX_CLUSTER + 1] = {0};                                  // resident clusters per cluster size
    for (int csz = min_csz; csz <= GR_MAX_CLUSTER; ++csz) {
        if (dbg_host && csz > 1) break;                              // the parity probe runs on one CTA
        if (hp->graph_cluster > 0 && csz != hp->graph_cluster) continue;
        if (hp->graph_cluster == 0 && csz > min_csz && (NB + csz - 1) / csz < 4) continue;   // fixed costs dominate tiny shares
        int slots = 0;
        if (launch_graph(DP, HP, pol, a, -1, csz, smem, ctx->stream, &slots) != cudaSuccess || slots < 1) {
            cudaGetLastError();
            continue;
        }
        slots_of[csz] = slots;
    }
    // Time of one instance on a cluster of csz CTAs, in units of "one round of 16 row blocks on a full SM":
    // a warp owns a row block, a CTA works through ceil(rb/16) rounds, and a round with w < 16 busy warps is
    // shorter because the warps share the SM's pipes - measured on B200 (tools/bench_graph_cluster.py):
    // w = 16/12/8/6/4 -> 1.00/0.875/0.70/0.69/0.55, i.e. about 0.4 + 0.6 w/16; a cluster adds ~6 % (two
    // cluster barriers per step, replicated generator forward).
    struct Choice { int csz, slots; double t_inst, cost; };
    auto choose = [&](int n) {
        Choice best{0, 0, 0.0, 1e300};
        for (int csz = 1; csz <= GR_MAX_CLUSTER; ++csz) {
            if (slots_of[csz] < 1) continue;
            const int rb_per_cta = (NB + csz - 1) / csz;
            const int tail_w = rb_per_cta % GR_NW;
            const double t_inst = double(rb_per_cta / GR_NW) + (tail_w ? 0.4 + 0.6 * tail_w / GR_NW : 0.0) + (csz > 1 ? 0.06 : 0.0);
            const double cost = std::ceil(double(n) / slots_of[csz]) * t_inst;
            if (cost < best.cost * 0.999) best = Choice{csz, slots_of[csz], t_inst, cost};
        }
        return best;
    };
    Choice mainc = choose(n_inst);
    if (mainc.csz == 0) return fail(ctx, CGNN_ERR_CUDA, "graph kernel cannot be launched (cluster size %d, %zu B smem)", hp->graph_cluster, smem);
    // A batch just above a multiple of the resident count pays a whole wave for its remainder.  So the batch may be cut
    // in two launches: k full waves on clusters of one size, the rest on the cluster size that suits IT (results do not
    // depend on the cluster size).  Every (size, k) is priced with the model above; the single launch wins ties.
    int full = n_inst, rem = 0;
    Choice tailc{0, 0, 0.0, 1e300};
    if (hp->graph_cluster == 0 && !dbg_host) {
        double best_cost = mainc.cost;
        for (int csz = 1; csz <= GR_MAX_CLUSTER; ++csz) {
            if (slots_of[csz] < 1) continue;
            const int rb_per_cta = (NB + csz - 1) / csz, tail_w = rb_per_cta % GR_NW;
            const double t1 = double(rb_per_cta / GR_NW) + (tail_w ? 0.4 + 0.6 * tail_w / GR_NW : 0.0) + (csz > 1 ? 0.06 : 0.0);
            const int kmax = n_inst / slots_of[csz];
            for (int k = kmax; k >= 1 && k >= kmax - 2; --k) {
                const int r = n_inst - k * slots_of[csz];
                if (r <= 0) continue;
                const Choice t = choose(r);
                if (t.csz == 0 || t.csz == csz) continue;
                const double cost = k * t1 + t.cost + 0.02;           // + a launch boundary
                if (cost < 0.97 * best_cost) {
                    best_cost = cost;
                    mainc = Choice{csz, slots_of[csz], t1, k * t1};
                    tailc = t;
                    full = k * slots_of[csz];
                    rem = r;
                }
            }
        }
    }
    ctx->last_cluster = mainc.csz;

    if (getenv("CGNN_DEBUG"))
        fprintf(stderr, "[cgnn_b200] graph batch: %d instances, main cluster %d (%d resident), full %d, remainder %d on cluster %d (%d resident)\n",
                n_inst, mainc.csz, mainc.slots, full, rem, tailc.csz, tailc.slots);
    CU(cudaEventRecord(ctx->ev0, ctx->stream));
    cudaError_t le;
    if (tailc.csz) {
        GraphArgs a1 = a, a2 = a;
        a1.n_inst = full;
        a2.n_inst = rem; a2.inst = a.inst + full; a2.inst_uid = a.inst_uid + full; a2.counter = a.counter + 32;
        le = launch_graph(DP, HP, pol, a1, std::min(full, mainc.slots), mainc.csz, smem, ctx->stream, nullptr);
        if (le == cudaSuccess) le = launch_graph(DP, HP, pol, a2, std::min(rem, tailc.slots), tailc.csz, smem, ctx->stream, nullptr);
        ctx->last_launches = 3;
    } else {
        le = launch_graph(DP, HP, pol, a, std::min(n_inst, mainc.slots), mainc.csz, smem, ctx->stream, nullptr);
        ctx->last_launches = 2;
    }
    if (le != cudaSuccess) return fail(ctx, CGNN_ERR_CUDA, "graph kernel launch: %s", cudaGetErrorString(le));
    CU(cudaEventRecord(ctx->ev1, ctx->stream));

    if ((rc = copy_back(ctx, scores, a.scores, sizeof(float) * n_inst))) return rc;
    if ((rc = copy_back(ctx, loss_trace, a.trace, sizeof(float) * size_t(n_inst) * epochs))) return rc;
    if ((rc = copy_back(ctx, params_out, a.params_out, sizeof(float) * size_t(n_inst) * ld_params))) return rc;
    if (dbg_host)
        for (int k = 0; k < 5; ++k)
            if ((rc = copy_back(ctx, dbg_host[k], dbg_dev[k], dbg_bytes[k]))) return rc;
    CU(cudaStreamSynchronize(ctx->stream));
    CU(cudaEventElapsedTime(&ctx->last_ms, ctx->ev0, ctx->ev1));
    return CGNN_OK;
}

extern "C" {

int cgnn_b200_last_graph_mode(const cgnn_ctx* ctx) { return ctx ? ctx->last_graph_mode : -1; }

int cgnn_b200_run_graph(cgnn_ctx* ctx, const cgnn_hparams* hp, const float* data, int n_rows_total,
                        const cgnn_graph_batch* graphs, int n_runs, const int* row_subsets, int n_rows,
                        unsigned long long uid_base, const unsigned long long* graph_uids, const float* noise,
                        const float* init_params, int ld_params, float* scores, float* loss_trace, float* params_out) {
    return run_graph_impl(ctx, hp, data, n_rows_total, graphs, n_runs, row_subsets, n_rows, uid_base, graph_uids, noise, init_params,
                          ld_params, scores, loss_trace, params_out, nullptr, 0);
}

int cgnn_b200_graph_step_debug(cgnn_ctx* ctx, const cgnn_hparams* hp, const float* data, int n_rows,
                               const cgnn_graph_batch* graph, int n_steps, const float* noise, const float* init_params,
                               int n_params, const float* fourier_wz, float* out_loss, float* out_pred, float* out_dldp,
                               float* out_grad, float* out_params) {
    if (!ctx) return CGNN_ERR_ARG;
    if (!hp) return fail(ctx, CGNN_ERR_ARG, "hparams is NULL");
    if (!noise || !init_params) return fail(ctx, CGNN_ERR_ARG, "step_debug needs caller-supplied noise and parameters");
    if (n_steps < 1) return fail(ctx, CGNN_ERR_ARG, "n_steps must be >= 1");
    cgnn_hparams h2 = *hp;
    h2.train_epochs = n_steps;
    h2.test_epochs = 0;
    h2.trace_every = 0;
    float* dbg[5] = {out_loss, out_pred, out_dldp, out_grad, out_params};
    return run_graph_impl(ctx, &h2, data, n_rows, graph, 1, nullptr, n_rows, 0ULL, nullptr, noise, init_params, n_params, nullptr,
                          nullptr, nullptr, dbg, n_steps, fourier_wz);
}

}  // extern "C"
