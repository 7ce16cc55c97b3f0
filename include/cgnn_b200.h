/* cgnn_b200 - C ABI of the B200-native CGNN hot-path engine.
 *
 * The reference (CaoAo/CGNN) has no native code: its seam for this path is a Python
 * function pointer selected by a backend string,
 *     backend_alg_dic = {"TensorFlow": tf_run_instance}        Code/cgnn/GNN.py:182,193
 *     self.infer_graph = run_CGNN_tf                           Code/cgnn/CGNN.py:344-345
 *     self.infer_graph = run_CGNN_confounders_tf               Code/cgnn/CGNN_confounders.py:475-476
 * called once per (pair | candidate graph, run) through joblib.  One call per instance cannot
 * fill 148 SMs, so the replacement boundary is BATCHED: one call scores every
 * (pair | graph, run[, direction]) instance of a batch.  Each entry point below names the
 * reference function(s) it replaces.  Plain C types only; no exceptions cross the boundary;
 * every function returns 0 on success or a negative cgnn_status, and the message for the
 * last failure on a context is available from cgnn_b200_last_error().
 *
 * Ownership: the caller owns every buffer passed in; buffers may be host memory (pageable or
 * pinned) or device memory on the context's device (detected with cudaPointerGetAttributes).
 * Calls are synchronous (results are in the caller's buffer on return).  A context is not
 * re-entrant: one call at a time per context; separate contexts are independent.
 * NaN/Inf scores are returned as they are, never masked (the reference filters them in the
 * graph paths only, CGNN.py:217,240 - that stays the caller's job).
 */
#ifndef CGNN_B200_H
#define CGNN_B200_H

#ifdef __cplusplus
extern "C" {
#endif

#define CGNN_B200_VERSION      100      /* major*10000 + minor*100 + patch */
#define CGNN_N_GAMMAS          7        /* Loss.py:10 - fixed at compile time, like the reference */
#define CGNN_MAX_ROWS_PAIRWISE 2048     /* reference subsamples to SETTINGS.max_nb_points = 1500 (GNN.py:150-153) */
#define CGNN_MAX_H             64       /* hidden width limit (reference default 20, scripts use 30) */
#define CGNN_MAX_H_GRAPH       32       /* hidden width limit of the graph path (reference default 20) */
#define CGNN_MAX_VARS          32       /* variables (columns) per graph */
#define CGNN_MAX_FAST_VECTORS  256      /* nb_vectors_approx_MMD limit of the random-feature loss (reference default 100) */
#define CGNN_MAX_ROWS_GRAPH    1536     /* rows per graph instance after subsampling (SETTINGS.max_nb_points = 1500).
                                           One graph instance must also fit one SM's shared memory (227 KB):
                                           rows x vars x 4 B + 2 KB per noise column + 20 B per (padded) parameter;
                                           e.g. 1536 x 20, h = 20 fits, 1536 x 32 does not (CGNN_ERR_LIMIT with the sizes) */

typedef enum cgnn_status {
    CGNN_OK            =  0,
    CGNN_ERR_ARG       = -1,   /* bad argument (null pointer, size out of range ...) */
    CGNN_ERR_CUDA      = -2,   /* a CUDA runtime call failed; see last_error */
    CGNN_ERR_NOMEM     = -3,   /* device allocation failed */
    CGNN_ERR_LIMIT     = -4,   /* problem exceeds an engine limit (rows, hidden width, inputs) */
    CGNN_ERR_NODEVICE  = -5    /* no usable sm_100 device */
} cgnn_status;

typedef struct cgnn_ctx cgnn_ctx;

#define CGNN_EXP_POLICY_AUTO     (-1)
#define CGNN_EXP_POLICY_PAIRWISE 3
#define CGNN_EXP_POLICY_GRAPH    1

/* Hyper-parameters.  Mirrors the knobs of Code/cgnn/utils/Settings.py:26-41 that reach the
 * device code, plus the engine's own (seed, exp_policy, trace_every). */
typedef struct cgnn_hparams {
    int   h_layer_dim;        /* SETTINGS.h_layer_dim (Settings.py:36) */
    int   train_epochs;       /* SETTINGS.train_epochs (Settings.py:37) */
    int   test_epochs;        /* SETTINGS.test_epochs  (Settings.py:38) */
    float learning_rate;      /* SETTINGS.learning_rate (Settings.py:31) */
    float init_std;           /* SETTINGS.init_weights  (Settings.py:32) */
    int   exp_policy;         /* how the 7 Gaussian kernels of one pair evaluation are produced (all within the
                                 1e-5 parity bound, measured 2e-6): 0 = seven MUFU ex2; 1 = five (gamma 0.5, 1 by
                                 squaring gamma 0.25); 2 = four (+ gamma 0.05 = 10th power of 0.005);
                                 3 = three (+ gamma 5 = 5th power of gamma 1);
                                 CGNN_EXP_POLICY_AUTO (-1) = the fastest measured per path: 3 on the pairwise path
                                 (bound by the MUFU pipe), 1 on the graph path (bound by the FMA pipe, where an
                                 ex2 is cheaper than the multiplies that replace it).  The graph kernel is built for
                                 0, 1 and 3; it runs a request for 2 as 1 (one ex2 more, smaller error) */
    int   trace_every;        /* >0: also compute the (pre-update) training loss on every trace_every-th
                                 train step into loss_trace (GNN.py:98-102 prints every 100th); 0 = never */
    int   fast_mmd_vectors;   /* 0: Gaussian-kernel MMD (SETTINGS.use_Fast_MMD = False).  k > 0: the random-feature loss
                                 of Loss.py:35-56 with k = SETTINGS.nb_vectors_approx_MMD vectors per bandwidth
                                 (k <= CGNN_MAX_FAST_VECTORS; on the graph path n_rows <= 2048) */
    unsigned long long seed;  /* engine Philox seed.  The reference is unseeded (no hit for "seed" in its tree). */
    int   graph_cluster;      /* graph path: CTAs (SMs) per instance, 1..8; 0 = chosen per batch (a thread-block cluster
                                 when the batch has too few instances to fill the GPU otherwise) */
    int   graph_noise_regen;  /* graph path: 0 = buffer a row block's noise columns when they fit shared memory (default),
                                 1 = always re-derive them at every use (the plan the engine falls back to when they do
                                 not fit; same random streams, bit-identical results - this switch exists for testing) */
} cgnn_hparams;

/* ---- life cycle ------------------------------------------------------------------------- */

int  cgnn_b200_version(void);

/* Creates a context bound to ONE device (one process per GPU is the multi-GPU model; the
 * reference's "/gpu:(offset + run % nb_gpu)" placement, GNN.py:159, becomes the caller's
 * sharding of instances over contexts).  Replaces GNN_tf.__init__'s session creation
 * (GNN.py:77-80) - done once, not once per instance. */
int  cgnn_b200_create(cgnn_ctx** out, int device_id);
void cgnn_b200_destroy(cgnn_ctx* ctx);
const char* cgnn_b200_last_error(const cgnn_ctx* ctx);   /* owned by ctx; valid until the next call on ctx */

/* Run the engine's kernels on an existing CUDA stream (e.g. torch's current stream) instead
 * of the context's own.  stream is a cudaStream_t passed as void*. */
int  cgnn_b200_set_stream(cgnn_ctx* ctx, void* stream);

/* Device time (ms, CUDA events on the launching stream) of the main kernel of the last
 * run_* call on this context; number of kernels that call launched. */
float cgnn_b200_last_kernel_ms(const cgnn_ctx* ctx);
int   cgnn_b200_last_launch_count(const cgnn_ctx* ctx);

/* ---- pairwise GNN ----------------------------------------------------------------------- */

/* Scores every (pair, run, direction) instance of a batch of cause-effect pairs.
 * Replaces tf_run_instance + tf_evalcausalscore_pairwise + GNN_tf.{__init__,train,evaluate}
 * (Code/cgnn/GNN.py:135-166, 129-132, 33-126) for all runs of all pairs at once.
 *
 *   xy            [n_pairs][ld_rows][2] fp32, column 0 = A, column 1 = B (already standardised and,
 *                 if needed, subsampled by the caller as GNN.py:150-153 does)
 *   n_rows        [n_pairs] valid rows of each pair, 1 <= n_rows[p] <= min(ld_rows, CGNN_MAX_ROWS_PAIRWISE)
 *   n_runs        SETTINGS.NB_RUNS
 *   uid_base      global index of this batch's first instance: instance (pair p, run r, direction d)
 *                 draws its random numbers from Philox stream uid_base + (p*n_runs + r)*2 + d, so
 *                 results do not depend on how a sweep is cut into batches or sharded over GPUs
 *   noise         NULL (in-kernel Philox N(0,1)) or [n_inst][train_epochs+test_epochs][ld_rows] fp32:
 *                 the e of GNN.py:64 for every sess.run, supplied by the caller (parity mode)
 *   init_params   NULL (in-kernel N(0, init_std^2), GNN.py:21-29) or [n_inst][4h+1] fp32, layout
 *                 [W_in row X (h) | W_in row e (h) | b_in (h) | W_out (h) | b_out]
 *   scores        out [n_pairs][n_runs][2]: mean test MMD of A->B (d=0) and B->A (d=1)  (GNN.py:161-163)
 *   loss_trace    NULL or out [n_inst][train_epochs+test_epochs]: training losses on traced steps
 *                 (NaN elsewhere) followed by every evaluation loss
 *   params_out    NULL or out [n_inst][4h+1]: parameters after training
 * with n_inst = n_pairs*n_runs*2 and instance index (p*n_runs + r)*2 + d. */
int  cgnn_b200_run_pairwise(cgnn_ctx* ctx, const cgnn_hparams* hp,
                            const float* xy, const int* n_rows, int n_pairs, int ld_rows, int n_runs,
                            unsigned long long uid_base,
                            const float* noise, const float* init_params,
                            float* scores, float* loss_trace, float* params_out);

/* Parity probe: n_steps consecutive TRAIN steps of one pairwise instance (column 0 of xy = cause)
 * with caller-supplied parameters and noise, dumping what the reference's graph would hold at
 * each sess.run (GNN.py:93-96): the loss before the update, the generated column, dLoss/d out_y,
 * the parameter gradient and the parameters after the Adam update.  Runs the production kernel.
 *   xy [n_rows][2], noise [n_steps][n_rows] or NULL (in-kernel Philox, stream uid 0), init_params [4h+1] or NULL
 *   fourier_wz  NULL, or - with hp->fast_mmd_vectors = k - the random matrix of Loss.py:35-37 for every step,
 *               [n_steps][3][7k] (rows: frequency on the cause, frequency on the effect, phase)
 *   out_loss [n_steps], out_yhat [n_steps][n_rows], out_dldy [n_steps][n_rows],
 *   out_grad [n_steps][4h+1], out_params [n_steps][4h+1]    (any of them may be NULL) */
int  cgnn_b200_pairwise_step_debug(cgnn_ctx* ctx, const cgnn_hparams* hp,
                                   const float* xy, int n_rows, int n_steps,
                                   const float* noise, const float* init_params, const float* fourier_wz,
                                   float* out_loss, float* out_yhat, float* out_dldy,
                                   float* out_grad, float* out_params);

/* ---- graph CGNN and CGNN with hidden variables --------------------------------------------- */

/* A batch of candidate graphs over the same n_var variables, each given as its nodes in the
 * reference's BUILD ORDER (repeated scan of graph.get_list_nodes(), a node is emitted once all its
 * parents are, Code/cgnn/CGNN.py:63-69).  Node k of graph g
 *   writes data column          node_col[g*n_var + k]
 *   has n_parents[g*n_var + k]  generated parent columns (graph.get_parents order, CGNN.py:67,76)
 *   has n_inputs[g*n_var + k]   inputs in total = parents, then its own noise column, then - hidden-variable
 *                               model only - one shared noise column per skeleton neighbour
 *                               (graph.skeleton.get_neighbors order, CGNN_confounders.py:82-96)
 * in_idx holds, graph after graph (graph g occupies [in_ptr[g], in_ptr[g+1])) and node after node, the
 * parent COLUMN indices followed by the NOISE-COLUMN ids.  Noise column ids: c < n_var is the own noise
 * of the node writing column c; n_var + e is the noise shared over skeleton edge e
 * (skeleton.get_list_edges_without_duplicate order, CGNN_confounders.py:64-72).
 * Parameters of a graph, flat: per node in build order [W_in (n_inputs x h) | b_in (h) | W_out (h) | b_out]. */
typedef struct cgnn_graph_batch {
    int n_graphs;
    int n_var;
    int n_noise_cols;         /* n_var for CGNN; n_var + number of skeleton edges for the hidden-variable model */
    const int* node_col;      /* [n_graphs][n_var]  (host memory) */
    const int* n_parents;     /* [n_graphs][n_var] */
    const int* n_inputs;      /* [n_graphs][n_var] */
    const int* in_ptr;        /* [n_graphs + 1] */
    const int* in_idx;        /* [in_ptr[n_graphs]] */
} cgnn_graph_batch;

/* Scores every (candidate graph, run) instance: mean test MMD after training (no complexity penalty,
 * no finite-filter: CGNN.py:217 and CGNN_confounders.py:236-237 stay with the caller).
 * Replaces run_CGNN_tf + CGNN_tf.{__init__,train,evaluate} (Code/cgnn/CGNN.py:35-195) and
 * run_CGNN_confounders_tf + CGNN_confounders_tf.* (Code/cgnn/CGNN_confounders.py:37-216) for all runs of
 * all candidates at once - the unit a hill-climbing sweep evaluates (CGNN.py:214-215,237-238).
 *
 *   data          [n_rows_total][n_var] fp32 row-major, columns in the order node_col refers to
 *   row_subsets   NULL (n_rows == n_rows_total, every instance sees all rows) or [n_inst][n_rows] row indices:
 *                 the per-call random subsample of CGNN.py:183-185, drawn by the caller
 *   graph_uids    NULL, or [n_graphs] (host): Philox stream base of each graph; run r of graph g then uses
 *                 stream graph_uids[g] + r instead of uid_base + g*n_runs + r.  The search drivers pass a
 *                 hash of the graph structure, which makes a candidate's score independent of the batch
 *                 it was evaluated in.
 *   noise         NULL (in-kernel Philox) or [n_inst][train+test epochs][n_noise_cols][n_rows]
 *   init_params   NULL (in-kernel N(0, init_std^2)) or [n_inst][ld_params]
 *   scores        out [n_graphs][n_runs]
 *   loss_trace    NULL or out [n_inst][train+test epochs];  params_out  NULL or out [n_inst][ld_params]
 * with n_inst = n_graphs*n_runs, instance index g*n_runs + r, Philox stream uid_base + that index. */
int  cgnn_b200_run_graph(cgnn_ctx* ctx, const cgnn_hparams* hp,
                         const float* data, int n_rows_total,
                         const cgnn_graph_batch* graphs, int n_runs,
                         const int* row_subsets, int n_rows,
                         unsigned long long uid_base, const unsigned long long* graph_uids,
                         const float* noise, const float* init_params, int ld_params,
                         float* scores, float* loss_trace, float* params_out);

/* Parity probe for the graph path: n_steps consecutive train steps of graph 0 of the batch with
 * caller-supplied noise [n_steps][n_noise_cols][n_rows] and parameters [n_params]; dumps per step the
 * pre-update loss, the generated matrix [n_rows][n_var], dLoss/dPred [n_rows][n_var], the parameter
 * gradient and the parameters after Adam (one sess.run of CGNN.py:118-121).  Host buffers. */
int  cgnn_b200_graph_step_debug(cgnn_ctx* ctx, const cgnn_hparams* hp,
                                const float* data, int n_rows, const cgnn_graph_batch* graph, int n_steps,
                                const float* noise, const float* init_params, int n_params,
                                const float* fourier_wz,   /* NULL, or [n_steps][n_var+1][7k] with hp->fast_mmd_vectors = k */
                                float* out_loss, float* out_pred, float* out_dldp,
                                float* out_grad, float* out_params);

/* ---- pairwise GNN for callers that must not allocate device memory ------------------------ */

/* No counterpart in the reference: this entry point exists only in this library.
 *
 * cgnn_b200_run_pairwise() grows staging buffers with cudaMalloc when a batch is larger than any
 * seen before.  Some hosts cannot afford that: a service whose device memory has been handed over
 * in full to another framework's pool (cudaMalloc would fail or evict), a loop with a latency
 * budget (cudaMalloc synchronises the device and may take milliseconds), or a process that freezes
 * its memory map after start-up.  For them this variant scores a batch with NO allocation of any
 * kind: no cudaMalloc/cudaFree, no context object, no host heap.  It is the same instance kernel;
 * the scratch it needs besides your buffers (work-queue head, row counts, instance table) is a
 * fixed 135,424-byte array compiled into the library, resident on every GPU that loads the
 * library for as long as the process lives.
 *
 * Limits (CGNN_ERR_LIMIT with a message otherwise): n_pairs <= 1024 and n_pairs*n_runs*2 <= 8192
 * instances per call; all pairs have the same n_rows (<= CGNN_MAX_ROWS_PAIRWISE).  Split larger
 * sweeps into several calls (advance uid_base by n_pairs*n_runs*2 each time).
 *
 *   device, stream  CUDA device ordinal and a cudaStream_t on it (as void*, NULL = default stream)
 *   d_xy            DEVICE memory, [n_pairs][n_rows][2] fp32, column 0 = A (see cgnn_b200_run_pairwise)
 *   d_scores        DEVICE memory, out [n_pairs][n_runs][2]
 *   errbuf          caller-owned buffer (may be NULL) that receives the message of a failure
 * Noise and initial parameters always come from the in-kernel Philox streams (seed, uid_base), so
 * the scores are bit-identical to cgnn_b200_run_pairwise() with noise = init_params = NULL.
 *
 * Things to know:
 *  - The first call on each device resolves the device address of the static array
 *    (cudaGetSymbolAddress) and loads the kernels; make that first call during start-up, before
 *    any stream capture begins.
 *  - The array is shared by every call in the process and keeps the previous call's contents; each
 *    call re-initialises the parts that must start from zero on `stream` before its main launch.
 *    Two calls whose GPU work overlaps (different streams or threads) race on it: serialise them.
 *  - Synchronous like the other entry points, on EVERY call and not only the first: it inspects the two pointers
 *    (cudaPointerGetAttributes), enqueues its work on `stream` and waits (cudaStreamSynchronize) until `stream`
 *    has drained before it returns.
 *  - `device` is made current while the call works; the calling thread gets back the device it had before. */
int  cgnn_b200_run_pairwise_noalloc(int device, void* stream, const cgnn_hparams* hp,
                                    const float* d_xy, int n_pairs, int n_rows, int n_runs,
                                    unsigned long long uid_base, float* d_scores,
                                    char* errbuf, int errbuf_len);

/* ---- measurement support ---------------------------------------------------------------- */

/* Micro-benchmarks that measure this chip's own pipe peaks (MEASURED_PEAKS.json holds only HBM
 * and bf16 tensor numbers): kind 0 = MUFU ex2.approx.ftz.f32 (result in exp/s),
 * kind 1 = FP32 FFMA (result in flop/s, 2 per FFMA), kind 2 = packed FFMA2 / fma.rn.f32x2 (flop/s, 4 per instruction),
 * kind 3 = packed FADD2 and kind 4 = the FMA-pipe instruction mix of the graph kernel's gradient loop (both in
 * FMA-pipe lane-slots per second: 2 per packed instruction and lane).  Best of `reps` launches. */
int  cgnn_b200_probe_peak(cgnn_ctx* ctx, int kind, int reps, double* out_per_second);

/* Number of SMs of the context's device and persistent CTAs per SM the engine launches. */
int  cgnn_b200_device_info(const cgnn_ctx* ctx, int* n_sm, int* ctas_per_sm_pairwise);

/* CTAs per instance (cluster size) the last cgnn_b200_run_graph call on this context used. */
int  cgnn_b200_last_graph_cluster(const cgnn_ctx* ctx);

/* Shared-memory plan the last graph call used (diagnostic; results do not depend on it): low two bits = where the
 * observed matrix is read from (0 shared memory, 1 streamed through the warps' buffers, 2 global/L2); +4 = the noise
 * columns are re-derived from their Philox counters at every use instead of being buffered per row block - the plan
 * for hidden-variable models (CGNN_confounders.py:66-96: one extra noise column per skeleton edge) at 1500 rows. */
int  cgnn_b200_last_graph_mode(const cgnn_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* CGNN_B200_H */
