// Arguments of the pairwise instance kernel: ONE type shared by every build of the kernel (512 / 256 / 128 threads per
// CTA, see pairwise.cuh) and by the C-ABI translation unit that fills it.
#pragma once
#include <cuda_runtime.h>

namespace cgnn {

constexpr int PW_MAX_ROWS = 2048;

struct PairwiseArgs {
    const float* xy;          // [n_pairs][ld_rows][2]
    const int*   n_rows;      // [n_pairs]
    const int4*  inst;        // [n_inst] (pair, run, dir, out_index), sorted by decreasing cost
    int n_inst, ld_rows, np_max;
    int h, train_epochs, test_epochs, trace_every;
    float lr, init_std;
    uint2 key;                  // Philox key = seed; the high half of the 64-bit instance uid is xor-ed into key.y
    unsigned long long uid_base;
    const float* noise;       // NULL or [n_inst][epochs][ld_rows]   (indexed by out_index)
    const float* init_params; // NULL or [n_inst][4h+1]
    float* scores;            // [n_inst] by out_index
    float* trace;             // NULL or [n_inst][train+test]
    float* params_out;        // NULL or [n_inst][4h+1]
    int*   counter;           // work queue head (scratch)
    // parity probe (single instance): any may be NULL
    float *dbg_loss, *dbg_yhat, *dbg_dldy, *dbg_grad, *dbg_params;
    // random-feature ("Fourier") MMD, Loss.py:35-56 (SETTINGS.use_Fast_MMD): fast_k = nb_vectors_approx_MMD, 0 = off
    int fast_k;
    const float* wz_ext;      // NULL (in-kernel Philox) or [n_inst][epochs][3][7*fast_k]: the matrix of Loss.py:35-37
};

}  // namespace cgnn
