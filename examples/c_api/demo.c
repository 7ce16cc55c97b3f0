/* Plain-C caller of the cgnn_b200 C ABI: scores one synthetic cause-effect pair (both directions, 4 runs)
 * and prints the causal score of Code/cgnn/GNN.py:196-207.  No Python, no C++ on the caller's side.
 *
 *   gcc -I include examples/c_api/demo.c -o demo -L cgnn_b200 -lcgnn_b200 -lm -Wl,-rpath,$PWD/cgnn_b200
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include "cgnn_b200.h"

int main(void) {
    enum { N = 400, RUNS = 4 };
    static float xy[N][2];
    double sx = 0, sy = 0, qx = 0, qy = 0;
    unsigned int s = 12345u;
    for (int i = 0; i < N; ++i) {                       /* y = x^2 + 0.5 x + noise, x uniform: not invertible */
        s = s * 1664525u + 1013904223u; double u = (s >> 8) / 16777216.0;
        s = s * 1664525u + 1013904223u; double v = (s >> 8) / 16777216.0;
        double x = 2.0 * u - 1.0, y = x * x + 0.5 * x + 0.3 * (v - 0.5);
        xy[i][0] = (float)x; xy[i][1] = (float)y;
        sx += x; sy += y; qx += x * x; qy += y * y;
    }
    double mx = sx / N, my = sy / N, dx = sqrt(qx / N - mx * mx), dy = sqrt(qy / N - my * my);
    for (int i = 0; i < N; ++i) { xy[i][0] = (float)((xy[i][0] - mx) / dx); xy[i][1] = (float)((xy[i][1] - my) / dy); }

    cgnn_ctx* ctx = NULL;
    int rc = cgnn_b200_create(&ctx, 0);
    if (rc != CGNN_OK) { fprintf(stderr, "cgnn_b200_create failed: %d (no sm_100 GPU?)\n", rc); return 2; }

    cgnn_hparams hp = {0};
    hp.h_layer_dim = 20; hp.train_epochs = 300; hp.test_epochs = 100;
    hp.learning_rate = 0.01f; hp.init_std = 0.05f; hp.exp_policy = 3; hp.seed = 7;
    int n_rows = N;
    float scores[RUNS][2];
    rc = cgnn_b200_run_pairwise(ctx, &hp, &xy[0][0], &n_rows, 1, N, RUNS, 0ULL, NULL, NULL, &scores[0][0], NULL, NULL);
    if (rc != CGNN_OK) { fprintf(stderr, "run_pairwise failed: %s\n", cgnn_b200_last_error(ctx)); return 3; }
    double ab = 0, ba = 0;
    for (int r = 0; r < RUNS; ++r) { ab += scores[r][0]; ba += scores[r][1]; }
    ab /= RUNS; ba /= RUNS;
    printf("MMD A->B %.5f  B->A %.5f  causal score %.4f (%s)  kernel %.2f ms\n", ab, ba, (ba - ab) / (ba + ab),
           ba > ab ? "A causes B" : "B causes A", cgnn_b200_last_kernel_ms(ctx));
    cgnn_b200_destroy(ctx);
    return (isfinite(ab) && isfinite(ba) && ab > 0 && ba > 0) ? 0 : 1;
}
