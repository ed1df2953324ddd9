// Phase timing of K-B (mask -> min-area rectangle): compiles csrc/mask_rect.cu with its clock64() hooks enabled and runs
// it on synthetic ellipse masks of BASELINE config-2 shape.  Diagnostic only; not part of libtdod.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -DTDOD_KB_TIMING tools/kb_phases.cu 3dod_b200/csrc/api.cu -o /tmp/kb_phases
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include "../3dod_b200/csrc/mask_rect.cu"

int main(int argc, char** argv) {
    const int H = 530, W = 730, N = argc > 1 ? atoi(argv[1]) : 2048;
    std::vector<uint8_t> h((size_t)N * H * W, 0);
    srand(1);
    for (int m = 0; m < N; ++m) {
        const double bw = (0.08 + 0.42 * rand() / RAND_MAX) * W, bh = (0.08 + 0.52 * rand() / RAND_MAX) * H;
        const double x0 = (double)rand() / RAND_MAX * (W - 1 - bw), y0 = (double)rand() / RAND_MAX * (H - 1 - bh);
        const double cx = x0 + bw / 2, cy = y0 + bh / 2;
        for (int y = (int)y0; y <= (int)(y0 + bh); ++y)
            for (int x = (int)x0; x <= (int)(x0 + bw); ++x) {
                const double ex = (x - cx) / (bw / 2), ey = (y - cy) / (bh / 2);
                if (ex * ex + ey * ey <= 1.0) h[(size_t)m * H * W + (size_t)y * W + x] = 1;
            }
    }
    uint8_t *d_m, *d_has; float* d_r; void* ws;
    cudaMalloc(&d_m, h.size()); cudaMemcpy(d_m, h.data(), h.size(), cudaMemcpyHostToDevice);
    cudaMalloc(&d_has, N); cudaMalloc(&d_r, N * 32);
    const int64_t wsb = tdod_mask_rect_workspace_bytes(N, H, W);
    cudaMalloc(&ws, wsb);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int it = 0; it < 3; ++it) {
        cudaEventRecord(e0);
        int rc = tdod_mask_rect(d_m, N, H, W, d_r, d_has, ws, wsb, 0);
        cudaEventRecord(e1); cudaDeviceSynchronize();
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        printf("rc=%d  %d masks  %.3f ms  (%.1f GB/s)\n", rc, N, ms, (double)h.size() / ms / 1e6);
    }
    std::vector<long long> t(32 * 8192);
    cudaMemcpyFromSymbol(t.data(), g_kb_t, sizeof(long long) * 32 * 8192);
    const int M = N < 8192 ? N : 8192;
    const char* names[9] = {"row range", "label_runs", "select_component", "rowL/R+extent", "chain sieve", "scan+compact",
                            "stacks+assemble", "calipers pre", "calipers walk"};
    double tot = 0, acc[9] = {0};
    double hv = 0, rows = 0, runs = 0;
    for (int m = 0; m < M; ++m) {
        for (int k = 0; k < 9; ++k) acc[k] += (double)(t[m * 32 + k + 1] - t[m * 32 + k]);
        tot += (double)(t[m * 32 + 9] - t[m * 16]);
        hv += t[m * 32 + 10]; rows += t[m * 32 + 11]; runs += t[m * 32 + 12];
    }
    printf("mean hull vertices %.1f, rows %.1f, runs %.1f; mean cycles per mask (CTA view) %.0f\n", hv / M, rows / M, runs / M, tot / M);
    for (int k = 0; k < 9; ++k) printf("  %-18s %9.0f cycles  %5.1f%%\n", names[k], acc[k] / M, 100 * acc[k] / tot);
    const char* ln[6] = {"  label: count pass", "  label: scan", "  label: emit+init", "  label: union", "  label: find", "  label: area"};
    const int a[6] = {1, 16, 17, 18, 19, 20}, b[6] = {16, 17, 18, 19, 20, 2};
    for (int k = 0; k < 6; ++k) {
        double d = 0;
        for (int m = 0; m < M; ++m) d += (double)(t[m * 32 + b[k]] - t[m * 32 + a[k]]);
        printf("  %-22s %9.0f cycles\n", ln[k], d / M);
    }
    return 0;
}
