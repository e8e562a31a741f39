// Stand-alone check + timing of the tcgen05 intersection kernel against a CPU popcount reference.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../../beyond-fixed-forms_b200/csrc/iou_umma.cuh"
int main(int argc, char **argv) {
    const int n = argc > 1 ? atoi(argv[1]) : 200, m = argc > 2 ? atoi(argv[2]) : 150;
    const long long bits = argc > 3 ? atoll(argv[3]) : 10007, words = (bits + 31) / 32;
    std::vector<uint32_t> ha((size_t)n * words), hb((size_t)m * words);
    srand(1);
    for (auto &w : ha) w = (uint32_t)rand() & (uint32_t)rand();
    for (auto &w : hb) w = (uint32_t)rand() & (uint32_t)rand() & (uint32_t)rand();
    uint32_t *da, *db; int32_t *di;
    cudaMalloc(&da, ha.size() * 4); cudaMalloc(&db, hb.size() * 4); cudaMalloc(&di, (size_t)n * m * 4);
    cudaMemcpy(da, ha.data(), ha.size() * 4, cudaMemcpyHostToDevice); cudaMemcpy(db, hb.data(), hb.size() * 4, cudaMemcpyHostToDevice);
    cudaMemset(di, 0, (size_t)n * m * 4);
    cudaError_t e = bff::umma::launch_intersect(da, n, db, m, words, 0, di, nullptr, nullptr, 148, 0);
    printf("launch: %s\n", cudaGetErrorString(e));
    e = cudaDeviceSynchronize();
    printf("sync: %s\n", cudaGetErrorString(e));
    if (e != cudaSuccess) return 1;
    std::vector<int32_t> hi((size_t)n * m);
    cudaMemcpy(hi.data(), di, hi.size() * 4, cudaMemcpyDeviceToHost);
    long long bad = 0;
    for (int i = 0; i < n && i < 64; ++i)
        for (int j = 0; j < m; ++j) {
            int s = 0;
            for (long long w = 0; w < words; ++w) s += __builtin_popcount(ha[i * words + w] & hb[j * words + w]);
            if (s != hi[(size_t)i * m + j]) { if (bad < 5) printf("mismatch (%d,%d): got %d want %d\n", i, j, hi[(size_t)i * m + j], s); ++bad; }
        }
    printf("rectangular %dx%d x %lld bits: %lld mismatches (first 64 rows checked)\n", n, m, bits, bad);
    // symmetric + timing
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int rep = 0; rep < 3; ++rep) {
        cudaMemset(di, 0, (size_t)n * m * 4);
        cudaEventRecord(e0);
        bff::umma::launch_intersect(da, n, da, n, words, 1, di, nullptr, nullptr, 148, 0);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        printf("symmetric %dx%d: %.4f ms  %.3e word-ops/s (%s)\n", n, n, ms, (double)n * n * words / (ms * 1e-3), cudaGetErrorString(cudaGetLastError()));
    }
    return bad != 0;
}
