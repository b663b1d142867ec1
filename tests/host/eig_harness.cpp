// Host harness for openpbe_b200/csrc/pbe_eig.cuh (tests/test_eig_host.py): reads
// [2*order][n] FP64 moments from a file, writes nodes [order][n] and weights [order][n], using the
// HOST instantiation of the same wheeler_dev<N> the CUDA kernels run.  Test infrastructure only.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>
#define PBE_EIG_COUNT_FALLBACKS 1
#include "pbe_eig.cuh"

template <int N>
static void run(long n, const std::vector<double> &m, std::vector<double> &x, std::vector<double> &w) {
    for (long c = 0; c < n; ++c) {
        double mm[2 * N], xx[N], ww[N];
        for (int k = 0; k < 2 * N; ++k) mm[k] = m[(size_t)k * n + c];
        pbe::wheeler_dev<N>(mm, xx, ww);
        for (int i = 0; i < N; ++i) {
            x[(size_t)i * n + c] = xx[i];
            w[(size_t)i * n + c] = ww[i];
        }
    }
}

int main(int argc, char **argv) {
    if (argc < 5) return 2;
    const int order = std::atoi(argv[1]);
    const long n = std::atol(argv[2]);
    std::vector<double> m((size_t)2 * order * n), x((size_t)order * n), w((size_t)order * n);
    FILE *f = std::fopen(argv[3], "rb");
    if (!f || std::fread(m.data(), 8, m.size(), f) != m.size()) return 3;
    std::fclose(f);
    switch (order) {
    case 1: run<1>(n, m, x, w); break;
    case 2: run<2>(n, m, x, w); break;
    case 3: run<3>(n, m, x, w); break;
    case 4: run<4>(n, m, x, w); break;
    default: return 4;
    }
    f = std::fopen(argv[4], "wb");
    std::fwrite(x.data(), 8, x.size(), f);
    std::fwrite(w.data(), 8, w.size(), f);
    std::fclose(f);
    std::printf("fallbacks %ld\n", pbe::jacobi3_fallbacks);
    return 0;
}
