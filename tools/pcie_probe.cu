// Host <-> device copy ceiling of the box, with 1 .. N GPUs active at once.
//
// What dc_convert_host moves for bearing x4 is 2 bytes in and 8 bytes out per
// pixel; this probe moves the same 1 : 4 mix with nothing but pinned buffers and
// one cudaMemcpyAsync per buffer (no batched-copy API), one host thread and two
// streams (one per direction) per GPU, all GPUs started together. If the
// aggregate stops scaling here, the box (PCIe switches / host memory), not the
// library, bounds the end-to-end number (VERDICT r1 "what's weak" 4).
//
//   nvcc -O2 -o pcie_probe tools/pcie_probe.cu -lpthread
//   pcie_probe [max_gpus] [MiB per H2D buffer] [repetitions]
// Prints one line per (active GPUs, mode): aggregate GB/s.
#include <cuda_runtime.h>

#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <thread>
#include <vector>

#define CK(x)                                                                            \
    do {                                                                                 \
        cudaError_t e_ = (x);                                                            \
        if (e_ != cudaSuccess) {                                                         \
            std::fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_));                \
            std::exit(2);                                                                \
        }                                                                                \
    } while (0)

struct gpu_buffers {
    void *h_in = nullptr, *h_out = nullptr, *d_in = nullptr, *d_out = nullptr;
    cudaStream_t up = nullptr, down = nullptr;
};

enum mode { H2D_ONLY, D2H_ONLY, BOTH_1_TO_4 };

int main(int argc, char** argv) {
    int n_dev = 0;
    CK(cudaGetDeviceCount(&n_dev));
    const int    max_gpus = argc > 1 ? std::min(std::atoi(argv[1]), n_dev) : n_dev;
    const size_t in_bytes = size_t(argc > 2 ? std::atoi(argv[2]) : 64) << 20;
    const int    reps     = argc > 3 ? std::atoi(argv[3]) : 20;
    const size_t out_bytes = 4 * in_bytes;
    std::vector<gpu_buffers> g(max_gpus);
    for (int d = 0; d < max_gpus; ++d) {
        CK(cudaSetDevice(d));
        CK(cudaHostAlloc(&g[d].h_in, in_bytes, cudaHostAllocDefault));
        CK(cudaHostAlloc(&g[d].h_out, out_bytes, cudaHostAllocDefault));
        CK(cudaMalloc(&g[d].d_in, in_bytes));
        CK(cudaMalloc(&g[d].d_out, out_bytes));
        CK(cudaMemset(g[d].d_out, 1, out_bytes));
        CK(cudaStreamCreateWithFlags(&g[d].up, cudaStreamNonBlocking));
        CK(cudaStreamCreateWithFlags(&g[d].down, cudaStreamNonBlocking));
        CK(cudaDeviceSynchronize());
    }
    std::printf("# pinned cudaMemcpyAsync, %zu MiB H2D + %zu MiB D2H per repetition per GPU, %d repetitions\n",
                in_bytes >> 20, out_bytes >> 20, reps);
    const char* names[3] = {"h2d", "d2h", "h2d+d2h(1:4)"};
    for (int active = 1; active <= max_gpus; active *= 2) {
        for (int m = 0; m < 3; ++m) {
            std::atomic<int> ready{0};
            std::atomic<bool> go{false};
            std::vector<std::thread> th;
            std::vector<double>      secs(active, 0.);
            for (int d = 0; d < active; ++d)
                th.emplace_back([&, d] {
                    CK(cudaSetDevice(d));
                    auto run = [&](int n) {
                        for (int r = 0; r < n; ++r) {
                            if (m != D2H_ONLY)
                                CK(cudaMemcpyAsync(g[d].d_in, g[d].h_in, in_bytes, cudaMemcpyHostToDevice, g[d].up));
                            if (m != H2D_ONLY)
                                CK(cudaMemcpyAsync(g[d].h_out, g[d].d_out, out_bytes, cudaMemcpyDeviceToHost, g[d].down));
                        }
                        CK(cudaStreamSynchronize(g[d].up));
                        CK(cudaStreamSynchronize(g[d].down));
                    };
                    run(2);
                    ready.fetch_add(1);
                    while (!go.load())
                        std::this_thread::yield();
                    const auto t0 = std::chrono::steady_clock::now();
                    run(reps);
                    secs[d] = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
                });
            while (ready.load() < active)
                std::this_thread::yield();
            go.store(true);
            for (auto& t : th)
                t.join();
            double worst = 0.;
            for (double s : secs)
                worst = std::max(worst, s);
            const double bytes = double(reps) * active *
                                 ((m != D2H_ONLY ? double(in_bytes) : 0.) + (m != H2D_ONLY ? double(out_bytes) : 0.));
            std::printf("gpus %d  %-14s %8.1f GB/s aggregate  (%.1f per GPU)\n", active, names[m],
                        bytes / worst / 1e9, bytes / worst / 1e9 / active);
        }
    }
    return 0;
}
