// h2d_ceiling.cu -- what can this box move from page-locked host memory into its GPUs at once?  The e2e leg of bench.py is bounded by it.
// One host thread per GPU, each copying `chunk` MiB pieces of its own pinned buffer to its device with cudaMemcpyAsync on two streams,
// all GPUs concurrently, for G = 1, 2, 4, ... up to the GPUs present.  Also the D2H direction and both at once.
// build: nvcc -O2 -o tools/h2d_ceiling tools/h2d_ceiling.cu     usage: h2d_ceiling [MiB per copy = 20] [seconds = 1.0]   -> one JSON object
#include <cuda_runtime.h>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <thread>
#include <vector>

static double run(int G, size_t bytes, double seconds, int mode)   // mode 0: H2D, 1: D2H, 2: both
{
    std::vector<std::thread> th;
    std::vector<double> moved(G, 0.0);
    std::atomic<int> ready{0};
    std::atomic<bool> go{false};
    for (int g = 0; g < G; g++)
        th.emplace_back([&, g] {
            cudaSetDevice(g);
            void *h[2], *d[2];
            cudaStream_t s[2];
            for (int k = 0; k < 2; k++) { cudaHostAlloc(&h[k], bytes, cudaHostAllocPortable); cudaMalloc(&d[k], bytes); cudaStreamCreate(&s[k]); }
            cudaMemcpy(d[0], h[0], bytes, cudaMemcpyHostToDevice);
            ready++;
            while (!go) std::this_thread::yield();
            const auto t0 = std::chrono::steady_clock::now();
            double n = 0;
            while (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() < seconds) {
                for (int k = 0; k < 2; k++) {
                    if (mode == 0 || (mode == 2 && k == 0)) cudaMemcpyAsync(d[k], h[k], bytes, cudaMemcpyHostToDevice, s[k]);
                    else cudaMemcpyAsync(h[k], d[k], bytes, cudaMemcpyDeviceToHost, s[k]);
                }
                for (int k = 0; k < 2; k++) cudaStreamSynchronize(s[k]);
                n += 2.0 * bytes;
            }
            moved[g] = n / std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
            for (int k = 0; k < 2; k++) { cudaFreeHost(h[k]); cudaFree(d[k]); cudaStreamDestroy(s[k]); }
        });
    while (ready < G) std::this_thread::yield();
    go = true;
    for (auto& t : th) t.join();
    double sum = 0;
    for (double v : moved) sum += v;
    return sum / 1e9;
}

int main(int argc, char** argv)
{
    const size_t bytes = (size_t)(argc > 1 ? atoi(argv[1]) : 20) << 20;
    const double seconds = argc > 2 ? atof(argv[2]) : 1.0;
    int present = 0;
    cudaGetDeviceCount(&present);
    printf("{\"gpus_present\": %d, \"host_threads\": %u, \"bytes_per_copy\": %zu, \"GBps\": {", present, std::thread::hardware_concurrency(), bytes);
    bool first = true;
    for (int G = 1; G <= present; G *= 2) {
        const double a = run(G, bytes, seconds, 0), b = run(G, bytes, seconds, 1), c = run(G, bytes, seconds, 2);
        printf("%s\"%d\": {\"h2d\": %.1f, \"d2h\": %.1f, \"both\": %.1f}", first ? "" : ", ", G, a, b, c);
        first = false;
    }
    printf("}}\n");
    return 0;
}
